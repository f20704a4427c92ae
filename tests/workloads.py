"""Shared synthetic workloads for tests and bench.  Host-side, numpy only.

Only the *description* of a workload lives here (seeds, sizes, distributions, how `hard` is drawn);
the pod columns themselves come from the counter-based generator, which exists twice — in the
oracle (oracle/quota_ref.c, gen_ref_*) and in the product (csrc/bgc_kernels.cuh, k_synth_pods).
"""
from __future__ import annotations

import numpy as np

SEEDS = {"C1": 0xB2000001, "C2": 0xB2000002, "C3": 0xB2000003, "C4": 0xB2000004, "C5": 0xB2000005}
KINDS = {2: [0, 1], 3: [0, 1, 2], 1: [1], 4: [0, 1, 2, 3]}         # gpu, cpu, memory, mig
UNITS = {2: [0, 1], 3: [0, 1, 2], 1: [1], 4: [0, 1, 2, 0]}         # COUNT, MILLI, BYTES


def _sm64(x):
    x = np.asarray(x, dtype=np.uint64)
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


def make_hard(used: np.ndarray, seed: int):
    """`hard[r][ns]` = floor(used * u), u in {0.5, 0.9, 1.0, 1.1, 2.0} chosen per (ns, r) by hash, so
    that roughly half of the verdicts deny; 5 % of namespaces get hard_mask = 0 (no quota object,
    reference src/controller.rs:90), the rest constrain every resource (the synchronizer writes all
    keys together, src/synchronizer.rs:249-281)."""
    R, NS = used.shape
    ns = np.arange(NS, dtype=np.uint64)
    num = np.array([5, 9, 10, 11, 20], dtype=np.int64)
    hard = np.empty_like(used)
    for r in range(R):
        h = _sm64(ns ^ np.uint64((seed + 977 * (r + 1)) & 0xFFFFFFFFFFFFFFFF))
        u = num[(h % np.uint64(5)).astype(np.int64)]
        hard[r] = (used[r] // 10) * u + ((used[r] % 10) * u) // 10
    hm = _sm64(ns ^ np.uint64(seed ^ 0x5151))
    hard_mask = np.where((hm % np.uint64(100)) < np.uint64(5), 0, (1 << R) - 1).astype(np.uint8)
    return hard, hard_mask
