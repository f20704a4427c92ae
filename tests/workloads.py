"""Shared synthetic workloads for tests and bench.  Host-side, numpy only.

Only the *description* of a workload lives here (seeds, sizes, distributions, how `hard` is drawn);
the pod columns themselves come from the counter-based generator, which exists twice — in the
oracle (oracle/quota_ref.c, gen_ref_*) and in the product (csrc/bgc_kernels.cuh, k_synth_pods).
"""
from __future__ import annotations

import numpy as np

SEEDS = {"C1": 0xB2000001, "C2": 0xB2000002, "C3": 0xB2000003, "C4": 0xB2000004, "C5": 0xB2000005}
# generator kinds: 0 gpu, 1 cpu millicores, 2 memory bytes, 3 mig; units: 0 COUNT, 1 MILLI, 2 BYTES.
# R = 6 is the pod-relevant key set of src/synchronizer.rs:249-281 (requests.* and limits.* for cpu/memory)
KINDS = {1: [1], 2: [0, 1], 3: [0, 1, 2], 4: [0, 1, 2, 3], 5: [0, 1, 2, 3, 1], 6: [0, 1, 2, 3, 1, 2],
         7: [0, 1, 2, 3, 1, 2, 0], 8: [0, 1, 2, 3, 1, 2, 0, 3]}
UNITS = {1: [1], 2: [0, 1], 3: [0, 1, 2], 4: [0, 1, 2, 0], 5: [0, 1, 2, 0, 1], 6: [0, 1, 2, 0, 1, 2],
         7: [0, 1, 2, 0, 1, 2, 0], 8: [0, 1, 2, 0, 1, 2, 0, 0]}
C1_KEYS6 = ["requests.nvidia.com/gpu", "requests.cpu", "requests.memory", "requests.nvidia.com/mig-1g.10gb", "limits.cpu", "limits.memory"]


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


# ---------------------------------------------------------------- C1 (BASELINE.json configs[0])
C1_KEYS = ["requests.nvidia.com/gpu", "requests.cpu", "requests.memory", "requests.nvidia.com/mig-1g.10gb"]


def c1_namespaces(seed: int = SEEDS["C1"], count: int = 100):
    """100 namespaces `ns-000..099`, each with the quota a sheet row would give it
    (reference src/synchronizer.rs:249-281): cpu in {8,16,32,64} cores, memory in {32,64,128,256} Gi,
    gpu in {1,2,4,8}, mig in {0,1,2}, storage in {100,500} Gi.  5 namespaces have no quota object."""
    rng = np.random.default_rng(seed)
    rows = []
    for i in range(count):
        rows.append(dict(name=f"ns-{i:03d}", cpu=int(rng.choice([8, 16, 32, 64])), memory=int(rng.choice([32, 64, 128, 256])),
                         gpu=int(rng.choice([1, 2, 4, 8])), mig=int(rng.choice([0, 1, 2])), storage=int(rng.choice([100, 500])),
                         has_quota=bool(i % 20 != 7)))
    return rows


def c1_pods(n: int, namespaces, seed: int = SEEDS["C1"], running: bool = True):
    """Synthetic Pod JSON objects (the shape kube-apiserver would send in a pod-create AdmissionReview)."""
    rng = np.random.default_rng(seed + (1 if running else 2))
    pods = []
    cpu_choices = ["250m", "500m", "1", "2", "4000m", "0.5", "5e-1", "1.5"]
    mem_choices = ["256Mi", "1Gi", "2Gi", "4096Mi", "8Gi", "1G", "1500M"]
    for i in range(n):
        ns = namespaces[int(rng.integers(0, len(namespaces)))]["name"]
        nc = int(rng.integers(1, 4))
        containers = []
        for c in range(nc):
            req = {}
            if rng.random() < 0.97:
                req["cpu"] = str(rng.choice(cpu_choices))
            if rng.random() < 0.97:
                req["memory"] = str(rng.choice(mem_choices))
            if rng.random() < 0.3:
                req["nvidia.com/gpu"] = str(int(rng.choice([1, 1, 2, 4])))
            if rng.random() < 0.05:
                req["nvidia.com/mig-1g.10gb"] = "1"
            containers.append({"name": f"c{c}", "resources": {"requests": req, "limits": dict(req)}})
        spec = {"containers": containers}
        if rng.random() < 0.2:
            spec["initContainers"] = [{"name": "init", "resources": {"requests": {"cpu": str(rng.choice(["2", "6"])), "memory": "512Mi"}}}]
        if rng.random() < 0.1:
            spec["overhead"] = {"cpu": "100m", "memory": "64Mi"}
        phase = "Running" if running else "Pending"
        if running and rng.random() < 0.1:
            phase = str(rng.choice(["Succeeded", "Failed"]))
        pods.append({"apiVersion": "v1", "kind": "Pod", "metadata": {"name": f"pod-{i}", "namespace": ns},
                     "spec": spec, "status": {"phase": phase}})
    return pods
