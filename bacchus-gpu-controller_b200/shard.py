"""Row E: namespace-hash sharding across GPUs (north_star).

`shard = mix32(ns_id) % G`: every pod of a namespace lands on one rank, so each rank owns a
disjoint slice of the usage/hard tables and produces final verdicts for its own pods with no
data-path collective; the host only gathers.  Used by bench.py (one process per GPU under
torchrun) and tested with world_size-2 gloo on CPU (tests/test_shard.py).
"""
from __future__ import annotations

import numpy as np


def mix32(x: np.ndarray) -> np.ndarray:
    """Integer finaliser (murmur3 fmix32) — spreads dense ids over ranks."""
    x = x.astype(np.uint32, copy=True)
    x ^= x >> np.uint32(16)
    x *= np.uint32(0x85EBCA6B)
    x ^= x >> np.uint32(13)
    x *= np.uint32(0xC2B2AE35)
    x ^= x >> np.uint32(16)
    return x


def shard_of(ns_id: np.ndarray, world: int) -> np.ndarray:
    return (mix32(ns_id) % np.uint32(world)).astype(np.uint32)


def partition(ns_id: np.ndarray, cols: list, world: int):
    """Host partition pass: returns, per rank, (index array into the batch, ns_id slice, column slices).
    This is itself a full pass over the batch on the CPU — the cost SURVEY.md §8(e) says to state."""
    s = shard_of(ns_id, world)
    out = []
    for g in range(world):
        idx = np.nonzero(s == g)[0]
        out.append((idx, ns_id[idx], [c[idx] for c in cols]))
    return out


def gather_verdicts(n: int, parts: list) -> np.ndarray:
    """Host gather: parts = [(index array, verdict array)] from every rank."""
    out = np.empty(n, np.uint8)
    for idx, v in parts:
        out[idx] = v
    return out


def owned_namespaces(NS: int, world: int, rank: int) -> np.ndarray:
    ids = np.arange(NS, dtype=np.uint32)
    return ids[shard_of(ids, world) == rank]


def local_ids(NS: int, world: int, rank: int):
    """(owned global ids in increasing order, local_of[NS]): the dense per-shard id of every owned
    namespace (0xFFFFFFFF for namespaces other ranks own).  A shard's tables then have
    len(owned) ~ NS/world rows instead of NS."""
    owned = owned_namespaces(NS, world, rank)
    local_of = np.full(NS, 0xFFFFFFFF, dtype=np.uint32)
    local_of[owned] = np.arange(owned.shape[0], dtype=np.uint32)
    return owned, local_of


# ---- row N5: the alternative with one exchange step ----------------------------
def index_slice(n_total: int, world: int, rank: int) -> tuple:
    """Pods sharded by INDEX: no host partition pass and perfect balance even under Zipf; every rank
    reduces its slice into a full R x NS usage table, the tables are summed with one int64
    all-reduce (8*R*NS bytes — 24 MB at 10^6 namespaces x 3 resources — over NVLink/NVSwitch), and each
    rank then evaluates its own slice against the global usage.  Integer sums: still bit-exact."""
    per = (n_total + world - 1) // world
    lo = min(n_total, rank * per)
    return lo, min(n_total, lo + per)


def allreduce_usage(used, group=None):
    """Sum the per-rank usage tables in place (torch tensor, int64; NCCL on GPUs, gloo on CPU)."""
    import torch.distributed as dist
    dist.all_reduce(used, op=dist.ReduceOp.SUM, group=group)
    return used
