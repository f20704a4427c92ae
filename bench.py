#!/usr/bin/env python
"""bench.py — pods evaluated/s and achieved HBM GB/s of the batched quota-evaluation path.

A step = one pass of the hot path over one batch: K1 usage_reduce over the running set
(n pods) followed by K2 quota_eval over the candidate set (n pods), both through the C ABI of
libbgc.so.  `value` = candidate pods evaluated per second with inputs resident in HBM (int64
columns); `e2e` = the same step through the host-pointer entry points with pinned HOST buffers
(H2D of both sets and D2H of the verdicts inside the timed region), on the compact column types
the narrow-column ABI takes (u16 namespace ids, u8 gpu, u16 cpu millicores, u32 memory MiB:
9 bytes per pod instead of 28) — `e2e_int64_columns` is the same with plain int64 columns.
Default workload = BASELINE.json configs[2] (10^8 pods, Zipf(1.1) over 10^4 namespaces, R=3: gpu,
cpu, memory) — the configuration the north_star quotes its HBM target on; it fits one B200.  With
--gpus N every rank owns an independent namespace shard of the same size (weak scaling, no
data-path collective) for the top-level metric, and the line carries a `c5` sub-record: ONE
10^9-pod x 10^6-namespace Zipf job split over the N GPUs (strong scaling) by namespace hash and
by pod index + one int64 all-reduce.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload c3|c2|c5|c5full]
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOADS = {
    # name: (pods per GPU, namespaces per GPU, R, dist, zipf s, description)
    "c3": (100_000_000, 10_000, 3, 1, 1.1, "C3: 1e8 pods, Zipf(1.1) over 1e4 namespaces, R=3 (gpu,cpu,memory), arrival order"),
    "c2": (1_048_576, 10_000, 2, 0, 0.0, "C2: 2^20 pods x 1e4 namespaces, uniform, R=2 (gpu,cpu)"),
    "c5": (125_000_000, 125_000, 3, 1, 1.1, "C5 shard: 1.25e8 pods x 1.25e5 namespaces per GPU (1e9 x 1e6 over 8), Zipf(1.1), R=3"),
    "c5full": (1_000_000_000, 1_000_000, 3, 1, 1.1, "C5 on one GPU: 1e9 pods x 1e6 namespaces, Zipf(1.1), R=3 (the 1-GPU reference point; 57 GB of inputs)"),
}
SEED = {"c3": 0xB2000003, "c2": 0xB2000002, "c5": 0xB2000005, "c5full": 0xB2000005}
C5_JOB = (1_000_000_000, 1_000_000, 3, 1.1)      # BASELINE.json configs[4]: pods, namespaces, R, zipf s


def ncu_traffic(version: str, workload: str, n: int):
    """dram__bytes_read.sum + dram__bytes_write.sum per K1 launch from the committed `ncu --set full` capture
    (profiles/r02_ncu_traffic.json) — used only when that capture was taken on THIS library version, workload
    and size; otherwise null (the figure is not something this run measured)."""
    p = os.path.join(ROOT, "profiles", "r02_ncu_traffic.json")
    try:
        d = json.load(open(p))
        if d["library_version"] != version or d["workload"] != workload or int(d["pods"]) != n:
            return None
        return d["k1"]["dram_bytes_read"] + d["k1"]["dram_bytes_write"]
    except Exception:
        return None


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        self.device, self.rows, self.proc = device, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.device}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
            except Exception:
                continue
            for nm, val in zip(names, r[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        busy = sorted(x for x in sm if mx is None or x > 0.3 * mx) or sorted(sm)
        med = busy[len(busy) // 2] if busy else None
        return {"sm_mhz": med, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def _parse_cpulist(text: str):
    cpus = []
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus += list(range(int(lo), int(hi or lo) + 1))
    return cpus


def bind_to_gpu_numa_node(device: int) -> dict:
    """Pin this rank's host threads to the CPUs next to its GPU BEFORE any pinned buffer is allocated, so the
    e2e leg's host buffers are local to the GPU's PCIe root.  sysfs first; when it reports node -1 (a VM with
    one flat node) fall back to the CPU affinity column of `nvidia-smi topo -m`.  Returns what was found."""
    info = {"source": None, "node": None, "cpus": None}
    try:
        import torch
        pr = torch.cuda.get_device_properties(device)
        path = f"/sys/bus/pci/devices/{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0/numa_node"
        node = int(open(path).read().strip())
        if node >= 0:
            cpus = _parse_cpulist(open(f"/sys/devices/system/node/node{node}/cpulist").read())
            info.update(source="sysfs", node=node)
        else:
            out = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
            cpus, node = None, None
            for line in out.splitlines():
                cells = line.split()
                if cells and cells[0] == f"GPU{device}":
                    # ... <links per GPU/NIC> <CPU Affinity> <NUMA Affinity> [GPU NUMA ID]
                    aff = [c for c in cells[1:] if c[0].isdigit() and ("-" in c or "," in c or c.isdigit())]
                    if aff:
                        cpus = _parse_cpulist(aff[0])
                        node = int(aff[1]) if len(aff) > 1 and aff[1].isdigit() else None
                    break
            info.update(source="nvidia-smi topo -m" if cpus else "single-node VM (sysfs numa_node = -1, no affinity reported)", node=node)
        if cpus:
            allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
            if allowed:
                os.sched_setaffinity(0, allowed)
                info["cpus"] = f"{allowed[0]}-{allowed[-1]} ({len(allowed)})"
    except Exception as e:  # noqa: BLE001
        info["source"] = f"unavailable ({type(e).__name__})"
    return info


def cpu_info():
    model = "unknown"
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                model = line.split(":", 1)[1].strip()
                break
    except Exception:
        pass
    return len(os.sched_getaffinity(0)) or 1, model


def make_config(desc, n, NS, R, grouped=False, by_index=False):
    """The workload description both arms print (identical keys and values for the same command line)."""
    _, _, bpp = narrow_plan(NS, R)
    k1_bytes, k2_bytes = n * (4 + 8 * R), n * (4 + 8 * R + 1)
    return {"workload": desc, "pods_per_gpu": n, "namespaces_per_gpu": NS, "resources": R,
            "step": "K1 usage_reduce(running set) + K2 quota_eval(candidate set)", "running_set": "grouped" if grouped else "arrival order",
            "l2": f"inputs {(k1_bytes + k2_bytes) / 1e6:.0f} MB per step >> 126 MB L2 (no flush needed)" if k1_bytes > 4e8 else "inputs fit in L2: treat GB/s as cache bandwidth",
            "e2e_host_columns": f"compact, {bpp} B/pod (int64 columns: {4 + 8 * R} B/pod, reported beside)",
            "sharding": ("pods sharded by index, one NCCL int64 all-reduce of the %d-byte usage table per step" % (8 * R * NS)) if by_index
            else "namespace-hash shards, independent per rank (replicas of the workload), no collective; the partitioned one-job run is the c5 sub-record"}


def narrow_plan(NS: int, R: int):
    """The compact host representation of the C3/C5 key set (what the row-A5 flattener would emit):
    namespace id u16 up to 65,536 namespaces, gpu count u8, cpu millicores u16, memory u32 MiB (shift 20),
    mig count u8.  Returns (ns numpy dtype, [(numpy dtype, shift)] per column, bytes per pod)."""
    from workloads import KINDS
    per_kind = {0: (np.uint8, 0), 1: (np.uint16, 0), 2: (np.uint32, 20), 3: (np.uint8, 0)}
    ns_t = np.uint16 if NS <= 65536 else np.uint32
    cols = [per_kind[k] for k in KINDS[R]]
    return ns_t, cols, np.dtype(ns_t).itemsize + sum(np.dtype(t).itemsize for t, _ in cols)


# ----------------------------------------------------------------------------------- the CPU arm
class CpuArm:
    """The reference side of the comparison.  The reference (Rust) has no implementation of this path and
    cannot be built here, so the CPU arm is the oracle port (oracle/quota_ref.c, kind = "port"): OpenMP over all
    host cores, per-thread private usage tables in a workspace allocated once (not inside the timed call),
    compiled -march=native on the box it runs on.  It runs the WHOLE workload — the same running and candidate
    sets as the GPU arm — on the same compact host columns the GPU arm's e2e leg ships over PCIe, and on plain
    int64 columns beside it."""

    def __init__(self, oracle, NS, R, threads):
        self.o, self.NS, self.R, self.T = oracle, NS, R, threads
        self.ws = oracle.usage_workspace(NS, R, threads)
        self.cm = 0b110 & ((1 << R) - 1)

    def prepare(self, run_ns, run_cols, shifts, seed):
        from workloads import make_hard
        self.used0 = self.o.usage_reduce_typed(run_ns, run_cols, shifts, self.NS, threads=self.T, workspace=self.ws)
        self.hard, self.hm = make_hard(self.used0, seed)

    def step_typed(self, run_ns, run_cols, cand_ns, cand_cols, shifts):
        t0 = time.perf_counter()
        u = self.o.usage_reduce_typed(run_ns, run_cols, shifts, self.NS, threads=self.T, workspace=self.ws)
        v = self.o.quota_eval_typed(cand_ns, cand_cols, shifts, None, self.NS, u, self.hard, self.hm, self.cm, threads=self.T)
        return time.perf_counter() - t0, u, v

    def step_i64(self, run_ns, run_req, cand_ns, cand_req):
        t0 = time.perf_counter()
        u = self.o.usage_reduce_ws(run_ns, run_req, self.NS, self.T, self.ws)
        v = self.o.quota_eval(cand_ns, cand_req, None, self.NS, u, self.hard, self.hm, self.cm, threads=self.T)
        return time.perf_counter() - t0, u, v


def narrow_numpy(ns, req, NS, R):
    """int64 generator columns -> the compact columns of narrow_plan(); checks the narrowing is lossless."""
    ns_t, plan, _ = narrow_plan(NS, R)
    cols, shifts = [], []
    for c, (t, sh) in zip(req, plan):
        d = (c >> sh).astype(t)
        if not np.array_equal(d.astype(np.int64) << sh, c):
            raise SystemExit("the synthetic column does not fit its compact type: narrow_plan() is wrong for this workload")
        cols.append(d); shifts.append(sh)
    return ns.astype(ns_t), cols, shifts


def run_reference(args, wl, rank):
    """--impl reference: the CPU arm alone, on the GPU arm's config (the whole workload, every step)."""
    if rank != 0:
        return
    import oracle
    from oracle import quota_ref as Q
    from workloads import KINDS
    oracle.build()
    native = True
    try:
        oracle.use_native_build()
    except Exception:
        native = False
    cores, model = cpu_info()
    n, NS, R, dist, s, desc = wl
    seed = SEED[args.workload]
    a, b = Q.scramble_params(NS)
    cdf = Q.zipf_cdf(NS, s) if dist else None
    t_gen = time.perf_counter()
    run = oracle.gen_fill(oracle.gen_params(seed, 0, NS, dist, 0, n, a, b, cdf, KINDS[R]), 0, n, threads=cores)
    cand = oracle.gen_fill(oracle.gen_params(seed, 1, NS, dist, 0, n, a, b, cdf, KINDS[R]), 0, n, threads=cores)
    r_ns, r_cols, shifts = narrow_numpy(run[0], run[1], NS, R)
    c_ns, c_cols, _ = narrow_numpy(cand[0], cand[1], NS, R)
    t_gen = time.perf_counter() - t_gen
    arm = CpuArm(oracle, NS, R, cores)
    arm.prepare(r_ns, r_cols, shifts, seed)
    for _ in range(max(1, min(args.warmup, 3))):
        arm.step_typed(r_ns, r_cols, c_ns, c_cols, shifts)
    ts = [arm.step_typed(r_ns, r_cols, c_ns, c_cols, shifts)[0] for _ in range(args.steps)]
    t = sum(ts) / len(ts)
    ti = [arm.step_i64(run[0], run[1], cand[0], cand[1])[0] for _ in range(max(2, min(args.steps, 5)))]
    ti = sum(ti[1:]) / len(ti[1:])
    v = n / t
    line = {
        "impl": "reference", "metric": "pods_evaluated_per_s", "value": v, "unit": "pods/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int64", "data": "synthetic",
        "config": make_config(desc, n, NS, R),
        "cpu_baseline": {"value": v, "unit": "pods/s", "cores": cores, "kind": "port", "cpu_model": model,
                         "sample": f"the whole workload every step: {n} running + {n} candidate pods (oracle/quota_ref.c, OpenMP, {cores} threads, "
                                   f"{'-march=native' if native else '-march=x86-64-v3'}, workspace allocated once)",
                         "int64_columns_value": n / ti, "generate_s": t_gen},
        "e2e": {"value": v, "unit": "pods/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------- GPU arm: sub-records
def _median_ms(torch, fn, reps, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def c2_record(torch, bgc, eng, oracle, Q):
    """BASELINE.json configs[1] under the driver's clock: 2^20 pods x 1e4 namespaces, uniform, R=2 (gpu, cpu), one
    K1+K2 step on device-resident columns.  20 MB of input: L2-resident and launch-bound, so the step is timed two
    ways — launched call by call through the C ABI, and replayed from a CUDA graph captured around the same calls
    (four nodes: clear, K1, headroom rebuild, K2)."""
    from workloads import KINDS, UNITS, make_hard
    n, NS, R, seed = 1 << 20, 10_000, 2, SEED["c2"]
    a, b = Q.scramble_params(NS)
    run_ns = torch.empty(n, dtype=torch.int32, device="cuda"); run_req = [torch.empty(n, dtype=torch.int64, device="cuda") for _ in range(R)]
    c_ns = torch.empty(n, dtype=torch.int32, device="cuda"); c_req = [torch.empty(n, dtype=torch.int64, device="cuda") for _ in range(R)]
    v = torch.empty(n, dtype=torch.uint8, device="cuda")
    eng.synth_pods_dev(bgc.synth_params(seed, 0, NS, 0, 0, n, a, b, None, KINDS[R]), 0, n, run_ns, run_req)
    eng.synth_pods_dev(bgc.synth_params(seed, 1, NS, 0, 0, n, a, b, None, KINDS[R]), 0, n, c_ns, c_req)
    eng.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
    eng.usage_reduce_dev(run_ns, run_req, 0)
    torch.cuda.synchronize()
    used0 = eng.usage_get()
    hard, hm = make_hard(used0, seed)
    eng.quota_table_set(hard, hm, UNITS[R])

    def step():
        eng.usage_reduce_dev(run_ns, run_req, 0)
        eng.quota_eval_dev(c_ns, c_req, None, v)
    for _ in range(5):
        step()
    torch.cuda.synchronize()
    reps = 200
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        step()
    e1.record(); torch.cuda.synchronize()
    eager_us = e0.elapsed_time(e1) / reps * 1e3
    rns, rreq, _ = oracle.gen_fill(oracle.gen_params(seed, 0, NS, 0, 0, n, a, b, None, KINDS[R]), 0, n, threads=8)
    cns, creq, _ = oracle.gen_fill(oracle.gen_params(seed, 1, NS, 0, 0, n, a, b, None, KINDS[R]), 0, n, threads=8)
    want_u = oracle.usage_reduce(rns, rreq, NS, threads=8)
    want_v = oracle.quota_eval(cns, creq, None, NS, want_u, hard, hm, 0b10, threads=8)
    ok = bool(np.array_equal(eng.usage_get(), want_u)) and bool(np.array_equal(v.cpu().numpy(), want_v))
    graph_us = None
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            eng.use_torch_stream()
            step()                                   # every kernel already loaded and configured before capture
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                step()
            for _ in range(5):
                g.replay()
            side.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(side)
            for _ in range(reps):
                g.replay()
            e1.record(side); side.synchronize()
            graph_us = e0.elapsed_time(e1) / reps * 1e3
            ok = ok and bool(np.array_equal(v.cpu().numpy(), want_v))
    except Exception as e:  # noqa: BLE001
        graph_us = f"capture failed: {type(e).__name__}: {e}"[:200]
    finally:
        eng.use_torch_stream()
    return {"config": "C2: 2^20 pods x 1e4 namespaces, uniform, R=2 (gpu, cpu); 21 MB per step, L2-resident, launch-bound",
            "step_us_call_by_call": eager_us, "pods_per_s_call_by_call": n / eager_us * 1e6,
            "step_us_cuda_graph_replay": graph_us, "pods_per_s_cuda_graph_replay": (n / graph_us * 1e6) if isinstance(graph_us, float) else None,
            "verified_vs_oracle": ok, "verified_scope": "full usage table and every verdict vs the oracle"}


def k0_record(torch, bgc, eng, oracle, peak):
    """Row A1 under the driver's clock: 2^30 requests through bgc_policy_eval_dev (5 B per request: 3 read, 2 written)."""
    m = 1 << 30
    op = torch.empty(m, dtype=torch.uint8, device="cuda"); fl = torch.empty(m, dtype=torch.int16, device="cuda")
    code = torch.empty(m, dtype=torch.uint8, device="cuda"); pm = torch.empty(m, dtype=torch.uint8, device="cuda")
    eng.synth_policy_dev(0xA1, 0, m, op, fl)
    ms = _median_ms(torch, lambda: eng.policy_eval_dev(op, fl, code, pm), 7, 3)
    cnt = 1 << 20
    rop, rfl = oracle.gen_policy_fill(0xA1, m - cnt, cnt, threads=8)
    rc, rp = oracle.policy_bits(rop, rfl, threads=8)
    ok = oracle.checksum_u8(rc, m - cnt) == eng.checksum_u8_dev(code[m - cnt:], m - cnt) and \
        oracle.checksum_u8(rp, m - cnt) == eng.checksum_u8_dev(pm[m - cnt:], m - cnt)
    gbs = m * 5 / ms / 1e6
    return {"kernel": "k0_policy_eval", "requests": m, "ms": ms, "requests_per_s": m / ms * 1e3, "achieved_GBps": gbs, "frac": gbs / peak,
            "algorithmic_bytes_per_request": 5, "verified_vs_oracle": bool(ok), "verified_scope": "code and patch_mask checksums of the last 2^20 requests"}


def k3_record(torch, bgc, eng, oracle, Q, wl, seed, peak):
    """Row N1 under the driver's clock: greedy in-batch admission of the workload's candidate set (uniform and Zipf)."""
    from workloads import KINDS, UNITS, make_hard
    n, NS, R = min(wl[0], 100_000_000), wl[1], wl[2]
    out = {}
    for name, dist, s in (("uniform", 0, 0.0), ("zipf1.1", 1, 1.1)):
        a, b = Q.scramble_params(NS)
        cdf = Q.zipf_cdf(NS, s) if dist else None
        d_cdf = torch.from_numpy(cdf.view(np.int64)).cuda() if dist else None
        ns = torch.empty(n, dtype=torch.int32, device="cuda")
        req = [torch.empty(n, dtype=torch.int64, device="cuda") for _ in range(R)]
        v = torch.empty(n, dtype=torch.uint8, device="cuda")
        eng.synth_pods_dev(bgc.synth_params(seed, 1, NS, dist, 0, n, a, b, d_cdf, KINDS[R]), 0, n, ns, req)
        # limits = the batch's own total demand x {0.5..2}: about half the pods fit
        eng.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
        eng.usage_reduce_dev(ns, req, 0)
        torch.cuda.synchronize()
        hard, hm = make_hard(eng.usage_get(), seed)
        eng.quota_table_set(hard, hm, UNITS[R])
        zero = np.zeros((R, NS), np.int64)

        def once():
            eng.usage_set(zero)
            eng.admit_greedy_dev(ns, req, None, v)
        once(); torch.cuda.synchronize()
        ts = []
        for _ in range(3):
            eng.usage_set(zero)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); eng.admit_greedy_dev(ns, req, None, v); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = sorted(ts)[1]
        gbs = n * (4 + 8 * R + 1) / ms / 1e6
        out[name] = {"pods": n, "ms": ms, "pods_per_s": n / ms * 1e3, "achieved_GBps": gbs, "frac": gbs / peak,
                     "admitted_fraction": float((v[::101] == 0).float().mean())}
        del ns, req, v
    return {"kernel": "k3_admit_greedy", "algorithmic_bytes_per_pod": 4 + 8 * R + 1, **out,
            "long_segments": os.environ.get("BGC_K3_LONG", "chain"),
            "parity": "bit-exact vs the literal loop in tests/test_gpu_greedy.py; at this size by tools/greedy_probe.py --verify (profiles/r02_greedy_n1_chain.jsonl), not re-checked in this run"}


def crossover_record(torch, bgc, eng, oracle, Q, seed, cores):
    """C4 in this run: bgc_quota_eval from pinned host buffers (copies, launch and sync inside) against the CPU
    port on all cores, batch sizes 2^10..2^18, tables resident on both sides, medians of 50 calls; the
    crossover is the smallest batch from which the GPU's median stays below the CPU's."""
    from workloads import KINDS, UNITS, make_hard
    NS, R, nmax = 10_000, 2, 1 << 18
    a, b = Q.scramble_params(NS)
    run = oracle.gen_fill(oracle.gen_params(seed, 0, NS, 0, 0, nmax, a, b, None, KINDS[R]), 0, nmax, threads=cores)
    used = oracle.usage_reduce(run[0], run[1], NS, threads=cores)
    hard, hm = make_hard(used, seed)
    eng.quota_table_set(hard, hm, UNITS[R])
    eng.usage_set(used)
    h_ns = torch.empty(nmax, dtype=torch.int32, pin_memory=True); h_ns.numpy()[:] = run[0].view(np.int32)
    h_req = [torch.empty(nmax, dtype=torch.int64, pin_memory=True) for _ in range(R)]
    for r in range(R):
        h_req[r].numpy()[:] = run[1][r]
    h_v = torch.empty(nmax, dtype=torch.uint8, pin_memory=True)
    ns_np, req_np = h_ns.numpy().view(np.uint32), [t.numpy() for t in h_req]
    rows, cross = [], None
    for lg in range(10, 19):
        m = 1 << lg
        tg, tc = [], []
        for _ in range(53):
            t0 = time.perf_counter()
            eng.quota_eval_ptr(h_ns.data_ptr(), [t.data_ptr() for t in h_req], 0, m, h_v.data_ptr())
            tg.append(time.perf_counter() - t0)
        for _ in range(53):
            t0 = time.perf_counter()
            oracle.quota_eval(ns_np[:m], [c[:m] for c in req_np], None, NS, used, hard, hm, 0b10, threads=cores)
            tc.append(time.perf_counter() - t0)
        g, c = sorted(tg[3:])[25] * 1e6, sorted(tc[3:])[25] * 1e6
        rows.append({"pods": m, "gpu_us": round(g, 1), "cpu_us": round(c, 1)})
    for i in range(len(rows)):
        if all(r["gpu_us"] < r["cpu_us"] for r in rows[i:]):
            cross = rows[i]["pods"]
            break
    return {"measured": "in this run: bgc_quota_eval (host pointers, pinned; H2D + launch + D2H + sync inside) vs oracle quota_eval on all cores; "
                        "NS=1e4, R=2, tables resident, median of 50", "cpu_threads": cores, "sweep": rows,
            "gpu_wins_from_pinned_host_at_pods": cross if cross else "never up to 262144",
            "from_pageable_host": "see profiles/r02_crossover_c4.md (tools/crossover.cpp sweep)"}


def c5_record(torch, dist, bgc, eng, oracle, Q, rank, world, steps, n_total, NS, R, zs, peak, verify_full=False):
    """BASELINE.json configs[4] as ONE job: a single global stream of n_total pods over NS namespaces, Zipf(zs),
    one seed, split over `world` GPUs two ways: (a) by namespace hash — rank g keeps the pods with
    mix32(ns) % world == g, owns that slice of the tables with dense local ids, no collective, shards unequal
    under Zipf (imbalance = max/mean shard size); (b) by pod index — equal slices, every rank reduces into a full
    table, ONE int64 all-reduce of 8*R*NS bytes, then evaluates its slice.  Strong scaling: whole-job pods/s =
    n_total / (max over ranks of the device time of one K1+K2 step)."""
    from workloads import KINDS, UNITS, make_hard
    seed = SEED["c5"]
    dev = torch.device("cuda", torch.cuda.current_device())
    a, b = Q.scramble_params(NS)
    cdf = Q.zipf_cdf(NS, zs)
    d_cdf = torch.from_numpy(cdf.view(np.int64)).to(dev)
    cm = 0b110 & ((1 << R) - 1)
    TAIL = min(n_total, 1 << 22)

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step_k1, step_k2):
        for _ in range(2):
            step_k1(); step_k2()
        sync_all()
        ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(steps)]
        for k in range(steps):
            ev[k][0].record(); step_k1(); ev[k][1].record(); step_k2(); ev[k][2].record()
        sync_all()
        t = torch.tensor([sum(e[0].elapsed_time(e[2]) for e in ev) / steps, sum(e[0].elapsed_time(e[1]) for e in ev) / steps,
                          sum(e[1].elapsed_time(e[2]) for e in ev) / steps], device=dev, dtype=torch.float64)
        mine = t.clone()
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.tolist(), mine.tolist()

    def gather_counts(x):
        t = torch.tensor([x], device=dev, dtype=torch.int64)
        if world == 1:
            return [int(x)]
        out = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return [int(o.item()) for o in out]

    def all_ok(ok):
        t = torch.tensor([1 if ok else 0], device=dev, dtype=torch.int64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return bool(t.item())

    rec = {"job": f"C5 as one job: {n_total} pods x {NS} namespaces, Zipf({zs}), R={R}, one global stream (seed {seed:#x}) over {world} GPU(s); strong scaling",
           "pods": n_total, "namespaces": NS, "steps": steps}

    # ---- (a) namespace-hash shards of the one stream -------------------------------------------------
    t0 = time.perf_counter()
    owned, local_of = bgc.shard.local_ids(NS, world, rank)
    NSg = int(owned.shape[0])
    d_local = torch.from_numpy(local_of.view(np.int32)).to(dev) if world > 1 else None
    sets = []
    for set_ in (0, 1):
        sp = bgc.synth_params(seed, set_, NS, 1, 0, n_total, a, b, d_cdf, KINDS[R])
        m = n_total if world == 1 else eng.synth_pods_shard_dev(sp, 0, n_total, world, rank)
        ns = torch.empty(max(m, 1), dtype=torch.int32, device=dev)
        req = [torch.empty(max(m, 1), dtype=torch.int64, device=dev) for _ in range(R)]
        if world == 1:
            eng.synth_pods_dev(sp, 0, n_total, ns, req)
        else:
            eng.synth_pods_shard_dev(sp, 0, n_total, world, rank, d_local, 0, ns, req)
        sets.append((m, ns, req))
    torch.cuda.synchronize()
    gen_s = time.perf_counter() - t0
    (m_run, r_ns, r_req), (m_cand, c_ns, c_req) = sets
    d_v = torch.empty(max(m_cand, 1), dtype=torch.uint8, device=dev)
    eng.quota_table_set(np.zeros((R, NSg), np.int64), np.zeros(NSg, np.uint8), UNITS[R])
    eng.usage_reduce_dev(r_ns, r_req, 0, None, n=m_run)
    torch.cuda.synchronize()
    used0 = eng.usage_get()
    hard, hm = make_hard(used0, seed)
    eng.quota_table_set(hard, hm, UNITS[R])
    (tot, k1, k2), mine = timed(lambda: eng.usage_reduce_dev(r_ns, r_req, 0, None, n=m_run),
                                lambda: eng.quota_eval_dev(c_ns, c_req, None, d_v, n=m_cand))
    variant = eng.last_k1_variant()
    # checks (outside the timed region): K1 by linearity (sum of the table = sum of the column), K2 by the verdict
    # checksum of the shard's tail against the oracle on the regenerated, filtered tail of the global stream
    used_now = eng.usage_get()
    ok = bool(np.array_equal(used_now, used0))
    with np.errstate(over="ignore"):
        for r in range(R):
            ok = ok and int(used_now[r].sum()) == int(r_req[r][:m_run].sum().item())
    p1 = oracle.gen_params(seed, 1, NS, 1, 0, n_total, a, b, cdf, KINDS[R])
    tns, treq, _ = oracle.gen_fill(p1, n_total - TAIL, TAIL, threads=8)
    keep = np.nonzero(bgc.shard.shard_of(tns, world) == rank)[0] if world > 1 else np.arange(TAIL)
    lns = local_of[tns[keep]] if world > 1 else tns
    tv = oracle.quota_eval(lns, [c[keep] for c in treq], None, NSg, used_now, hard, hm, cm, threads=8)
    mt = int(keep.shape[0])
    ok = ok and oracle.checksum_u8(tv, 0) == eng.checksum_u8_dev(d_v[m_cand - mt:m_cand], 0, mt)
    if verify_full:   # the oracle streams the WHOLE global running set, filtered to this shard
        p0 = oracle.gen_params(seed, 0, NS, 1, 0, n_total, a, b, cdf, KINDS[R])
        acc = np.zeros((R, NSg), np.int64)
        cores_v = os.cpu_count() or 1
        for lo in range(0, n_total, 1 << 24):
            rns, rreq, _ = oracle.gen_fill(p0, lo, min(1 << 24, n_total - lo), threads=cores_v)
            kp = np.nonzero(bgc.shard.shard_of(rns, world) == rank)[0] if world > 1 else slice(None)
            l2 = local_of[rns[kp]] if world > 1 else rns
            with np.errstate(over="ignore"):
                acc += oracle.usage_reduce(l2, [c[kp] for c in rreq], NSg, threads=cores_v)
        ok = ok and bool(np.array_equal(acc, used_now))
    run_counts, cand_counts = gather_counts(m_run), gather_counts(m_cand)
    step_ms_per_rank = gather_counts(int(mine[0] * 1e6))
    rec["namespace_hash"] = {
        "sharding": "rank = mix32(ns_id) % G on the one global stream; dense per-shard namespace ids; no collective; host gathers verdicts",
        "running_pods_per_rank": run_counts, "candidate_pods_per_rank": cand_counts, "namespaces_on_rank0": NSg,
        "imbalance_max_over_mean": max(cand_counts) / (sum(cand_counts) / world),
        "ms_per_step": tot, "k1_ms": k1, "k2_ms": k2, "step_ms_per_rank": [x / 1e6 for x in step_ms_per_rank],
        "pods_per_s": n_total / (tot * 1e-3), "k1_variant": variant,
        "k1_frac_of_hbm_peak_on_slowest_rank": max(run_counts) * (4 + 8 * R) / (k1 * 1e-3) / 1e9 / peak,
        "generate_and_partition_s_on_device": gen_s,
        "verified_vs_oracle": all_ok(ok),
        "verified_scope": ("full shard usage table vs the oracle streamed over the whole global stream + " if verify_full else "usage table sums = column sums + ")
                          + f"verdict checksum of each shard's share of the last {TAIL} candidates",
    }
    del sets, r_ns, r_req, c_ns, c_req, d_v, d_local
    torch.cuda.empty_cache()

    # ---- (b) pod-index slices + one int64 all-reduce of the usage table ---------------------------------
    if world > 1:
        lo, hi = bgc.shard.index_slice(n_total, world, rank)
        m = hi - lo
        r_ns = torch.empty(m, dtype=torch.int32, device=dev); r_req = [torch.empty(m, dtype=torch.int64, device=dev) for _ in range(R)]
        c_ns = torch.empty(m, dtype=torch.int32, device=dev); c_req = [torch.empty(m, dtype=torch.int64, device=dev) for _ in range(R)]
        d_v = torch.empty(m, dtype=torch.uint8, device=dev)
        d_used = torch.empty(R * NS, dtype=torch.int64, device=dev)
        eng.synth_pods_dev(bgc.synth_params(seed, 0, NS, 1, 0, n_total, a, b, d_cdf, KINDS[R]), lo, m, r_ns, r_req)
        eng.synth_pods_dev(bgc.synth_params(seed, 1, NS, 1, 0, n_total, a, b, d_cdf, KINDS[R]), lo, m, c_ns, c_req)
        eng.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])

        def k1_exchange():
            eng.usage_reduce_dev(r_ns, r_req, 0)
            eng.usage_get_dev(d_used)
            dist.all_reduce(d_used, op=dist.ReduceOp.SUM)      # the one exchange step: 8*R*NS bytes over NVLink/NVSwitch
            eng.usage_set_dev(d_used)
        k1_exchange()
        torch.cuda.synchronize()
        used0 = eng.usage_get()
        hard, hm = make_hard(used0, seed)
        eng.quota_table_set(hard, hm, UNITS[R])
        (tot, k1, k2), mine = timed(k1_exchange, lambda: eng.quota_eval_dev(c_ns, c_req, None, d_v))
        # the all-reduce alone
        ar_ms = _median_ms(torch, lambda: dist.all_reduce(d_used, op=dist.ReduceOp.SUM), 5, 2)
        k1_exchange(); torch.cuda.synchronize()
        used_now = eng.usage_get()
        ok = bool(np.array_equal(used_now, used0))
        T2 = min(m, TAIL)
        tns, treq, _ = oracle.gen_fill(p1, hi - T2, T2, threads=8)
        tv = oracle.quota_eval(tns, treq, None, NS, used_now, hard, hm, cm, threads=8)
        ok = ok and oracle.checksum_u8(tv, 0) == eng.checksum_u8_dev(d_v[m - T2:], 0, T2)
        if verify_full:   # the summed table against the oracle streamed over the WHOLE global running set
            p0 = oracle.gen_params(seed, 0, NS, 1, 0, n_total, a, b, cdf, KINDS[R])
            acc = np.zeros((R, NS), np.int64)
            cores_v = os.cpu_count() or 1
            for lo2 in range(0, n_total, 1 << 24):
                rns, rreq, _ = oracle.gen_fill(p0, lo2, min(1 << 24, n_total - lo2), threads=cores_v)
                with np.errstate(over="ignore"):
                    acc += oracle.usage_reduce(rns, rreq, NS, threads=cores_v)
            ok = ok and bool(np.array_equal(acc, used_now))
        rec["index_allreduce"] = {
            "sharding": f"rank g takes pods [g*n/G, (g+1)*n/G) of the same stream; every rank reduces into a full {R}x{NS} table; "
                        f"ONE NCCL int64 all-reduce of {8 * R * NS} bytes per step; each rank evaluates its own slice",
            "pods_per_rank": gather_counts(m), "imbalance_max_over_mean": 1.0,
            "ms_per_step": tot, "k1_plus_exchange_ms": k1, "k2_ms": k2, "allreduce_alone_ms": ar_ms, "pods_per_s": n_total / (tot * 1e-3),
            "k1_variant": eng.last_k1_variant(), "verified_vs_oracle": all_ok(ok),
            "verified_scope": ("summed usage table vs the oracle streamed over the whole global stream + " if verify_full else "summed usage table stable across steps + ")
                              + f"verdict checksum of the last {T2} candidates of each slice vs the oracle on the global usage",
        }
        del r_ns, r_req, c_ns, c_req, d_v, d_used
        torch.cuda.empty_cache()
    else:
        rec["index_allreduce"] = "identical to namespace_hash on one GPU (one slice, no exchange)"
    return rec


# ----------------------------------------------------------------------------------- GPU arm: main
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--pods", type=int, default=0, help="override pods per GPU")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-sub", action="store_true", help="skip the k0 / k3 / crossover sub-records (1-GPU runs)")
    ap.add_argument("--no-c5", action="store_true", help="skip the c5 sub-record (one 1e9 x 1e6 job over the N GPUs)")
    ap.add_argument("--c5-pods", type=int, default=C5_JOB[0])
    ap.add_argument("--c5-namespaces", type=int, default=C5_JOB[1])
    ap.add_argument("--k1", type=int, default=0, help="force a K1 variant (BGC_K1_*)")
    ap.add_argument("--grouped", action="store_true", help="running set grouped by namespace (segmented path)")
    ap.add_argument("--verify-full", action="store_true",
                    help="after timing, stream the whole running set through the CPU oracle in chunks and compare the full usage table")
    ap.add_argument("--sharding", default="namespace", choices=["namespace", "index"],
                    help="top-level metric on N GPUs: independent namespace-hash shards (no collective, default) or pods sharded by index "
                         "with one int64 all-reduce of the usage table per step (SURVEY.md §8(e)(b), row N5)")
    args = ap.parse_args()
    wl = list(WORKLOADS[args.workload])
    if args.pods:
        wl[0] = args.pods
    wl = tuple(wl)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, wl, rank)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local)
    numa = bind_to_gpu_numa_node(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    from workloads import make_hard, KINDS, UNITS
    from oracle import quota_ref as Q  # zipf table + scramble parameters are INPUTS shared with the checker
    import oracle
    oracle.build()
    bgc = importlib.import_module("bacchus-gpu-controller_b200")
    eng = bgc.Engine(local)
    eng.use_torch_stream()
    if args.k1:
        eng.set_k1_variant(args.k1)

    n, NS, R, dist_kind, zs, desc = wl
    by_index = args.sharding == "index" and world > 1
    # namespace sharding: every rank owns an independent namespace shard (own seed, no exchange);
    # index sharding: all ranks see the same namespaces, rank r takes pods [r*n, (r+1)*n) of one global stream
    seed = SEED[args.workload] + (0 if by_index else 7919 * rank)
    first = rank * n if by_index else 0               # position of this rank's slice in the global stream
    n_total = world * n if by_index else n
    a, b = Q.scramble_params(NS)
    cdf = Q.zipf_cdf(NS, zs) if dist_kind else None
    d_cdf = torch.from_numpy(cdf.view(np.int64)).cuda() if dist_kind else None
    dev = torch.device("cuda", local)
    peak, peak_src = peaks()

    def alloc_set():
        return (torch.empty(n, dtype=torch.int32, device=dev), [torch.empty(n, dtype=torch.int64, device=dev) for _ in range(R)])
    run_ns, run_req = alloc_set()
    cand_ns, cand_req = alloc_set()
    d_verdict = torch.empty(n, dtype=torch.uint8, device=dev)
    eng.synth_pods_dev(bgc.synth_params(seed, 0, NS, dist_kind, int(args.grouped), n_total, a, b, d_cdf, KINDS[R]), first, n, run_ns, run_req)
    eng.synth_pods_dev(bgc.synth_params(seed, 1, NS, dist_kind, 0, n_total, a, b, d_cdf, KINDS[R]), first, n, cand_ns, cand_req)
    # quota table: hard drawn from the running set's own usage so that about half the verdicts deny
    eng.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
    flags = bgc.REDUCE_GROUPED if args.grouped else 0
    d_used = torch.empty(R * NS, dtype=torch.int64, device=dev)

    def reduce_usage():
        """K1 over this rank's running pods; with index sharding, plus the one exchange step: an int64
        all-reduce (NCCL over NVLink/NVSwitch) of the R x NS usage table, written back for K2."""
        eng.usage_reduce_dev(run_ns, run_req, flags)
        if by_index:
            eng.usage_get_dev(d_used)
            dist.all_reduce(d_used, op=dist.ReduceOp.SUM)
            eng.usage_set_dev(d_used)

    reduce_usage()
    torch.cuda.synchronize()
    used0 = eng.usage_get()
    hard, hard_mask = make_hard(used0, seed)
    eng.quota_table_set(hard, hard_mask, UNITS[R])

    def step():
        reduce_usage()
        eng.quota_eval_dev(cand_ns, cand_req, None, d_verdict)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # clocks are sampled from just before the warm-up to just after the timed region: the timed
    # region alone (K steps of ~1 ms) is shorter than nvidia-smi's sampling period
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    # keep the GPU under the same load for ~0.5 s so that the clock sampler sees it.  The number of
    # extra steps is decided by rank 0 and broadcast: with index sharding every step contains a
    # collective, so all ranks must run exactly the same number of steps.
    t_w = time.perf_counter()
    step()
    torch.cuda.synchronize()
    extra = torch.tensor([int(min(2000, max(0, 0.5 / max(time.perf_counter() - t_w, 1e-5))))], device=dev, dtype=torch.int64)
    if world > 1:
        dist.broadcast(extra, src=0)
    for _ in range(int(extra.item())):
        step()
    barrier()
    launches0 = eng.timings()["launches"]
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    barrier()
    t_all0 = torch.cuda.Event(enable_timing=True); t_all1 = torch.cuda.Event(enable_timing=True)
    t_all0.record()
    for k in range(args.steps):
        ev[k][0].record()
        reduce_usage()
        ev[k][1].record()
        eng.quota_eval_dev(cand_ns, cand_req, None, d_verdict)
        ev[k][2].record()
    t_all1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = eng.timings()["launches"] - launches0
    total_ms = t_all0.elapsed_time(t_all1)
    k1_ms = sum(e[0].elapsed_time(e[1]) for e in ev) / args.steps
    k2_ms = sum(e[1].elapsed_time(e[2]) for e in ev) / args.steps
    if world > 1:
        t = torch.tensor([total_ms, k1_ms, k2_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, k1_ms, k2_ms = t.tolist()
    ms_per_step = total_ms / args.steps
    value = world * n / (ms_per_step * 1e-3)
    k1_variant = eng.last_k1_variant()

    # ---- parity spot-check of what was just timed (checker only; not in any timed region)
    verified = None
    verified_full = False
    cnt = min(n, 1 << 20)
    v_tail = None
    if rank == 0:
        p1 = oracle.gen_params(seed, 1, NS, dist_kind, 0, n_total, a, b, cdf, KINDS[R])
        cns, creq, _ = oracle.gen_fill(p1, first + n - cnt, cnt, threads=8)
        used_now = eng.usage_get()
        v_tail = oracle.quota_eval(cns, creq, None, NS, used_now, hard, hard_mask, 0b110 & ((1 << R) - 1), threads=8)
        verified = bool(np.array_equal(used_now, used0)) and oracle.checksum_u8(v_tail, n - cnt) == eng.checksum_u8_dev(d_verdict[n - cnt:], n - cnt)
        if n_total <= (1 << 25) or args.verify_full:
            # the oracle regenerates the same counter-based stream in chunks (no 28 GB host buffer)
            p0 = oracle.gen_params(seed, 0, NS, dist_kind, int(args.grouped), n_total, a, b, cdf, KINDS[R])
            acc = np.zeros((R, NS), np.int64)
            cores_v = os.cpu_count() or 1
            for lo in range(0, n_total, 1 << 24):
                rns, rreq, _ = oracle.gen_fill(p0, lo, min(1 << 24, n_total - lo), threads=cores_v)
                with np.errstate(over="ignore"):
                    acc += oracle.usage_reduce(rns, rreq, NS, threads=cores_v)
            verified = verified and bool(np.array_equal(acc, used_now))
            verified_full = True

    # ---- end to end through the host-pointer entry points (pinned host buffers)
    e2e = e2e_i64 = None
    host = None
    ns_t, plan, bpp = narrow_plan(NS, R)
    if not args.no_e2e and not by_index:
        tt = {np.dtype(np.uint8): (torch.uint8, bgc.T_U8, 0xFF), np.dtype(np.uint16): (torch.int16, bgc.T_U16, 0xFFFF),
              np.dtype(np.uint32): (torch.int32, bgc.T_U32, 0xFFFFFFFF)}

        def to_host_narrow(ns, req):
            """device int64 columns -> pinned host columns of narrow_plan() (lossless, checked on the device)"""
            tdt, tcode, msk = tt[np.dtype(ns_t)]
            h_ns = torch.empty(n, dtype=tdt, pin_memory=True); h_ns.copy_(ns.to(tdt))
            cols, spec = [], []
            for c, (t, sh) in zip(req, plan):
                tdt2, code2, msk2 = tt[np.dtype(t)]
                d = (c >> sh).to(tdt2)
                if not bool((((d.to(torch.int64) & msk2) << sh) == c).all()):
                    raise SystemExit("the synthetic column does not fit its compact type: narrow_plan() is wrong for this workload")
                h = torch.empty(n, dtype=tdt2, pin_memory=True); h.copy_(d)
                cols.append(h); spec.append((h.data_ptr(), code2, sh))
                del d
            return h_ns, tcode, cols, bgc.make_cols(spec)
        hr_ns, ns_code, hr_cols, hr_spec = to_host_narrow(run_ns, run_req)
        hc_ns, _, hc_cols, hc_spec = to_host_narrow(cand_ns, cand_req)
        want_i64 = world == 1
        if want_i64:
            h_run_ns = torch.empty(n, dtype=torch.int32, pin_memory=True); h_run_ns.copy_(run_ns)
            h_run_req = [torch.empty(n, dtype=torch.int64, pin_memory=True) for _ in range(R)]
            h_cand_ns = torch.empty(n, dtype=torch.int32, pin_memory=True); h_cand_ns.copy_(cand_ns)
            h_cand_req = [torch.empty(n, dtype=torch.int64, pin_memory=True) for _ in range(R)]
            for r in range(R):
                h_run_req[r].copy_(run_req[r]); h_cand_req[r].copy_(cand_req[r])
        h_verdict = torch.empty(n, dtype=torch.uint8, pin_memory=True)
        torch.cuda.synchronize()
        del run_ns, run_req, cand_ns, cand_req
        torch.cuda.empty_cache()
        eng.set_stream(None)

        def time_e2e(fn, steps):
            for _ in range(2):
                fn()
            barrier()
            t0 = time.perf_counter()
            for _ in range(steps):
                fn()              # each host-pointer call returns after its last D2H has completed
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            if world > 1:
                t = torch.tensor([dt], device=dev, dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                dt = t.item()
            return dt

        def step_typed():
            eng.usage_reduce_typed_ptr(hr_ns.data_ptr(), ns_code, hr_spec, n, 0, 0)
            eng.quota_eval_typed_ptr(hc_ns.data_ptr(), ns_code, hc_spec, 0, n, h_verdict.data_ptr())
        dt = time_e2e(step_typed, args.steps)
        tm = eng.timings()
        if rank == 0 and verified is not None:
            verified = verified and bool(np.array_equal(eng.usage_get(), used0)) and \
                oracle.checksum_u8(v_tail, n - cnt) == oracle.checksum_u8(h_verdict[n - cnt:].numpy(), n - cnt)
        e2e = {"value": world * n * args.steps / dt, "unit": "pods/s", "h2d_bytes_per_step": world * 2 * n * bpp,
               "d2h_bytes_per_step": world * n, "ms_per_step": dt / args.steps * 1e3, "host_memory": "pinned",
               "host_columns": f"compact, {bpp} B/pod ({np.dtype(ns_t).name} namespace id, " + ", ".join(f"{np.dtype(t).name}<<{s}" for t, s in plan) + "), widened inside the K1/K2 loads",
               "entry_points": "bgc_usage_reduce_typed + bgc_quota_eval_typed", "numa": numa,
               "last_quota_eval_call": {"h2d_ms": tm["h2d_ns"] / 1e6, "kernel_ms": tm["kernel_ns"] / 1e6, "d2h_ms": tm["d2h_ns"] / 1e6}}
        if want_i64:
            def step_i64():
                eng.usage_reduce_ptr(h_run_ns.data_ptr(), [t.data_ptr() for t in h_run_req], n, flags, 0)
                eng.quota_eval_ptr(h_cand_ns.data_ptr(), [t.data_ptr() for t in h_cand_req], 0, n, h_verdict.data_ptr())
            st = max(2, min(args.steps, 5))
            dt = time_e2e(step_i64, st)
            if rank == 0 and verified is not None:
                verified = verified and oracle.checksum_u8(v_tail, n - cnt) == oracle.checksum_u8(h_verdict[n - cnt:].numpy(), n - cnt)
            e2e_i64 = {"value": n * st / dt, "unit": "pods/s", "h2d_bytes_per_step": 2 * n * (4 + 8 * R), "d2h_bytes_per_step": n,
                       "ms_per_step": dt / st * 1e3, "steps": st, "entry_points": "bgc_usage_reduce + bgc_quota_eval (int64 columns, as in round 1)"}
            host = (h_run_ns, h_run_req, h_cand_ns, h_cand_req)
    else:
        del run_ns, run_req, cand_ns, cand_req
        torch.cuda.empty_cache()

    # ---- CPU baseline beside it (rank 0, the WHOLE workload on all host cores, same host columns as the e2e leg)
    cpu = None
    if rank == 0 and not args.no_cpu and world == 1 and e2e is not None:
        native = True
        try:
            oracle.use_native_build()
        except Exception:
            native = False
        cores, model = cpu_info()
        arm = CpuArm(oracle, NS, R, cores)
        npv = lambda t, dt: t.numpy().view(dt)   # noqa: E731
        r_ns, c_ns = npv(hr_ns, ns_t), npv(hc_ns, ns_t)
        r_cols = [npv(h, t) for h, (t, _) in zip(hr_cols, plan)]
        c_cols = [npv(h, t) for h, (t, _) in zip(hc_cols, plan)]
        shifts = [s for _, s in plan]
        arm.hard, arm.hm = hard, hard_mask
        t1, u, v = arm.step_typed(r_ns, r_cols, c_ns, c_cols, shifts)
        cpu_agrees = bool(np.array_equal(u, used0)) and oracle.checksum_u8(v[n - cnt:], n - cnt) == oracle.checksum_u8(v_tail, n - cnt)
        reps = int(max(2, min(10, 10.0 / max(t1, 1e-3))))
        ts = [arm.step_typed(r_ns, r_cols, c_ns, c_cols, shifts)[0] for _ in range(reps)]
        cpu = {"value": n / (sum(ts) / len(ts)), "unit": "pods/s", "cores": cores, "kind": "port", "cpu_model": model,
               "sample": f"the whole workload: {n} running + {n} candidate pods per step x {reps} steps, on the e2e leg's compact host columns "
                         f"(oracle/quota_ref.c, OpenMP, {'-march=native' if native else '-march=x86-64-v3'}, workspace allocated once)",
               "agrees_with_gpu": cpu_agrees}
        if host is not None:
            hn = [host[0].numpy().view(np.uint32), [t.numpy() for t in host[1]], host[2].numpy().view(np.uint32), [t.numpy() for t in host[3]]]
            arm.step_i64(*hn)
            ti = [arm.step_i64(*hn)[0] for _ in range(3)]
            cpu["int64_columns_value"] = n / (sum(ti) / len(ti))
        one = CpuArm(oracle, NS, R, 1)
        one.hard, one.hm = hard, hard_mask
        ms = min(n, 1 << 22)
        one.step_typed(r_ns[:ms], [c[:ms] for c in r_cols], c_ns[:ms], [c[:ms] for c in c_cols], shifts)
        cpu["single_thread_pods_per_s"] = ms / min(one.step_typed(r_ns[:ms], [c[:ms] for c in r_cols], c_ns[:ms], [c[:ms] for c in c_cols], shifts)[0] for _ in range(2))
    host = None

    # ---- sub-records: the other kernels under the same clock, the in-run crossover, and C5 as one job
    eng.use_torch_stream()
    subs = {}
    if rank == 0 and world == 1 and not args.no_sub and e2e is not None:
        # the same step on the compact columns RESIDENT in HBM (the typed _dev entry points): what the kernels do
        # when the bytes per pod drop from 28 to 9 — no longer HBM-bound but bound by the shared-memory atomics
        try:
            d_rns, d_cns = hr_ns.to(dev), hc_ns.to(dev)
            d_rc, d_cc = [h.to(dev) for h in hr_cols], [h.to(dev) for h in hc_cols]
            types = [tt[np.dtype(t)][1] for t, _ in plan]
            shifts = [sft for _, sft in plan]
            d_v2 = torch.empty(n, dtype=torch.uint8, device=dev)
            eng.quota_table_set(hard, hard_mask, UNITS[R])
            k1t = _median_ms(torch, lambda: eng.usage_reduce_typed_dev(d_rns, d_rc, shifts, 0, None, types=types), 7, 2)
            k2t = _median_ms(torch, lambda: eng.quota_eval_typed_dev(d_cns, d_cc, shifts, None, d_v2, types=types), 7, 2)
            same = bool(np.array_equal(eng.usage_get(), used0)) and eng.checksum_u8_dev(d_v2[n - cnt:], n - cnt) == oracle.checksum_u8(v_tail, n - cnt)
            subs["compact_columns_device_resident"] = {
                "entry_points": "bgc_usage_reduce_typed_dev + bgc_quota_eval_typed_dev", "bytes_per_pod": {"k1": bpp, "k2": bpp + 1},
                "k1_ms": k1t, "k2_ms": k2t, "pods_per_s": n / ((k1t + k2t) * 1e-3),
                "k1_GBps": n * bpp / k1t / 1e6, "k1_frac_of_hbm_peak": n * bpp / k1t / 1e6 / peak,
                "k2_GBps": n * (bpp + 1) / k2t / 1e6, "k2_frac_of_hbm_peak": n * (bpp + 1) / k2t / 1e6 / peak,
                "verified_vs_oracle": same}
            del d_rns, d_cns, d_rc, d_cc, d_v2
        except Exception as e:  # noqa: BLE001
            subs["compact_columns_device_resident"] = {"error": f"{type(e).__name__}: {e}"[:300]}
        torch.cuda.empty_cache()
    if rank == 0 and world == 1 and not args.no_sub:
        for name, fn in (("c2", lambda: c2_record(torch, bgc, eng, oracle, Q)),
                         ("k0", lambda: k0_record(torch, bgc, eng, oracle, peak)),
                         ("k3", lambda: k3_record(torch, bgc, eng, oracle, Q, wl, seed, peak)),
                         ("crossover", lambda: crossover_record(torch, bgc, eng, oracle, Q, SEED["c2"] + 2, cpu_info()[0]))):
            try:
                subs[name] = fn()
            except Exception as e:  # noqa: BLE001  (a failing sub-record must not lose the line)
                subs[name] = {"error": f"{type(e).__name__}: {e}"[:300]}
            torch.cuda.empty_cache()
    c5 = None
    if not args.no_c5:
        try:
            c5 = c5_record(torch, dist, bgc, eng, oracle, Q, rank, world, max(3, min(args.steps, 10)), args.c5_pods, args.c5_namespaces,
                           C5_JOB[2], C5_JOB[3], peak, verify_full=args.verify_full)
        except Exception as e:  # noqa: BLE001
            c5 = {"error": f"{type(e).__name__}: {e}"[:300]}

    if rank == 0:
        k1_bytes = n * (4 + 8 * R)
        achieved = k1_bytes / (k1_ms * 1e-3) / 1e9
        k2_bytes = n * (4 + 8 * R + 1)
        traffic = None if args.grouped or args.k1 else ncu_traffic(bgc.version(), args.workload, n)
        line = {
            "metric": "pods_evaluated_per_s", "value": value, "unit": "pods/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int64", "data": "synthetic",
            "config": make_config(desc, n, NS, R, args.grouped, by_index), "k1_variant": k1_variant, "library": bgc.version(),
            "roofline": {"bound": "hbm", "kernel": "k1_scatter (usage_reduce)", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic,
                         "traffic_source": "profiles/r02_ncu_traffic.json (ncu --set full of this library version, workload and size)" if traffic else None,
                         "algorithmic_bytes_per_launch": k1_bytes, "peak_source": peak_src, "frac_of_8000_datasheet": achieved / 8000.0,
                         "algorithmic_bytes_per_pod": 4 + 8 * R, "k1_ms": k1_ms,
                         "k2": {"ms": k2_ms, "achieved": k2_bytes / (k2_ms * 1e-3) / 1e9, "frac": k2_bytes / (k2_ms * 1e-3) / 1e9 / peak,
                                "algorithmic_bytes_per_pod": 4 + 8 * R + 1}},
            "cpu_baseline": cpu, "e2e": e2e, "e2e_int64_columns": e2e_i64, "clocks": clocks, "gpu_launches": int(launches), "verified_vs_oracle": verified,
            "verified_scope": ("full usage table (oracle streamed over all %d running pods) + verdict checksum of the last 2^20 candidates" % n_total) if verified_full
                              else "usage table stable across steps + verdict checksum of the last 2^20 candidates vs oracle (device-resident, typed e2e and int64 e2e legs)",
            **subs, "c5": c5,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    eng.close()


if __name__ == "__main__":
    main()
