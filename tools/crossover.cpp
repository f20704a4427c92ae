// tools/crossover.cpp — C4 latency sweep without interpreter overhead (BASELINE.json configs[3]).
//
// One quota evaluation end to end through the host-pointer C ABI of libbgc.so (H2D + launch + D2H
// + sync inside the timed region, quota tables resident), from pinned and from pageable host
// memory, next to the CPU oracle port (oracle/quota_ref.c; 1 thread and all cores) on the same
// batch.  Also the policy decision (K0 vs the restated mutate()).  Benchmark harness: like
// bench.py's cpu_baseline leg it may link the oracle; the product never does.
//
//   build: make -C tools     run: tools/crossover_bench [max_log2] > gpurun_out/crossover_c.jsonl
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include <cuda_runtime.h>

#include "../include/bgc.h"
#include "../oracle/policy_ref.h"
#include "../oracle/quota_ref.h"

using clk = std::chrono::steady_clock;

template <class F>
static void stats(F&& fn, int reps, double* med, double* p99) {
  for (int i = 0; i < std::min(reps, 20); ++i) fn();
  std::vector<double> t(reps);
  for (int i = 0; i < reps; ++i) {
    auto a = clk::now();
    fn();
    t[i] = std::chrono::duration<double, std::micro>(clk::now() - a).count();
  }
  std::sort(t.begin(), t.end());
  *med = t[reps / 2];
  *p99 = t[std::min(reps - 1, (int)(reps * 0.99))];
}

#define CK(x) do { int rc_ = (x); if (rc_) { fprintf(stderr, "%s failed: %d %s\n", #x, rc_, bgc_last_error(ctx)); return 1; } } while (0)

int main(int argc, char** argv) {
  const int max_log2 = argc > 1 ? atoi(argv[1]) : 24;
  const int cores = (int)std::thread::hardware_concurrency();
  const size_t nmax = (size_t)1 << max_log2;
  const uint32_t R = 2;
  bgc_ctx* ctx = nullptr;
  if (bgc_ctx_create(0, &ctx)) { fprintf(stderr, "no CUDA device\n"); return 1; }

  for (uint32_t NS : {100u, 10000u}) {
    gen_ref_params gp{};
    gp.seed = 0xB2000004; gp.ns_count = NS; gp.dist = GEN_DIST_UNIFORM; gp.n_total = nmax;
    gp.scramble_a = 1; gp.scramble_b = 0; gp.R = R; gp.kind[0] = GEN_KIND_GPU; gp.kind[1] = GEN_KIND_CPU_MILLI;
    // running set → usage; hard = used * {0.5..2} roughly: here simply used (+/-) so verdicts split
    std::vector<uint32_t> ns(nmax);
    std::vector<int64_t> q0(nmax), q1(nmax);
    int64_t* qs[2] = {q0.data(), q1.data()};
    gp.set = 0;
    gen_ref_fill(&gp, 0, (size_t)1 << 20, ns.data(), qs, nullptr, cores);
    std::vector<int64_t> used((size_t)R * NS), hard((size_t)R * NS);
    const int64_t* cq[2] = {q0.data(), q1.data()};
    quota_ref_usage_reduce(ns.data(), cq, (size_t)1 << 20, NS, R, used.data(), cores);
    std::vector<uint8_t> hm(NS, 3), unit = {BGC_UNIT_COUNT, BGC_UNIT_MILLI};
    for (size_t i = 0; i < hard.size(); ++i) hard[i] = used[i] + (int64_t)((i * 2654435761u) % 5) - 1;
    CK(bgc_quota_table_set(ctx, NS, R, hard.data(), hm.data(), unit.data()));
    CK(bgc_usage_set(ctx, used.data()));
    gp.set = 1;
    gen_ref_fill(&gp, 0, nmax, ns.data(), qs, nullptr, cores);
    // pinned copies
    uint32_t* p_ns; int64_t *p_q0, *p_q1; uint8_t* p_v;
    cudaHostAlloc(&p_ns, nmax * 4, cudaHostAllocDefault); cudaHostAlloc(&p_q0, nmax * 8, cudaHostAllocDefault);
    cudaHostAlloc(&p_q1, nmax * 8, cudaHostAllocDefault); cudaHostAlloc(&p_v, nmax, cudaHostAllocDefault);
    memcpy(p_ns, ns.data(), nmax * 4); memcpy(p_q0, q0.data(), nmax * 8); memcpy(p_q1, q1.data(), nmax * 8);
    const int64_t* pq[2] = {p_q0, p_q1};
    std::vector<uint8_t> v_gpu(nmax), v_cpu(nmax);
    long long cross_pin = -1, cross_page = -1;
    for (int lg = 0; lg <= max_log2; ++lg) {
      const size_t n = (size_t)1 << lg;
      const int reps = n <= 4096 ? 2000 : n <= 65536 ? 400 : n <= (1u << 20) ? 50 : 10;
      quota_ref_quota_eval(ns.data(), cq, nullptr, n, NS, R, used.data(), hard.data(), hm.data(), 2, v_cpu.data(), 1);
      CK(bgc_quota_eval(ctx, ns.data(), cq, nullptr, n, v_gpu.data()));
      if (memcmp(v_gpu.data(), v_cpu.data(), n)) { fprintf(stderr, "MISMATCH pageable n=%zu\n", n); return 2; }
      CK(bgc_quota_eval(ctx, p_ns, pq, nullptr, n, p_v));
      if (memcmp(p_v, v_cpu.data(), n)) { fprintf(stderr, "MISMATCH pinned n=%zu\n", n); return 2; }
      double gp_m, gp_9, gg_m, gg_9, c1_m, c1_9, ca_m, ca_9;
      stats([&] { bgc_quota_eval(ctx, p_ns, pq, nullptr, n, p_v); }, reps, &gp_m, &gp_9);
      stats([&] { bgc_quota_eval(ctx, ns.data(), cq, nullptr, n, v_gpu.data()); }, reps, &gg_m, &gg_9);
      stats([&] { quota_ref_quota_eval(ns.data(), cq, nullptr, n, NS, R, used.data(), hard.data(), hm.data(), 2, v_cpu.data(), 1); }, reps, &c1_m, &c1_9);
      stats([&] { quota_ref_quota_eval(ns.data(), cq, nullptr, n, NS, R, used.data(), hard.data(), hm.data(), 2, v_cpu.data(), cores); }, reps, &ca_m, &ca_9);
      const double best = std::min(c1_m, ca_m);
      if (cross_pin < 0 && gp_m < best) cross_pin = (long long)n;
      if (cross_page < 0 && gg_m < best) cross_page = (long long)n;
      printf("{\"kind\":\"quota_eval\",\"NS\":%u,\"n\":%zu,\"gpu_pinned_med\":%.2f,\"gpu_pinned_p99\":%.2f,\"gpu_pageable_med\":%.2f,\"gpu_pageable_p99\":%.2f,"
             "\"cpu_1t_med\":%.2f,\"cpu_1t_p99\":%.2f,\"cpu_all_med\":%.2f,\"cpu_all_p99\":%.2f,\"cores\":%d}\n",
             NS, n, gp_m, gp_9, gg_m, gg_9, c1_m, c1_9, ca_m, ca_9, cores);
      fflush(stdout);
    }
    printf("{\"kind\":\"crossover\",\"NS\":%u,\"pinned\":%lld,\"pageable\":%lld}\n", NS, cross_pin, cross_page);
    cudaFreeHost(p_ns); cudaFreeHost(p_q0); cudaFreeHost(p_q1); cudaFreeHost(p_v);
  }
  // policy decision
  {
    const size_t pm = (size_t)1 << std::min(max_log2, 24);
    std::vector<uint8_t> op(pm), code(pm), pmask(pm), rc(pm), rp(pm);
    std::vector<uint16_t> fl(pm);
    gen_ref_policy_fill(0xB2000001, 0, pm, op.data(), fl.data(), cores);
    long long cross = -1;
    for (int lg = 0; (size_t)1 << lg <= pm; ++lg) {
      const size_t n = (size_t)1 << lg;
      const int reps = n <= 4096 ? 2000 : n <= 65536 ? 400 : n <= (1u << 20) ? 50 : 10;
      policy_ref_bits_batch(op.data(), fl.data(), n, rc.data(), rp.data(), 1);
      CK(bgc_policy_eval(ctx, op.data(), fl.data(), n, code.data(), pmask.data()));
      if (memcmp(code.data(), rc.data(), n) || memcmp(pmask.data(), rp.data(), n)) { fprintf(stderr, "MISMATCH policy n=%zu\n", n); return 2; }
      double g_m, g_9, c1_m, c1_9, ca_m, ca_9;
      stats([&] { bgc_policy_eval(ctx, op.data(), fl.data(), n, code.data(), pmask.data()); }, reps, &g_m, &g_9);
      stats([&] { policy_ref_bits_batch(op.data(), fl.data(), n, rc.data(), rp.data(), 1); }, reps, &c1_m, &c1_9);
      stats([&] { policy_ref_bits_batch(op.data(), fl.data(), n, rc.data(), rp.data(), cores); }, reps, &ca_m, &ca_9);
      if (cross < 0 && g_m < std::min(c1_m, ca_m)) cross = (long long)n;
      printf("{\"kind\":\"policy_eval\",\"n\":%zu,\"gpu_pageable_med\":%.2f,\"gpu_pageable_p99\":%.2f,\"cpu_1t_med\":%.2f,\"cpu_all_med\":%.2f,\"cores\":%d}\n",
             n, g_m, g_9, c1_m, ca_m, cores);
      fflush(stdout);
    }
    printf("{\"kind\":\"crossover\",\"policy\":%lld}\n", cross);
  }
  bgc_ctx_destroy(ctx);
  return 0;
}
