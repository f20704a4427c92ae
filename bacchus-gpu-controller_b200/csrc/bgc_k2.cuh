// bgc_k2.cuh — K2 quota_eval (row A4).  Part of libbgc.so's device code; included through bgc_kernels.cuh.
#pragma once
#include "bgc_common.cuh"

namespace bgc {

// ------------------------------------------------------------------ K2 quota_eval
// K2 gathers one 32-byte sector per pod from a compact table derived from the usage and limit tables:
// head[ns].h[r] = hard[r] - used[r], saturated to the i64 range, and for R <= 3 h[3] = hard_mask.
// "exceeded" is the overflow-free comparison used + req > hard (as Kubernetes' arbitrary-precision
// Quantity arithmetic would decide it), i.e. req > hard - used with the difference saturated.
// The table is NS rows of HEADW(R) words: 4 (one 32-byte sector) for R <= 3, 8 (two sectors) for
// R <= 8; the last word holds hard_mask when there is room (R < HEADW), else the mask is gathered
// from its own byte array.
__host__ __device__ constexpr int head_words(int R) { return R <= 3 ? 4 : 8; }

__device__ __forceinline__ long long sat_sub_i64(long long a, long long b) {
  const long long d = (long long)((unsigned long long)a - (unsigned long long)b);
  if (((a ^ b) & (a ^ d)) < 0) return a < 0 ? (long long)0x8000000000000000ull : 0x7FFFFFFFFFFFFFFFll;
  return d;
}

static __global__ void k_build_headroom(const long long* __restrict__ used, const long long* __restrict__ hard,
                                        const uint8_t* __restrict__ hard_mask, long long* head, uint32_t NS, int R) {
  const int W = head_words(R);
  for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < (size_t)NS * W;
       t += (size_t)gridDim.x * blockDim.x) {
    const uint32_t ns = (uint32_t)(t / W);
    const int r = (int)(t % W);
    long long v = 0;
    if (r < R) v = sat_sub_i64(hard[(size_t)r * NS + ns], used[(size_t)r * NS + ns]);
    else if (r == W - 1) v = hard_mask[ns];
    head[t] = v;
  }
}

// verdict = missing << SHIFT | exceeded, SHIFT = 4 for the u8 form (R <= 4), 8 for the u16 form
template <int R, int SHIFT>
__device__ __forceinline__ uint32_t quota_verdict(uint32_t ns, const long long (&q)[R], uint32_t rmask,
                                                  const long long* __restrict__ head,
                                                  const uint8_t* __restrict__ hard_mask, uint32_t NS,
                                                  uint32_t compute_mask) {
  if (ns >= NS) return SHIFT == 4 ? 0xFFu : 0xFFFFu;
  constexpr int W = head_words(R);
  long long h[W];
  {
    long long t[4];
    ldg_row_half(head + (size_t)ns * W, t);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = t[i];
    if constexpr (W == 8) {
      ldg_row_half(head + (size_t)ns * W + 4, t);
#pragma unroll
      for (int i = 0; i < 4; ++i) h[4 + i] = t[i];
    }
  }
  const uint32_t hm = (R < W) ? (uint32_t)h[W - 1] : (uint32_t)hard_mask[ns];
  uint32_t missing = 0, exceeded = 0;
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const bool constrained = (hm >> r) & 1u;
    const bool present = (rmask >> r) & 1u;
    if (constrained && !present && ((compute_mask >> r) & 1u)) missing |= 1u << r;
    if (constrained && present && q[r] > 0 && q[r] > h[r]) exceeded |= 1u << r;
  }
  return (missing << SHIFT) | exceeded;
}

constexpr int kK2Threads = 256;

// PODS pods per thread (4 for R <= 4, 2 above): one vector ns load, R vector request loads, one
// HeadRow gather per pod (one or two 32-byte sectors), one packed verdict store.
// (min 4 CTAs per SM = 64 registers: at 68 the kernel ran three CTAs per SM; the fourth one's gathers in flight are
// worth 82 % -> 92 % of the HBM roofline at 10^6 namespaces and 99 % -> 101 % on C3)
template <int R, bool HAS_MASK, bool VEC, typename VT>
__global__ void __launch_bounds__(kK2Threads, 4)
k2_quota_eval(const uint32_t* __restrict__ ns_id, ReqCols<R> req, const uint8_t* __restrict__ req_mask,
              size_t n, const long long* __restrict__ head, const uint8_t* __restrict__ hard_mask,
              uint32_t NS, uint32_t compute_mask, VT* __restrict__ verdict, uint32_t pf_dist) {
  constexpr int PODS = R <= 4 ? 4 : 2;
  constexpr int SHIFT = sizeof(VT) == 1 ? 4 : 8;
  const size_t groups = VEC ? n / PODS : 0;
  for (size_t g = (size_t)blockIdx.x * kK2Threads + threadIdx.x; g < groups;
       g += (size_t)gridDim.x * kK2Threads) {
    const size_t base = g * PODS;
    if (pf_dist && (threadIdx.x & 31u) == 0) {  // L2 bulk prefetch of the warp's tile pf_dist trips ahead
      const size_t gp = g + (size_t)pf_dist * gridDim.x * kK2Threads;
      if (gp + 32 <= groups) {
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(ns_id + gp * PODS), "r"(32u * PODS * 4u) : "memory");
#pragma unroll
        for (int r = 0; r < R; ++r)
          asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(req.p[r] + gp * PODS), "r"(32u * PODS * 8u) : "memory");
      }
    }
    uint32_t ns[PODS];
    long long q[R][PODS];
    ldg_ns<PODS>(ns_id + base, ns);
#pragma unroll
    for (int r = 0; r < R; ++r) ldg_req<PODS>(req.p[r] + base, q[r]);
    uint32_t rm = 0xFFFFFFFFu;
    if (HAS_MASK) {
      if constexpr (PODS == 4) rm = __ldg(reinterpret_cast<const uint32_t*>(req_mask + base));
      else rm = __ldg(reinterpret_cast<const unsigned short*>(req_mask + base));
    }
    VT out[PODS];
#pragma unroll
    for (int k = 0; k < PODS; ++k) {
      long long qk[R];
#pragma unroll
      for (int r = 0; r < R; ++r) qk[r] = q[r][k];
      out[k] = (VT)quota_verdict<R, SHIFT>(ns[k], qk, (rm >> (8 * k)) & 0xFFu, head, hard_mask, NS, compute_mask);
    }
    if constexpr (sizeof(VT) * PODS == 8) {
      *reinterpret_cast<uint2*>(verdict + base) =
          make_uint2((uint32_t)out[0] | ((uint32_t)out[1] << 16), (uint32_t)out[2] | ((uint32_t)out[3] << 16));
    } else if constexpr (sizeof(VT) * PODS == 4) {
      uint32_t w = 0;
#pragma unroll
      for (int k = 0; k < PODS; ++k) w |= (uint32_t)out[k] << (8 * sizeof(VT) * k);
      *reinterpret_cast<uint32_t*>(verdict + base) = w;
    } else {
      *reinterpret_cast<unsigned short*>(verdict + base) = (unsigned short)((uint32_t)out[0] | ((uint32_t)out[1] << 8));
    }
  }
  for (size_t i = groups * PODS + (size_t)blockIdx.x * kK2Threads + threadIdx.x; i < n;
       i += (size_t)gridDim.x * kK2Threads) {
    long long qk[R];
#pragma unroll
    for (int r = 0; r < R; ++r) qk[r] = req.p[r][i];
    const uint32_t rm = HAS_MASK ? req_mask[i] : 0xFFu;
    verdict[i] = (VT)quota_verdict<R, SHIFT>(ns_id[i], qk, rm, head, hard_mask, NS, compute_mask);
  }
}

// ------------------------------------------------------------------ K2 with the headroom table in shared memory
// With one or two resources a pod is 12 or 20 bytes of stream but still one 32-byte sector of table gather:
// the L2 sector rate (~210 G random sectors/s), not HBM, bounds the kernel at 0.47 ms per 10^8 pods (68 % of
// the roofline at R=2).  When the table is small enough — NS * (8R + 1) bytes <= 227 KB, i.e. 13,600
// namespaces at R=2 — every CTA keeps its own copy in shared memory instead: one persistent 1024-thread
// CTA per SM, h[r][ns] as 64-bit words plus the mask bytes, gathers become LDS and only the pod
// columns touch L2/HBM.
template <int R, bool HAS_MASK, bool VEC>
__global__ void __launch_bounds__(1024, 1)
k2_quota_eval_smem(const uint32_t* __restrict__ ns_id, ReqCols<R> req, const uint8_t* __restrict__ req_mask, size_t n,
                   const long long* __restrict__ head, const uint8_t* __restrict__ hard_mask, uint32_t NS,
                   uint32_t compute_mask, uint8_t* __restrict__ verdict, uint32_t pf_dist) {
  extern __shared__ __align__(16) long long k2_smem[];
  constexpr int THREADS = 1024;
  constexpr int W = head_words(R);
  long long* s_h = k2_smem;                                        // [R][NS]
  uint8_t* s_m = reinterpret_cast<uint8_t*>(k2_smem + (size_t)R * NS);  // [NS]
  for (uint32_t i = threadIdx.x; i < NS; i += THREADS) {
#pragma unroll
    for (int r = 0; r < R; ++r) s_h[(size_t)r * NS + i] = head[(size_t)i * W + r];
    s_m[i] = hard_mask[i];
  }
  __syncthreads();
  auto decide = [&](uint32_t ns, const long long (&q)[R], uint32_t rmask) -> uint32_t {
    if (ns >= NS) return 0xFFu;
    const uint32_t hm = s_m[ns];
    uint32_t missing = 0, exceeded = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const long long h = s_h[(size_t)r * NS + ns];
      const bool constrained = (hm >> r) & 1u;
      const bool present = (rmask >> r) & 1u;
      if (constrained && !present && ((compute_mask >> r) & 1u)) missing |= 1u << r;
      if (constrained && present && q[r] > 0 && q[r] > h) exceeded |= 1u << r;
    }
    return (missing << 4) | exceeded;
  };
  constexpr int PODS = 4;
  const size_t groups = VEC ? n / PODS : 0;
  const size_t stride = (size_t)gridDim.x * THREADS;
  for (size_t g = (size_t)blockIdx.x * THREADS + threadIdx.x; g < groups; g += stride) {
    const size_t base = g * PODS;
    if (pf_dist && (threadIdx.x & 31u) == 0) {
      const size_t gp = g + (size_t)pf_dist * stride;
      if (gp + 32 <= groups) {
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(ns_id + gp * PODS), "r"(32u * PODS * 4u) : "memory");
#pragma unroll
        for (int r = 0; r < R; ++r)
          asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(req.p[r] + gp * PODS), "r"(32u * PODS * 8u) : "memory");
      }
    }
    uint32_t ns[PODS];
    long long q[R][PODS];
    ldg_ns<PODS>(ns_id + base, ns);
#pragma unroll
    for (int r = 0; r < R; ++r) ldg_req<PODS>(req.p[r] + base, q[r]);
    uint32_t rm = 0xFFFFFFFFu;
    if (HAS_MASK) rm = __ldg(reinterpret_cast<const uint32_t*>(req_mask + base));
    uint32_t w = 0;
#pragma unroll
    for (int k = 0; k < PODS; ++k) {
      long long qk[R];
#pragma unroll
      for (int r = 0; r < R; ++r) qk[r] = q[r][k];
      w |= decide(ns[k], qk, (rm >> (8 * k)) & 0xFFu) << (8 * k);
    }
    *reinterpret_cast<uint32_t*>(verdict + base) = w;
  }
  for (size_t i = groups * PODS + (size_t)blockIdx.x * THREADS + threadIdx.x; i < n; i += stride) {
    long long qk[R];
#pragma unroll
    for (int r = 0; r < R; ++r) qk[r] = req.p[r][i];
    verdict[i] = (uint8_t)decide(ns_id[i], qk, HAS_MASK ? req_mask[i] : 0xFFu);
  }
}

}  // namespace bgc
