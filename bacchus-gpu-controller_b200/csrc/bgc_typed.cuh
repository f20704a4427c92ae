// bgc_typed.cuh — K1/K2 over NARROW request columns (rows A3/A4 behind bgc_*_typed).  Part of libbgc.so's
// device code; included through bgc_kernels.cuh.
//
// Why: end to end the path is PCIe-bound (57 bytes per pod-step cross the bus with int64 columns, the
// kernels are ~100x faster than the wire), and most request columns do not need 64 bits: a GPU count fits
// u8, cpu millicores u16/u32, memory is a multiple of MiB, a namespace id fits u16 up to 65,536
// namespaces.  The typed entry points take each column as {pointer, element type, shift} and widen on
// the fly inside the loads, value = (int64)element << shift (wrapping), so narrow columns cross PCIe —
// and HBM — as they are.  Results are bit-identical to the int64 entry points on the widened columns.
#pragma once
#include "bgc_k1.cuh"
#include "bgc_k2.cuh"

namespace bgc {

enum { T_I64 = 0, T_U32 = 1, T_U16 = 2, T_U8 = 3, T_I32 = 4 };  // = BGC_T_* of include/bgc.h

__host__ __device__ __forceinline__ uint32_t type_width(uint32_t t) {
  return t == T_I64 ? 8u : (t == T_U32 || t == T_I32) ? 4u : t == T_U16 ? 2u : 1u;
}

struct TCol {
  const void* p;
  uint32_t type, shift;
};
template <int R>
struct TypedCols {
  TCol c[R];
  const void* ns;
  uint32_t ns_type;  // T_U32 or T_U16
};

__device__ __forceinline__ uint32_t ldg_nc_b32(const void* p) {
  uint32_t v;
  asm(BGC_LD_STREAM ".b32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ uint2 ldg_nc_b64(const void* p) {
  uint2 v;
  asm(BGC_LD_STREAM ".v2.b32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
  return v;
}
__device__ __forceinline__ uint32_t ldg_nc_b16(const void* p) {
  unsigned short v;
  asm(BGC_LD_STREAM ".b16 %0, [%1];" : "=h"(v) : "l"(p));
  return v;
}

// PODS consecutive elements of a narrow unsigned column as u32 (one vector load; caller checked alignment)
template <int PODS>
__device__ __forceinline__ void ld_u16xN(const void* p, size_t base, uint32_t (&v)[PODS]) {
  const unsigned short* q = static_cast<const unsigned short*>(p) + base;
  if constexpr (PODS == 4) {
    const uint2 w = ldg_nc_b64(q);
    v[0] = w.x & 0xFFFFu; v[1] = w.x >> 16; v[2] = w.y & 0xFFFFu; v[3] = w.y >> 16;
  } else {
    const uint32_t w = ldg_nc_b32(q);
    v[0] = w & 0xFFFFu; v[1] = w >> 16;
  }
}
template <int PODS>
__device__ __forceinline__ void ld_u8xN(const void* p, size_t base, uint32_t (&v)[PODS]) {
  const uint8_t* q = static_cast<const uint8_t*>(p) + base;
  if constexpr (PODS == 4) {
    const uint32_t w = ldg_nc_b32(q);
    v[0] = w & 0xFFu; v[1] = (w >> 8) & 0xFFu; v[2] = (w >> 16) & 0xFFu; v[3] = w >> 24;
  } else {
    const uint32_t w = ldg_nc_b16(q);
    v[0] = w & 0xFFu; v[1] = w >> 8;
  }
}

template <int PODS>
__device__ __forceinline__ void ld_ns_typed(const void* p, uint32_t type, size_t base, uint32_t (&v)[PODS]) {
  if (type == T_U16) ld_u16xN<PODS>(p, base, v);
  else ldg_ns<PODS>(static_cast<const uint32_t*>(p) + base, v);
}

// the type switch is uniform over the grid: no divergence, a handful of extra instructions per column
template <int PODS>
__device__ __forceinline__ void ld_col_typed(const TCol& c, size_t base, long long (&v)[PODS]) {
  if (c.type == T_I64) {
    ldg_req<PODS>(static_cast<const long long*>(c.p) + base, v);
  } else {
    uint32_t t[PODS];
    if (c.type == T_U16) ld_u16xN<PODS>(c.p, base, t);
    else if (c.type == T_U8) ld_u8xN<PODS>(c.p, base, t);
    else ldg_ns<PODS>(static_cast<const uint32_t*>(c.p) + base, t);
    // element << shift as ONE wide multiply by 2^shift (a 64-bit shift by a run-time amount is five instructions:
    // ncu counted 105 thread instructions per pod in the typed K2 against 66 in the int64 one, SHF / SEL / ISETP on top).
    // The branches are uniform over the grid.
    if (c.type != T_I32 && c.shift < 32u) {
      const uint32_t m = 1u << c.shift;
#pragma unroll
      for (int k = 0; k < PODS; ++k) v[k] = (long long)((unsigned long long)t[k] * m);
      return;
    }
    if (c.type == T_I32 && c.shift < 31u) {
      const int m = 1 << c.shift;
#pragma unroll
      for (int k = 0; k < PODS; ++k) v[k] = (long long)(int)t[k] * m;
      return;
    }
#pragma unroll
    for (int k = 0; k < PODS; ++k) v[k] = c.type == T_I32 ? (long long)(int)t[k] : (long long)t[k];
  }
#pragma unroll
  for (int k = 0; k < PODS; ++k) v[k] = (long long)((unsigned long long)v[k] << c.shift);
}

__device__ __forceinline__ uint32_t ld_ns_one(const void* p, uint32_t type, size_t i) {
  return type == T_U16 ? (uint32_t) static_cast<const unsigned short*>(p)[i] : static_cast<const uint32_t*>(p)[i];
}
__device__ __forceinline__ long long ld_col_one(const TCol& c, size_t i) {
  long long v;
  switch (c.type) {
    case T_U32: v = (long long)static_cast<const uint32_t*>(c.p)[i]; break;
    case T_U16: v = (long long)static_cast<const unsigned short*>(c.p)[i]; break;
    case T_U8: v = (long long)static_cast<const uint8_t*>(c.p)[i]; break;
    case T_I32: v = (long long)static_cast<const int*>(c.p)[i]; break;
    default: v = static_cast<const long long*>(c.p)[i]; break;
  }
  return (long long)((unsigned long long)v << c.shift);
}

// ------------------------------------------------------------------ K1 over typed columns
// The arrival-order scatter kernel (bgc_k1.cuh: CTA-private shared-memory table, direct or hash
// sink, one persistent 1024-thread CTA per SM) with the typed loads in front.  vec = every column
// pointer is aligned for its vector load; otherwise everything takes the scalar loop.
template <int R, int SINK>
__global__ void __launch_bounds__(1024, 1)
k1_scatter_typed(TypedCols<R> in, size_t n, long long* used, uint32_t NS, uint32_t cap, uint32_t wide_mask,
                 uint32_t pf_dist, uint32_t vec) {
  extern __shared__ __align__(16) uint32_t k1_smem[];
  constexpr int THREADS = 1024;
  const uint32_t nwide = __popc(wide_mask & ((1u << R) - 1u));
  const uint32_t words = (R + nwide) * cap;
  for (uint32_t i = threadIdx.x; i < words; i += THREADS) k1_smem[i] = 0;
  if (SINK == SINK_HASH)
    for (uint32_t i = threadIdx.x; i < cap; i += THREADS) k1_smem[words + i] = kHashEmpty;
  ScatterCtx<R, SINK> C;
  C.used = used; C.NS = NS; C.cap = cap; C.wide_mask = wide_mask;
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(k1_smem);
#pragma unroll
  for (int r = 0; r < R; ++r) {
    C.lo_base[r] = sbase + (uint32_t)r * cap * 4u;
    C.hi_base[r] = sbase + (R + __popc(wide_mask & ((1u << r) - 1u))) * cap * 4u;
  }
  C.key_base = sbase + words * 4u;
  __syncthreads();

  constexpr int PODS = R <= 4 ? 4 : 2;
  const size_t groups = vec ? n / PODS : 0;
  const size_t stride = (size_t)gridDim.x * THREADS;
  const uint32_t nsw = type_width(in.ns_type);
  for (size_t g = (size_t)blockIdx.x * THREADS + threadIdx.x; g < groups; g += stride) {
    if (pf_dist && (threadIdx.x & 31u) == 0) {  // the warp's tile pf_dist trips ahead (host checked 16-byte alignment)
      const size_t gp = g + (size_t)pf_dist * stride;
      if (gp + 32 <= groups) {
        prefetch_l2_bulk(static_cast<const char*>(in.ns) + gp * PODS * nsw, 32u * PODS * nsw);
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const uint32_t w = type_width(in.c[r].type);
          prefetch_l2_bulk(static_cast<const char*>(in.c[r].p) + gp * PODS * w, 32u * PODS * w);
        }
      }
    }
    uint32_t ns[PODS];
    long long q[R][PODS];
    const size_t base = g * PODS;
    ld_ns_typed<PODS>(in.ns, in.ns_type, base, ns);
#pragma unroll
    for (int r = 0; r < R; ++r) ld_col_typed<PODS>(in.c[r], base, q[r]);
    scatter_group<R, SINK, PODS>(C, ns, q);
  }
  for (size_t i = groups * PODS + (size_t)blockIdx.x * THREADS + threadIdx.x; i < n; i += stride) {
    uint32_t ns[1] = {ld_ns_one(in.ns, in.ns_type, i)};
    long long q[R][1];
#pragma unroll
    for (int r = 0; r < R; ++r) q[r][0] = ld_col_one(in.c[r], i);
    scatter_pods<R, SINK, 1>(C, ns, q);
  }

  __syncthreads();
  for (uint32_t i = threadIdx.x; i < R * cap; i += THREADS) {
    const uint32_t r = i / cap, sl = i - r * cap;
    const uint32_t hi = ((wide_mask >> r) & 1u) ? k1_smem[(R + __popc(wide_mask & ((1u << r) - 1u))) * cap + sl] : 0u;
    const unsigned long long v = ((unsigned long long)hi << 32) | k1_smem[i];
    if (v) {
      const uint32_t ns = (SINK == SINK_HASH) ? k1_smem[words + sl] : sl;
      red_add_i64(used_at(used, NS, r, ns), (long long)v);
    }
  }
}

// ------------------------------------------------------------------ K2 over typed columns
// (min 4 CTAs per SM: left alone the generic widening code takes 114 registers and two CTAs per SM — ncu showed
// 23 % of the warp slots occupied and the kernel at half the speed of the int64 one)
template <int R, bool HAS_MASK, typename VT>
__global__ void __launch_bounds__(kK2Threads, 4)
k2_quota_eval_typed(TypedCols<R> in, const uint8_t* __restrict__ req_mask, size_t n, const long long* __restrict__ head,
                    const uint8_t* __restrict__ hard_mask, uint32_t NS, uint32_t compute_mask, VT* __restrict__ verdict,
                    uint32_t pf_dist, uint32_t vec) {
  constexpr int PODS = R <= 4 ? 4 : 2;
  constexpr int SHIFT = sizeof(VT) == 1 ? 4 : 8;
  const size_t groups = vec ? n / PODS : 0;
  const uint32_t nsw = type_width(in.ns_type);
  for (size_t g = (size_t)blockIdx.x * kK2Threads + threadIdx.x; g < groups; g += (size_t)gridDim.x * kK2Threads) {
    const size_t base = g * PODS;
    if (pf_dist && (threadIdx.x & 31u) == 0) {
      const size_t gp = g + (size_t)pf_dist * gridDim.x * kK2Threads;
      if (gp + 32 <= groups) {
        prefetch_l2_bulk(static_cast<const char*>(in.ns) + gp * PODS * nsw, 32u * PODS * nsw);
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const uint32_t w = type_width(in.c[r].type);
          prefetch_l2_bulk(static_cast<const char*>(in.c[r].p) + gp * PODS * w, 32u * PODS * w);
        }
      }
    }
    uint32_t ns[PODS];
    long long q[R][PODS];
    ld_ns_typed<PODS>(in.ns, in.ns_type, base, ns);
#pragma unroll
    for (int r = 0; r < R; ++r) ld_col_typed<PODS>(in.c[r], base, q[r]);
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
  for (size_t i = groups * PODS + (size_t)blockIdx.x * kK2Threads + threadIdx.x; i < n; i += (size_t)gridDim.x * kK2Threads) {
    long long qk[R];
#pragma unroll
    for (int r = 0; r < R; ++r) qk[r] = ld_col_one(in.c[r], i);
    const uint32_t rm = HAS_MASK ? req_mask[i] : 0xFFu;
    verdict[i] = (VT)quota_verdict<R, SHIFT>(ld_ns_one(in.ns, in.ns_type, i), qk, rm, head, hard_mask, NS, compute_mask);
  }
}

}  // namespace bgc
