// bgc_common.cuh — shared prelude of libbgc.so's device code: constants, column structs and the
// streaming load/store helpers (whole-sector 128/256-bit read-only loads that bypass L1, global reds).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace bgc {

constexpr int kMaxR = 8;   // u8 verdicts up to R = 4, u16 verdicts up to R = 8
constexpr int kMaxR8 = 4;  // widest R of the u8 verdict encoding
constexpr uint32_t kBadNs = 0xFFFFFFFFu;

// Quota state: two [R][NS] int64 tables, `used` (K1 accumulates into it with global reds; the L2
// atomic rate does not depend on the layout, see tools/micro/red_bench.cu) and `hard`, plus
// hard_mask[NS].  K2 reads the compact per-namespace HeadRow derived from them (k_build_headroom).
__device__ __forceinline__ long long* used_at(long long* used, uint32_t NS, int r, uint32_t ns) {
  return used + (size_t)r * NS + ns;
}

template <int R>
struct ReqCols {
  const long long* p[R];
};
template <int R>
struct ReqColsOut {
  long long* p[R];
};

// ------------------------------------------------------------------ load/store helpers
// Pod columns are read once and never again: besides bypassing L1 they are marked evict-first in L2
// (LDG.E.NA.EFL2), so that 5.7 GB of stream per step does not push the usage/headroom tables — tens of
// MB at 10^6 namespaces, re-read by every CTA — out of the 126 MB L2.  -DBGC_NO_L2_HINTS builds without.
// (ptxas accepts the L2 eviction priority on 256-bit loads only — the request columns, 6/7 of the stream.)
#define BGC_LD_STREAM "ld.global.nc.L1::no_allocate"
#ifndef BGC_NO_L2_HINTS
#define BGC_LD_STREAM256 "ld.global.nc.L1::no_allocate.L2::evict_first"
#define BGC_LD_TABLE256 "ld.global.L2::evict_last"
#else
#define BGC_LD_STREAM256 "ld.global.nc.L1::no_allocate"
#define BGC_LD_TABLE256 "ld.global"
#endif
__device__ __forceinline__ uint4 ldg_nc_u32x4(const uint32_t* p) {
  uint4 v;
  asm(BGC_LD_STREAM ".v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}
// 256-bit streaming load (sm_100: LDG.E.256)
__device__ __forceinline__ void ldg_nc_i64x4(const long long* p, long long (&v)[4]) {
  asm(BGC_LD_STREAM256 ".v4.u64 {%0,%1,%2,%3}, [%4];"
               : "=l"(v[0]), "=l"(v[1]), "=l"(v[2]), "=l"(v[3])
               : "l"(p));
}
// 256-bit gather of one HeadRow; table lines are re-used, so let them allocate in L1/L2
__device__ __forceinline__ void ldg_row_half(const long long* p, long long (&v)[4]) {
  asm volatile(BGC_LD_TABLE256 ".v4.u64 {%0,%1,%2,%3}, [%4];"
               : "=l"(v[0]), "=l"(v[1]), "=l"(v[2]), "=l"(v[3])
               : "l"(p));
}
// PODS consecutive pods per thread: 4 (R <= 4: 128-bit ns + 256-bit request loads) or 2 (R > 4, to
// stay inside 64 registers: 64-bit ns + 128-bit request loads)
template <int PODS>
__device__ __forceinline__ void ldg_ns(const uint32_t* p, uint32_t (&v)[PODS]) {
  if constexpr (PODS == 4) {
    const uint4 x = ldg_nc_u32x4(p);
    v[0] = x.x; v[1] = x.y; v[2] = x.z; v[3] = x.w;
  } else {
    asm(BGC_LD_STREAM ".v2.u32 {%0,%1}, [%2];" : "=r"(v[0]), "=r"(v[1]) : "l"(p));
  }
}
template <int PODS>
__device__ __forceinline__ void ldg_req(const long long* p, long long (&v)[PODS]) {
  if constexpr (PODS == 4) {
    ldg_nc_i64x4(p, v);
  } else {
    asm(BGC_LD_STREAM ".v2.u64 {%0,%1}, [%2];" : "=l"(v[0]), "=l"(v[1]) : "l"(p));
  }
}

// 256-bit streaming load / store of 8 words (one full 32-byte sector per thread).  No L2 eviction hint here: K0,
// the only user, has no table to protect, and with evict-first loads it ran at 78 % instead of 105 % of the copy peak.
__device__ __forceinline__ void ldg_nc_u32x8(const void* p, uint32_t (&v)[8]) {
  asm(BGC_LD_STREAM ".v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
      : "l"(p));
}
template <int MODE = 0>
__device__ __forceinline__ void stg_u32x8(void* p, const uint32_t (&v)[8]) {
  if constexpr (MODE == 0)
    asm volatile("st.global.L1::no_allocate.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(v[0]), "r"(v[1]),
                 "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
  else if constexpr (MODE == 1)
    asm volatile("st.global.cs.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(v[0]), "r"(v[1]),
                 "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
  else
    asm volatile("st.global.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(v[0]), "r"(v[1]),
                 "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
}
__device__ __forceinline__ void red_add_i64(long long* p, long long v) {
  asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}


}  // namespace bgc
