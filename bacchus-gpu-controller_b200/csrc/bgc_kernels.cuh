// bgc_kernels.cuh — sm_100a device code of libbgc.so.
//
// K0 policy_eval  : batched form of mutate(), reference src/admission.rs:241-431   (row A1)
// K1 usage_reduce : per-namespace wrapping-i64 usage, no reference function        (row A3)
// K2 quota_eval   : snapshot quota verdict per candidate pod, no reference function (row A4)
//
// All three are HBM-streaming integer kernels: no tensor cores, no TMEM.  What matters is
// 128/256-bit coalesced read-only loads that bypass L1, enough loads in flight per thread, a
// grid that is a multiple of the 148 SMs, and — for K1 — keeping the atomic traffic per pod low.
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
__device__ __forceinline__ uint4 ldg_nc_u32x4(const uint32_t* p) {
  uint4 v;
  asm("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}
// 256-bit streaming load (sm_100: LDG.E.256)
__device__ __forceinline__ void ldg_nc_i64x4(const long long* p, long long (&v)[4]) {
  asm("ld.global.nc.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];"
               : "=l"(v[0]), "=l"(v[1]), "=l"(v[2]), "=l"(v[3])
               : "l"(p));
}
// 256-bit gather of one HeadRow; table lines are re-used, so let them allocate in L1/L2
__device__ __forceinline__ void ldg_row_half(const long long* p, long long (&v)[4]) {
  asm volatile("ld.global.v4.u64 {%0,%1,%2,%3}, [%4];"
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
    asm("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(v[0]), "=r"(v[1]) : "l"(p));
  }
}
template <int PODS>
__device__ __forceinline__ void ldg_req(const long long* p, long long (&v)[PODS]) {
  if constexpr (PODS == 4) {
    ldg_nc_i64x4(p, v);
  } else {
    asm("ld.global.nc.L1::no_allocate.v2.u64 {%0,%1}, [%2];" : "=l"(v[0]), "=l"(v[1]) : "l"(p));
  }
}

// 256-bit streaming load / store of 8 words (one full 32-byte sector per thread)
__device__ __forceinline__ void ldg_nc_u32x8(const void* p, uint32_t (&v)[8]) {
  asm("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
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

// ------------------------------------------------------------------ K0 policy_eval
// Decision of one flattened request; restates src/admission.rs:241-431 in evaluation order.
// Returns code | patch_mask << 8.
__device__ __forceinline__ uint32_t policy_decide(uint32_t op, uint32_t f) {
  const bool has_username = f & 1u, normal = f & 2u, in_group = f & 4u, has_object = f & 8u;
  const bool has_name = f & 16u, name_eq = f & 32u, parse_ok = f & 64u, ku_nonempty = f & 128u;
  const bool has_quota = f & 256u, has_rb = f & 512u;
  auto ret = [](uint32_t klass, uint32_t reason, uint32_t patch) {
    return (klass << 4) | reason | (patch << 8);
  };
  if (!has_username) return ret(3, 1, 0);                      // :248-255 invalid
  if (op == 0) {                                               // Create :273-283
    if (normal && !in_group) return ret(2, 2, 0);
  } else if (op == 2) {                                        // Delete :284-294
    return normal ? ret(2, 3, 0) : ret(0, 12, 0);
  } else if (op == 1) {                                        // Update :295-304
    if (normal) return ret(2, 4, 0);
  } else {
    return ret(3, 5, 0);                                       // :305-309 invalid operation
  }
  if (!has_object) return ret(0, 13, 0);                       // :312-318
  if (!has_name) return ret(3, 6, 0);                          // :321-328
  if (normal && !name_eq) return ret(2, 7, 0);                 // :331-338
  if (!parse_ok) return ret(3, 8, 0);                          // :341-347
  uint32_t patch = 0;
  if (normal) patch |= 1;                                      // :352-358 P_USER
  else if (!ku_nonempty) return ret(2, 9, 0);                  // :359-373
  if (has_quota && normal) return ret(2, 10, 0);               // :376-383
  if (!has_rb) patch |= normal ? (2u | 4u) : 2u;               // :386-416 P_RB (+ SUBJ_ORIG)
  else if (normal) return ret(2, 11, 0);                       // :417-424
  return ret(patch ? 1u : 0u, 0, patch);                       // :426-430
}

// The decision is a function of 12 bits, so each CTA tabulates it once (4096 x u16 = 8 KB of
// shared memory) and the streaming loop is one LDS per request: 3 B in, 2 B out per request.
constexpr int kK0Threads = 256;
constexpr int kK0PerThread = 32;

template <int ST>
static __global__ void __launch_bounds__(kK0Threads)
k0_policy_eval(const uint8_t* __restrict__ op, const uint16_t* __restrict__ flags, size_t n,
               uint8_t* __restrict__ code, uint8_t* __restrict__ patch_mask, int vec_ok) {
  __shared__ uint16_t lut[4096];
  for (uint32_t i = threadIdx.x; i < 4096; i += kK0Threads)
    lut[i] = (uint16_t)policy_decide(i >> 10, i & 1023u);
  __syncthreads();

  // 32 requests per thread, every access a whole 32-byte sector: one 256-bit op load, two 256-bit
  // flags loads, one 256-bit store each for code and patch_mask
  const size_t groups = vec_ok ? n / kK0PerThread : 0;
  for (size_t g = (size_t)blockIdx.x * kK0Threads + threadIdx.x; g < groups;
       g += (size_t)gridDim.x * kK0Threads) {
    const size_t base = g * kK0PerThread;
    uint32_t ow[8], fw[16], cw[8], pw[8];
    ldg_nc_u32x8(op + base, ow);
    ldg_nc_u32x8(flags + base, *reinterpret_cast<uint32_t(*)[8]>(&fw[0]));
    ldg_nc_u32x8(flags + base + 16, *reinterpret_cast<uint32_t(*)[8]>(&fw[8]));
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      uint32_t c = 0, p = 0;
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const uint32_t raw = (ow[w] >> (8 * b)) & 0xFFu;   // anything outside 0..2 is "other"
        const uint32_t opv = raw > 2u ? 3u : raw;
        const uint32_t fl = (fw[w * 2 + (b >> 1)] >> (16 * (b & 1))) & 1023u;
        const uint32_t e = lut[(opv << 10) | fl];
        c |= (e & 0xFFu) << (8 * b);
        p |= (e >> 8) << (8 * b);
      }
      cw[w] = c;
      pw[w] = p;
    }
    stg_u32x8<ST>(code + base, cw);
    stg_u32x8<ST>(patch_mask + base, pw);
  }
  // scalar tail (and the whole batch when the pointers are not 32-byte aligned)
  for (size_t i = groups * kK0PerThread + (size_t)blockIdx.x * kK0Threads + threadIdx.x; i < n;
       i += (size_t)gridDim.x * kK0Threads) {
    const uint32_t raw = op[i];
    const uint32_t e = lut[((raw > 2u ? 3u : raw) << 10) | (flags[i] & 1023u)];
    code[i] = (uint8_t)(e & 0xFFu);
    patch_mask[i] = (uint8_t)(e >> 8);
  }
}

// ------------------------------------------------------------------ K1 usage_reduce
// One warp owns a contiguous chunk of pods and walks it in tiles of 128 (4 consecutive pods per
// lane, so every warp-level load is a fully coalesced 512 B / 1 KB request).  Runs of equal
// namespace are summed before anything is published:
//   1. inside a lane (4 pods),
//   2. across lanes with a segmented shuffle scan keyed on run boundaries,
//   3. across tiles with a carry held by lane 0,
// so pods grouped by namespace cost ~R atomics per 4096 pods, while fully shuffled input
// degenerates to R updates per pod.  Updates go to one of three sinks:
//   SINK_GLOBAL : red.global.add.u64 straight into the global usage table (L2 atomics);
//   SINK_DIRECT : CTA-private shared-memory table indexed by ns (NS*R*8 B must fit), flushed once;
//   SINK_HASH   : CTA-private shared-memory table of S claim-once slots (hot namespaces claim
//                 a slot first), cold namespaces fall through to global reds.
// Shared-memory accumulators are split lo/hi u32 because ATOMS.ADD is native only for 32 bits
// (a 64-bit shared atomicAdd compiles to a CAS spin loop): the carry out of the low word is
// detected from the returned old value and folded into the high-word add, which is exact
// modulo 2^64 under any interleaving.
enum { SINK_GLOBAL = 0, SINK_DIRECT = 1, SINK_HASH = 2 };

// slot hash of the claim-once table (murmur3 finaliser): namespace ids handed out by affine maps
// collide systematically under a bare multiplicative hash
__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x85EBCA6Bu; x ^= x >> 13; x *= 0xC2B2AE35u; x ^= x >> 16;
  return x;
}
__device__ __forceinline__ uint32_t slot_hash(uint32_t x, uint32_t shift) { return mix32(x) >> shift; }

constexpr int kK1Tile = 128;            // pods per warp iteration
constexpr uint32_t kHashEmpty = 0xFFFFFFFFu;

template <int R, int SINK>
struct K1Sink {
  long long* used;
  uint32_t NS;
  uint32_t* s_lo;   // [R][cap]      low words of every column
  uint32_t* s_hi;   // [nwide][cap]  high words of the wide columns only
  uint32_t* s_key;  // [cap] (hash)
  uint32_t cap;     // NS (direct) or slots (hash, power of two)
  uint32_t hash_shift;
  uint32_t wide_mask;  // bit r: column r keeps its high word in shared memory (byte-valued columns)

  __device__ __forceinline__ uint32_t hi_slot(int r) const { return __popc(wide_mask & ((1u << r) - 1u)); }

  // Narrow columns (counts, millicores) almost never carry into bit 32, so they get a low word
  // only and the rare carry / high part goes straight to the global table.
  __device__ __forceinline__ void smem_add(int r, uint32_t idx, uint32_t ns, long long v) {
    uint32_t lo = (uint32_t)(unsigned long long)v, hi = (uint32_t)((unsigned long long)v >> 32);
    if (lo) {
      const uint32_t old = atomicAdd(&s_lo[r * cap + idx], lo);
      hi += (old + lo < old);  // carry out of the low word
    }
    if (hi) {
      if ((wide_mask >> r) & 1u) atomicAdd(&s_hi[hi_slot(r) * cap + idx], hi);
      else red_add_i64(used_at(used, NS, r, ns), (long long)((unsigned long long)hi << 32));
    }
  }

  __device__ __forceinline__ void add(uint32_t ns, const long long (&v)[R]) {
    if (ns >= NS) return;  // kBadNs padding and out-of-range ids contribute nothing
    if (SINK == SINK_GLOBAL) {
#pragma unroll
      for (int r = 0; r < R; ++r)
        if (v[r]) red_add_i64(used_at(used, NS, r, ns), v[r]);
    } else if (SINK == SINK_DIRECT) {
#pragma unroll
      for (int r = 0; r < R; ++r)
        if (v[r]) smem_add(r, ns, ns, v[r]);
    } else {
      const uint32_t s = slot_hash(ns, hash_shift);
      uint32_t k = s_key[s];
      if (k == kHashEmpty) {
        k = atomicCAS(&s_key[s], kHashEmpty, ns);
        if (k == kHashEmpty) k = ns;
      }
      if (k == ns) {
#pragma unroll
        for (int r = 0; r < R; ++r)
          if (v[r]) smem_add(r, s, ns, v[r]);
      } else {
#pragma unroll
        for (int r = 0; r < R; ++r)
          if (v[r]) red_add_i64(used_at(used, NS, r, ns), v[r]);
      }
    }
  }
};

template <int R, int SINK, bool VEC>
__global__ void __launch_bounds__(SINK == SINK_GLOBAL ? 512 : 1024, 1)
k1_usage_reduce(const uint32_t* __restrict__ ns_id, ReqCols<R> req, size_t n, long long* used,
                uint32_t NS, uint32_t cap, uint32_t chunk_tiles, uint32_t wide_mask) {
  extern __shared__ __align__(16) uint32_t k1_smem[];
  K1Sink<R, SINK> sink;
  const uint32_t nwide = __popc(wide_mask & ((1u << R) - 1u));
  sink.used = used;
  sink.NS = NS;
  sink.cap = cap;
  sink.wide_mask = wide_mask;
  sink.s_lo = k1_smem;
  sink.s_hi = k1_smem + (size_t)R * cap;
  sink.s_key = k1_smem + (size_t)(R + nwide) * cap;
  sink.hash_shift = 32 - (31 - __clz(cap | 1));
  if (SINK != SINK_GLOBAL) {
    for (uint32_t i = threadIdx.x; i < (R + nwide) * cap; i += blockDim.x) k1_smem[i] = 0;
    if (SINK == SINK_HASH)
      for (uint32_t i = threadIdx.x; i < cap; i += blockDim.x) sink.s_key[i] = kHashEmpty;
    __syncthreads();
  }

  const uint32_t lane = threadIdx.x & 31u;
  const size_t warps_total = (size_t)gridDim.x * (blockDim.x >> 5);
  const size_t warp_global = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const size_t tiles = (n + kK1Tile - 1) / kK1Tile;
  const size_t chunks = (tiles + chunk_tiles - 1) / chunk_tiles;

  for (size_t c = warp_global; c < chunks; c += warps_total) {
    const size_t t0 = c * chunk_tiles;
    const size_t t1 = (t0 + chunk_tiles < tiles) ? t0 + chunk_tiles : tiles;
    // carry = the still-open run at the end of the previous tile (meaningful on lane 0)
    uint32_t carry_ns = kBadNs;
    long long carry_v[R];
#pragma unroll
    for (int r = 0; r < R; ++r) carry_v[r] = 0;

    for (size_t t = t0; t < t1; ++t) {
      const size_t base = t * kK1Tile + lane * 4;
      uint32_t ns[4];
      long long q[R][4];
      if (VEC && base + 4 <= n) {
        const uint4 v = ldg_nc_u32x4(ns_id + base);
        ns[0] = v.x; ns[1] = v.y; ns[2] = v.z; ns[3] = v.w;
#pragma unroll
        for (int r = 0; r < R; ++r) ldg_nc_i64x4(req.p[r] + base, q[r]);
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const bool ok = base + k < n;
          ns[k] = ok ? ns_id[base + k] : kBadNs;
#pragma unroll
          for (int r = 0; r < R; ++r) q[r][k] = ok ? req.p[r][base + k] : 0;
        }
      }
      // 1. runs inside the lane; interior runs are published by the lane itself
      const uint32_t first_ns = ns[0];
      uint32_t cur = ns[0];
      long long acc[R];
#pragma unroll
      for (int r = 0; r < R; ++r) acc[r] = q[r][0];
      bool uniform = true;
#pragma unroll
      for (int k = 1; k < 4; ++k) {
        if (ns[k] != cur) {
          // the first run of a lane may continue the previous lane's run only when the lane is
          // uniform; otherwise it is published here, which is still exact (sums commute)
          sink.add(cur, acc);
          cur = ns[k];
          uniform = false;
#pragma unroll
          for (int r = 0; r < R; ++r) acc[r] = q[r][k];
        } else {
#pragma unroll
          for (int r = 0; r < R; ++r) acc[r] += q[r][k];
        }
      }
      // 2. segmented scan over the lanes' open (tail) runs
      uint32_t prev_tail = __shfl_up_sync(0xFFFFFFFFu, cur, 1);
      if (lane == 0) prev_tail = carry_ns;
      const bool head = !uniform || first_ns != prev_tail;
      const uint32_t heads = __ballot_sync(0xFFFFFFFFu, head);
      if (lane == 0) {
        if (head) {
          sink.add(carry_ns, carry_v);  // the carried run ended at the tile boundary
        } else {
#pragma unroll
          for (int r = 0; r < R; ++r) acc[r] += carry_v[r];
        }
      }
      if (heads != 0xFFFFFFFFu) {  // at least one lane continues its predecessor
        const uint32_t below = (heads | 1u) & (0xFFFFFFFFu >> (31u - lane));
        const uint32_t dist = lane - (31u - __clz(below));
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
#pragma unroll
          for (int r = 0; r < R; ++r) {
            const long long o = __shfl_up_sync(0xFFFFFFFFu, acc[r], d);
            if (dist >= (uint32_t)d) acc[r] += o;
          }
        }
      }
      // a lane publishes its scanned run when the next lane starts a new one
      const bool next_is_head = lane < 31u && ((heads >> (lane + 1u)) & 1u);
      if (next_is_head) sink.add(cur, acc);
      // 3. lane 31's run stays open and is carried to the next tile on lane 0
      carry_ns = __shfl_sync(0xFFFFFFFFu, cur, 31);
#pragma unroll
      for (int r = 0; r < R; ++r) carry_v[r] = __shfl_sync(0xFFFFFFFFu, acc[r], 31);
    }
    if (lane == 0) sink.add(carry_ns, carry_v);
  }

  if (SINK != SINK_GLOBAL) {
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < R * cap; i += blockDim.x) {
      const uint32_t r = i / cap, s = i - r * cap;
      const uint32_t hi = ((wide_mask >> r) & 1u) ? sink.s_hi[sink.hi_slot(r) * cap + s] : 0u;
      const unsigned long long v = ((unsigned long long)hi << 32) | sink.s_lo[i];
      if (v) {
        const uint32_t ns = (SINK == SINK_HASH) ? sink.s_key[s] : s;
        red_add_i64(used_at(used, NS, r, ns), (long long)v);
      }
    }
  }
}

// ------------------------------------------------------------------ K1, arrival-order input
// When the caller does not promise grouped input, merging runs buys nothing (two neighbouring pods
// share a namespace with probability sum p_i^2), and the segmented machinery above costs ~120
// issued instructions per pod slot.  k1_scatter is the lean form for that case: every thread takes
// U groups of 4 consecutive pods, issues all of its loads, then all of its shared-memory atomics
// (native 32-bit ATOMS on shared-window addresses, so no generic-address conversion per access),
// and only afterwards looks at the returned old values to propagate the rare carries.
__device__ __forceinline__ uint32_t atoms_add_u32(uint32_t saddr, uint32_t v) {
  uint32_t old;
  asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(old) : "r"(saddr), "r"(v));
  return old;
}
__device__ __forceinline__ void reds_add_u32(uint32_t saddr, uint32_t v) {
  asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(saddr), "r"(v));
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t saddr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(saddr));
  return v;
}
__device__ __forceinline__ uint32_t atoms_cas_u32(uint32_t saddr, uint32_t cmp, uint32_t val) {
  uint32_t old;
  asm volatile("atom.shared.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "r"(saddr), "r"(cmp), "r"(val));
  return old;
}

template <int R, int SINK>
struct ScatterCtx {
  long long* used;
  uint32_t NS, cap, wide_mask, hash_shift;
  uint32_t lo_base[R];  // shared-window byte address of column r's low words
  uint32_t hi_base[R];  // ... of its high words (wide columns only)
  uint32_t key_base;

  // slot of a namespace in the CTA-private table, or kHashEmpty when it has none (hash sink, cold)
  __device__ __forceinline__ uint32_t slot_of(uint32_t ns) const {
    if (SINK == SINK_DIRECT) return ns;
    const uint32_t s = slot_hash(ns, hash_shift);
    const uint32_t kaddr = key_base + s * 4u;
    uint32_t k = lds_u32(kaddr);
    if (k == kHashEmpty) {
      k = atoms_cas_u32(kaddr, kHashEmpty, ns);
      if (k == kHashEmpty) k = ns;
    }
    return k == ns ? s : kHashEmpty;
  }
};

template <int R, int SINK, int PODS>
__device__ __forceinline__ void scatter_pods(const ScatterCtx<R, SINK>& C, const uint32_t (&ns)[PODS],
                                             const long long (&q)[R][PODS]) {
  uint32_t slot[PODS];
  uint32_t old[R][PODS];
  if (SINK == SINK_HASH) {
    // Two candidate slots per namespace (top and bottom bits of one mixed word): with a single
    // slot, two HOT namespaces landing on the same slot send the loser's millions of pods to one
    // global address.  All key probes of the group are issued before any is examined.
    uint32_t s1[PODS], s2[PODS], k1[PODS], k2[PODS];
#pragma unroll
    for (int k = 0; k < PODS; ++k) {
      const uint32_t h = mix32(ns[k]);
      s1[k] = h >> C.hash_shift;
      s2[k] = h & (C.cap - 1u);
      k1[k] = lds_u32(C.key_base + s1[k] * 4u);
      k2[k] = lds_u32(C.key_base + s2[k] * 4u);
    }
#pragma unroll
    for (int k = 0; k < PODS; ++k) {
      uint32_t got = kHashEmpty;
      if (ns[k] < C.NS) {
        if (k1[k] == ns[k]) got = s1[k];
        else if (k2[k] == ns[k]) got = s2[k];
        else {
          if (k1[k] == kHashEmpty) {  // claim-once: first namespace to arrive owns the slot
            const uint32_t o = atoms_cas_u32(C.key_base + s1[k] * 4u, kHashEmpty, ns[k]);
            if (o == kHashEmpty || o == ns[k]) got = s1[k];
          }
          if (got == kHashEmpty && k2[k] == kHashEmpty && s2[k] != s1[k]) {
            // a namespace never holds two slots: only one thread per namespace can reach this CAS
            // with s1 refused, because s1's refusal is permanent once its owner is somebody else
            const uint32_t o = atoms_cas_u32(C.key_base + s2[k] * 4u, kHashEmpty, ns[k]);
            if (o == kHashEmpty || o == ns[k]) got = s2[k];
          }
        }
      }
      slot[k] = got;
    }
  } else {
#pragma unroll
    for (int k = 0; k < PODS; ++k) slot[k] = ns[k] < C.NS ? ns[k] : kHashEmpty;
  }
  // phase 1: every low-word atomic of the group is in flight before any result is consumed
#pragma unroll
  for (int k = 0; k < PODS; ++k) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const uint32_t lo = (uint32_t)(unsigned long long)q[r][k];
      old[r][k] = 0;
      if (slot[k] != kHashEmpty && lo) old[r][k] = atoms_add_u32(C.lo_base[r] + slot[k] * 4u, lo);
    }
  }
  // phase 2: high words and carries
#pragma unroll
  for (int k = 0; k < PODS; ++k) {
    if (ns[k] >= C.NS) continue;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const unsigned long long v = (unsigned long long)q[r][k];
      if (slot[k] == kHashEmpty) {  // cold namespace of the hash sink: straight to the global table
        if (v) red_add_i64(used_at(C.used, C.NS, r, ns[k]), (long long)v);
        continue;
      }
      const uint32_t lo = (uint32_t)v;
      const uint32_t hi = (uint32_t)(v >> 32) + (uint32_t)(old[r][k] + lo < old[r][k]);
      if (hi) {
        if ((C.wide_mask >> r) & 1u) reds_add_u32(C.hi_base[r] + slot[k] * 4u, hi);
        else red_add_i64(used_at(C.used, C.NS, r, ns[k]), (long long)((unsigned long long)hi << 32));
      }
    }
  }
}

template <int R, int SINK, bool VEC, int THREADS, int U>
__global__ void __launch_bounds__(THREADS, 1)
k1_scatter(const uint32_t* __restrict__ ns_id, ReqCols<R> req, size_t n, long long* used, uint32_t NS,
           uint32_t cap, uint32_t wide_mask) {
  extern __shared__ __align__(16) uint32_t k1_smem[];
  const uint32_t nwide = __popc(wide_mask & ((1u << R) - 1u));
  const uint32_t words = (R + nwide) * cap;
  for (uint32_t i = threadIdx.x; i < words; i += THREADS) k1_smem[i] = 0;
  if (SINK == SINK_HASH)
    for (uint32_t i = threadIdx.x; i < cap; i += THREADS) k1_smem[words + i] = kHashEmpty;
  ScatterCtx<R, SINK> C;
  C.used = used; C.NS = NS; C.cap = cap; C.wide_mask = wide_mask;
  C.hash_shift = 32 - (31 - __clz(cap | 1));
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(k1_smem);
#pragma unroll
  for (int r = 0; r < R; ++r) {
    C.lo_base[r] = sbase + (uint32_t)r * cap * 4u;
    C.hi_base[r] = sbase + (R + __popc(wide_mask & ((1u << r) - 1u))) * cap * 4u;
  }
  C.key_base = sbase + words * 4u;
  __syncthreads();

  constexpr int PODS = R <= 4 ? 4 : 2;
  const size_t groups = VEC ? n / PODS : 0;
  const size_t stride = (size_t)gridDim.x * THREADS;
  size_t g = (size_t)blockIdx.x * THREADS + threadIdx.x;
  // U groups per trip: all loads first (U*(1+R) independent vector requests per thread)
  for (; g + (U - 1) * stride < groups; g += U * stride) {
    uint32_t ns[U][PODS];
    long long q[U][R][PODS];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t base = (g + u * stride) * PODS;
      ldg_ns<PODS>(ns_id + base, ns[u]);
#pragma unroll
      for (int r = 0; r < R; ++r) ldg_req<PODS>(req.p[r] + base, q[u][r]);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) scatter_pods<R, SINK, PODS>(C, ns[u], q[u]);
  }
  for (; g < groups; g += stride) {
    uint32_t ns[PODS];
    long long q[R][PODS];
    const size_t base = g * PODS;
    ldg_ns<PODS>(ns_id + base, ns);
#pragma unroll
    for (int r = 0; r < R; ++r) ldg_req<PODS>(req.p[r] + base, q[r]);
    scatter_pods<R, SINK, PODS>(C, ns, q);
  }
  // scalar tail (and everything when the columns are not 16/32-byte aligned)
  for (size_t i = groups * PODS + (size_t)blockIdx.x * THREADS + threadIdx.x; i < n; i += stride) {
    uint32_t ns[1] = {ns_id[i]};
    long long q[R][1];
#pragma unroll
    for (int r = 0; r < R; ++r) q[r][0] = req.p[r][i];
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

// ------------------------------------------------------------------ K1, cluster-distributed table
// Between ~14k and ~230k namespaces a CTA-private table no longer fits one SM, and the hash sink
// pays a global red (L2 atomic, ~230 G/s chip-wide) for every cold namespace.  Thread-block
// clusters give CSIZE neighbouring SMs one table between them: CTA `ns % CSIZE` of the cluster owns
// namespace ns, and every CTA scatters its own slice of pods into whichever SM owns the slot with
// shared::cluster atomics over distributed shared memory — no misses, no global traffic until the
// single flush at the end.
__device__ __forceinline__ uint32_t mapa_u32(uint32_t saddr, uint32_t cta_rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(cta_rank));
  return r;
}
__device__ __forceinline__ uint32_t atomc_add_u32(uint32_t caddr, uint32_t v) {
  uint32_t old;
  asm volatile("atom.shared::cluster.add.u32 %0, [%1], %2;" : "=r"(old) : "r"(caddr), "r"(v) : "memory");
  return old;
}
__device__ __forceinline__ void redc_add_u32(uint32_t caddr, uint32_t v) {
  asm volatile("red.shared::cluster.add.u32 [%0], %1;" ::"r"(caddr), "r"(v) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}

template <int R, int PODS, int CSIZE>
__device__ __forceinline__ void scatter_pods_cluster(long long* used, uint32_t NS, uint32_t wide_mask, uint32_t sbase,
                                                     uint32_t cap, const uint32_t (&hi_slot)[R], const uint32_t (&ns)[PODS],
                                                     const long long (&q)[R][PODS]) {
  constexpr uint32_t kLog = CSIZE == 16 ? 4 : CSIZE == 8 ? 3 : CSIZE == 4 ? 2 : 1;
  uint32_t base[PODS];   // shared::cluster address of slot 0 column 0 in the owning CTA
  uint32_t local[PODS];
  uint32_t old[R][PODS];
#pragma unroll
  for (int k = 0; k < PODS; ++k) {
    local[k] = ns[k] >> kLog;
    base[k] = mapa_u32(sbase, ns[k] & (CSIZE - 1));
  }
#pragma unroll
  for (int k = 0; k < PODS; ++k) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const uint32_t lo = (uint32_t)(unsigned long long)q[r][k];
      old[r][k] = 0;
      if (ns[k] < NS && lo) old[r][k] = atomc_add_u32(base[k] + ((uint32_t)r * cap + local[k]) * 4u, lo);
    }
  }
#pragma unroll
  for (int k = 0; k < PODS; ++k) {
    if (ns[k] >= NS) continue;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const unsigned long long v = (unsigned long long)q[r][k];
      const uint32_t lo = (uint32_t)v;
      const uint32_t hi = (uint32_t)(v >> 32) + (uint32_t)(old[r][k] + lo < old[r][k]);
      if (hi) {
        if ((wide_mask >> r) & 1u) redc_add_u32(base[k] + ((R + hi_slot[r]) * cap + local[k]) * 4u, hi);
        else red_add_i64(used_at(used, NS, r, ns[k]), (long long)((unsigned long long)hi << 32));
      }
    }
  }
}

// cap = slots per CTA (>= ceil(NS / CSIZE)); launched with cluster dimension CSIZE
template <int R, bool VEC, int CSIZE>
__global__ void __launch_bounds__(1024, 1)
k1_scatter_cluster(const uint32_t* __restrict__ ns_id, ReqCols<R> req, size_t n, long long* used, uint32_t NS,
                   uint32_t cap, uint32_t wide_mask) {
  extern __shared__ __align__(16) uint32_t k1_smem[];
  constexpr int THREADS = 1024;
  constexpr uint32_t kLog = CSIZE == 16 ? 4 : CSIZE == 8 ? 3 : CSIZE == 4 ? 2 : 1;
  const uint32_t nwide = __popc(wide_mask & ((1u << R) - 1u));
  const uint32_t words = (R + nwide) * cap;
  for (uint32_t i = threadIdx.x; i < words; i += THREADS) k1_smem[i] = 0;
  uint32_t hi_slot[R];
#pragma unroll
  for (int r = 0; r < R; ++r) hi_slot[r] = __popc(wide_mask & ((1u << r) - 1u));
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(k1_smem);
  cluster_sync_all();  // every table of the cluster is zeroed before anyone scatters into it

  constexpr int PODS = R <= 4 ? 4 : 2;
  const size_t groups = VEC ? n / PODS : 0;
  const size_t stride = (size_t)gridDim.x * THREADS;
  for (size_t g = (size_t)blockIdx.x * THREADS + threadIdx.x; g < groups; g += stride) {
    uint32_t ns[PODS];
    long long q[R][PODS];
    const size_t base = g * PODS;
    ldg_ns<PODS>(ns_id + base, ns);
#pragma unroll
    for (int r = 0; r < R; ++r) ldg_req<PODS>(req.p[r] + base, q[r]);
    scatter_pods_cluster<R, PODS, CSIZE>(used, NS, wide_mask, sbase, cap, hi_slot, ns, q);
  }
  for (size_t i = groups * PODS + (size_t)blockIdx.x * THREADS + threadIdx.x; i < n; i += stride) {
    uint32_t ns[1] = {ns_id[i]};
    long long q[R][1];
#pragma unroll
    for (int r = 0; r < R; ++r) q[r][0] = req.p[r][i];
    scatter_pods_cluster<R, 1, CSIZE>(used, NS, wide_mask, sbase, cap, hi_slot, ns, q);
  }

  cluster_sync_all();  // all remote atomics into this CTA's table have landed
  const uint32_t me = cluster_ctarank();
  for (uint32_t i = threadIdx.x; i < R * cap; i += THREADS) {
    const uint32_t r = i / cap, sl = i - r * cap;
    const uint32_t hi = ((wide_mask >> r) & 1u) ? k1_smem[(R + hi_slot[r]) * cap + sl] : 0u;
    const unsigned long long v = ((unsigned long long)hi << 32) | k1_smem[i];
    const uint32_t ns = (sl << kLog) | me;
    if (v && ns < NS) red_add_i64(used_at(used, NS, r, ns), (long long)v);
  }
}

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
template <int R, bool HAS_MASK, bool VEC, typename VT>
__global__ void __launch_bounds__(kK2Threads)
k2_quota_eval(const uint32_t* __restrict__ ns_id, ReqCols<R> req, const uint8_t* __restrict__ req_mask,
              size_t n, const long long* __restrict__ head, const uint8_t* __restrict__ hard_mask,
              uint32_t NS, uint32_t compute_mask, VT* __restrict__ verdict) {
  constexpr int PODS = R <= 4 ? 4 : 2;
  constexpr int SHIFT = sizeof(VT) == 1 ? 4 : 8;
  const size_t groups = VEC ? n / PODS : 0;
  for (size_t g = (size_t)blockIdx.x * kK2Threads + threadIdx.x; g < groups;
       g += (size_t)gridDim.x * kK2Threads) {
    const size_t base = g * PODS;
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

// ------------------------------------------------------------------ K3 admit_greedy (row N1)
// Kubernetes-faithful sequential admission inside a batch: per namespace, in arrival order, a pod
// is admitted iff its verdict against the CURRENT usage is 0, and only admitted pods add to usage.
// The recurrence is order-dependent inside a namespace and independent across namespaces, so:
//   1. pod indices are grouped by namespace with a stable radix sort (arrival order preserved);
//   2. one warp walks one namespace, 32 pods per tile, alternating two warp-wide steps:
//      admit-run : exclusive prefix sums of the requests assume "everyone so far is admitted"; the
//                  FIRST lane whose check fails under that prefix is the true first rejection
//                  (everything before it really was admitted), found with one ballot;
//      reject-run: while usage stands still, every following pod that does not fit on its own is
//                  rejected, up to the first one that fits — again one ballot.
// A tile costs one scan per admit->reject alternation, not one step per pod.
static __global__ void k_greedy_keys(const uint32_t* __restrict__ ns_id, size_t n, uint32_t NS,
                                     uint32_t* keys, uint32_t* idx) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const uint32_t ns = ns_id[i];
    keys[i] = ns < NS ? ns : NS;  // unknown namespaces sort behind every real one
    idx[i] = (uint32_t)i;
  }
}

// seg[k] = first position in the sorted key array with key >= k, for k = 0..NS+1
static __global__ void k_greedy_segments(const uint32_t* __restrict__ sorted_keys, size_t n, uint32_t NS,
                                         uint32_t* seg) {
  for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k <= (size_t)NS + 1;
       k += (size_t)gridDim.x * blockDim.x) {
    size_t lo = 0, hi = n;
    while (lo < hi) {
      const size_t mid = lo + ((hi - lo) >> 1);
      if (sorted_keys[mid] < (uint32_t)k) lo = mid + 1; else hi = mid;
    }
    seg[k] = (uint32_t)lo;
  }
}

static __global__ void k_greedy_bad_ns(const uint32_t* __restrict__ order, const uint32_t* __restrict__ seg,
                                       uint32_t NS, uint8_t* verdict) {
  const size_t beg = seg[NS], end = seg[NS + 1];
  for (size_t i = beg + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < end; i += (size_t)gridDim.x * blockDim.x)
    verdict[order[i]] = 0xFFu;
}

// requests permuted into grouped order, so the sequential walkers below stream contiguous memory
template <int R, bool HAS_MASK>
static __global__ void __launch_bounds__(256)
k_greedy_gather(const uint32_t* __restrict__ order, ReqCols<R> req, const uint8_t* __restrict__ req_mask, size_t n,
                ReqColsOut<R> out, uint8_t* __restrict__ mask_out) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const uint32_t idx = order[i];
#pragma unroll
    for (int r = 0; r < R; ++r) out.p[r][i] = req.p[r][idx];
    if (HAS_MASK) mask_out[i] = req_mask[idx];
  }
}

template <int R>
struct GreedyPod {
  long long q[R];
  uint32_t mask;
};

// verdict of one pod against usage T
template <int R>
__device__ __forceinline__ uint32_t greedy_verdict(const GreedyPod<R>& p, const long long (&T)[R],
                                                   const long long (&H)[R], uint32_t hm, uint32_t compute_mask) {
  uint32_t missing = 0, exceeded = 0;
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const bool constrained = (hm >> r) & 1u, present = (p.mask >> r) & 1u;
    if (constrained && !present && ((compute_mask >> r) & 1u)) missing |= 1u << r;
    if (constrained && present && p.q[r] > 0 && p.q[r] > sat_sub_i64(H[r], T[r])) exceeded |= 1u << r;
  }
  return (missing << 4) | exceeded;
}

template <int R, bool HAS_MASK>
__global__ void __launch_bounds__(256)
k3_admit_greedy(const uint32_t* __restrict__ order, const uint32_t* __restrict__ seg, ReqCols<R> req,
                const uint8_t* __restrict__ req_mask, long long* used, const long long* __restrict__ hard,
                const uint8_t* __restrict__ hard_mask, uint32_t NS, uint32_t compute_mask,
                uint8_t* __restrict__ verdict, uint32_t team_min) {
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t warps_total = gridDim.x * (blockDim.x >> 5);
  // heavy namespaces first would be ideal; a strided walk at least spreads neighbours over warps
  for (uint32_t ns = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); ns < NS; ns += warps_total) {
    const uint32_t beg = seg[ns], end = seg[ns + 1];
    if (beg == end || end - beg >= team_min) continue;  // long segments belong to the CTA-wide kernel
    long long S[R], H[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      S[r] = used[(size_t)r * NS + ns];
      H[r] = hard[(size_t)r * NS + ns];
    }
    const uint32_t hm = hard_mask[ns];

    // `req` / `req_mask` are the GROUPED columns (k_greedy_gather): tile t+1 is in flight while tile t is decided
    GreedyPod<R> p0;
#pragma unroll
    for (int r = 0; r < R; ++r) p0.q[r] = (beg + lane < end) ? req.p[r][beg + lane] : 0;
    p0.mask = (HAS_MASK && beg + lane < end) ? req_mask[beg + lane] : 0xFFu;
    for (uint32_t t = beg; t < end; t += 32) {
      const bool valid = t + lane < end;
      const GreedyPod<R> p = p0;
      const bool valid1 = t + 32 + lane < end;
#pragma unroll
      for (int r = 0; r < R; ++r) p0.q[r] = valid1 ? req.p[r][t + 32 + lane] : 0;
      p0.mask = (HAS_MASK && valid1) ? req_mask[t + 32 + lane] : 0xFFu;

      uint32_t my_verdict = 0;
      uint32_t pos = 0;  // lanes < pos are decided
      // Monotone fast path.  When every request of the tile is in [0, 2^40] and usage is in
      // [0, 2^62], usage only grows and cannot overflow, so a pod that does not fit on its own NOW
      // can never fit later: all such pods are struck out at once, and only pods that lose their
      // place because of admissions inside this very tile cost an extra round.
      bool tame = true;
#pragma unroll
      for (int r = 0; r < R; ++r) tame = tame && p.q[r] >= 0 && p.q[r] <= (1ll << 40) && S[r] >= 0 && S[r] <= (1ll << 62);
      if (__all_sync(0xFFFFFFFFu, tame)) {
        long long S0[R];
#pragma unroll
        for (int r = 0; r < R; ++r) S0[r] = S[r];
        uint32_t live = __ballot_sync(0xFFFFFFFFu, valid), admitted = 0;
        while (live) {
          const uint32_t v_alone = greedy_verdict<R>(p, S, H, hm, compute_mask);
          live &= ~__ballot_sync(0xFFFFFFFFu, v_alone != 0);
          if (!live) break;
          const bool mine = (live >> lane) & 1u;
          long long E[R], incl[R];
#pragma unroll
          for (int r = 0; r < R; ++r) incl[r] = mine ? p.q[r] : 0;
#pragma unroll
          for (int d = 1; d < 32; d <<= 1) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
              const long long o = __shfl_up_sync(0xFFFFFFFFu, incl[r], d);
              if (lane >= (uint32_t)d) incl[r] += o;
            }
          }
          long long T[R];
#pragma unroll
          for (int r = 0; r < R; ++r) {
            E[r] = incl[r] - (mine ? p.q[r] : 0);
            T[r] = S[r] + E[r];
          }
          const uint32_t fail = __ballot_sync(0xFFFFFFFFu, mine && greedy_verdict<R>(p, T, H, hm, compute_mask) != 0);
          if (!fail) {
            admitted |= live;
#pragma unroll
            for (int r = 0; r < R; ++r) S[r] += __shfl_sync(0xFFFFFFFFu, incl[r], 31);
            break;
          }
          const uint32_t v = __ffs(fail) - 1u;  // first pod squeezed out by admissions of this tile
          admitted |= live & ((1u << v) - 1u);
#pragma unroll
          for (int r = 0; r < R; ++r) S[r] += __shfl_sync(0xFFFFFFFFu, E[r], v);
          live &= ~((2u << v) - 1u);
        }
        // reason bits of the rejected pods are taken against the usage they actually met
        long long incl[R], T[R];
        const bool adm = (admitted >> lane) & 1u;
#pragma unroll
        for (int r = 0; r < R; ++r) incl[r] = adm ? p.q[r] : 0;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
#pragma unroll
          for (int r = 0; r < R; ++r) {
            const long long o = __shfl_up_sync(0xFFFFFFFFu, incl[r], d);
            if (lane >= (uint32_t)d) incl[r] += o;
          }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) T[r] = S0[r] + incl[r] - (adm ? p.q[r] : 0);
        my_verdict = adm ? 0u : greedy_verdict<R>(p, T, H, hm, compute_mask);
        pos = 32;
      }
      while (pos < 32) {
        // ---- admit-run: exclusive prefix of the adds of lanes >= pos
        long long E[R], incl[R];
#pragma unroll
        for (int r = 0; r < R; ++r) incl[r] = (valid && lane >= pos) ? p.q[r] : 0;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
#pragma unroll
          for (int r = 0; r < R; ++r) {
            const long long o = __shfl_up_sync(0xFFFFFFFFu, incl[r], d);
            if (lane >= (uint32_t)d) incl[r] = (long long)((unsigned long long)incl[r] + (unsigned long long)o);
          }
        }
        long long T[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
          E[r] = (long long)((unsigned long long)incl[r] - (unsigned long long)((valid && lane >= pos) ? p.q[r] : 0));
          T[r] = (long long)((unsigned long long)S[r] + (unsigned long long)E[r]);
        }
        const uint32_t v_run = greedy_verdict<R>(p, T, H, hm, compute_mask);
        const uint32_t fail = __ballot_sync(0xFFFFFFFFu, valid && lane >= pos && v_run != 0);
        if (fail == 0) {  // everyone left in the tile is admitted
#pragma unroll
          for (int r = 0; r < R; ++r)
            S[r] = (long long)((unsigned long long)S[r] + (unsigned long long)__shfl_sync(0xFFFFFFFFu, incl[r], 31));
          break;
        }
        const uint32_t v = __ffs(fail) - 1u;  // first true rejection
        if (lane == v) my_verdict = v_run;
#pragma unroll
        for (int r = 0; r < R; ++r)
          S[r] = (long long)((unsigned long long)S[r] + (unsigned long long)__shfl_sync(0xFFFFFFFFu, E[r], v));
        pos = v + 1;
        if (pos >= 32) break;
        // ---- reject-run: usage stands still at S until somebody fits on their own
        const uint32_t v_alone = greedy_verdict<R>(p, S, H, hm, compute_mask);
        const uint32_t later = 0xFFFFFFFFu << pos;
        const uint32_t fits = __ballot_sync(0xFFFFFFFFu, valid && v_alone == 0) & later;
        const uint32_t stop = fits ? (__ffs(fits) - 1u) : 32u;
        if (lane >= pos && lane < stop && valid) my_verdict = v_alone;
        pos = stop;
        if (pos >= 32) break;
      }
      if (valid) verdict[order[t + lane]] = (uint8_t)my_verdict;
    }
    if (lane == 0) {
#pragma unroll
      for (int r = 0; r < R; ++r) used[(size_t)r * NS + ns] = S[r];
    }
  }
}

// Hot namespaces (Zipf heads) would keep one warp busy for hundreds of milliseconds, so segments of
// kGreedyTeamMin pods or more are walked by a whole 1024-thread CTA instead: the same admit-run /
// reject-run alternation on tiles of 1024 pods, with a two-level (warp shuffle + shared memory)
// exclusive scan and a shared-memory atomicMin to find the first rejection / first fit.
constexpr uint32_t kGreedyTeamMin = 32768;
constexpr int kGreedyTeam = 1024;
constexpr int kGreedyDepth = 4;  // tiles of requests in flight per CTA (ring of R * 8 KB per tile)

template <int R, bool HAS_MASK>
__global__ void __launch_bounds__(kGreedyTeam, 1)
k3_admit_greedy_team(const uint32_t* __restrict__ order, const uint32_t* __restrict__ seg, ReqCols<R> req,
                     const uint8_t* __restrict__ req_mask, long long* used, const long long* __restrict__ hard,
                     const uint8_t* __restrict__ hard_mask, uint32_t NS, uint32_t compute_mask,
                     uint8_t* __restrict__ verdict) {
  extern __shared__ __align__(16) long long g_ring[];  // [kGreedyDepth][R][kGreedyTeam]
  __shared__ long long s_wtot[R][32];    // per-warp totals, then their exclusive prefix
  __shared__ long long s_total[R];
  __shared__ long long s_E[R][kGreedyTeam];
  __shared__ uint32_t s_first[2], s_fit[2];
  const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  constexpr uint32_t BIG = 0xFFFFFFFFu;

  for (uint32_t ns = blockIdx.x; ns < NS; ns += gridDim.x) {
    const uint32_t beg = seg[ns], end = seg[ns + 1];
    if (end - beg < kGreedyTeamMin) continue;  // short segments belong to the warp kernel
    long long S[R], H[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      S[r] = used[(size_t)r * NS + ns];
      H[r] = hard[(size_t)r * NS + ns];
    }
    const uint32_t hm = hard_mask[ns];
    __syncthreads();  // previous namespace's shared state is no longer read

    // A single CTA walks this namespace, so its own loads are all the memory parallelism there is:
    // kGreedyDepth tiles of requests are kept in flight with cp.async into a shared-memory ring
    // (every thread only ever reads back the elements it copied itself, so no barrier is needed).
    auto issue_tile = [&](uint32_t t, uint32_t slot) {
      if (t + tid < end) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&g_ring[((size_t)slot * R + r) * kGreedyTeam + tid]);
          asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(req.p[r] + t + tid) : "memory");
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    uint32_t mask_q[kGreedyDepth];
#pragma unroll
    for (int d = 0; d < kGreedyDepth; ++d) {
      const uint32_t t = beg + d * kGreedyTeam;
      issue_tile(t, d);
      mask_q[d] = (HAS_MASK && t < end && t + tid < end) ? req_mask[t + tid] : 0xFFu;
    }
    uint32_t iter = 0, slot = 0;
    for (uint32_t t = beg; t < end; t += kGreedyTeam) {
      const bool valid = t + tid < end;
      asm volatile("cp.async.wait_group %0;" ::"n"(kGreedyDepth - 1) : "memory");
      GreedyPod<R> p;
#pragma unroll
      for (int r = 0; r < R; ++r) p.q[r] = valid ? g_ring[((size_t)slot * R + r) * kGreedyTeam + tid] : 0;
      p.mask = mask_q[0];
#pragma unroll
      for (int d = 0; d + 1 < kGreedyDepth; ++d) mask_q[d] = mask_q[d + 1];
      {
        const uint32_t tn = t + kGreedyDepth * kGreedyTeam;
        issue_tile(tn < end ? tn : end, slot);  // refills the slot just consumed (or commits an empty group)
        mask_q[kGreedyDepth - 1] = (HAS_MASK && tn < end && tn + tid < end) ? req_mask[tn + tid] : 0xFFu;
      }
      slot = slot + 1 == kGreedyDepth ? 0 : slot + 1;

      uint32_t my_verdict = 0, pos = 0;
      // block-wide exclusive prefix sum of x[] (0 for threads that do not take part); also leaves the
      // per-thread prefixes in s_E and the block total in s_total.  Two barriers.
      auto block_exclusive = [&](const long long (&x)[R], long long (&E)[R]) {
        long long incl[R];
#pragma unroll
        for (int r = 0; r < R; ++r) incl[r] = x[r];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
#pragma unroll
          for (int r = 0; r < R; ++r) {
            const long long o = __shfl_up_sync(0xFFFFFFFFu, incl[r], d);
            if (lane >= (uint32_t)d) incl[r] = (long long)((unsigned long long)incl[r] + (unsigned long long)o);
          }
        }
        if (lane == 31) {
#pragma unroll
          for (int r = 0; r < R; ++r) s_wtot[r][warp] = incl[r];
        }
        __syncthreads();
        if (warp == 0) {
#pragma unroll
          for (int r = 0; r < R; ++r) {
            const long long mine = s_wtot[r][lane];
            long long y = mine;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
              const long long o = __shfl_up_sync(0xFFFFFFFFu, y, d);
              if (lane >= (uint32_t)d) y = (long long)((unsigned long long)y + (unsigned long long)o);
            }
            s_wtot[r][lane] = (long long)((unsigned long long)y - (unsigned long long)mine);
            if (lane == 31) s_total[r] = y;
          }
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < R; ++r) {
          E[r] = (long long)((unsigned long long)s_wtot[r][warp] + (unsigned long long)incl[r] - (unsigned long long)x[r]);
          s_E[r][tid] = E[r];
        }
      };
      // Monotone fast path (see the warp kernel): requests in [0, 2^40], usage in [0, 2^62].
      bool tame = true;
#pragma unroll
      for (int r = 0; r < R; ++r) tame = tame && p.q[r] >= 0 && p.q[r] <= (1ll << 40) && S[r] >= 0 && S[r] <= (1ll << 62);
      if (__syncthreads_and(tame)) {
        long long S0[R];
#pragma unroll
        for (int r = 0; r < R; ++r) S0[r] = S[r];
        bool live = valid, adm = false;
        while (true) {
          if (live && greedy_verdict<R>(p, S, H, hm, compute_mask) != 0) live = false;  // can never fit again
          if (!__syncthreads_or(live)) break;  // saturated: nobody left to consider
          const uint32_t par = iter & 1u;
          ++iter;
          if (tid == 0) s_first[par] = BIG;
          long long x[R], E[R], T[R];
#pragma unroll
          for (int r = 0; r < R; ++r) x[r] = live ? p.q[r] : 0;
          block_exclusive(x, E);  // its barriers also publish the reset of s_first[par]
#pragma unroll
          for (int r = 0; r < R; ++r) T[r] = S[r] + E[r];
          const uint32_t fail = __ballot_sync(0xFFFFFFFFu, live && greedy_verdict<R>(p, T, H, hm, compute_mask) != 0);
          if (fail && lane == (uint32_t)(__ffs(fail) - 1)) atomicMin(&s_first[par], tid);
          __syncthreads();
          const uint32_t v = s_first[par];
          if (v == BIG) {
            adm = adm || live;
#pragma unroll
            for (int r = 0; r < R; ++r) S[r] += s_total[r];
            break;
          }
          if (live && tid < v) adm = true;
#pragma unroll
          for (int r = 0; r < R; ++r) S[r] += s_E[r][v];
          if (tid <= v) live = false;
        }
        if (__syncthreads_or(adm)) {  // reason bits against the usage each rejected pod actually met
          long long x[R], E[R], T[R];
#pragma unroll
          for (int r = 0; r < R; ++r) x[r] = adm ? p.q[r] : 0;
          block_exclusive(x, E);
#pragma unroll
          for (int r = 0; r < R; ++r) T[r] = S0[r] + E[r];
          my_verdict = adm ? 0u : greedy_verdict<R>(p, T, H, hm, compute_mask);
        } else {                      // nothing admitted in this tile: usage never moved
          my_verdict = valid ? greedy_verdict<R>(p, S0, H, hm, compute_mask) : 0u;
        }
        pos = kGreedyTeam;
        __syncthreads();  // s_E / s_wtot are reused by the next tile
      }
      while (pos < (uint32_t)kGreedyTeam) {
        const uint32_t par = iter & 1u;
        ++iter;
        // ---- admit-run: block-wide exclusive prefix of the adds of threads >= pos
        long long a[R], incl[R];
#pragma unroll
        for (int r = 0; r < R; ++r) incl[r] = a[r] = (valid && tid >= pos) ? p.q[r] : 0;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
#pragma unroll
          for (int r = 0; r < R; ++r) {
            const long long o = __shfl_up_sync(0xFFFFFFFFu, incl[r], d);
            if (lane >= (uint32_t)d) incl[r] = (long long)((unsigned long long)incl[r] + (unsigned long long)o);
          }
        }
        if (lane == 31) {
#pragma unroll
          for (int r = 0; r < R; ++r) s_wtot[r][warp] = incl[r];
        }
        if (tid == 0) { s_first[par] = BIG; s_fit[par] = BIG; }
        __syncthreads();
        if (warp == 0) {  // exclusive scan of the 32 warp totals
#pragma unroll
          for (int r = 0; r < R; ++r) {
            const long long mine = s_wtot[r][lane];
            long long x = mine;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
              const long long o = __shfl_up_sync(0xFFFFFFFFu, x, d);
              if (lane >= (uint32_t)d) x = (long long)((unsigned long long)x + (unsigned long long)o);
            }
            s_wtot[r][lane] = (long long)((unsigned long long)x - (unsigned long long)mine);
            if (lane == 31) s_total[r] = x;
          }
        }
        __syncthreads();
        long long E[R], T[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
          E[r] = (long long)((unsigned long long)s_wtot[r][warp] + (unsigned long long)incl[r] - (unsigned long long)a[r]);
          T[r] = (long long)((unsigned long long)S[r] + (unsigned long long)E[r]);
          s_E[r][tid] = E[r];
        }
        const uint32_t v_run = greedy_verdict<R>(p, T, H, hm, compute_mask);
        const uint32_t fail = __ballot_sync(0xFFFFFFFFu, valid && tid >= pos && v_run != 0);
        if (fail && lane == (uint32_t)(__ffs(fail) - 1)) atomicMin(&s_first[par], tid);
        __syncthreads();
        const uint32_t v = s_first[par];
        if (v == BIG) {  // everyone left in the tile is admitted
#pragma unroll
          for (int r = 0; r < R; ++r) S[r] = (long long)((unsigned long long)S[r] + (unsigned long long)s_total[r]);
          break;
        }
        if (tid == v) my_verdict = v_run;
#pragma unroll
        for (int r = 0; r < R; ++r) S[r] = (long long)((unsigned long long)S[r] + (unsigned long long)s_E[r][v]);
        pos = v + 1;
        if (pos >= (uint32_t)kGreedyTeam) break;
        // ---- reject-run: usage stands still at S until somebody fits on their own
        const uint32_t v_alone = greedy_verdict<R>(p, S, H, hm, compute_mask);
        const uint32_t fits = __ballot_sync(0xFFFFFFFFu, valid && tid >= pos && v_alone == 0);
        if (fits && lane == (uint32_t)(__ffs(fits) - 1)) atomicMin(&s_fit[par], tid);
        __syncthreads();
        const uint32_t stop = s_fit[par] == BIG ? (uint32_t)kGreedyTeam : s_fit[par];
        if (tid >= pos && tid < stop && valid) my_verdict = v_alone;
        pos = stop;
        if (pos >= (uint32_t)kGreedyTeam) break;
      }
      if (valid) verdict[order[t + tid]] = (uint8_t)my_verdict;
    }
    if (tid == 0) {
#pragma unroll
      for (int r = 0; r < R; ++r) used[(size_t)r * NS + ns] = S[r];
    }
  }
}

// ------------------------------------------------------------------ synthetic generator
// Independent re-implementation of the generator spec of SURVEY.md §8(d) (integer only, so the
// CPU oracle reproduces it bit for bit).
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
__host__ __device__ __forceinline__ uint64_t stream_key(uint64_t seed, uint32_t stream) {
  return mix64(seed ^ ((uint64_t)(stream + 1) * 0xD1B54A32D192ED03ull));
}
__host__ __device__ __forceinline__ uint64_t draw(uint64_t key, uint64_t index) {
  return mix64(key ^ (index * 0x9E3779B97F4A7C15ull));
}
__device__ __forceinline__ long long synth_value(uint32_t kind, uint64_t h) {
  const uint32_t t = (uint32_t)(h >> 32);
  if (kind == 0) {
    const uint32_t c = t % 100u;
    return c < 50 ? 0 : c < 80 ? 1 : c < 90 ? 2 : c < 97 ? 4 : 8;
  }
  if (kind == 1) return 250ll << (t % 7u);
  if (kind == 2) return 1ll << (28 + (t % 9u));
  const uint32_t c = t % 100u;
  return c < 90 ? 0 : c < 97 ? 1 : 2;
}

struct SynthArgs {
  uint64_t key_ns, key_mask, key_req[kMaxR];
  uint32_t kind[kMaxR];
  uint32_t R, ns_count, dist, grouped;
  uint64_t step, a, b;
  const uint64_t* cdf;
  uint32_t* ns_id;
  long long* req[kMaxR];
  uint8_t* req_mask;
};

static __global__ void __launch_bounds__(256) k_synth_pods(SynthArgs A, uint64_t first, size_t count) {
  for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < count;
       k += (size_t)gridDim.x * blockDim.x) {
    const uint64_t i = first + k;
    const uint64_t u = A.grouped ? i * A.step : draw(A.key_ns, i);
    uint32_t rank;
    if (A.dist == 1) {
      uint32_t lo = 0, hi = A.ns_count;
      while (lo < hi) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        if (A.cdf[mid] > u) hi = mid; else lo = mid + 1;
      }
      rank = lo < A.ns_count ? lo : A.ns_count - 1;
    } else {
      rank = (uint32_t)(((u >> 32) * (uint64_t)A.ns_count) >> 32);
    }
    A.ns_id[k] = (uint32_t)(((uint64_t)rank * A.a + A.b) % A.ns_count);
    for (uint32_t r = 0; r < A.R; ++r) A.req[r][k] = synth_value(A.kind[r], draw(A.key_req[r], i));
    if (A.req_mask) {
      const uint64_t h = draw(A.key_mask, i);
      uint32_t m = 0xFFu;
      for (uint32_t r = 0; r < A.R; ++r)
        if (((h >> (8 * r)) & 0xFFu) < 8u) m &= ~(1u << r);
      A.req_mask[k] = (uint8_t)m;
    }
  }
}

static __global__ void __launch_bounds__(256)
k_synth_policy(uint64_t key, uint64_t first, size_t count, uint8_t* op, uint16_t* flags) {
  for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < count;
       k += (size_t)gridDim.x * blockDim.x) {
    const uint64_t h = draw(key, first + k);
    op[k] = (uint8_t)(h & 3u);
    flags[k] = (uint16_t)((h >> 8) & 0x3FFu);
  }
}

static __global__ void __launch_bounds__(256)
k_checksum_u8(const uint8_t* __restrict__ v, uint64_t first, size_t count, unsigned long long* out) {
  unsigned long long s = 0;
  for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < count;
       k += (size_t)gridDim.x * blockDim.x)
    s += mix64((first + k) ^ ((uint64_t)v[k] << 56));
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, d);
  if ((threadIdx.x & 31) == 0 && s) atomicAdd(out, s);
}

static __global__ void __launch_bounds__(256)
k_histogram_u8(const uint8_t* __restrict__ v, size_t count, unsigned long long* out256) {
  __shared__ unsigned int h[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  // uniform trip count per CTA, so the periodic flush (u32 bins cannot overflow) may sync
  unsigned int done = 0;
  for (size_t base = (size_t)blockIdx.x * blockDim.x; base < count;
       base += (size_t)gridDim.x * blockDim.x) {
    const size_t k = base + threadIdx.x;
    if (k < count) atomicAdd(&h[v[k]], 1u);
    if (++done == (1u << 20)) {
      __syncthreads();
      if (h[threadIdx.x]) atomicAdd(&out256[threadIdx.x], (unsigned long long)h[threadIdx.x]);
      h[threadIdx.x] = 0;
      done = 0;
      __syncthreads();
    }
  }
  __syncthreads();
  if (h[threadIdx.x]) atomicAdd(&out256[threadIdx.x], (unsigned long long)h[threadIdx.x]);
}

}  // namespace bgc
