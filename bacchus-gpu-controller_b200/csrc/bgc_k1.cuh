// bgc_k1.cuh — K1 usage_reduce (row A3): segmented, scatter and cluster kernels.  Part of libbgc.so's device code; included through bgc_kernels.cuh.
#pragma once
#include "bgc_common.cuh"

namespace bgc {

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
// One lane asks L2 for a contiguous range (TMA-style bulk prefetch, no registers, no shared memory):
// keeps HBM requests in flight while the warp is busy with its shared-memory atomics.
__device__ __forceinline__ void prefetch_l2_bulk(const void* p, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
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
  uint32_t NS, cap, wide_mask;
  uint32_t lo_base[R];  // shared-window byte address of column r's low words
  uint32_t hi_base[R];  // ... of its high words (wide columns only)
  uint32_t key_base;
};

template <int R, int SINK, int PODS>
__device__ __forceinline__ void scatter_pods(const ScatterCtx<R, SINK>& C, const uint32_t (&ns)[PODS],
                                             const long long (&q)[R][PODS]) {
  uint32_t slot[PODS];
  uint32_t old[R][PODS];
  if constexpr (SINK == SINK_DIRECT) {
    // The direct sink is HBM-bound as it is (and its few branches are rarely divergent: a zero GPU
    // count is the only common skip); the predicated form below costs it registers for nothing.
#pragma unroll
    for (int k = 0; k < PODS; ++k) slot[k] = ns[k] < C.NS ? ns[k] : kHashEmpty;
#pragma unroll
    for (int k = 0; k < PODS; ++k) {
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const uint32_t lo = (uint32_t)(unsigned long long)q[r][k];
        old[r][k] = 0;
        if (slot[k] != kHashEmpty && lo) old[r][k] = atoms_add_u32(C.lo_base[r] + slot[k] * 4u, lo);
      }
    }
#pragma unroll
    for (int k = 0; k < PODS; ++k) {
      if (ns[k] >= C.NS) continue;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const unsigned long long v = (unsigned long long)q[r][k];
        const uint32_t lo = (uint32_t)v;
        const uint32_t hi = (uint32_t)(v >> 32) + (uint32_t)(old[r][k] + lo < old[r][k]);
        if (hi) {
          if ((C.wide_mask >> r) & 1u) reds_add_u32(C.hi_base[r] + slot[k] * 4u, hi);
          else red_add_i64(used_at(C.used, C.NS, r, ns[k]), (long long)((unsigned long long)hi << 32));
        }
      }
    }
  } else {
  bool valid[PODS], hit[PODS];
#pragma unroll
  for (int k = 0; k < PODS; ++k) valid[k] = ns[k] < C.NS;  // kBadNs padding and out-of-range ids contribute nothing
  if (SINK == SINK_HASH) {
    // Two candidate slots per namespace (two multiply-shift reductions of one mixed word): with a single
    // slot, two HOT namespaces landing on the same slot send the loser's millions of pods to one
    // global address.  All key probes of the group are issued before any is examined.
    uint32_t s1[PODS], s2[PODS], k1[PODS], k2[PODS];
#pragma unroll
    for (int k = 0; k < PODS; ++k) {
      const uint32_t h = mix32(ns[k]);
      s1[k] = __umulhi(h, C.cap);                 // multiply-shift range reduction: any capacity,
      s2[k] = __umulhi(h * 0x9E3779B1u, C.cap);   // so the table uses all of shared memory
      k1[k] = lds_u32(C.key_base + s1[k] * 4u);
    }
    // The second candidate is probed only when the first slot is somebody else's: the kernel is bound
    // by shared-memory wavefronts (ncu: L1TEX the busiest unit at 70 %), and most pods of a skewed
    // stream sit in their first slot, so this saves most of one LDS per pod (measured +7..12 %).
#pragma unroll
    for (int k = 0; k < PODS; ++k) {
      k2[k] = 0;  // placeholder, never a key that matters: k1 already matched
      if (k1[k] != ns[k]) k2[k] = lds_u32(C.key_base + s2[k] * 4u);
    }
#pragma unroll
    for (int k = 0; k < PODS; ++k) {
      uint32_t got = k1[k] == ns[k] ? s1[k] : (k2[k] == ns[k] ? s2[k] : kHashEmpty);
      // claim-once: the first namespace to arrive owns a slot.  Rare once the table has filled up,
      // so this is the one real branch of the body.
      if (got == kHashEmpty && valid[k] && (k1[k] == kHashEmpty || k2[k] == kHashEmpty)) {
        if (k1[k] == kHashEmpty) {
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
      slot[k] = got;
      hit[k] = valid[k] && got != kHashEmpty;
    }
  } else {
#pragma unroll
    for (int k = 0; k < PODS; ++k) { slot[k] = ns[k]; hit[k] = valid[k]; }
  }
  // (Tried and measured slower, round 2: PTX-predicated atomics — ptxas turns them back into
  // BSSY/BRA/op/BSYNC — and unconditional atomics with slot-less pods adding into per-lane dummy
  // words: the kernel is bound by shared-memory atomic wavefronts (ncu: L1TEX 70 %, the busiest
  // unit), so every skipped zero addend counts for more than the branch that skips it.)
  // phase 1: every low-word atomic of the group is in flight before any result is consumed
#pragma unroll
  for (int k = 0; k < PODS; ++k) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const uint32_t lo = (uint32_t)(unsigned long long)q[r][k];
      old[r][k] = 0;
      if (hit[k] && lo) old[r][k] = atoms_add_u32(C.lo_base[r] + slot[k] * 4u, lo);
    }
  }
  // phase 2: high words and carries
#pragma unroll
  for (int k = 0; k < PODS; ++k) {
    if (!valid[k]) continue;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const unsigned long long v = (unsigned long long)q[r][k];
      if (!hit[k]) {  // cold namespace of the hash sink: straight to the global table
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
  }  // hash sink
}

// The hash sink keeps two candidate slots, two keys, the chosen slot and R old values per pod live on
// top of the loaded columns: with four pods at once a 64-register thread spills.  Two pods at a time
// fit; the loads were all issued before, so nothing is lost on the memory side.
template <int R, int SINK, int PODS>
__device__ __forceinline__ void scatter_group(const ScatterCtx<R, SINK>& C, const uint32_t (&ns)[PODS],
                                              const long long (&q)[R][PODS]) {
  if constexpr (SINK == SINK_HASH && PODS == 4) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      uint32_t ns2[2] = {ns[2 * h], ns[2 * h + 1]};
      long long q2[R][2];
#pragma unroll
      for (int r = 0; r < R; ++r) { q2[r][0] = q[r][2 * h]; q2[r][1] = q[r][2 * h + 1]; }
      scatter_pods<R, SINK, 2>(C, ns2, q2);
    }
  } else {
    scatter_pods<R, SINK, PODS>(C, ns, q);
  }
}

template <int R, int SINK, bool VEC, int THREADS, int U>
__global__ void __launch_bounds__(THREADS, 1)
k1_scatter(const uint32_t* __restrict__ ns_id, ReqCols<R> req, size_t n, long long* used, uint32_t NS,
           uint32_t cap, uint32_t wide_mask, uint32_t pf_dist) {
  extern __shared__ __align__(16) uint32_t k1_smem[];
  // layout: R low-word columns, then one high-word column per wide resource, then (hash) the keys.
  const uint32_t colw = cap;
  const uint32_t nwide = __popc(wide_mask & ((1u << R) - 1u));
  const uint32_t words = (R + nwide) * colw;
  for (uint32_t i = threadIdx.x; i < words; i += THREADS) k1_smem[i] = 0;
  if (SINK == SINK_HASH)
    for (uint32_t i = threadIdx.x; i < cap; i += THREADS) k1_smem[words + i] = kHashEmpty;
  ScatterCtx<R, SINK> C;
  C.used = used; C.NS = NS; C.cap = cap; C.wide_mask = wide_mask;
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(k1_smem);
#pragma unroll
  for (int r = 0; r < R; ++r) {
    C.lo_base[r] = sbase + (uint32_t)r * colw * 4u;
    C.hi_base[r] = sbase + (R + __popc(wide_mask & ((1u << r) - 1u))) * colw * 4u;
  }
  C.key_base = sbase + words * 4u;
  __syncthreads();

  constexpr int PODS = R <= 4 ? 4 : 2;
  const size_t groups = VEC ? n / PODS : 0;
  const size_t stride = (size_t)gridDim.x * THREADS;
  size_t g = (size_t)blockIdx.x * THREADS + threadIdx.x;
  // U groups per trip: all loads first (U*(1+R) independent vector requests per thread)
  for (; g + (U - 1) * stride < groups; g += U * stride) {
    if (pf_dist && (threadIdx.x & 31u) == 0) {  // the warp's tiles pf_dist trips ahead, whole tiles only
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const size_t gp = g + ((size_t)pf_dist * U + u) * stride;
        if (gp + 32 <= groups) {
          prefetch_l2_bulk(ns_id + gp * PODS, 32u * PODS * 4u);
#pragma unroll
          for (int r = 0; r < R; ++r) prefetch_l2_bulk(req.p[r] + gp * PODS, 32u * PODS * 8u);
        }
      }
    }
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
    for (int u = 0; u < U; ++u) scatter_group<R, SINK, PODS>(C, ns[u], q[u]);
  }
  for (; g < groups; g += stride) {
    uint32_t ns[PODS];
    long long q[R][PODS];
    const size_t base = g * PODS;
    ldg_ns<PODS>(ns_id + base, ns);
#pragma unroll
    for (int r = 0; r < R; ++r) ldg_req<PODS>(req.p[r] + base, q[r]);
    scatter_group<R, SINK, PODS>(C, ns, q);
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
    const uint32_t hi = ((wide_mask >> r) & 1u) ? k1_smem[(R + __popc(wide_mask & ((1u << r) - 1u))) * colw + sl] : 0u;
    const unsigned long long v = ((unsigned long long)hi << 32) | k1_smem[r * colw + sl];
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

}  // namespace bgc
