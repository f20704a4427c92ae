// bgc_k3.cuh — K3 admit_greedy (row N1).  Part of libbgc.so's device code; included through bgc_kernels.cuh.
#pragma once
#include "bgc_common.cuh"

namespace bgc {

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
// ------------------------------------------------------------------ stable grouping by namespace
// Hand-written LSD radix sort of (key = namespace id clamped to NS, value = pod index) pairs, 8 bits per
// pass, only ceil(log2(NS+1)) key bits (2 passes up to 65,535 namespaces, 3 up to 16.7 M).  Stability —
// pods of one namespace stay in arrival order, which the greedy recurrence depends on — comes from
// (i) fixed tiles of kSortTile consecutive elements per CTA, (ii) a digit-major histogram
// hist[digit][tile] whose exclusive scan gives every (digit, tile) its output range in tile order, and
// (iii) an in-order rank inside the tile: rounds of 256 consecutive elements, __match_any_sync for the
// rank among equal digits inside a warp, a per-digit prefix over the 8 warps, a running count per digit
// across rounds.  The first pass reads the namespace ids directly (keys and indices are implicit).
constexpr int kSortThreads = 256;
constexpr int kSortRounds = 32;
constexpr int kSortTile = kSortThreads * kSortRounds;  // 8192 elements per CTA tile

__device__ __forceinline__ uint32_t sort_key_at(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ ns_id,
                                               uint32_t NS, size_t i) {
  if (keys_in) return keys_in[i];
  const uint32_t ns = ns_id[i];
  return ns < NS ? ns : NS;  // unknown namespaces sort behind every real one
}

static __global__ void __launch_bounds__(kSortThreads)
k_rsort_hist(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ ns_id, uint32_t NS, size_t n, int shift,
             uint32_t* __restrict__ hist, uint32_t ntile) {
  __shared__ uint32_t s_cnt[256];
  for (uint32_t t = blockIdx.x; t < ntile; t += gridDim.x) {
    s_cnt[threadIdx.x] = 0;
    __syncthreads();
    const size_t base = (size_t)t * kSortTile;
#pragma unroll 4
    for (int r = 0; r < kSortRounds; ++r) {
      const size_t i = base + (size_t)r * kSortThreads + threadIdx.x;
      if (i < n) atomicAdd(&s_cnt[(sort_key_at(keys_in, ns_id, NS, i) >> shift) & 0xFFu], 1u);
    }
    __syncthreads();
    hist[(size_t)threadIdx.x * ntile + t] = s_cnt[threadIdx.x];
    __syncthreads();
  }
}

// exclusive scan of a u32 array in place, one CTA (the digit-major histogram: a few MB, L2-resident)
static __global__ void __launch_bounds__(1024) k_scan_u32(uint32_t* a, size_t count) {
  __shared__ uint32_t s_part[1024];
  const size_t per = (count + 1023) / 1024;
  const size_t lo = (size_t)threadIdx.x * per, hi = lo + per < count ? lo + per : count;
  uint32_t sum = 0;
  for (size_t k = lo; k < hi; ++k) sum += a[k];
  s_part[threadIdx.x] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t run = 0;
    for (int i = 0; i < 1024; ++i) { const uint32_t v = s_part[i]; s_part[i] = run; run += v; }
  }
  __syncthreads();
  uint32_t run = s_part[threadIdx.x];
  for (size_t k = lo; k < hi; ++k) { const uint32_t v = a[k]; a[k] = run; run += v; }
}

// The histogram scan in three small launches (a single CTA took 3 ms per pass on the 3 M entries of a 10^8-pair
// sort): every CTA scans 8192 consecutive entries in place and publishes its total, one CTA scans the
// totals (k_scan_u32), every CTA adds its base.
constexpr int kScanPerThread = 8;
constexpr int kScanBlock = 1024 * kScanPerThread;

static __global__ void __launch_bounds__(1024) k_scan_blocks(uint32_t* a, size_t count, uint32_t* block_total) {
  __shared__ uint32_t s_w[32];
  const size_t base = (size_t)blockIdx.x * kScanBlock + (size_t)threadIdx.x * kScanPerThread;
  uint32_t v[kScanPerThread], sum = 0;
#pragma unroll
  for (int k = 0; k < kScanPerThread; ++k) { v[k] = base + k < count ? a[base + k] : 0u; sum += v[k]; }
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  uint32_t incl = sum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
    if (lane >= (uint32_t)d) incl += o;
  }
  if (lane == 31) s_w[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = s_w[lane], wi = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, wi, d);
      if (lane >= (uint32_t)d) wi += o;
    }
    s_w[lane] = wi - w;
    if (lane == 31) block_total[blockIdx.x] = wi;
  }
  __syncthreads();
  uint32_t run = s_w[warp] + incl - sum;
#pragma unroll
  for (int k = 0; k < kScanPerThread; ++k) {
    if (base + k < count) a[base + k] = run;
    run += v[k];
  }
}

static __global__ void __launch_bounds__(1024) k_scan_add(uint32_t* a, size_t count, const uint32_t* __restrict__ block_base) {
  const uint32_t add = block_base[blockIdx.x];
  const size_t base = (size_t)blockIdx.x * kScanBlock + (size_t)threadIdx.x * kScanPerThread;
#pragma unroll
  for (int k = 0; k < kScanPerThread; ++k)
    if (base + k < count) a[base + k] += add;
}

// The request columns (and masks) travel with the pairs through every pass, so that the last pass leaves
// them in grouped order for the walkers: a separate gather by the final permutation reads one 32-byte
// sector per 8-byte value from all over a multi-GB array (6.7 ms per 10^8 pods, measured), whereas here
// every read stays inside the tile's own 64 KB window of a column (L2 hits) and every write is a run.
struct SortPayload {
  const long long* in[kMaxR];
  long long* out[kMaxR];
  const uint8_t* mask_in;
  uint8_t* mask_out;
  int R;
};

// (Tried and measured slower, round 2: warp-owned contiguous chunks ranked with two passes of __match_any_sync and
// only four block barriers per tile — 12.5 ms per 10^8-pod greedy call against 9.8 ms for the rounds below: the
// second match per element costs more than the 124 barriers it saves.)
// (Also tried, round 2: keeping only (digit, place in the tile) per element while the tile is ranked — 32 KB of shared
// memory and 64 registers instead of 80 KB and 123, four CTAs per SM instead of two — and fetching key and index
// again at write-out like the payload columns: 9.8 ms per 10^8-pod greedy call against 8.2 ms, the scattered 4-byte
// re-reads cost a whole L2 sector each.  And an L2 bulk prefetch, at the start of a tile, of its payload columns and of
// the next tile's keys and indices: 8.4 ms against 8.1 ms.)
// One tile per trip: the tile is sorted by digit INSIDE shared memory first (in-order ranks as described
// above), then written out — every digit's elements of the tile are consecutive in shared memory and go to
// consecutive global positions, so the writes are whole 32-byte sectors instead of one 4-byte element per
// digit per round (the first version scattered straight to global memory: 9 ms per pass on 10^8 pairs,
// all of it partial-sector read-modify-write traffic).
static __global__ void __launch_bounds__(kSortThreads)
k_rsort_scatter(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ idx_in, const uint32_t* __restrict__ ns_id,
                uint32_t NS, size_t n, int shift, const uint32_t* __restrict__ offs, uint32_t ntile,
                uint32_t* __restrict__ keys_out, uint32_t* __restrict__ idx_out, SortPayload P) {
  extern __shared__ __align__(16) uint32_t rs_smem[];
  uint32_t* s_key = rs_smem;                    // [kSortTile] the tile, grouped by digit
  uint32_t* s_idx = rs_smem + kSortTile;        // [kSortTile]
  unsigned short* s_src = reinterpret_cast<unsigned short*>(rs_smem + 2 * kSortTile);  // [kSortTile] where in the tile it came from
  __shared__ uint32_t s_warp[kSortThreads / 32][256];  // per-warp digit counts of the current round, then bases
  __shared__ uint32_t s_run[256];                      // cursor of each digit inside the tile
  __shared__ uint32_t s_lbase[256];                    // first tile position of each digit
  __shared__ uint32_t s_goff[256];                     // first global position of each digit's run of this tile
  __shared__ uint32_t s_scan[kSortThreads / 32];
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  for (uint32_t t = blockIdx.x; t < ntile; t += gridDim.x) {
    const size_t base = (size_t)t * kSortTile;
    const uint32_t live_n = (uint32_t)(n - base < (size_t)kSortTile ? n - base : (size_t)kSortTile);
    {  // this tile's digit counts = differences of the scanned digit-major histogram; exclusive scan over the 256 digits
      const size_t at = (size_t)threadIdx.x * ntile + t;
      const uint32_t off = offs[at];
      // count of (digit d, tile t) = next entry in scan order - this one; the very last entry ends at n
      const uint32_t nxt = (threadIdx.x == 255 && t + 1 == ntile) ? (uint32_t)n : offs[at + 1];
      const uint32_t cnt = nxt - off;
      uint32_t incl = cnt;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= (uint32_t)d) incl += o;
      }
      if (lane == 31) s_scan[warp] = incl;
      __syncthreads();
      uint32_t before = incl - cnt;
      for (uint32_t w = 0; w < warp; ++w) before += s_scan[w];
      s_lbase[threadIdx.x] = before;
      s_run[threadIdx.x] = before;
      s_goff[threadIdx.x] = off;
    }
    // every load of the tile is issued before the first round (ncu: with one load per thread per round the kernel
    // sat in long-scoreboard stalls at 13 % issue utilisation, a tile taking ~100 us of exposed latency)
    uint32_t keys[kSortRounds], idxs[kSortRounds];
#pragma unroll
    for (int r = 0; r < kSortRounds; ++r) {
      const size_t i = base + (size_t)r * kSortThreads + threadIdx.x;
      keys[r] = i < n ? sort_key_at(keys_in, ns_id, NS, i) : 0u;
      idxs[r] = i < n ? (idx_in ? idx_in[i] : (uint32_t)i) : 0u;
    }
#pragma unroll
    for (int r = 0; r < kSortRounds; ++r) {
#pragma unroll
      for (int w = 0; w < kSortThreads / 32; ++w) s_warp[w][threadIdx.x] = 0;
      __syncthreads();
      const size_t i = base + (size_t)r * kSortThreads + threadIdx.x;
      const bool live = i < n;
      const uint32_t key = keys[r];
      uint32_t digit = 0x100u + lane;  // dead lanes get digits nobody else has
      if (live) digit = (key >> shift) & 0xFFu;
      const uint32_t peers = __match_any_sync(0xFFFFFFFFu, digit);
      const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
      if (live && rank == 0) s_warp[warp][digit] = __popc(peers);  // the first lane of each group publishes its size
      __syncthreads();
      {  // thread d: turn the 8 per-warp counts of digit d into exclusive bases, advance the digit's cursor
        uint32_t run = s_run[threadIdx.x];
#pragma unroll
        for (int w = 0; w < kSortThreads / 32; ++w) {
          const uint32_t c = s_warp[w][threadIdx.x];
          s_warp[w][threadIdx.x] = run;
          run += c;
        }
        s_run[threadIdx.x] = run;
      }
      __syncthreads();
      if (live) {
        const uint32_t pos = s_warp[warp][digit] + rank;
        s_key[pos] = key;
        s_idx[pos] = idxs[r];
        s_src[pos] = (unsigned short)(r * kSortThreads + threadIdx.x);
      }
      __syncthreads();
    }
    // write-out, kWB elements per thread at a time: all of a batch's payload loads are issued before its stores (the
    // compiler cannot hoist them itself, input and output columns may alias as far as it knows)
    constexpr int kWB = 8;
    for (uint32_t j0 = threadIdx.x; j0 < live_n; j0 += kSortThreads * kWB) {
      uint32_t gpos[kWB];
      size_t src[kWB];
      bool on[kWB];
#pragma unroll
      for (int u = 0; u < kWB; ++u) {
        const uint32_t j = j0 + (uint32_t)u * kSortThreads;
        on[u] = j < live_n;
        const uint32_t key = on[u] ? s_key[j] : 0u;
        const uint32_t d = (key >> shift) & 0xFFu;
        gpos[u] = on[u] ? s_goff[d] + (j - s_lbase[d]) : 0u;
        src[u] = base + (on[u] ? s_src[j] : 0u);
        if (on[u]) {
          keys_out[gpos[u]] = key;
          idx_out[gpos[u]] = s_idx[j];
        }
      }
      for (int q = 0; q < P.R; ++q) {
        long long v[kWB];
#pragma unroll
        for (int u = 0; u < kWB; ++u) v[u] = on[u] ? P.in[q][src[u]] : 0;
#pragma unroll
        for (int u = 0; u < kWB; ++u)
          if (on[u]) P.out[q][gpos[u]] = v[u];
      }
      if (P.mask_in) {
        uint8_t m[kWB];
#pragma unroll
        for (int u = 0; u < kWB; ++u) m[u] = on[u] ? P.mask_in[src[u]] : (uint8_t)0;
#pragma unroll
        for (int u = 0; u < kWB; ++u)
          if (on[u]) P.mask_out[gpos[u]] = m[u];
      }
    }
    __syncthreads();
  }
}

// The shipped form of the pass.  Two things changed against k_rsort_scatter above (kept behind BGC_K3_SORT=1 for
// comparison), each measured on the 10^8-pod greedy call:
//  * WARP-PRIVATE ranking: every warp owns a contiguous chunk of the tile and ranks it round by round against its own
//    row of digit counters, with two __syncwarp per round and no block barrier; one block-wide step then turns the
//    per-warp counts into bases.  3 barriers per tile instead of 128, and only (digit, rank) live in registers while
//    ranking — 24 registers instead of 64 — so that 512 threads fit twice on an SM (32 warps instead of 16).  Keys and
//    indices are read a second time when the tile is placed in shared memory: coalesced, and L2 hits.
//  * the payload columns go THROUGH shared memory at write-out: the tile's slice of a column is read coalesced into
//    the space keys and indices have left, and gathered from there.  Gathering it from global memory by source index
//    cost a 32-byte L2 sector per 8-byte value (L1 cannot hold a 64 KB window per column next to the shared memory).
// Alone, the first made the call slower (9.7 against 8.1 ms: twice the threads thrashing L1 with those gathers) and
// the second took 8.1 to 7.4 ms; together 5.9 ms (scatter passes 3.46 + 2.50 -> 2.03 + 1.79 ms).  Staging keys and
// indices the same way as the payload changed nothing (5.97 ms).
constexpr int kSort2Threads = 512;
constexpr int kSort2Warps = kSort2Threads / 32;
constexpr int kSort2Chunk = kSortTile / kSort2Warps;  // 512 consecutive elements per warp
constexpr int kSort2Rounds = kSort2Chunk / 32;        // 16
static __global__ void __launch_bounds__(kSort2Threads, 2)
k_rsort_scatter_wp(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ idx_in, const uint32_t* __restrict__ ns_id,
                 uint32_t NS, size_t n, int shift, const uint32_t* __restrict__ offs, uint32_t ntile,
                 uint32_t* __restrict__ keys_out, uint32_t* __restrict__ idx_out, SortPayload P) {
  extern __shared__ __align__(16) uint32_t rs_smem[];
  uint32_t* s_key = rs_smem;
  uint32_t* s_idx = rs_smem + kSortTile;
  unsigned short* s_src = reinterpret_cast<unsigned short*>(rs_smem + 2 * kSortTile);
  long long* stage = reinterpret_cast<long long*>(rs_smem);  // one payload column of the tile, over s_key + s_idx once they are out
  __shared__ uint32_t s_warp[kSort2Warps][256];  // per-warp digit counts, then the warp's base for each digit
  __shared__ uint32_t s_lbase[256];
  __shared__ uint32_t s_goff[256];
  __shared__ uint32_t s_scan[8];
  const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  for (uint32_t t = blockIdx.x; t < ntile; t += gridDim.x) {
    const size_t base = (size_t)t * kSortTile;
    const uint32_t live_n = (uint32_t)(n - base < (size_t)kSortTile ? n - base : (size_t)kSortTile);
    uint32_t cnt = 0, off = 0, incl = 0;
    if (tid < 256) {  // this tile's digit counts = differences of the scanned digit-major histogram
      const size_t at = (size_t)tid * ntile + t;
      off = offs[at];
      const uint32_t nxt = (tid == 255 && t + 1 == ntile) ? (uint32_t)n : offs[at + 1];
      cnt = nxt - off;
      incl = cnt;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= (uint32_t)d) incl += o;
      }
      if (lane == 31) s_scan[warp] = incl;
    }
#pragma unroll
    for (int d = 0; d < 256 / 32; ++d) s_warp[warp][d * 32 + lane] = 0;
    __syncthreads();
    if (tid < 256) {
      uint32_t before = incl - cnt;
      for (uint32_t w = 0; w < warp; ++w) before += s_scan[w];
      s_lbase[tid] = before;
      s_goff[tid] = off;
    }
    // ---- phase A: the warp ranks its own chunk (element e = warp * 512 + k * 32 + lane: memory order)
    const uint32_t e0 = warp * kSort2Chunk + lane;
    uint32_t dig[kSort2Rounds / 4], rk[kSort2Rounds / 2];
#pragma unroll
    for (int k4 = 0; k4 < kSort2Rounds / 4; ++k4) {
      uint32_t kk[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t e = e0 + (uint32_t)(k4 * 4 + u) * 32u;
        kk[u] = e < live_n ? sort_key_at(keys_in, ns_id, NS, base + e) : 0u;
      }
      dig[k4] = ((kk[0] >> shift) & 0xFFu) | (((kk[1] >> shift) & 0xFFu) << 8) | (((kk[2] >> shift) & 0xFFu) << 16) | (((kk[3] >> shift) & 0xFFu) << 24);
    }
#pragma unroll
    for (int k = 0; k < kSort2Rounds; ++k) {
      const bool live = e0 + (uint32_t)k * 32u < live_n;
      const uint32_t d = (dig[k >> 2] >> (8 * (k & 3))) & 0xFFu;
      const uint32_t digit = live ? d : 0x100u + lane;  // dead lanes get digits nobody else has
      const uint32_t peers = __match_any_sync(0xFFFFFFFFu, digit);
      const uint32_t ahead = __popc(peers & ((1u << lane) - 1u));
      const uint32_t seen = live ? s_warp[warp][d] : 0u;  // same-digit elements of the earlier rounds of this chunk
      __syncwarp();
      if (live && ahead == 0) s_warp[warp][d] = seen + __popc(peers);
      __syncwarp();
      const uint32_t r = seen + ahead;  // < 512
      if (k & 1) rk[k >> 1] |= r << 16; else rk[k >> 1] = r;
    }
    __syncthreads();
    // ---- phase B: digit d's counts of the 16 warps become the warps' first positions inside the tile
    if (tid < 256) {
      uint32_t run = s_lbase[tid];
#pragma unroll
      for (int w = 0; w < kSort2Warps; ++w) {
        const uint32_t c = s_warp[w][tid];
        s_warp[w][tid] = run;
        run += c;
      }
    }
    __syncthreads();
    // ---- phase C: place the chunk (keys and indices read again, coalesced)
#pragma unroll
    for (int k8 = 0; k8 < kSort2Rounds / 8; ++k8) {
      uint32_t kk[8], ii[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const uint32_t e = e0 + (uint32_t)(k8 * 8 + u) * 32u;
        const bool live = e < live_n;
        kk[u] = live ? sort_key_at(keys_in, ns_id, NS, base + e) : 0u;
        ii[u] = (live && idx_in) ? idx_in[base + e] : (uint32_t)(base + e);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int k = k8 * 8 + u;
        const uint32_t e = e0 + (uint32_t)k * 32u;
        if (e < live_n) {
          const uint32_t d = (dig[k >> 2] >> (8 * (k & 3))) & 0xFFu;
          const uint32_t pos = s_warp[warp][d] + ((rk[k >> 1] >> (16 * (k & 1))) & 0xFFFFu);
          s_key[pos] = kk[u];
          s_idx[pos] = ii[u];
          s_src[pos] = (unsigned short)e;
        }
      }
    }
    __syncthreads();
    // write-out.  Keys and indices first, straight from shared memory; then the payload, one column at a time THROUGH
    // shared memory: the tile's slice of the column is read coalesced into the space keys and indices have left, and
    // gathered from there.  (Gathering it from global memory by source index costs a 32-byte L2 sector per 8-byte value,
    // and L1 cannot hold a 64 KB window per column next to 160 KB of shared memory.)  A thread places the same 16
    // positions for every column: where they go and where they came from stay in registers.
    constexpr int kWB = 8;
    constexpr int kPer = kSortTile / kSort2Threads;  // 16
    uint32_t gp[kPer], sr[kPer / 2];
#pragma unroll
    for (int m = 0; m < kPer; ++m) {
      const uint32_t j = tid + (uint32_t)m * kSort2Threads;
      uint32_t from = 0;
      gp[m] = 0xFFFFFFFFu;
      if (j < live_n) {
        const uint32_t key = s_key[j];
        const uint32_t d = (key >> shift) & 0xFFu;
        gp[m] = s_goff[d] + (j - s_lbase[d]);
        from = s_src[j];
        keys_out[gp[m]] = key;
        idx_out[gp[m]] = s_idx[j];
      }
      if (m & 1) sr[m >> 1] |= from << 16; else sr[m >> 1] = from;
    }
    __syncthreads();
    for (int q = 0; q < P.R; ++q) {
      const long long* __restrict__ col = P.in[q] + base;
      for (uint32_t i0 = tid; i0 < live_n; i0 += kSort2Threads * kWB) {
        long long v[kWB];
#pragma unroll
        for (int u = 0; u < kWB; ++u) {
          const uint32_t i = i0 + (uint32_t)u * kSort2Threads;
          v[u] = i < live_n ? col[i] : 0;
        }
#pragma unroll
        for (int u = 0; u < kWB; ++u) {
          const uint32_t i = i0 + (uint32_t)u * kSort2Threads;
          if (i < live_n) stage[i] = v[u];
        }
      }
      __syncthreads();
      long long* __restrict__ dst = P.out[q];
#pragma unroll
      for (int m = 0; m < kPer; ++m)
        if (gp[m] != 0xFFFFFFFFu) dst[gp[m]] = stage[(sr[m >> 1] >> (16 * (m & 1))) & 0xFFFFu];
      __syncthreads();
    }
    if (P.mask_in) {
      uint8_t* stage8 = reinterpret_cast<uint8_t*>(stage);
      for (uint32_t i = tid; i < live_n; i += kSort2Threads) stage8[i] = P.mask_in[base + i];
      __syncthreads();
#pragma unroll
      for (int m = 0; m < kPer; ++m)
        if (gp[m] != 0xFFFFFFFFu) P.mask_out[gp[m]] = stage8[(sr[m >> 1] >> (16 * (m & 1))) & 0xFFFFu];
    }
    __syncthreads();
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

// VT = uint8_t: verdict = missing << 4 | exceeded (R <= 4); VT = uint16_t: missing << 8 | exceeded (R <= 8)
template <typename VT>
static __global__ void k_greedy_bad_ns(const uint32_t* __restrict__ order, const uint32_t* __restrict__ seg,
                                       uint32_t NS, VT* verdict) {
  const size_t beg = seg[NS], end = seg[NS + 1];
  for (size_t i = beg + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < end; i += (size_t)gridDim.x * blockDim.x)
    verdict[order[i]] = (VT)~(VT)0;
}

template <int R>
struct GreedyPod {
  long long q[R];
  uint32_t mask;
};

// verdict of one pod against usage T
template <int R, int SHIFT>
__device__ __forceinline__ uint32_t greedy_verdict(const GreedyPod<R>& p, const long long (&T)[R],
                                                   const long long (&H)[R], uint32_t hm, uint32_t compute_mask) {
  uint32_t missing = 0, exceeded = 0;
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const bool constrained = (hm >> r) & 1u, present = (p.mask >> r) & 1u;
    if (constrained && !present && ((compute_mask >> r) & 1u)) missing |= 1u << r;
    if (constrained && present && p.q[r] > 0 && p.q[r] > sat_sub_i64(H[r], T[r])) exceeded |= 1u << r;
  }
  return (missing << SHIFT) | exceeded;
}

template <int R, bool HAS_MASK, typename VT>
__global__ void __launch_bounds__(256)
k3_admit_greedy(const uint32_t* __restrict__ order, const uint32_t* __restrict__ seg, ReqCols<R> req,
                const uint8_t* __restrict__ req_mask, long long* used, const long long* __restrict__ hard,
                const uint8_t* __restrict__ hard_mask, uint32_t NS, uint32_t compute_mask,
                VT* __restrict__ verdict, uint32_t team_min) {
  constexpr int SHIFT = sizeof(VT) == 1 ? 4 : 8;
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
    uint32_t ord_next = beg + lane < end ? order[beg + lane] : 0u;  // verdict destination, fetched one tile ahead
    for (uint32_t t = beg; t < end; t += 32) {
      const uint32_t ord = ord_next;
      ord_next = t + 32 + lane < end ? order[t + 32 + lane] : 0u;
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
          const uint32_t v_alone = greedy_verdict<R, SHIFT>(p, S, H, hm, compute_mask);
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
          const uint32_t fail = __ballot_sync(0xFFFFFFFFu, mine && greedy_verdict<R, SHIFT>(p, T, H, hm, compute_mask) != 0);
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
        // reason bits of the rejected pods are taken against the usage they actually met: a second scan, over the
        // admitted pods only — not needed when the whole tile got in, nor when nobody did (usage never moved)
        const bool adm = (admitted >> lane) & 1u;
        const uint32_t everyone = __ballot_sync(0xFFFFFFFFu, valid);
        if (admitted == everyone) {
          my_verdict = 0u;
        } else if (admitted == 0u) {
          my_verdict = valid ? greedy_verdict<R, SHIFT>(p, S0, H, hm, compute_mask) : 0u;
        } else {
          long long incl[R], T[R];
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
          my_verdict = adm ? 0u : greedy_verdict<R, SHIFT>(p, T, H, hm, compute_mask);
        }
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
        const uint32_t v_run = greedy_verdict<R, SHIFT>(p, T, H, hm, compute_mask);
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
        const uint32_t v_alone = greedy_verdict<R, SHIFT>(p, S, H, hm, compute_mask);
        const uint32_t later = 0xFFFFFFFFu << pos;
        const uint32_t fits = __ballot_sync(0xFFFFFFFFu, valid && v_alone == 0) & later;
        const uint32_t stop = fits ? (__ffs(fits) - 1u) : 32u;
        if (lane >= pos && lane < stop && valid) my_verdict = v_alone;
        pos = stop;
        if (pos >= 32) break;
      }
      if (valid) verdict[ord] = (VT)my_verdict;
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
constexpr uint32_t kGreedyTeamMin = 32768;  // measured on 10^8 pods: 16,384 the same, 65,536 +0.5 ms under Zipf, 8,192 doubles the uniform case
                                            // (10^4-pod segments are walked faster by one warp each than as chained tiles)
constexpr int kGreedyTeam = 1024;
// tiles of requests in flight per CTA (ring of R * 8 KB per tile; the ring and the R * 8 KB scan scratch share
// the 227 KB of dynamic shared memory)
__host__ __device__ constexpr int greedy_depth(int R) { return R <= 4 ? 4 : R <= 6 ? 3 : 2; }
__host__ __device__ constexpr size_t greedy_team_smem(int R) { return (size_t)(greedy_depth(R) + 1) * R * kGreedyTeam * 8; }

template <int R, bool HAS_MASK, typename VT>
__global__ void __launch_bounds__(kGreedyTeam, 1)
k3_admit_greedy_team(const uint32_t* __restrict__ order, const uint32_t* __restrict__ seg, ReqCols<R> req,
                     const uint8_t* __restrict__ req_mask, long long* used, const long long* __restrict__ hard,
                     const uint8_t* __restrict__ hard_mask, uint32_t NS, uint32_t compute_mask,
                     VT* __restrict__ verdict) {
  constexpr int SHIFT = sizeof(VT) == 1 ? 4 : 8;
  constexpr int kGreedyDepth = greedy_depth(R);
  extern __shared__ __align__(16) long long g_ring[];  // [kGreedyDepth][R][kGreedyTeam], then s_E[R][kGreedyTeam]
  __shared__ long long s_wtot[R][32];    // per-warp totals, then their exclusive prefix
  __shared__ long long s_total[R];
  long long (*s_E)[kGreedyTeam] = reinterpret_cast<long long (*)[kGreedyTeam]>(g_ring + (size_t)kGreedyDepth * R * kGreedyTeam);
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
    // (the verdict's destination, order[t + tid], is loaded where it is stored: fetching it a tile ahead — it is the
    // top stall in ncu — costs two more spilled registers at the 64-register cap and measured 8 % slower)
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
          if (live && greedy_verdict<R, SHIFT>(p, S, H, hm, compute_mask) != 0) live = false;  // can never fit again
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
          const uint32_t fail = __ballot_sync(0xFFFFFFFFu, live && greedy_verdict<R, SHIFT>(p, T, H, hm, compute_mask) != 0);
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
          my_verdict = adm ? 0u : greedy_verdict<R, SHIFT>(p, T, H, hm, compute_mask);
        } else {                      // nothing admitted in this tile: usage never moved
          my_verdict = valid ? greedy_verdict<R, SHIFT>(p, S0, H, hm, compute_mask) : 0u;
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
        const uint32_t v_run = greedy_verdict<R, SHIFT>(p, T, H, hm, compute_mask);
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
        const uint32_t v_alone = greedy_verdict<R, SHIFT>(p, S, H, hm, compute_mask);
        const uint32_t fits = __ballot_sync(0xFFFFFFFFu, valid && tid >= pos && v_alone == 0);
        if (fits && lane == (uint32_t)(__ffs(fits) - 1)) atomicMin(&s_fit[par], tid);
        __syncthreads();
        const uint32_t stop = s_fit[par] == BIG ? (uint32_t)kGreedyTeam : s_fit[par];
        if (tid >= pos && tid < stop && valid) my_verdict = v_alone;
        pos = stop;
        if (pos >= (uint32_t)kGreedyTeam) break;
      }
      if (valid) verdict[order[t + tid]] = (VT)my_verdict;
    }
    if (tid == 0) {
#pragma unroll
      for (int r = 0; r < R; ++r) used[(size_t)r * NS + ns] = S[r];
    }
  }
}

// ------------------------------------------------------------------ long segments on MANY CTAs: chained tiles
// k3_admit_greedy_team gives a long segment to ONE CTA, so a Zipf head of 10^7 pods is 10^4 tiles walked one after
// the other (4 us each: 40-70 ms, the whole greedy call waits for it).  Here the long segments are cut into tiles
// of 1024 pods that ANY CTA may take (a ticket counter hands them out in order, so a tile only ever waits for
// tiles held by CTAs that are already running), and the only thing a tile needs from its predecessors — the usage
// S_in it starts from — travels through per-tile descriptors, decoupled look-back style:
//   AGG(X)   "if S_in + X fits the quota, all my eligible pods are admitted and S_out = S_in + X" — published as
//            soon as the tile's own requests are summed, before S_in is known;
//   IDENT    "S_out = S_in" — published as soon as ANY exact usage of an earlier tile shows that no pod with a
//            non-zero request fits on its own: requests are >= 0 in these tiles, so usage only grows and a pod that
//            does not fit against a lower bound of S_in does not fit against S_in;
//   AGG2(X') the same lower bound strikes out the pods that can never fit; X' sums the ones left: "if S_in + X' fits,
//            exactly those are admitted and S_out = S_in + X'" (one resource full, the others still filling);
//   EXACT(S) the tile's S_out, final;
//   OPAQUE   a tile with a request outside [0, 2^40]: nothing can be said before S_in is known.
// A tile resolves S_in by walking back to the nearest EXACT and adding the AGGs in between (valid when the sum
// still fits: every intermediate prefix is smaller), IDENTs adding nothing.  While everything fits, and again once
// the namespace is saturated, no tile waits for its predecessor's walk: the chain is only sequential where pods
// are still slipping into the last of the headroom.  Verdict bits need the exact S_in, so an IDENT tile still
// waits for it, but nobody waits for that tile.
constexpr uint32_t kChainPending = 0, kChainAgg = 1, kChainOpaque = 2, kChainIdent = 3, kChainExact = 4, kChainAgg2 = 5;
constexpr uint32_t kChainMaxBack = 2048;  // tiles: 2^62 + 2048 * 1024 * 2^40 < 2^63, so a walked-over stretch cannot wrap

struct ChainBufs {
  uint32_t* tile_base;    // [NS + 1] exclusive prefix of the tile counts (0 tiles for a short segment); [NS] = all tiles
  uint32_t* long_ns;      // [long segments] their namespaces, in no particular order
  uint32_t* long_nt;      // [long segments] their tile counts
  uint32_t* tile_li;      // [tiles] index into long_ns of the tile's segment
  uint4* ticket_rec;      // [tiles] what a ticket stands for — tile 0 of every long segment, then tile 1 of every ... —
                          // as one 16-byte record: tile, namespace, first tile of the namespace, first pod of the tile
  uint32_t* status;       // [tiles] kChain*
  long long* agg;         // [tiles][R]
  long long* agg2;        // [tiles][R]
  long long* exact;       // [tiles][R]
  uint32_t* ticket;       // [0] next ticket, [1..5] statistics, [6] number of long segments
};

__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// tiles per namespace: ceil(len / 1024) for a segment the warp kernel leaves alone, else 0; the long ones are listed
static __global__ void k_chain_count(const uint32_t* __restrict__ seg, uint32_t NS, uint32_t team_min, ChainBufs C) {
  for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k <= (size_t)NS; k += (size_t)gridDim.x * blockDim.x) {
    const uint32_t len = k < NS ? seg[k + 1] - seg[k] : 0u;
    const uint32_t nt = len >= team_min ? (len + 1023u) / 1024u : 0u;
    C.tile_base[k] = nt;
    if (nt) {
      const uint32_t li = atomicAdd(C.ticket + 6, 1u);
      C.long_ns[li] = (uint32_t)k;
      C.long_nt[li] = nt;
    }
  }
}

// tile_li[]: one warp per long segment
static __global__ void k_chain_fill(ChainBufs C) {
  const uint32_t lane = threadIdx.x & 31u, L = C.ticket[6];
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  for (uint32_t li = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; li < L; li += warps) {
    const uint32_t ns = C.long_ns[li];
    const uint32_t b = C.tile_base[ns], e = C.tile_base[ns + 1];
    for (uint32_t t = b + lane; t < e; t += 32) C.tile_li[t] = li;
  }
}

// Ticket order: tile t of EVERY long segment before tile t + 1 of any, so that the CTAs in flight work on as many
// segments as there are (a tile only ever waits for lower tiles of its own segment: the fewer of those are in
// flight, the fresher the usage it looks back to, and the sequential stretches of different segments overlap).
// Position of tile (segment li, t) = sum over segments of min(tiles, t) + the number of segments before li that
// have a tile t.
static __global__ void k_chain_order(const uint32_t* __restrict__ seg, uint32_t NS, ChainBufs C) {
  const uint32_t L = C.ticket[6], tiles = C.tile_base[NS];
  for (uint32_t d = blockIdx.x * blockDim.x + threadIdx.x; d < tiles; d += gridDim.x * blockDim.x) {
    const uint32_t li = C.tile_li[d];
    const uint32_t ns = C.long_ns[li], first = C.tile_base[ns];
    const uint32_t t = d - first;
    uint32_t pos = 0;
    for (uint32_t j = 0; j < L; ++j) {
      const uint32_t nt = __ldg(C.long_nt + j);
      pos += (nt < t ? nt : t) + ((j < li && nt > t) ? 1u : 0u);
    }
    C.ticket_rec[pos] = make_uint4(d, ns, first, seg[ns] + t * (uint32_t)kGreedyTeam);
  }
}

// shared memory of one 1024-pod tile step (the admit-run / reject-run alternation of the team kernel)
template <int R>
struct TileSmem {
  long long (*E)[kGreedyTeam];  // [R][1024] exclusive prefixes (dynamic shared memory)
  long long (*wtot)[32];        // [R][32]
  long long* total;             // [R]
  uint32_t* first;              // [2]
  uint32_t* fit;                // [2]
};

// block-wide exclusive prefix sum of x[] (0 for threads that do not take part); leaves the per-thread prefixes in
// M.E and the block total in M.total.  Two barriers.
template <int R>
__device__ __forceinline__ void tile_exclusive(const TileSmem<R>& M, const long long (&x)[R], long long (&E)[R]) {
  const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
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
    for (int r = 0; r < R; ++r) M.wtot[r][warp] = incl[r];
  }
  __syncthreads();
  if (warp == 0) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const long long mine = M.wtot[r][lane];
      long long y = mine;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const long long o = __shfl_up_sync(0xFFFFFFFFu, y, d);
        if (lane >= (uint32_t)d) y = (long long)((unsigned long long)y + (unsigned long long)o);
      }
      M.wtot[r][lane] = (long long)((unsigned long long)y - (unsigned long long)mine);
      if (lane == 31) M.total[r] = y;
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < R; ++r) {
    E[r] = (long long)((unsigned long long)M.wtot[r][warp] + (unsigned long long)incl[r] - (unsigned long long)x[r]);
    M.E[r][tid] = E[r];
  }
}

// One tile of 1024 consecutive pods of one namespace, usage S on entry: returns this thread's verdict and leaves the
// usage after the tile in S (the same in every thread).  Ends on a barrier: the shared state may be reused at once.
// `done(S)` is called by every thread as soon as the usage after the tile is final (before the reason bits of the
// rejected pods are worked out).
template <int R, int SHIFT, typename Done>
__device__ __forceinline__ uint32_t greedy_tile(const TileSmem<R>& M, const GreedyPod<R>& p, bool valid, long long (&S)[R],
                                               const long long (&H)[R], uint32_t hm, uint32_t compute_mask, uint32_t& iter,
                                               Done done) {
  const uint32_t tid = threadIdx.x, lane = tid & 31u;
  constexpr uint32_t BIG = 0xFFFFFFFFu;
  uint32_t my_verdict = 0, pos = 0;
  bool tame = true;  // monotone fast path: requests in [0, 2^40], usage in [0, 2^62] (see the warp kernel)
#pragma unroll
  for (int r = 0; r < R; ++r) tame = tame && p.q[r] >= 0 && p.q[r] <= (1ll << 40) && S[r] >= 0 && S[r] <= (1ll << 62);
  if (__syncthreads_and(tame)) {
    long long S0[R];
#pragma unroll
    for (int r = 0; r < R; ++r) S0[r] = S[r];
    bool live = valid, adm = false;
    while (true) {
      if (live && greedy_verdict<R, SHIFT>(p, S, H, hm, compute_mask) != 0) live = false;  // can never fit again
      if (!__syncthreads_or(live)) break;
      const uint32_t par = iter & 1u;
      ++iter;
      if (tid == 0) M.first[par] = BIG;
      long long x[R], E[R], T[R];
#pragma unroll
      for (int r = 0; r < R; ++r) x[r] = live ? p.q[r] : 0;
      tile_exclusive<R>(M, x, E);  // its barriers also publish the reset of first[par]
#pragma unroll
      for (int r = 0; r < R; ++r) T[r] = S[r] + E[r];
      const uint32_t fail = __ballot_sync(0xFFFFFFFFu, live && greedy_verdict<R, SHIFT>(p, T, H, hm, compute_mask) != 0);
      if (fail && lane == (uint32_t)(__ffs(fail) - 1)) atomicMin(&M.first[par], tid);
      __syncthreads();
      const uint32_t v = M.first[par];
      if (v == BIG) {
        adm = adm || live;
#pragma unroll
        for (int r = 0; r < R; ++r) S[r] += M.total[r];
        break;
      }
      if (live && tid < v) adm = true;
#pragma unroll
      for (int r = 0; r < R; ++r) S[r] += M.E[r][v];
      if (tid <= v) live = false;
    }
    done(S);
    if (__syncthreads_or(adm)) {  // reason bits against the usage each rejected pod actually met
      long long x[R], E[R], T[R];
#pragma unroll
      for (int r = 0; r < R; ++r) x[r] = adm ? p.q[r] : 0;
      tile_exclusive<R>(M, x, E);
#pragma unroll
      for (int r = 0; r < R; ++r) T[r] = S0[r] + E[r];
      my_verdict = adm ? 0u : greedy_verdict<R, SHIFT>(p, T, H, hm, compute_mask);
    } else {                      // nothing admitted in this tile: usage never moved
      my_verdict = valid ? greedy_verdict<R, SHIFT>(p, S0, H, hm, compute_mask) : 0u;
    }
    __syncthreads();
    return my_verdict;
  }
  while (pos < (uint32_t)kGreedyTeam) {
    const uint32_t par = iter & 1u;
    ++iter;
    // ---- admit-run: block-wide exclusive prefix of the adds of threads >= pos
    if (tid == 0) { M.first[par] = BIG; M.fit[par] = BIG; }
    long long a[R], E[R], T[R];
#pragma unroll
    for (int r = 0; r < R; ++r) a[r] = (valid && tid >= pos) ? p.q[r] : 0;
    tile_exclusive<R>(M, a, E);
#pragma unroll
    for (int r = 0; r < R; ++r) T[r] = (long long)((unsigned long long)S[r] + (unsigned long long)E[r]);
    const uint32_t v_run = greedy_verdict<R, SHIFT>(p, T, H, hm, compute_mask);
    const uint32_t fail = __ballot_sync(0xFFFFFFFFu, valid && tid >= pos && v_run != 0);
    if (fail && lane == (uint32_t)(__ffs(fail) - 1)) atomicMin(&M.first[par], tid);
    __syncthreads();
    const uint32_t v = M.first[par];
    if (v == BIG) {  // everyone left in the tile is admitted
#pragma unroll
      for (int r = 0; r < R; ++r) S[r] = (long long)((unsigned long long)S[r] + (unsigned long long)M.total[r]);
      break;
    }
    if (tid == v) my_verdict = v_run;
#pragma unroll
    for (int r = 0; r < R; ++r) S[r] = (long long)((unsigned long long)S[r] + (unsigned long long)M.E[r][v]);
    pos = v + 1;
    if (pos >= (uint32_t)kGreedyTeam) break;
    // ---- reject-run: usage stands still at S until somebody fits on their own
    const uint32_t v_alone = greedy_verdict<R, SHIFT>(p, S, H, hm, compute_mask);
    const uint32_t fits = __ballot_sync(0xFFFFFFFFu, valid && tid >= pos && v_alone == 0);
    if (fits && lane == (uint32_t)(__ffs(fits) - 1)) atomicMin(&M.fit[par], tid);
    __syncthreads();
    const uint32_t stop = M.fit[par] == BIG ? (uint32_t)kGreedyTeam : M.fit[par];
    if (tid >= pos && tid < stop && valid) my_verdict = v_alone;
    pos = stop;
  }
  done(S);
  __syncthreads();
  return my_verdict;
}

// block-wide sum of x[] into out[] (shared); one barrier inside, the caller adds the one that publishes out[]
template <int R>
__device__ __forceinline__ void tile_sum(long long (&x)[R], long long (*wtot)[32], long long* out) {
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
#pragma unroll
  for (int r = 0; r < R; ++r) {
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) x[r] += __shfl_xor_sync(0xFFFFFFFFu, x[r], d);
  }
  if (lane == 0) {
#pragma unroll
    for (int r = 0; r < R; ++r) wtot[r][warp] = x[r];
  }
  __syncthreads();
  if (warp == 0) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      long long y = wtot[r][lane];
#pragma unroll
      for (int d = 16; d >= 1; d >>= 1) y += __shfl_xor_sync(0xFFFFFFFFu, y, d);
      if (lane == 0) out[r] = y;  // <= 1024 * 2^40
    }
  }
}

// does usage S + X still fit?  (only resources somebody asks for can be exceeded; sums stay below 2^63)
template <int R>
__device__ __forceinline__ bool chain_fits(const long long (&S)[R], const long long (&X)[R], const long long (&H)[R], uint32_t hm) {
  bool ok = true;
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const long long sum = S[r] + X[r];  // S in [0, 2^62], X in [0, 2^61]
    ok = ok && sum <= (1ll << 62) && !(((hm >> r) & 1u) && X[r] > 0 && sum > H[r]);
  }
  return ok;
}

// Warp 0 of the CTA that holds tile c walks back over the descriptors of tiles first..c-1 of the same namespace.
// Returns 0 with S = S_in when it is known, 1 with S = a lower bound of S_in (the nearest exact usage) when the
// aggregates in between do not fit, 2 when nothing can be said yet.
template <int R>
__device__ __forceinline__ int chain_lookback(const ChainBufs& C, uint32_t c, uint32_t first, const long long (&H)[R], uint32_t hm,
                                              long long (&S)[R]) {
  const uint32_t lane = threadIdx.x & 31u;
  long long A[R];
#pragma unroll
  for (int r = 0; r < R; ++r) A[r] = 0;
  uint32_t pos = c;  // tiles pos..c-1 are summed in A
  for (uint32_t hop = 0; hop < kChainMaxBack / 32; ++hop) {
    const bool in = pos >= first + lane + 1;
    const uint32_t j = pos - 1u - lane;
    const uint32_t st = in ? ld_acquire_u32(C.status + j) : kChainPending;
    const uint32_t m_exact = __ballot_sync(0xFFFFFFFFu, st == kChainExact);
    const uint32_t m_block = __ballot_sync(0xFFFFFFFFu, st == kChainPending || st == kChainOpaque);
    const uint32_t e = m_exact ? (uint32_t)__ffs(m_exact) - 1u : 32u;
    const uint32_t nearer = e == 32u ? 0xFFFFFFFFu : ((1u << e) - 1u);
    if (m_block & nearer) return 2;
    const bool take = lane < e && (st == kChainAgg || st == kChainAgg2);
    const long long* from = st == kChainAgg2 ? C.agg2 : C.agg;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      long long a = take ? __ldcg(from + (size_t)j * R + r) : 0;
#pragma unroll
      for (int d = 16; d >= 1; d >>= 1) a += __shfl_xor_sync(0xFFFFFFFFu, a, d);
      A[r] += a;                                 // <= 2^61 + 32 * 2^50
      if (A[r] > (1ll << 61)) A[r] = 1ll << 61;  // too much to fit anyway
    }
    if (e < 32u) {
      bool sane = true, zero = true;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        long long b = lane == e ? __ldcg(C.exact + (size_t)j * R + r) : 0;
        S[r] = __shfl_sync(0xFFFFFFFFu, b, e);
        sane = sane && S[r] >= 0 && S[r] <= (1ll << 62);
        zero = zero && A[r] == 0;
      }
      if (zero) return 0;  // nothing but IDENT tiles (and tiles of all-zero requests) in between
      if (!sane) return 2;
      if (chain_fits<R>(S, A, H, hm)) {
#pragma unroll
        for (int r = 0; r < R; ++r) S[r] += A[r];
        return 0;
      }
      return 1;
    }
    pos -= 32u;
  }
  return 2;
}

template <int R, bool HAS_MASK, typename VT>
__global__ void __launch_bounds__(kGreedyTeam, 1)
k3_admit_greedy_chain(const uint32_t* __restrict__ order, const uint32_t* __restrict__ seg, ReqCols<R> req,
                      const uint8_t* __restrict__ req_mask, long long* used, const long long* __restrict__ hard,
                      const uint8_t* __restrict__ hard_mask, uint32_t NS, uint32_t compute_mask,
                      VT* __restrict__ verdict, ChainBufs C) {
  constexpr int SHIFT = sizeof(VT) == 1 ? 4 : 8;
  extern __shared__ __align__(16) long long g_chain_E[];  // [R][kGreedyTeam]
  __shared__ long long s_wtot[R][32];
  __shared__ long long s_total[R];
  __shared__ long long s_X[R], s_in[R];
  __shared__ uint32_t s_first[2], s_fit[2], s_tile, s_res;
  TileSmem<R> M;
  M.E = reinterpret_cast<long long (*)[kGreedyTeam]>(g_chain_E);
  M.wtot = s_wtot; M.total = s_total; M.first = s_first; M.fit = s_fit;
  const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const uint32_t tiles = C.tile_base[NS];
  uint32_t iter = 0;
  for (;;) {
    if (tid == 0) s_tile = atomicAdd(C.ticket, 1u);
    __syncthreads();
    if (s_tile >= tiles) break;
#ifdef BGC_K3_CLOCKS
    const long long ck0 = clock64();
#endif
    // (Taking the NEXT ticket once S_in is known, and prefetching that tile's requests into L2 while this one is
    // walked, measured slower — 11.9 against 11.6 ms: a tile is PENDING, opaque to its successors, from its ticket
    // until its CTA has summed it, and that window grew.)
    const uint4 rec = C.ticket_rec[s_tile];  // one load instead of a chain of five dependent ones
    const uint32_t c = rec.x, ns = rec.y, first = rec.z, t = rec.w;
    const uint32_t end = seg[ns + 1];
    const bool valid = t + tid < end;
    const bool is_last = t + (uint32_t)kGreedyTeam >= end;
    GreedyPod<R> p;
#pragma unroll
    for (int r = 0; r < R; ++r) p.q[r] = valid ? req.p[r][t + tid] : 0;
    p.mask = (HAS_MASK && valid) ? req_mask[t + tid] : 0xFFu;
    const uint32_t ord = valid ? order[t + tid] : 0u;
    long long H[R], S[R];
#pragma unroll
    for (int r = 0; r < R; ++r) H[r] = hard[(size_t)r * NS + ns];
    const uint32_t hm = hard_mask[ns];
    // what can be said before S_in is known
    bool tame_q = true;
#pragma unroll
    for (int r = 0; r < R; ++r) tame_q = tame_q && p.q[r] >= 0 && p.q[r] <= (1ll << 40);
    const bool tile_tame = __syncthreads_and(tame_q);
    uint32_t missing = 0;
#pragma unroll
    for (int r = 0; r < R; ++r)
      if (((hm >> r) & 1u) && !((p.mask >> r) & 1u) && ((compute_mask >> r) & 1u)) missing |= 1u << r;
    const bool eligible = valid && missing == 0;  // a pod that misses a required request is rejected whatever the usage
    if (tile_tame) {
      long long x[R];
#pragma unroll
      for (int r = 0; r < R; ++r) x[r] = eligible ? p.q[r] : 0;
      tile_sum<R>(x, s_wtot, s_X);
      if (tid == 0 && c != first) {
#pragma unroll
        for (int r = 0; r < R; ++r) __stcg(C.agg + (size_t)c * R + r, s_X[r]);
      }
    }
    if (tid == 0 && c != first) {
      __threadfence();
      st_release_u32(C.status + c, tile_tame ? kChainAgg : kChainOpaque);
    }
    __syncthreads();
#ifdef BGC_K3_CLOCKS
    const long long ck1 = clock64();
#endif
    // ---- S_in
    bool ident = false, agg2 = false, live = eligible;  // live: not yet struck out by a lower bound of S_in
    if (c == first) {
#pragma unroll
      for (int r = 0; r < R; ++r) S[r] = used[(size_t)r * NS + ns];
    } else {
      bool tried_ident = !tile_tame;
      unsigned long long t0 = 0;
      for (uint32_t spin = 0;;) {
        if (warp == 0) {
          // warp 0 polls on its own until there is something for the whole CTA to do: S_in is known, or a first lower
          // bound has turned up (the other 31 warps wait at the barrier below instead of taking part in every poll)
          long long W[R];
          int res;
          for (;;) {
            res = chain_lookback<R>(C, c, first, H, hm, W);
            if (res == 0 || (res == 1 && !tried_ident)) break;
            if (++spin > 64) __nanosleep(200);
            if ((spin & 1023u) == 1023u) {  // a tile never waits for a tile that is not running: this is a bug trap, not a time limit
              const unsigned long long now = global_ns();
              if (t0 == 0) t0 = now;
              else if (now - t0 > 30000000000ull) __trap();
            }
          }
          if (lane == 0) {
            s_res = (uint32_t)res;
#pragma unroll
            for (int r = 0; r < R; ++r) s_in[r] = W[r];
          }
        }
        __syncthreads();
        const uint32_t res = s_res;
#pragma unroll
        for (int r = 0; r < R; ++r) S[r] = s_in[r];
        if (res == 0) break;
        if (res == 1 && !tried_ident) {  // S is a lower bound of S_in: who does not fit against it never will
          tried_ident = true;
          live = eligible && greedy_verdict<R, SHIFT>(p, S, H, hm, compute_mask) == 0;
          bool moves = false;  // an admitted pod of all-zero requests leaves the usage where it is
#pragma unroll
          for (int r = 0; r < R; ++r) moves = moves || p.q[r] != 0;
          if (!__syncthreads_or(live && moves)) {
            ident = true;
            if (tid == 0) st_release_u32(C.status + c, kChainIdent);
          } else {
            agg2 = true;
            long long x[R];
#pragma unroll
            for (int r = 0; r < R; ++r) x[r] = live ? p.q[r] : 0;
            tile_sum<R>(x, s_wtot, s_X);  // replaces the sum over all eligible pods: the struck ones are out for good
            if (tid == 0) {
#pragma unroll
              for (int r = 0; r < R; ++r) __stcg(C.agg2 + (size_t)c * R + r, s_X[r]);
              __threadfence();
              st_release_u32(C.status + c, kChainAgg2);
            }
          }
        }
        __syncthreads();  // s_res / s_in are rewritten by the next walk
      }
    }
#ifdef BGC_K3_CLOCKS
    const long long ck2 = clock64();
#endif
    // ---- S_out known without walking the tile?  Tell the successors now.
    bool early = false;
    if (tile_tame) {
      long long X[R];
      bool sane = true;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        X[r] = ident ? 0 : s_X[r];
        sane = sane && S[r] >= 0 && S[r] <= (1ll << 62);
      }
      early = sane && (ident || chain_fits<R>(S, X, H, hm));
      if (early && tid == 0) {
#pragma unroll
        for (int r = 0; r < R; ++r) __stcg(C.exact + (size_t)c * R + r, S[r] + X[r]);
        __threadfence();
        st_release_u32(C.status + c, kChainExact);
      }
    }
    uint32_t my_verdict;
    if (early && ident) {  // nobody moves the usage: every pod met S_in
      my_verdict = valid ? greedy_verdict<R, SHIFT>(p, S, H, hm, compute_mask) : 0u;
    } else if (early) {    // the pods still live are admitted one and all; the others met S_in + what was admitted before them
      long long x[R], E[R], T[R];
#pragma unroll
      for (int r = 0; r < R; ++r) x[r] = live ? p.q[r] : 0;
      tile_exclusive<R>(M, x, E);
#pragma unroll
      for (int r = 0; r < R; ++r) T[r] = S[r] + E[r];
      my_verdict = (live || !valid) ? 0u : greedy_verdict<R, SHIFT>(p, T, H, hm, compute_mask);
#pragma unroll
      for (int r = 0; r < R; ++r) S[r] += M.total[r];
      __syncthreads();  // M is reused by the next tile
    } else {
      my_verdict = greedy_tile<R, SHIFT>(M, p, valid, S, H, hm, compute_mask, iter, [&](const long long (&So)[R]) {
        if (tid == 0) {
#pragma unroll
          for (int r = 0; r < R; ++r) __stcg(C.exact + (size_t)c * R + r, So[r]);
          __threadfence();
          st_release_u32(C.status + c, kChainExact);
        }
      });
    }
    if (valid) verdict[ord] = (VT)my_verdict;
#ifdef BGC_K3_CLOCKS
    if (tid == 0) {  // cycles: ticket -> AGG published, waiting for S_in, walking the tile
      unsigned long long* ck = reinterpret_cast<unsigned long long*>(C.ticket + 8);
      atomicAdd(ck + 0, (unsigned long long)(ck1 - ck0));
      atomicAdd(ck + 1, (unsigned long long)(ck2 - ck1));
      atomicAdd(ck + 2, (unsigned long long)(clock64() - ck2));
    }
#endif
    if (tid == 0)  // how the tile got its S_out, for BGC_K3_STATS: [1] first of a segment, [2] IDENT, [3] AGG2 fitted, [4] AGG fitted, [5] walked
      atomicAdd(C.ticket + (c == first ? 1 : ident ? 2 : early && agg2 ? 3 : early ? 4 : 5), 1u);
    if (is_last && tid == 0) {
#pragma unroll
      for (int r = 0; r < R; ++r) used[(size_t)r * NS + ns] = S[r];
    }
  }
}

}  // namespace bgc
