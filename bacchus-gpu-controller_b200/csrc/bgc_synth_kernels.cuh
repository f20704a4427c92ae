// bgc_synth_kernels.cuh — synthetic generator, checksum and histogram kernels.  Part of libbgc.so's device code; included through bgc_kernels.cuh.
#pragma once
#include "bgc_common.cuh"

namespace bgc {

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

// ------------------------------------------------------------------ one global stream, one shard of it
// Row E: rank g of G keeps the pods of the global counter stream whose
// namespace hashes to it, shard = mix32(ns_id) % G, in stream order.  Two passes over fixed tiles of
// kShardTile pods so that the output order does not depend on scheduling: count per tile, exclusive
// scan of the tile counts, then regenerate and write each kept pod at its final position.
constexpr int kShardTile = 2048;  // pods per CTA tile: 256 threads x 8 consecutive pods

__device__ __forceinline__ uint32_t synth_ns(const SynthArgs& A, uint64_t i) {
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
  return (uint32_t)(((uint64_t)rank * A.a + A.b) % A.ns_count);
}
__device__ __forceinline__ uint32_t shard_mix32(uint32_t x) {  // murmur3 fmix32, same as shard.py
  x ^= x >> 16; x *= 0x85EBCA6Bu; x ^= x >> 13; x *= 0xC2B2AE35u; x ^= x >> 16;
  return x;
}

// pass 1: tile_count[t] = pods of tile t owned by `rank`
static __global__ void __launch_bounds__(256)
k_synth_shard_count(SynthArgs A, uint64_t first, size_t count, uint32_t world, uint32_t rank, uint32_t* tile_count) {
  __shared__ uint32_t s_warp[8];
  const size_t tiles = (count + kShardTile - 1) / kShardTile;
  for (size_t t = blockIdx.x; t < tiles; t += gridDim.x) {
    uint32_t mine = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const size_t k = t * kShardTile + (size_t)threadIdx.x * 8 + j;
      if (k < count) mine += shard_mix32(synth_ns(A, first + k)) % world == rank;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) mine += __shfl_xor_sync(0xFFFFFFFFu, mine, d);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = mine;
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t tot = 0;
      for (int w = 0; w < 8; ++w) tot += s_warp[w];
      tile_count[t] = tot;
    }
    __syncthreads();
  }
}

// exclusive scan of the tile counts into 64-bit offsets starting at `base`; one CTA; total -> *out_total
static __global__ void __launch_bounds__(1024)
k_scan_tiles(const uint32_t* tile_count, size_t tiles, unsigned long long base, unsigned long long* tile_off,
             unsigned long long* out_total) {
  __shared__ unsigned long long s_part[1024];
  const size_t per = (tiles + 1023) / 1024;
  const size_t lo = (size_t)threadIdx.x * per, hi = lo + per < tiles ? lo + per : tiles;
  unsigned long long sum = 0;
  for (size_t t = lo; t < hi; ++t) sum += tile_count[t];
  s_part[threadIdx.x] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long run = base;
    for (int i = 0; i < 1024; ++i) { const unsigned long long v = s_part[i]; s_part[i] = run; run += v; }
    *out_total = run - base;
  }
  __syncthreads();
  unsigned long long run = s_part[threadIdx.x];
  for (size_t t = lo; t < hi; ++t) { tile_off[t] = run; run += tile_count[t]; }
}

// pass 2: regenerate, keep, write in stream order; ids are rewritten through local_of[] (global ->
// dense per-shard id) when given; src_index[] (optional) records each kept pod's global position
static __global__ void __launch_bounds__(256)
k_synth_shard_emit(SynthArgs A, uint64_t first, size_t count, uint32_t world, uint32_t rank,
                   const uint32_t* __restrict__ local_of, const unsigned long long* __restrict__ tile_off,
                   unsigned long long* src_index) {
  __shared__ uint32_t s_warp[8];
  const size_t tiles = (count + kShardTile - 1) / kShardTile;
  for (size_t t = blockIdx.x; t < tiles; t += gridDim.x) {
    uint32_t ns[8];
    uint32_t keep = 0, mine = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const size_t k = t * kShardTile + (size_t)threadIdx.x * 8 + j;
      ns[j] = 0;
      if (k < count) {
        ns[j] = synth_ns(A, first + k);
        if (shard_mix32(ns[j]) % world == rank) { keep |= 1u << j; ++mine; }
      }
    }
    // exclusive prefix of `mine` over the CTA in thread order (= stream order)
    uint32_t incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
      if ((threadIdx.x & 31) >= (uint32_t)d) incl += o;
    }
    if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = incl;
    __syncthreads();
    uint32_t before = incl - mine;
    for (uint32_t w = 0; w < (threadIdx.x >> 5); ++w) before += s_warp[w];
    unsigned long long pos = tile_off[t] + before;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (!((keep >> j) & 1u)) continue;
      const uint64_t i = first + t * kShardTile + (size_t)threadIdx.x * 8 + j;
      A.ns_id[pos] = local_of ? local_of[ns[j]] : ns[j];
      for (uint32_t r = 0; r < A.R; ++r) A.req[r][pos] = synth_value(A.kind[r], draw(A.key_req[r], i));
      if (A.req_mask) {
        const uint64_t h = draw(A.key_mask, i);
        uint32_t m = 0xFFu;
        for (uint32_t r = 0; r < A.R; ++r)
          if (((h >> (8 * r)) & 0xFFu) < 8u) m &= ~(1u << r);
        A.req_mask[pos] = (uint8_t)m;
      }
      if (src_index) src_index[pos] = i;
      ++pos;
    }
    __syncthreads();
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
