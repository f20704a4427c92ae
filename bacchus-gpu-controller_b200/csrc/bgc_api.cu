// bgc_api.cu — the C ABI of libbgc.so (include/bgc.h, include/bgc_synth.h).
//
// Host side of the boundary that sits behind the reference's `mutate` seam
// (reference src/admission.rs:241-244): a context per caller thread, caller-owned POD buffers,
// integer status codes, no exceptions across the boundary, no CPU fallback.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "../../include/bgc.h"
#include "../../include/bgc_synth.h"

#include "bgc_kernels.cuh"

using namespace bgc;

namespace {

constexpr size_t kChunkPods = size_t(4) << 20;      // host-pointer calls move 4 Mi pods per stage
constexpr size_t kSmemBudget = 224 * 1024;          // CTA-private K1 table budget (<= 227 KB opt-in)
constexpr size_t kSmemMaxDynamic = 227 * 1024;      // the opt-in maximum of dynamic shared memory per CTA on sm_100
constexpr int kSlots = 2;
constexpr size_t kZeroCopyMaxPods = 262144;         // host batches up to this size take the latency path
constexpr size_t kZeroCopyBytes = kZeroCopyMaxPods * (4 + 8 * kMaxR + 1 + 2 + 2) + 16 * 256;

struct Slot {
  uint32_t* ns = nullptr;
  long long* req[kMaxR] = {};
  uint8_t* mask = nullptr;
  uint8_t* out0 = nullptr;  // verdict / code
  uint8_t* out1 = nullptr;  // patch_mask
  uint16_t* flags = nullptr;
  cudaEvent_t h2d_done = nullptr, k_done = nullptr, free_ev = nullptr;
  bool used = false;
};

}  // namespace

struct bgc_ctx {
  int device = 0;
  int sm_count = 148;
  cudaStream_t own_stream = nullptr, stream = nullptr, h2d_stream = nullptr, d2h_stream = nullptr;
  std::string err;
  // quota table
  long long* d_used = nullptr;       // [R][NS]; K1 accumulates here
  long long* d_hard = nullptr;       // [R][NS]
  uint8_t* d_hard_mask = nullptr;
  uint32_t NS = 0;
  int R = 0;
  uint32_t compute_mask = 0;
  uint32_t wide_mask = 0;            // byte-valued columns: K1 keeps their high words in shared memory
  long long* head = nullptr;         // K2's compact table (hard - used, mask), rebuilt when stale
  bool head_stale = true;
  // K1
  int k1_forced = BGC_K1_AUTO;
  int k1_shape = 0;                  // scatter kernel launch shape (tuning knob, see bgc_ctx_set_k1_variant)
  // L2 bulk-prefetch distance in loop trips; -1 = automatic (measured, profiles/r02_probe_*): one trip ahead for
  // the hash-sink scatter kernel (+4..6 %) and for K2 at R >= 3 (86 % -> 99 % on C3), none for the direct sink,
  // which is HBM-bound already.  BGC_K1_PREFETCH / BGC_K2_PREFETCH override.
  int k1_prefetch = -1;
  int k2_prefetch = -1;
  const char* last_k1 = "none";
  // staging for host-pointer calls
  size_t zc_max_pods = kZeroCopyMaxPods;  // BGC_ZEROCOPY_MAX_PODS overrides (0 disables the latency path)
  Slot slot[kSlots];
  size_t slot_pods = 0;
  int slot_R = 0;
  std::vector<cudaEvent_t> ev_pool;
  size_t ev_next = 0;
  unsigned long long* d_scratch = nullptr;  // 256 x u64
  // greedy admission (row N1): grouping buffers, grown on demand
  uint32_t *g_keys = nullptr, *g_keys_sorted = nullptr, *g_idx = nullptr, *g_order = nullptr, *g_seg = nullptr;
  uint32_t* g_hist = nullptr;  // digit-major histogram of the grouping sort: 256 x tiles
  long long* g_req[kMaxR] = {};  // request columns in grouped order (the sort's last pass writes them)
  long long* g_req2[kMaxR] = {}; // ... and their ping-pong partners for the passes in between
  uint8_t *g_mask = nullptr, *g_mask2 = nullptr;
  int g_R = 0;
  // device copies of a host batch for bgc_admit_greedy (host-pointer form), grown on demand
  uint32_t* gh_ns = nullptr;
  long long* gh_req[kMaxR] = {};
  uint8_t *gh_mask = nullptr, *gh_v = nullptr;
  size_t gh_cap = 0;
  int gh_R = 0;
  size_t g_cap = 0, g_hist_cap = 0, g_seg_cap = 0;
  // long segments of the greedy walk, cut into chained tiles (bgc_k3.cuh): tile table and descriptors
  uint32_t *g_tile_base = nullptr, *g_tile_li = nullptr, *g_ticket_rec = nullptr, *g_long = nullptr, *g_chain_status = nullptr;
  long long *g_chain_agg = nullptr, *g_chain_agg2 = nullptr, *g_chain_exact = nullptr;
  size_t g_tile_base_cap = 0, g_chain_cap = 0;
  int g_chain_R = 0;
  int k3_long = -1;  // -1: read BGC_K3_LONG once (chain | team); 0 chain, 1 team
  // latency path: small host batches are read by the kernel straight over PCIe (zero-copy) from
  // the caller's pinned buffers, or from this pinned, device-mapped staging area
  uint8_t* zc_host = nullptr;
  uint8_t* zc_dev = nullptr;
  // pinned bounce ring for PAGEABLE callers of the host-pointer entry points (two slots, grown on demand)
  uint8_t* bounce[2] = {nullptr, nullptr};
  size_t bounce_bytes = 0;
  int bounce_threads = 0;
  // one-shard-of-a-global-stream generator (row E): per-tile counts and offsets of one chunk
  uint32_t* sh_tile_count = nullptr;
  unsigned long long* sh_tile_off = nullptr;  // [kShardChunkTiles] offsets + 1 total
  bgc_timings tm{};
};

namespace {

int fail(bgc_ctx* c, int code, const char* fmt, ...) {
  if (c) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    c->err = buf;
  }
  return code;
}

#define CU(call)                                                                              \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess)                                                                    \
      return fail(ctx, e_ == cudaErrorMemoryAllocation ? BGC_E_OOM : BGC_E_CUDA, "%s: %s",    \
                  #call, cudaGetErrorString(e_));                                             \
  } while (0)

#define NEED(cond, msg)                                   \
  do {                                                    \
    if (!(cond)) return fail(ctx, BGC_E_BADARG, "%s", msg); \
  } while (0)

inline bool aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; }

int grid_for(size_t work_items, int per_block, int max_blocks) {
  size_t b = (work_items + per_block - 1) / per_block;
  if (b < 1) b = 1;
  if (b > (size_t)max_blocks) b = max_blocks;
  return (int)b;
}

cudaEvent_t next_event(bgc_ctx* ctx) {
  if (ctx->ev_next == ctx->ev_pool.size()) {
    cudaEvent_t e;
    if (cudaEventCreate(&e) != cudaSuccess) return nullptr;
    ctx->ev_pool.push_back(e);
  }
  return ctx->ev_pool[ctx->ev_next++];
}

// ---- kernel launchers (device pointers, ctx->stream unless told otherwise) -------------------
int launch_k0(bgc_ctx* ctx, cudaStream_t st, const uint8_t* op, const uint16_t* flags, size_t n,
              uint8_t* code, uint8_t* pm) {
  if (n == 0) return BGC_OK;
  const int vec = aligned(op, 32) && aligned(flags, 32) && aligned(code, 32) && aligned(pm, 32);
  // tuning knobs; measured on 2^30 requests: 8 CTAs/SM of grid-stride looping 88 % of the copy peak, 256 CTAs/SM
  // (3-4 trips per thread) 105 %: a streaming kernel wants its resident CTAs to sweep memory as one front
  static const int st_mode = std::getenv("BGC_K0_STORE") ? std::atoi(std::getenv("BGC_K0_STORE")) : 0;
  static const int blocks_per_sm = std::getenv("BGC_K0_BLOCKS") ? std::atoi(std::getenv("BGC_K0_BLOCKS")) : 256;
  const int grid = grid_for((n + kK0PerThread - 1) / kK0PerThread, kK0Threads, ctx->sm_count * blocks_per_sm);
  if (st_mode == 1) k0_policy_eval<1><<<grid, kK0Threads, 0, st>>>(op, flags, n, code, pm, vec);
  else if (st_mode == 2) k0_policy_eval<2><<<grid, kK0Threads, 0, st>>>(op, flags, n, code, pm, vec);
  else k0_policy_eval<0><<<grid, kK0Threads, 0, st>>>(op, flags, n, code, pm, vec);
  ctx->tm.launches++;
  CU(cudaGetLastError());
  return BGC_OK;
}

// bytes of CTA-private table per slot: a low word per column, a high word per wide column
size_t k1_slot_bytes(const bgc_ctx* ctx, int R) {
  return 4 * (size_t)(R + __builtin_popcount(ctx->wide_mask & ((1u << R) - 1u)));
}

template <int R, int SINK, bool VEC>
int launch_k1_inst(bgc_ctx* ctx, cudaStream_t st, const uint32_t* ns, const long long* const* req,
                   size_t n, uint32_t cap) {
  ReqCols<R> cols;
  for (int r = 0; r < R; ++r) cols.p[r] = req[r];
  auto kern = k1_usage_reduce<R, SINK, VEC>;
  size_t smem = 0;
  int threads = 512, blocks_per_sm = 1;
  if (SINK == SINK_GLOBAL) {
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, kern, threads, 0));
    if (blocks_per_sm < 1) blocks_per_sm = 1;
  } else {
    threads = 1024;
    smem = k1_slot_bytes(ctx, R) * cap + (SINK == SINK_HASH ? (size_t)cap * 4 : 0);
    CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  const size_t tiles = (n + kK1Tile - 1) / kK1Tile;
  int grid = ctx->sm_count * blocks_per_sm;
  const size_t warps_per_block = threads / 32;
  if ((size_t)grid * warps_per_block > tiles) grid = (int)((tiles + warps_per_block - 1) / warps_per_block);
  if (grid < 1) grid = 1;
  // a warp's chunk: long enough to amortise the carry flush, short enough to balance the grid
  size_t chunk_tiles = tiles / ((size_t)grid * warps_per_block * 4);
  if (chunk_tiles < 1) chunk_tiles = 1;
  if (chunk_tiles > 32) chunk_tiles = 32;
  kern<<<grid, threads, smem, st>>>(ns, cols, n, ctx->d_used, ctx->NS, cap, (uint32_t)chunk_tiles, ctx->wide_mask);
  ctx->head_stale = true;
  ctx->tm.launches++;
  CU(cudaGetLastError());
  return BGC_OK;
}

// arrival-order input -> lean scatter kernel (one persistent CTA per SM, CTA-private table)
template <int R, int SINK, bool VEC>
int launch_k1_scatter(bgc_ctx* ctx, cudaStream_t st, const uint32_t* ns, const long long* const* req,
                      size_t n, uint32_t cap) {
  ReqCols<R> cols;
  for (int r = 0; r < R; ++r) cols.p[r] = req[r];
  const size_t smem = k1_slot_bytes(ctx, R) * cap + (SINK == SINK_HASH ? (size_t)cap * 4 : 0);
  const size_t groups = (n + 1) / 2;
  auto go = [&](auto kern, int threads) -> int {
    // always the opt-in maximum: the attribute is per function and process-wide, and the allocation-free
    // entry point (bgc_noalloc.cu) launches the same kernels with its own table sizes
    CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemMaxDynamic));
    int grid = ctx->sm_count;
    if ((size_t)grid * threads > groups) grid = (int)((groups + threads - 1) / threads);
    if (grid < 1) grid = 1;
    kern<<<grid, threads, smem, st>>>(ns, cols, n, ctx->d_used, ctx->NS, cap, ctx->wide_mask,
                                      (uint32_t)(ctx->k1_prefetch >= 0 ? ctx->k1_prefetch : (SINK == SINK_HASH ? 1 : 0)));
    ctx->head_stale = true;
    ctx->tm.launches++;
    CU(cudaGetLastError());
    return BGC_OK;
  };
  if constexpr (R <= 4) {
    switch (ctx->k1_shape) {
      case 1: return go(k1_scatter<R, SINK, VEC, 512, 2>, 512);
      case 2: return go(k1_scatter<R, SINK, VEC, 1024, 2>, 1024);
      case 3: return go(k1_scatter<R, SINK, VEC, 512, 1>, 512);
      default: break;
    }
  }
  return go(k1_scatter<R, SINK, VEC, 1024, 1>, 1024);
}

// cluster-distributed table: CSIZE SMs share one direct-indexed table over distributed shared memory
template <int R, bool VEC, int CSIZE>
int launch_k1_cluster(bgc_ctx* ctx, cudaStream_t st, const uint32_t* ns, const long long* const* req, size_t n) {
  ReqCols<R> cols;
  for (int r = 0; r < R; ++r) cols.p[r] = req[r];
  auto kern = k1_scatter_cluster<R, VEC, CSIZE>;
  const uint32_t cap = (ctx->NS + CSIZE - 1) / CSIZE;
  const size_t smem = k1_slot_bytes(ctx, R) * cap;
  CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (CSIZE > 8) CU(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  cudaLaunchConfig_t cfg{};
  cfg.blockDim = dim3(1024);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr{};
  attr.id = cudaLaunchAttributeClusterDimension;
  attr.val.clusterDim.x = CSIZE;
  attr.val.clusterDim.y = 1;
  attr.val.clusterDim.z = 1;
  cfg.attrs = &attr;
  cfg.numAttrs = 1;
  int clusters = 0;
  cfg.gridDim = dim3(CSIZE);
  CU(cudaOccupancyMaxActiveClusters(&clusters, kern, &cfg));
  if (clusters < 1) return fail(ctx, BGC_E_CUDA, "no room for a %d-CTA cluster with %zu B of shared memory", CSIZE, smem);
  cfg.gridDim = dim3((unsigned)(clusters * CSIZE));
  CU(cudaLaunchKernelEx(&cfg, kern, ns, cols, n, ctx->d_used, ctx->NS, cap, ctx->wide_mask));
  ctx->head_stale = true;
  ctx->tm.launches++;
  CU(cudaGetLastError());
  return BGC_OK;
}

template <int R>
int launch_k1_r(bgc_ctx* ctx, cudaStream_t st, const uint32_t* ns, const long long* const* req,
                size_t n, unsigned flags) {
  bool vec = aligned(ns, 16);
  for (int r = 0; r < R; ++r) vec = vec && aligned(req[r], 32);
  // variant choice
  int variant = ctx->k1_forced;
  const size_t direct_bytes = (size_t)ctx->NS * k1_slot_bytes(ctx, R);
  if (variant == BGC_K1_AUTO) {
    if (flags & BGC_REDUCE_GROUPED) variant = BGC_K1_GLOBAL;
    else if (direct_bytes <= kSmemBudget) variant = BGC_K1_SMEM_DIRECT;
    else variant = BGC_K1_SMEM_HASH;
  }
  if (variant == BGC_K1_SMEM_DIRECT && direct_bytes > kSmemBudget) variant = BGC_K1_SMEM_HASH;
  uint32_t cap = ctx->NS;
  if (variant == BGC_K1_SMEM_HASH) {
    if (flags & BGC_REDUCE_GROUPED) {  // the segmented kernel hashes by shift: power-of-two slot count
      cap = 1;
      while ((size_t)cap * 2 * (k1_slot_bytes(ctx, R) + 4) <= kSmemBudget) cap *= 2;
    } else {  // the scatter kernel reduces the hash by multiply-shift: every byte of shared memory is a slot
      cap = (uint32_t)(kSmemMaxDynamic / (k1_slot_bytes(ctx, R) + 4));
    }
  }
  if (variant == BGC_K1_CLUSTER8 || variant == BGC_K1_CLUSTER16) {
    const int cs = variant == BGC_K1_CLUSTER8 ? 8 : 16;
    if ((size_t)((ctx->NS + cs - 1) / cs) * k1_slot_bytes(ctx, R) > kSmemBudget)
      return fail(ctx, BGC_E_BADARG, "namespace table does not fit a %d-CTA cluster", cs);
    if (cs == 8) {
      ctx->last_k1 = "scatter_cluster8_dsmem";
      return vec ? launch_k1_cluster<R, true, 8>(ctx, st, ns, req, n) : launch_k1_cluster<R, false, 8>(ctx, st, ns, req, n);
    }
    ctx->last_k1 = "scatter_cluster16_dsmem";
    return vec ? launch_k1_cluster<R, true, 16>(ctx, st, ns, req, n) : launch_k1_cluster<R, false, 16>(ctx, st, ns, req, n);
  }
  switch (variant) {
    case BGC_K1_GLOBAL:
      ctx->last_k1 = "segmented_global_red";
      return vec ? launch_k1_inst<R, SINK_GLOBAL, true>(ctx, st, ns, req, n, cap)
                 : launch_k1_inst<R, SINK_GLOBAL, false>(ctx, st, ns, req, n, cap);
    case BGC_K1_SMEM_DIRECT:
      if (!(flags & BGC_REDUCE_GROUPED)) {
        ctx->last_k1 = "scatter_smem_direct";
        return vec ? launch_k1_scatter<R, SINK_DIRECT, true>(ctx, st, ns, req, n, cap)
                   : launch_k1_scatter<R, SINK_DIRECT, false>(ctx, st, ns, req, n, cap);
      }
      if constexpr (R > 4) {  // the segmented kernel keeps its shared-memory sinks for R <= 4 only
        ctx->last_k1 = "segmented_global_red";
        return vec ? launch_k1_inst<R, SINK_GLOBAL, true>(ctx, st, ns, req, n, cap)
                   : launch_k1_inst<R, SINK_GLOBAL, false>(ctx, st, ns, req, n, cap);
      } else {
      ctx->last_k1 = "segmented_smem_direct";
      return vec ? launch_k1_inst<R, SINK_DIRECT, true>(ctx, st, ns, req, n, cap)
                 : launch_k1_inst<R, SINK_DIRECT, false>(ctx, st, ns, req, n, cap);
      }
    default:
      if (!(flags & BGC_REDUCE_GROUPED)) {
        ctx->last_k1 = "scatter_smem_hash";
        return vec ? launch_k1_scatter<R, SINK_HASH, true>(ctx, st, ns, req, n, cap)
                   : launch_k1_scatter<R, SINK_HASH, false>(ctx, st, ns, req, n, cap);
      }
      if constexpr (R > 4) {
        ctx->last_k1 = "segmented_global_red";
        return vec ? launch_k1_inst<R, SINK_GLOBAL, true>(ctx, st, ns, req, n, cap)
                   : launch_k1_inst<R, SINK_GLOBAL, false>(ctx, st, ns, req, n, cap);
      } else {
      ctx->last_k1 = "segmented_smem_hash";
      return vec ? launch_k1_inst<R, SINK_HASH, true>(ctx, st, ns, req, n, cap)
                 : launch_k1_inst<R, SINK_HASH, false>(ctx, st, ns, req, n, cap);
      }
  }
}

int launch_k1(bgc_ctx* ctx, cudaStream_t st, const uint32_t* ns, const long long* const* req,
              size_t n, unsigned flags) {
  if (n == 0) return BGC_OK;
  switch (ctx->R) {
    case 1: return launch_k1_r<1>(ctx, st, ns, req, n, flags);
    case 2: return launch_k1_r<2>(ctx, st, ns, req, n, flags);
    case 3: return launch_k1_r<3>(ctx, st, ns, req, n, flags);
    case 4: return launch_k1_r<4>(ctx, st, ns, req, n, flags);
    case 5: return launch_k1_r<5>(ctx, st, ns, req, n, flags);
    case 6: return launch_k1_r<6>(ctx, st, ns, req, n, flags);
    case 7: return launch_k1_r<7>(ctx, st, ns, req, n, flags);
    default: return launch_k1_r<8>(ctx, st, ns, req, n, flags);
  }
}

template <int R, bool HAS_MASK, typename VT>
int launch_k2_inst(bgc_ctx* ctx, cudaStream_t st, const uint32_t* ns, const long long* const* req,
                   const uint8_t* mask, size_t n, VT* verdict) {
  ReqCols<R> cols;
  constexpr int PODS = R <= 4 ? 4 : 2;
  bool vec = aligned(ns, 16) && aligned(verdict, 8) && (!HAS_MASK || aligned(mask, 4));
  for (int r = 0; r < R; ++r) {
    cols.p[r] = req[r];
    vec = vec && aligned(req[r], 32);
  }
  if (ctx->head_stale) {  // usage or limits changed since the compact table was built
    const int g = grid_for((size_t)ctx->NS * head_words(R), 256, ctx->sm_count * 4);
    k_build_headroom<<<g, 256, 0, st>>>(ctx->d_used, ctx->d_hard, ctx->d_hard_mask, ctx->head, ctx->NS, R);
    ctx->tm.launches++;
    CU(cudaGetLastError());
    ctx->head_stale = false;
  }
  // one or two resources and a table that fits shared memory: CTA-private copy of the table (see bgc_k2.cuh)
  if constexpr (R <= 2 && sizeof(VT) == 1) {
    static const int use_smem = std::getenv("BGC_K2_SMEM") ? std::atoi(std::getenv("BGC_K2_SMEM")) : 1;
    const size_t smem = (size_t)ctx->NS * 8 * R + ctx->NS;
    if (use_smem && smem <= kSmemMaxDynamic && n >= (size_t)ctx->sm_count * 1024 * 8) {
      const uint32_t pf = (uint32_t)(ctx->k2_prefetch >= 0 ? ctx->k2_prefetch : 1);
      auto go = [&](auto kern) -> int {
        CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemMaxDynamic));
        kern<<<ctx->sm_count, 1024, smem, st>>>(ns, cols, mask, n, ctx->head, ctx->d_hard_mask, ctx->NS, ctx->compute_mask,
                                                reinterpret_cast<uint8_t*>(verdict), pf);
        ctx->tm.launches++;
        CU(cudaGetLastError());
        return BGC_OK;
      };
      return vec ? go(k2_quota_eval_smem<R, HAS_MASK, true>) : go(k2_quota_eval_smem<R, HAS_MASK, false>);
    }
  }
  static const int k2_blocks = std::getenv("BGC_K2_BLOCKS") ? std::atoi(std::getenv("BGC_K2_BLOCKS")) : 32;  // tuning knob
  const int grid = grid_for((n + PODS - 1) / PODS, kK2Threads, ctx->sm_count * k2_blocks);
  if (vec)
    k2_quota_eval<R, HAS_MASK, true, VT><<<grid, kK2Threads, 0, st>>>(
        ns, cols, mask, n, ctx->head, ctx->d_hard_mask, ctx->NS, ctx->compute_mask, verdict,
        (uint32_t)(ctx->k2_prefetch >= 0 ? ctx->k2_prefetch : (R >= 3 ? 1 : 0)));
  else
    k2_quota_eval<R, HAS_MASK, false, VT><<<grid, kK2Threads, 0, st>>>(
        ns, cols, mask, n, ctx->head, ctx->d_hard_mask, ctx->NS, ctx->compute_mask, verdict,
        (uint32_t)(ctx->k2_prefetch >= 0 ? ctx->k2_prefetch : (R >= 3 ? 1 : 0)));
  ctx->tm.launches++;
  CU(cudaGetLastError());
  return BGC_OK;
}

// VT = uint8_t: missing << 4 | exceeded (R <= 4);  VT = uint16_t: missing << 8 | exceeded (R <= 8)
template <typename VT>
int launch_k2(bgc_ctx* ctx, cudaStream_t st, const uint32_t* ns, const long long* const* req,
              const uint8_t* mask, size_t n, VT* verdict) {
  if (n == 0) return BGC_OK;
#define K2CASE(RR)                                                                       \
  case RR:                                                                               \
    return mask ? launch_k2_inst<RR, true, VT>(ctx, st, ns, req, mask, n, verdict)       \
                : launch_k2_inst<RR, false, VT>(ctx, st, ns, req, mask, n, verdict);
  if constexpr (sizeof(VT) == 1) {
    switch (ctx->R) {
      K2CASE(1) K2CASE(2) K2CASE(3) K2CASE(4)
      default: return fail(ctx, BGC_E_BADARG, "u8 verdicts hold at most 4 resources: use bgc_quota_eval16 for R = %d", ctx->R);
    }
  } else {
    switch (ctx->R) {
      K2CASE(1) K2CASE(2) K2CASE(3) K2CASE(4) K2CASE(5) K2CASE(6) K2CASE(7) K2CASE(8)
      default: return fail(ctx, BGC_E_BADARG, "R out of range");
    }
  }
#undef K2CASE
}

int clear_used(bgc_ctx* ctx, cudaStream_t st) {
  ctx->head_stale = true;
  CU(cudaMemsetAsync(ctx->d_used, 0, (size_t)ctx->NS * ctx->R * 8, st));
  return BGC_OK;
}

// the usage table already has the [R][NS] layout of the ABI: a plain device-to-device copy
int export_used(bgc_ctx* ctx, cudaStream_t st, long long* d_out) {
  CU(cudaMemcpyAsync(d_out, ctx->d_used, (size_t)ctx->NS * ctx->R * 8, cudaMemcpyDeviceToDevice, st));
  return BGC_OK;
}

void free_slots(bgc_ctx* ctx) {
  for (auto& s : ctx->slot) {
    cudaFree(s.ns);
    for (auto& p : s.req) { cudaFree(p); p = nullptr; }
    cudaFree(s.mask);
    cudaFree(s.out0);
    cudaFree(s.out1);
    cudaFree(s.flags);
    s.ns = nullptr; s.mask = s.out0 = s.out1 = nullptr; s.flags = nullptr;
  }
  ctx->slot_pods = 0;
  ctx->slot_R = 0;
}

// device staging for host-pointer calls, sized to the largest chunk seen so far
int ensure_slots(bgc_ctx* ctx, size_t pods, int R) {
  if (pods > kChunkPods) pods = kChunkPods;
  pods = (pods + 127) & ~size_t(127);
  if (pods <= ctx->slot_pods && R <= ctx->slot_R) return BGC_OK;
  if (pods < ctx->slot_pods) pods = ctx->slot_pods;
  if (R < ctx->slot_R) R = ctx->slot_R;
  CU(cudaDeviceSynchronize());
  free_slots(ctx);
  for (auto& s : ctx->slot) {
    CU(cudaMalloc(&s.ns, pods * 4));
    for (int r = 0; r < R; ++r) CU(cudaMalloc(&s.req[r], pods * 8));
    CU(cudaMalloc(&s.mask, pods));
    CU(cudaMalloc(&s.out0, pods * 2));
    CU(cudaMalloc(&s.out1, pods));
    CU(cudaMalloc(&s.flags, pods * 2));
    if (!s.h2d_done) {
      CU(cudaEventCreateWithFlags(&s.h2d_done, cudaEventDisableTiming));
      CU(cudaEventCreateWithFlags(&s.k_done, cudaEventDisableTiming));
      CU(cudaEventCreateWithFlags(&s.free_ev, cudaEventDisableTiming));
    }
    s.used = false;
  }
  ctx->slot_pods = pods;
  ctx->slot_R = R;
  return BGC_OK;
}

// Latency path helpers.  A column that already lives in pinned (page-locked) host memory is handed
// to the kernel as is; a pageable one is copied into the context's pinned staging area first.
struct ZcCursor {
  size_t off = 0;
};

const void* zc_in(bgc_ctx* ctx, ZcCursor& cur, const void* host, size_t bytes) {
  cudaPointerAttributes a{};
  if (cudaPointerGetAttributes(&a, host) == cudaSuccess && a.type == cudaMemoryTypeHost && a.devicePointer)
    return a.devicePointer;
  cudaGetLastError();
  const size_t at = cur.off;
  cur.off = (cur.off + bytes + 255) & ~size_t(255);
  std::memcpy(ctx->zc_host + at, host, bytes);
  return ctx->zc_dev + at;
}

// returns the device-visible address to write to; *bounce = staging address to copy back from (or null)
void* zc_out(bgc_ctx* ctx, ZcCursor& cur, void* host, size_t bytes, uint8_t** bounce) {
  cudaPointerAttributes a{};
  *bounce = nullptr;
  if (cudaPointerGetAttributes(&a, host) == cudaSuccess && a.type == cudaMemoryTypeHost && a.devicePointer)
    return a.devicePointer;
  cudaGetLastError();
  const size_t at = cur.off;
  cur.off = (cur.off + bytes + 255) & ~size_t(255);
  *bounce = ctx->zc_host + at;
  return ctx->zc_dev + at;
}

struct Spans {
  std::vector<cudaEvent_t> h2d, k, d2h;  // start/stop pairs
};

int finish_timings(bgc_ctx* ctx, const Spans& sp) {
  auto sum = [&](const std::vector<cudaEvent_t>& v, uint64_t* out) -> cudaError_t {
    double total = 0;
    for (size_t i = 0; i + 1 < v.size(); i += 2) {
      float ms = 0;
      cudaError_t e = cudaEventElapsedTime(&ms, v[i], v[i + 1]);
      if (e != cudaSuccess) return e;
      total += ms;
    }
    *out = (uint64_t)(total * 1e6);
    return cudaSuccess;
  };
  CU(sum(sp.h2d, &ctx->tm.h2d_ns));
  CU(sum(sp.k, &ctx->tm.kernel_ns));
  CU(sum(sp.d2h, &ctx->tm.d2h_ns));
  return BGC_OK;
}

// ---- pageable callers ------------------------------------------------------------------------------
// cudaMemcpyAsync from pageable memory is staged by the driver through its own small pinned buffer, on
// the calling thread, at a fraction of the PCIe rate (round 1: the GPU never beat the host cores from
// pageable memory).  A real caller at the mutate_handler seam (reference src/admission.rs:184-187) holds
// serde-owned heap buffers, i.e. pageable ones.  For them a large host batch is cut into 4 Mi-pod pieces;
// while the pinned pipeline works on piece j out of one pinned slot, a helper thread — itself fanning out
// over a few cores — memcpys piece j+1 from the caller's memory into the other slot.
struct BounceCol {
  const void* src;
  size_t width;
};
struct BounceOut {
  void* dst;
  size_t width;
};
constexpr size_t kBouncePods = size_t(4) << 20;
constexpr int kBounceMaxCols = kMaxR + 3;

bool is_pageable(const void* p) {
  cudaPointerAttributes a{};
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return true;
  }
  return a.type == cudaMemoryTypeUnregistered;
}

bool wants_bounce(const bgc_ctx* ctx, const void* first_column, size_t n) {
  // below ~1 Mi pods the driver's own staging is as fast (measured, profiles/r02_crossover_c4.md)
  return ctx->bounce_threads > 0 && n > ctx->zc_max_pods && n > kBouncePods / 4 && is_pageable(first_column);
}

template <class Call>
int bounced(bgc_ctx* ctx, const BounceCol* cols, int ncols, const BounceOut* outs, int nouts, size_t n, Call&& call) {
  size_t col_off[kBounceMaxCols], out_off[4], bytes = 0;
  for (int c = 0; c < ncols; ++c) { col_off[c] = bytes; bytes += (kBouncePods * cols[c].width + 255) & ~size_t(255); }
  for (int o = 0; o < nouts; ++o) { out_off[o] = bytes; bytes += (kBouncePods * outs[o].width + 255) & ~size_t(255); }
  if (bytes > ctx->bounce_bytes) {
    cudaFreeHost(ctx->bounce[0]); cudaFreeHost(ctx->bounce[1]);
    ctx->bounce[0] = ctx->bounce[1] = nullptr;
    ctx->bounce_bytes = 0;
    CU(cudaHostAlloc(&ctx->bounce[0], bytes, cudaHostAllocDefault));
    CU(cudaHostAlloc(&ctx->bounce[1], bytes, cudaHostAllocDefault));
    ctx->bounce_bytes = bytes;
  }
  const int T = ctx->bounce_threads;
  auto fill = [&](int slot, size_t off, size_t m) {
    // (column, slice) pairs over T threads: every thread streams a contiguous piece
    const int pieces = ncols * T;
#pragma omp parallel for num_threads(T) schedule(static)
    for (int k = 0; k < pieces; ++k) {
      const int c = k / T, part = k % T;
      const size_t total = m * cols[c].width;
      const size_t lo = total * part / T, hi = total * (part + 1) / T;
      std::memcpy(ctx->bounce[slot] + col_off[c] + lo, static_cast<const char*>(cols[c].src) + off * cols[c].width + lo, hi - lo);
    }
  };
  // fixed 4 Mi-pod pieces: smaller ones were measured slower (every piece pays a pipeline fill and the helper
  // thread's start-up), so a batch up to 4 Mi pods is one parallel copy followed by one pinned call
  const size_t piece = kBouncePods;
  fill(0, 0, n < piece ? n : piece);
  int rc = BGC_OK;
  size_t j = 0;
  for (size_t off = 0; off < n; off += piece, ++j) {
    const size_t m = n - off < piece ? n - off : piece;
    const int slot = (int)(j & 1);
    std::thread next;
    bool next_filled = true;
    if (off + piece < n) {
      const size_t off2 = off + piece, m2 = n - off2 < piece ? n - off2 : piece;
      try {  // no exception may cross the C boundary: if a thread cannot be started the copy simply is not overlapped
        next = std::thread(fill, slot ^ 1, off2, m2);
      } catch (...) {
        next_filled = false;
      }
    }
    const void* in[kBounceMaxCols];
    void* out[4];
    for (int c = 0; c < ncols; ++c) in[c] = ctx->bounce[slot] + col_off[c];
    for (int o = 0; o < nouts; ++o) out[o] = ctx->bounce[slot] + out_off[o];
    rc = call(in, out, m, j == 0);
    if (rc == BGC_OK)
      for (int o = 0; o < nouts; ++o) std::memcpy(static_cast<char*>(outs[o].dst) + off * outs[o].width, out[o], m * outs[o].width);
    if (next.joinable()) next.join();
    if (rc) break;
    if (!next_filled) {
      const size_t off2 = off + piece, m2 = n - off2 < piece ? n - off2 : piece;
      fill(slot ^ 1, off2, m2);
    }
  }
  return rc;
}

}  // namespace

// =============================================================================== C ABI
extern "C" {

const char* bgc_version(void) { return "bgc 0.5 (sm_100a)"; }

int bgc_ctx_create(int device, bgc_ctx** out) {
  if (!out) return BGC_E_BADARG;
  *out = nullptr;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) return BGC_E_CUDA;
  bgc_ctx* ctx = new (std::nothrow) bgc_ctx;
  if (!ctx) return BGC_E_OOM;
  ctx->device = device;
  cudaError_t e = cudaSetDevice(device);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->h2d_stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->d2h_stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaMalloc(&ctx->d_scratch, 256 * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaHostAlloc(&ctx->zc_host, kZeroCopyBytes, cudaHostAllocMapped);
  if (e == cudaSuccess) e = cudaHostGetDevicePointer(&ctx->zc_dev, ctx->zc_host, 0);
  if (e != cudaSuccess) {  // release whatever was created before the failure
    if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
    if (ctx->h2d_stream) cudaStreamDestroy(ctx->h2d_stream);
    if (ctx->d2h_stream) cudaStreamDestroy(ctx->d2h_stream);
    cudaFree(ctx->d_scratch);
    cudaFreeHost(ctx->zc_host);
    delete ctx;
    cudaGetLastError();
    return e == cudaErrorMemoryAllocation ? BGC_E_OOM : BGC_E_CUDA;
  }
  ctx->stream = ctx->own_stream;
  if (const char* v = std::getenv("BGC_K1_PREFETCH")) ctx->k1_prefetch = std::atoi(v);
  if (const char* v = std::getenv("BGC_K2_PREFETCH")) ctx->k2_prefetch = std::atoi(v);
  {
    const unsigned hw = std::thread::hardware_concurrency();
    ctx->bounce_threads = hw >= 16 ? 8 : hw >= 4 ? (int)hw / 2 : 1;
    if (const char* v = std::getenv("BGC_BOUNCE_THREADS")) ctx->bounce_threads = std::atoi(v) > 0 ? std::atoi(v) : 0;  // 0 disables the bounce pipeline
  }
  if (const char* v = std::getenv("BGC_ZEROCOPY_MAX_PODS")) {
    const long long want = std::atoll(v);
    ctx->zc_max_pods = want < 0 ? 0 : ((size_t)want > kZeroCopyMaxPods ? kZeroCopyMaxPods : (size_t)want);
  }
  *out = ctx;
  return BGC_OK;
}

int bgc_ctx_destroy(bgc_ctx* ctx) {
  if (!ctx) return BGC_E_BADARG;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  free_slots(ctx);
  for (auto& s : ctx->slot) {
    if (s.h2d_done) cudaEventDestroy(s.h2d_done);
    if (s.k_done) cudaEventDestroy(s.k_done);
    if (s.free_ev) cudaEventDestroy(s.free_ev);
  }
  for (auto e : ctx->ev_pool) cudaEventDestroy(e);
  cudaFree(ctx->d_used);
  cudaFree(ctx->d_hard);
  cudaFree(ctx->head);
  cudaFree(ctx->d_hard_mask);
  cudaFree(ctx->d_scratch);
  cudaFree(ctx->sh_tile_count); cudaFree(ctx->sh_tile_off);
  cudaFreeHost(ctx->bounce[0]); cudaFreeHost(ctx->bounce[1]);
  cudaFree(ctx->g_keys); cudaFree(ctx->g_keys_sorted); cudaFree(ctx->g_idx); cudaFree(ctx->g_order);
  cudaFree(ctx->g_seg); cudaFree(ctx->g_hist); cudaFree(ctx->g_mask); cudaFree(ctx->g_mask2);
  cudaFree(ctx->g_tile_base); cudaFree(ctx->g_tile_li); cudaFree(ctx->g_ticket_rec); cudaFree(ctx->g_long); cudaFree(ctx->g_chain_status); cudaFree(ctx->g_chain_agg); cudaFree(ctx->g_chain_agg2); cudaFree(ctx->g_chain_exact);
  for (auto q : ctx->g_req) cudaFree(q);
  for (auto q : ctx->g_req2) cudaFree(q);
  cudaFree(ctx->gh_ns); cudaFree(ctx->gh_mask); cudaFree(ctx->gh_v);
  for (auto q : ctx->gh_req) cudaFree(q);
  cudaFreeHost(ctx->zc_host);
  cudaStreamDestroy(ctx->own_stream);
  cudaStreamDestroy(ctx->h2d_stream);
  cudaStreamDestroy(ctx->d2h_stream);
  delete ctx;
  return BGC_OK;
}

const char* bgc_last_error(const bgc_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int bgc_ctx_set_stream(bgc_ctx* ctx, void* cuda_stream) {
  if (!ctx) return BGC_E_BADARG;
  ctx->stream = static_cast<cudaStream_t>(cuda_stream);
  return BGC_OK;
}

int bgc_ctx_reset_stream(bgc_ctx* ctx) {
  if (!ctx) return BGC_E_BADARG;
  ctx->stream = ctx->own_stream;
  return BGC_OK;
}

int bgc_ctx_sync(bgc_ctx* ctx) {
  if (!ctx) return BGC_E_BADARG;
  CU(cudaSetDevice(ctx->device));
  CU(cudaStreamSynchronize(ctx->stream));
  CU(cudaStreamSynchronize(ctx->h2d_stream));
  CU(cudaStreamSynchronize(ctx->d2h_stream));
  return BGC_OK;
}

int bgc_timings_get(bgc_ctx* ctx, bgc_timings* out) {
  if (!ctx || !out) return BGC_E_BADARG;
  *out = ctx->tm;
  return BGC_OK;
}

int bgc_ctx_set_k1_variant(bgc_ctx* ctx, int variant) {
  if (!ctx) return BGC_E_BADARG;
  NEED((variant & 0xFF) >= BGC_K1_AUTO && (variant & 0xFF) <= BGC_K1_CLUSTER16, "unknown K1 variant");
  ctx->k1_forced = variant & 0xFF;
  ctx->k1_shape = (variant >> 8) & 0xFF;  // bits 8..15: scatter launch shape (development knob)
  if ((variant >> 16) & 0xFF) ctx->k1_prefetch = ((variant >> 16) & 0xFF) - 1;  // bits 16..23: prefetch distance + 1
  if ((variant >> 24) & 0x7F) ctx->k2_prefetch = ((variant >> 24) & 0x7F) - 1;  // bits 24..30: K2 prefetch distance + 1
  return BGC_OK;
}

const char* bgc_last_k1_variant(const bgc_ctx* ctx) { return ctx ? ctx->last_k1 : "none"; }

// ---------------------------------------------------------------- K0
int bgc_policy_eval_dev(bgc_ctx* ctx, const uint8_t* d_op, const uint16_t* d_flags, size_t n,
                        uint8_t* d_code, uint8_t* d_patch_mask) {
  if (!ctx) return BGC_E_BADARG;
  NEED(n == 0 || (d_op && d_flags && d_code && d_patch_mask), "null column pointer");
  CU(cudaSetDevice(ctx->device));
  return launch_k0(ctx, ctx->stream, d_op, d_flags, n, d_code, d_patch_mask);
}

int bgc_policy_eval(bgc_ctx* ctx, const uint8_t* op, const uint16_t* flags, size_t n, uint8_t* code,
                    uint8_t* patch_mask) {
  if (!ctx) return BGC_E_BADARG;
  NEED(n == 0 || (op && flags && code && patch_mask), "null column pointer");
  CU(cudaSetDevice(ctx->device));
  ctx->tm.h2d_ns = ctx->tm.kernel_ns = ctx->tm.d2h_ns = 0;
  if (ctx->stream != ctx->own_stream) CU(cudaStreamSynchronize(ctx->stream));
  if (n == 0) return BGC_OK;
  if (wants_bounce(ctx, op, n)) {  // pageable caller: pinned bounce ring, copies overlapped with the pipeline
    const BounceCol cols[2] = {{op, 1}, {flags, 2}};
    const BounceOut outs[2] = {{code, 1}, {patch_mask, 1}};
    return bounced(ctx, cols, 2, outs, 2, n, [&](const void* const* in, void* const* out, size_t m, bool) {
      return bgc_policy_eval(ctx, static_cast<const uint8_t*>(in[0]), static_cast<const uint16_t*>(in[1]), m,
                             static_cast<uint8_t*>(out[0]), static_cast<uint8_t*>(out[1]));
    });
  }
  if (n <= ctx->zc_max_pods) {  // latency path: one launch, no staged copies
    ZcCursor cur;
    uint8_t *b0, *b1;
    const uint8_t* z_op = static_cast<const uint8_t*>(zc_in(ctx, cur, op, n));
    const uint16_t* z_fl = static_cast<const uint16_t*>(zc_in(ctx, cur, flags, n * 2));
    uint8_t* z_c = static_cast<uint8_t*>(zc_out(ctx, cur, code, n, &b0));
    uint8_t* z_p = static_cast<uint8_t*>(zc_out(ctx, cur, patch_mask, n, &b1));
    if (int rc0 = launch_k0(ctx, ctx->own_stream, z_op, z_fl, n, z_c, z_p)) return rc0;
    CU(cudaStreamSynchronize(ctx->own_stream));
    if (b0) std::memcpy(code, b0, n);
    if (b1) std::memcpy(patch_mask, b1, n);
    return BGC_OK;
  }
  int rc = ensure_slots(ctx, n, ctx->slot_R);
  if (rc) return rc;
  ctx->ev_next = 0;
  Spans sp;
  const size_t chunk = ctx->slot_pods;
  const bool single = n <= chunk;  // one chunk: copy-in, kernel, copy-out on one stream, no hand-offs
  cudaStream_t hs = single ? ctx->own_stream : ctx->h2d_stream;
  cudaStream_t ds = single ? ctx->own_stream : ctx->d2h_stream;
  (void)ds;
  size_t k = 0;
  for (size_t off = 0; off < n; off += chunk, ++k) {
    Slot& s = ctx->slot[k % kSlots];
    const size_t m = (n - off < chunk) ? n - off : chunk;
    cudaEvent_t a = next_event(ctx), b = next_event(ctx), c = next_event(ctx), d = next_event(ctx),
                e = next_event(ctx), f = next_event(ctx);
    if (!f) return fail(ctx, BGC_E_CUDA, "event pool exhausted");
    if (s.used && !single) CU(cudaStreamWaitEvent(hs, s.free_ev, 0));
    CU(cudaEventRecord(a, hs));
    CU(cudaMemcpyAsync(s.mask, op + off, m, cudaMemcpyHostToDevice, hs));  // op column staged in `mask`
    CU(cudaMemcpyAsync(s.flags, flags + off, m * 2, cudaMemcpyHostToDevice, hs));
    CU(cudaEventRecord(b, hs));
    if (!single) {
      CU(cudaEventRecord(s.h2d_done, hs));
      CU(cudaStreamWaitEvent(ctx->own_stream, s.h2d_done, 0));
    }
    CU(cudaEventRecord(c, ctx->own_stream));
    rc = launch_k0(ctx, ctx->own_stream, s.mask, s.flags, m, s.out0, s.out1);
    if (rc) return rc;
    CU(cudaEventRecord(d, ctx->own_stream));
    if (!single) {
      CU(cudaEventRecord(s.k_done, ctx->own_stream));
      CU(cudaStreamWaitEvent(ds, s.k_done, 0));
    }
    CU(cudaEventRecord(e, ds));
    CU(cudaMemcpyAsync(code + off, s.out0, m, cudaMemcpyDeviceToHost, ds));
    CU(cudaMemcpyAsync(patch_mask + off, s.out1, m, cudaMemcpyDeviceToHost, ds));
    CU(cudaEventRecord(f, ds));
    if (!single) CU(cudaEventRecord(s.free_ev, ds));
    s.used = !single;
    sp.h2d.push_back(a); sp.h2d.push_back(b);
    sp.k.push_back(c); sp.k.push_back(d);
    sp.d2h.push_back(e); sp.d2h.push_back(f);
  }
  CU(cudaStreamSynchronize(ds));
  return finish_timings(ctx, sp);
}

// ---------------------------------------------------------------- quota table
int bgc_quota_table_set(bgc_ctx* ctx, uint32_t ns_count, uint32_t R, const int64_t* hard,
                        const uint8_t* hard_mask, const uint8_t* unit) {
  if (!ctx) return BGC_E_BADARG;
  NEED(ns_count > 0 && ns_count < 0xFFFFFFFFu, "ns_count out of range");
  NEED(R >= 1 && R <= BGC_MAX_R, "R must be 1..BGC_MAX_R");
  NEED(hard && hard_mask && unit, "null table pointer");
  // validate everything before touching the context, so that a refused call leaves it as it was
  uint32_t compute_mask = 0, wide_mask = 0;
  for (uint32_t r = 0; r < R; ++r) {
    NEED(unit[r] <= BGC_UNIT_BYTES, "unknown unit");
    if (unit[r] != BGC_UNIT_COUNT) compute_mask |= 1u << r;
    if (unit[r] == BGC_UNIT_BYTES) wide_mask |= 1u << r;
  }
  CU(cudaSetDevice(ctx->device));
  CU(cudaDeviceSynchronize());
  // any failure from here on leaves the context WITHOUT a table (later calls return BGC_E_STATE)
  auto drop_tables = [&] {
    cudaFree(ctx->d_used); ctx->d_used = nullptr;
    cudaFree(ctx->d_hard); ctx->d_hard = nullptr;
    cudaFree(ctx->head); ctx->head = nullptr;
    cudaFree(ctx->d_hard_mask); ctx->d_hard_mask = nullptr;
    ctx->NS = 0; ctx->R = 0;
    ctx->head_stale = true;
  };
  auto step = [&](cudaError_t e, const char* what) -> int {
    if (e == cudaSuccess) return BGC_OK;
    drop_tables();
    cudaGetLastError();
    return fail(ctx, e == cudaErrorMemoryAllocation ? BGC_E_OOM : BGC_E_CUDA, "%s: %s", what, cudaGetErrorString(e));
  };
  int rc;
  if (ns_count != ctx->NS || (int)R != ctx->R) {
    drop_tables();
    if ((rc = step(cudaMalloc(&ctx->d_used, (size_t)ns_count * R * 8), "cudaMalloc(used)"))) return rc;
    if ((rc = step(cudaMalloc(&ctx->d_hard, (size_t)ns_count * R * 8), "cudaMalloc(hard)"))) return rc;
    if ((rc = step(cudaMalloc(&ctx->head, (size_t)ns_count * head_words((int)R) * 8), "cudaMalloc(headroom)"))) return rc;
    if ((rc = step(cudaMalloc(&ctx->d_hard_mask, ns_count), "cudaMalloc(hard_mask)"))) return rc;
  }
  if ((rc = step(cudaMemcpyAsync(ctx->d_hard, hard, (size_t)ns_count * R * 8, cudaMemcpyHostToDevice, ctx->own_stream), "copy(hard)"))) return rc;
  if ((rc = step(cudaMemcpyAsync(ctx->d_hard_mask, hard_mask, ns_count, cudaMemcpyHostToDevice, ctx->own_stream), "copy(hard_mask)"))) return rc;
  if ((rc = step(cudaMemsetAsync(ctx->d_used, 0, (size_t)ns_count * R * 8, ctx->own_stream), "clear(used)"))) return rc;
  if ((rc = step(cudaStreamSynchronize(ctx->own_stream), "sync"))) return rc;
  // commit
  ctx->NS = ns_count;
  ctx->R = (int)R;
  ctx->compute_mask = compute_mask;
  ctx->wide_mask = wide_mask;
  ctx->head_stale = true;
  return BGC_OK;
}

int bgc_usage_get(bgc_ctx* ctx, int64_t* used_out) {
  if (!ctx) return BGC_E_BADARG;
  if (!ctx->d_used) return fail(ctx, BGC_E_STATE, "quota table not set");
  NEED(used_out, "null output");
  CU(cudaSetDevice(ctx->device));
  CU(cudaStreamSynchronize(ctx->stream));
  CU(cudaMemcpyAsync(used_out, ctx->d_used, (size_t)ctx->NS * ctx->R * 8, cudaMemcpyDeviceToHost, ctx->own_stream));
  CU(cudaStreamSynchronize(ctx->own_stream));
  return BGC_OK;
}

int bgc_usage_set(bgc_ctx* ctx, const int64_t* used) {
  if (!ctx) return BGC_E_BADARG;
  if (!ctx->d_used) return fail(ctx, BGC_E_STATE, "quota table not set");
  NEED(used, "null input");
  CU(cudaSetDevice(ctx->device));
  CU(cudaStreamSynchronize(ctx->stream));
  CU(cudaMemcpyAsync(ctx->d_used, used, (size_t)ctx->NS * ctx->R * 8, cudaMemcpyHostToDevice, ctx->own_stream));
  ctx->head_stale = true;
  CU(cudaStreamSynchronize(ctx->own_stream));
  return BGC_OK;
}

// device-pointer forms, enqueued on the context's stream (row N5: the usage table is what the
// index-sharded multi-GPU variant all-reduces between K1 and K2)
int bgc_usage_get_dev(bgc_ctx* ctx, int64_t* d_used_out) {
  if (!ctx) return BGC_E_BADARG;
  if (!ctx->d_used) return fail(ctx, BGC_E_STATE, "quota table not set");
  NEED(d_used_out, "null output");
  CU(cudaSetDevice(ctx->device));
  return export_used(ctx, ctx->stream, reinterpret_cast<long long*>(d_used_out));
}

int bgc_usage_set_dev(bgc_ctx* ctx, const int64_t* d_used) {
  if (!ctx) return BGC_E_BADARG;
  if (!ctx->d_used) return fail(ctx, BGC_E_STATE, "quota table not set");
  NEED(d_used, "null input");
  CU(cudaSetDevice(ctx->device));
  CU(cudaMemcpyAsync(ctx->d_used, d_used, (size_t)ctx->NS * ctx->R * 8, cudaMemcpyDeviceToDevice, ctx->stream));
  ctx->head_stale = true;
  return BGC_OK;
}

// ---------------------------------------------------------------- K1
int bgc_usage_reduce_dev(bgc_ctx* ctx, const uint32_t* d_ns_id, const int64_t* const* d_req, size_t n,
                         unsigned flags, int64_t* d_used_out) {
  if (!ctx) return BGC_E_BADARG;
  if (!ctx->d_used) return fail(ctx, BGC_E_STATE, "quota table not set");
  NEED(n == 0 || (d_ns_id && d_req), "null column pointer");
  const long long* cols[kMaxR] = {};
  for (int r = 0; r < ctx->R && n; ++r) {
    NEED(d_req[r], "null request column");
    cols[r] = reinterpret_cast<const long long*>(d_req[r]);
  }
  CU(cudaSetDevice(ctx->device));
  int rc;
  if (!(flags & BGC_REDUCE_ACCUMULATE) && (rc = clear_used(ctx, ctx->stream))) return rc;
  if ((rc = launch_k1(ctx, ctx->stream, d_ns_id, cols, n, flags))) return rc;
  if (d_used_out) return export_used(ctx, ctx->stream, reinterpret_cast<long long*>(d_used_out));
  return BGC_OK;
}

int bgc_usage_reduce(bgc_ctx* ctx, const uint32_t* ns_id, const int64_t* const* req, size_t n,
                     unsigned flags, int64_t* used_out) {
  if (!ctx) return BGC_E_BADARG;
  if (!ctx->d_used) return fail(ctx, BGC_E_STATE, "quota table not set");
  NEED(n == 0 || (ns_id && req), "null column pointer");
  const int R = ctx->R;
  for (int r = 0; r < R && n; ++r) NEED(req[r], "null request column");
  CU(cudaSetDevice(ctx->device));
  ctx->tm.h2d_ns = ctx->tm.kernel_ns = ctx->tm.d2h_ns = 0;
  if (ctx->stream != ctx->own_stream) CU(cudaStreamSynchronize(ctx->stream));
  int rc;
  if (wants_bounce(ctx, ns_id, n)) {  // pageable caller: pinned bounce ring; pieces after the first accumulate
    BounceCol cols[kBounceMaxCols];
    cols[0] = {ns_id, 4};
    for (int r = 0; r < R; ++r) cols[1 + r] = {req[r], 8};
    rc = bounced(ctx, cols, 1 + R, nullptr, 0, n, [&](const void* const* in, void* const*, size_t m, bool first) {
      const int64_t* rq[kMaxR];
      for (int r = 0; r < R; ++r) rq[r] = static_cast<const int64_t*>(in[1 + r]);
      return bgc_usage_reduce(ctx, static_cast<const uint32_t*>(in[0]), rq, m, first ? flags : (flags | BGC_REDUCE_ACCUMULATE), nullptr);
    });
    if (rc) return rc;
    return used_out ? bgc_usage_get(ctx, used_out) : BGC_OK;
  }
  if (!(flags & BGC_REDUCE_ACCUMULATE) && (rc = clear_used(ctx, ctx->own_stream))) return rc;
  // The columns are in host memory here, so the grouping hint can be inferred for free: 1,024
  // evenly spaced neighbour pairs; if nearly all of them share a namespace the segmented kernel is
  // the right one.  Either kernel is exact for any input — this only picks the faster.
  if (!(flags & BGC_REDUCE_GROUPED) && ctx->k1_forced == BGC_K1_AUTO && n >= 4096) {
    const size_t step = (n - 1) / 1024;
    size_t same = 0;
    for (size_t k = 0; k < 1024; ++k) same += ns_id[k * step] == ns_id[k * step + 1];
    if (same >= 1000) flags |= BGC_REDUCE_GROUPED;
  }
  Spans sp;
  if (n) {
    if ((rc = ensure_slots(ctx, n, R))) return rc;
    ctx->ev_next = 0;
    const size_t chunk = ctx->slot_pods;
    const bool single = n <= chunk;
    cudaStream_t hs = single ? ctx->own_stream : ctx->h2d_stream;
    size_t k = 0;
    for (size_t off = 0; off < n; off += chunk, ++k) {
      Slot& s = ctx->slot[k % kSlots];
      const size_t m = (n - off < chunk) ? n - off : chunk;
      cudaEvent_t a = next_event(ctx), b = next_event(ctx), c = next_event(ctx), d = next_event(ctx);
      if (!d) return fail(ctx, BGC_E_CUDA, "event pool exhausted");
      if (s.used && !single) CU(cudaStreamWaitEvent(hs, s.free_ev, 0));
      CU(cudaEventRecord(a, hs));
      CU(cudaMemcpyAsync(s.ns, ns_id + off, m * 4, cudaMemcpyHostToDevice, hs));
      for (int r = 0; r < R; ++r)
        CU(cudaMemcpyAsync(s.req[r], req[r] + off, m * 8, cudaMemcpyHostToDevice, hs));
      CU(cudaEventRecord(b, hs));
      if (!single) {
        CU(cudaEventRecord(s.h2d_done, hs));
        CU(cudaStreamWaitEvent(ctx->own_stream, s.h2d_done, 0));
      }
      CU(cudaEventRecord(c, ctx->own_stream));
      if ((rc = launch_k1(ctx, ctx->own_stream, s.ns, s.req, m, flags))) return rc;
      CU(cudaEventRecord(d, ctx->own_stream));
      if (!single) CU(cudaEventRecord(s.free_ev, ctx->own_stream));
      s.used = !single;
      sp.h2d.push_back(a); sp.h2d.push_back(b);
      sp.k.push_back(c); sp.k.push_back(d);
    }
  }
  if (used_out) {
    cudaEvent_t e = next_event(ctx), f = next_event(ctx);
    if (!f) return fail(ctx, BGC_E_CUDA, "event pool exhausted");
    CU(cudaEventRecord(e, ctx->own_stream));
    CU(cudaMemcpyAsync(used_out, ctx->d_used, (size_t)ctx->NS * R * 8, cudaMemcpyDeviceToHost, ctx->own_stream));
    CU(cudaEventRecord(f, ctx->own_stream));
    sp.d2h.push_back(e); sp.d2h.push_back(f);
  }
  CU(cudaStreamSynchronize(ctx->own_stream));
  return finish_timings(ctx, sp);
}

// ---------------------------------------------------------------- K2
}  // extern "C"

namespace {

template <typename VT>
int quota_eval_dev_impl(bgc_ctx* ctx, const uint32_t* d_ns_id, const int64_t* const* d_req,
                        const uint8_t* d_req_mask, size_t n, VT* d_verdict) {
  if (!ctx) return BGC_E_BADARG;
  if (!ctx->d_used) return fail(ctx, BGC_E_STATE, "quota table not set");
  NEED(n == 0 || (d_ns_id && d_req && d_verdict), "null column pointer");
  const long long* cols[kMaxR] = {};
  for (int r = 0; r < ctx->R && n; ++r) {
    NEED(d_req[r], "null request column");
    cols[r] = reinterpret_cast<const long long*>(d_req[r]);
  }
  CU(cudaSetDevice(ctx->device));
  return launch_k2(ctx, ctx->stream, d_ns_id, cols, d_req_mask, n, d_verdict);
}

template <typename VT>
int quota_eval_host_impl(bgc_ctx* ctx, const uint32_t* ns_id, const int64_t* const* req,
                         const uint8_t* req_mask, size_t n, VT* verdict) {
  if (!ctx) return BGC_E_BADARG;
  if (!ctx->d_used) return fail(ctx, BGC_E_STATE, "quota table not set");
  NEED(n == 0 || (ns_id && req && verdict), "null column pointer");
  const int R = ctx->R;
  for (int r = 0; r < R && n; ++r) NEED(req[r], "null request column");
  CU(cudaSetDevice(ctx->device));
  ctx->tm.h2d_ns = ctx->tm.kernel_ns = ctx->tm.d2h_ns = 0;
  if (ctx->stream != ctx->own_stream) CU(cudaStreamSynchronize(ctx->stream));
  if (n == 0) return BGC_OK;
  int rc;
  if (wants_bounce(ctx, ns_id, n)) {  // pageable caller: pinned bounce ring
    BounceCol cols[kBounceMaxCols];
    int nc = 0;
    cols[nc++] = {ns_id, 4};
    for (int r = 0; r < R; ++r) cols[nc++] = {req[r], 8};
    if (req_mask) cols[nc++] = {req_mask, 1};
    const BounceOut outs[1] = {{verdict, sizeof(VT)}};
    return bounced(ctx, cols, nc, outs, 1, n, [&](const void* const* in, void* const* out, size_t m, bool) {
      const int64_t* rq[kMaxR];
      for (int r = 0; r < R; ++r) rq[r] = static_cast<const int64_t*>(in[1 + r]);
      return quota_eval_host_impl<VT>(ctx, static_cast<const uint32_t*>(in[0]), rq,
                                      req_mask ? static_cast<const uint8_t*>(in[1 + R]) : nullptr, m, static_cast<VT*>(out[0]));
    });
  }
  if (n <= ctx->zc_max_pods) {  // latency path: one launch, the kernel reads/writes host memory itself
    ZcCursor cur;
    uint8_t* bounce;
    const long long* cols[kMaxR] = {};
    const uint32_t* z_ns = static_cast<const uint32_t*>(zc_in(ctx, cur, ns_id, n * 4));
    for (int r = 0; r < R; ++r) cols[r] = static_cast<const long long*>(zc_in(ctx, cur, req[r], n * 8));
    const uint8_t* z_mask = req_mask ? static_cast<const uint8_t*>(zc_in(ctx, cur, req_mask, n)) : nullptr;
    VT* z_v = static_cast<VT*>(zc_out(ctx, cur, verdict, n * sizeof(VT), &bounce));
    if ((rc = launch_k2(ctx, ctx->own_stream, z_ns, cols, z_mask, n, z_v))) return rc;
    CU(cudaStreamSynchronize(ctx->own_stream));
    if (bounce) std::memcpy(verdict, bounce, n * sizeof(VT));
    return BGC_OK;
  }
  if ((rc = ensure_slots(ctx, n, R))) return rc;
  ctx->ev_next = 0;
  Spans sp;
  const size_t chunk = ctx->slot_pods;
  const bool single = n <= chunk;  // one chunk: copy-in, kernel, copy-out on one stream, no hand-offs
  cudaStream_t hs = single ? ctx->own_stream : ctx->h2d_stream;
  cudaStream_t ds = single ? ctx->own_stream : ctx->d2h_stream;
  (void)ds;
  size_t k = 0;
  for (size_t off = 0; off < n; off += chunk, ++k) {
    Slot& s = ctx->slot[k % kSlots];
    const size_t m = (n - off < chunk) ? n - off : chunk;
    cudaEvent_t a = next_event(ctx), b = next_event(ctx), c = next_event(ctx), d = next_event(ctx),
                e = next_event(ctx), f = next_event(ctx);
    if (!f) return fail(ctx, BGC_E_CUDA, "event pool exhausted");
    if (s.used && !single) CU(cudaStreamWaitEvent(hs, s.free_ev, 0));
    CU(cudaEventRecord(a, hs));
    CU(cudaMemcpyAsync(s.ns, ns_id + off, m * 4, cudaMemcpyHostToDevice, hs));
    for (int r = 0; r < R; ++r)
      CU(cudaMemcpyAsync(s.req[r], req[r] + off, m * 8, cudaMemcpyHostToDevice, hs));
    if (req_mask) CU(cudaMemcpyAsync(s.mask, req_mask + off, m, cudaMemcpyHostToDevice, hs));
    CU(cudaEventRecord(b, hs));
    if (!single) {
      CU(cudaEventRecord(s.h2d_done, hs));
      CU(cudaStreamWaitEvent(ctx->own_stream, s.h2d_done, 0));
    }
    CU(cudaEventRecord(c, ctx->own_stream));
    if ((rc = launch_k2(ctx, ctx->own_stream, s.ns, s.req, req_mask ? s.mask : nullptr, m, reinterpret_cast<VT*>(s.out0)))) return rc;
    CU(cudaEventRecord(d, ctx->own_stream));
    if (!single) {
      CU(cudaEventRecord(s.k_done, ctx->own_stream));
      CU(cudaStreamWaitEvent(ds, s.k_done, 0));
    }
    CU(cudaEventRecord(e, ds));
    CU(cudaMemcpyAsync(verdict + off, s.out0, m * sizeof(VT), cudaMemcpyDeviceToHost, ds));
    CU(cudaEventRecord(f, ds));
    if (!single) CU(cudaEventRecord(s.free_ev, ds));
    s.used = !single;
    sp.h2d.push_back(a); sp.h2d.push_back(b);
    sp.k.push_back(c); sp.k.push_back(d);
    sp.d2h.push_back(e); sp.d2h.push_back(f);
  }
  CU(cudaStreamSynchronize(ds));
  return finish_timings(ctx, sp);
}

}  // namespace

extern "C" {

int bgc_quota_eval_dev(bgc_ctx* ctx, const uint32_t* d_ns_id, const int64_t* const* d_req,
                       const uint8_t* d_req_mask, size_t n, uint8_t* d_verdict) {
  return quota_eval_dev_impl<uint8_t>(ctx, d_ns_id, d_req, d_req_mask, n, d_verdict);
}
int bgc_quota_eval16_dev(bgc_ctx* ctx, const uint32_t* d_ns_id, const int64_t* const* d_req,
                         const uint8_t* d_req_mask, size_t n, uint16_t* d_verdict) {
  return quota_eval_dev_impl<uint16_t>(ctx, d_ns_id, d_req, d_req_mask, n, d_verdict);
}
int bgc_quota_eval(bgc_ctx* ctx, const uint32_t* ns_id, const int64_t* const* req, const uint8_t* req_mask,
                   size_t n, uint8_t* verdict) {
  return quota_eval_host_impl<uint8_t>(ctx, ns_id, req, req_mask, n, verdict);
}
int bgc_quota_eval16(bgc_ctx* ctx, const uint32_t* ns_id, const int64_t* const* req, const uint8_t* req_mask,
                     size_t n, uint16_t* verdict) {
  return quota_eval_host_impl<uint16_t>(ctx, ns_id, req, req_mask, n, verdict);
}

}  // extern "C"

// ---------------------------------------------------------------- K3 (row N1): greedy in-batch admission
namespace {

// tiles of 1024 pods over the segments of kGreedyTeamMin pods or more: each has at most len / 1024 + 1
inline size_t chain_tile_bound(size_t n) { return (n / kGreedyTeam + n / kGreedyTeamMin + 3) & ~size_t(1); }  // even: 8-byte aligned counters behind it

template <int R, typename VT>
int launch_k3_r(bgc_ctx* ctx, cudaStream_t st, size_t n, const uint8_t* mask_in, VT* verdict) {
  // the grouping sort already left the request columns (and masks) in grouped order in the context's buffers
  ReqCols<R> cols;
  for (int r = 0; r < R; ++r) cols.p[r] = ctx->g_req[r];
  const uint8_t* mask = mask_in ? ctx->g_mask : nullptr;
  int grid = ctx->sm_count * 8;
  const int warps_per_block = 256 / 32;
  if ((size_t)grid * warps_per_block > ctx->NS) grid = (int)((ctx->NS + warps_per_block - 1) / warps_per_block);
  // short segments: one warp per namespace; long ones (Zipf heads): one 1024-thread CTA per namespace
  int team_grid = ctx->sm_count;
  if ((uint32_t)team_grid > ctx->NS) team_grid = (int)ctx->NS;
  const size_t ring = greedy_team_smem(R);
  if (ctx->k3_long < 0) {
    const char* e = std::getenv("BGC_K3_LONG");
    ctx->k3_long = e && !std::strcmp(e, "team") ? 1 : 0;
  }
  auto go = [&](auto warp_kern, auto team_kern, auto chain_kern) -> int {
    warp_kern<<<grid, 256, 0, st>>>(ctx->g_order, ctx->g_seg, cols, mask, ctx->d_used, ctx->d_hard, ctx->d_hard_mask, ctx->NS,
                                    ctx->compute_mask, verdict, kGreedyTeamMin);
    ctx->tm.launches++;
    CU(cudaGetLastError());
    if (n < kGreedyTeamMin) return BGC_OK;  // no segment can be long
    if (ctx->k3_long == 1) {  // one CTA per long segment (the round-2 form before the chained tiles; kept for comparison)
      CU(cudaFuncSetAttribute(team_kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ring));
      team_kern<<<team_grid, kGreedyTeam, ring, st>>>(ctx->g_order, ctx->g_seg, cols, mask, ctx->d_used, ctx->d_hard, ctx->d_hard_mask,
                                                      ctx->NS, ctx->compute_mask, verdict);
      ctx->tm.launches++;
      CU(cudaGetLastError());
      return BGC_OK;
    }
    // long segments as chained tiles over all SMs
    const size_t cnt = (size_t)ctx->NS + 1;
    const unsigned blocks = (unsigned)((cnt + kScanBlock - 1) / kScanBlock);
    uint32_t* totals = ctx->g_tile_base + cnt;
    const size_t bound = chain_tile_bound(n);
    ChainBufs C{ctx->g_tile_base, ctx->g_long, ctx->g_long + bound, ctx->g_tile_li, reinterpret_cast<uint4*>(ctx->g_ticket_rec), ctx->g_chain_status,
                ctx->g_chain_agg, ctx->g_chain_agg2, ctx->g_chain_exact, ctx->g_chain_status + bound};
    CU(cudaMemsetAsync(ctx->g_chain_status, 0, (bound + 16) * 4, st));
    k_chain_count<<<grid_for(cnt, 256, ctx->sm_count * 4), 256, 0, st>>>(ctx->g_seg, ctx->NS, kGreedyTeamMin, C);
    k_scan_blocks<<<blocks, 1024, 0, st>>>(ctx->g_tile_base, cnt, totals);
    k_scan_u32<<<1, 1024, 0, st>>>(totals, blocks);
    k_scan_add<<<blocks, 1024, 0, st>>>(ctx->g_tile_base, cnt, totals);
    k_chain_fill<<<ctx->sm_count, 256, 0, st>>>(C);
    k_chain_order<<<ctx->sm_count * 4, 256, 0, st>>>(ctx->g_seg, ctx->NS, C);
    const size_t smem = (size_t)R * kGreedyTeam * 8;
    CU(cudaFuncSetAttribute(chain_kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    chain_kern<<<ctx->sm_count, kGreedyTeam, smem, st>>>(ctx->g_order, ctx->g_seg, cols, mask, ctx->d_used, ctx->d_hard, ctx->d_hard_mask,
                                                         ctx->NS, ctx->compute_mask, verdict, C);
    ctx->tm.launches += 7;
    CU(cudaGetLastError());
    if (std::getenv("BGC_K3_STATS")) {  // debugging aid: synchronises
      uint32_t h[16];
      CU(cudaStreamSynchronize(st));
      CU(cudaMemcpy(h, ctx->g_chain_status + bound, sizeof h, cudaMemcpyDeviceToHost));
#ifdef BGC_K3_CLOCKS
      {
        unsigned long long ck[3];
        std::memcpy(ck, h + 8, sizeof ck);
        std::fprintf(stderr, "[bgc k3 chain] cycles per tile: load+sum %.0f, wait for S_in %.0f, walk %.0f\n", (double)ck[0] / (h[0] ? h[0] : 1),
                     (double)ck[1] / (h[0] ? h[0] : 1), (double)ck[2] / (h[0] ? h[0] : 1));
      }
#endif
      std::fprintf(stderr, "[bgc k3 chain] tickets %u: first-of-segment %u, IDENT %u, AGG2 fitted %u, AGG fitted %u, walked %u; long segments %u\n", h[0], h[1], h[2], h[3], h[4], h[5], h[6]);
    }
    return BGC_OK;
  };
  return mask ? go(k3_admit_greedy<R, true, VT>, k3_admit_greedy_team<R, true, VT>, k3_admit_greedy_chain<R, true, VT>)
              : go(k3_admit_greedy<R, false, VT>, k3_admit_greedy_team<R, false, VT>, k3_admit_greedy_chain<R, false, VT>);
}

int ensure_greedy_buffers(bgc_ctx* ctx, size_t n) {
  if (n > ctx->g_cap || ctx->R > ctx->g_R) {
    CU(cudaDeviceSynchronize());
    cudaFree(ctx->g_keys); cudaFree(ctx->g_keys_sorted); cudaFree(ctx->g_idx); cudaFree(ctx->g_order); cudaFree(ctx->g_mask); cudaFree(ctx->g_mask2);
    for (auto& q : ctx->g_req) { cudaFree(q); q = nullptr; }
    for (auto& q : ctx->g_req2) { cudaFree(q); q = nullptr; }
    ctx->g_mask2 = nullptr;
    ctx->g_keys = ctx->g_keys_sorted = ctx->g_idx = ctx->g_order = nullptr;
    ctx->g_mask = nullptr;
    if (n < ctx->g_cap) n = ctx->g_cap;
    ctx->g_cap = 0;
    ctx->g_R = 0;
    CU(cudaMalloc(&ctx->g_keys, n * 4));
    CU(cudaMalloc(&ctx->g_keys_sorted, n * 4));
    CU(cudaMalloc(&ctx->g_idx, n * 4));
    CU(cudaMalloc(&ctx->g_order, n * 4));
    CU(cudaMalloc(&ctx->g_mask, n));
    CU(cudaMalloc(&ctx->g_mask2, n));
    for (int r = 0; r < ctx->R; ++r) CU(cudaMalloc(&ctx->g_req[r], n * 8));
    for (int r = 0; r < ctx->R; ++r) CU(cudaMalloc(&ctx->g_req2[r], n * 8));
    ctx->g_cap = n;
    ctx->g_R = ctx->R;
  }
  if ((size_t)ctx->NS + 2 > ctx->g_seg_cap) {
    CU(cudaDeviceSynchronize());
    cudaFree(ctx->g_seg); ctx->g_seg = nullptr; ctx->g_seg_cap = 0;
    CU(cudaMalloc(&ctx->g_seg, ((size_t)ctx->NS + 2) * 4));
    ctx->g_seg_cap = (size_t)ctx->NS + 2;
  }
  {
    const size_t base_need = (size_t)ctx->NS + 1 + ((size_t)ctx->NS + 1 + kScanBlock - 1) / kScanBlock + 1;  // + the scan's block totals
    if (base_need > ctx->g_tile_base_cap) {
      CU(cudaDeviceSynchronize());
      cudaFree(ctx->g_tile_base); ctx->g_tile_base = nullptr; ctx->g_tile_base_cap = 0;
      CU(cudaMalloc(&ctx->g_tile_base, base_need * 4));
      ctx->g_tile_base_cap = base_need;
    }
    const size_t tiles = chain_tile_bound(n);
    if (tiles > ctx->g_chain_cap || ctx->R > ctx->g_chain_R) {
      CU(cudaDeviceSynchronize());
      cudaFree(ctx->g_tile_li); cudaFree(ctx->g_ticket_rec); cudaFree(ctx->g_long); cudaFree(ctx->g_chain_status); cudaFree(ctx->g_chain_agg); cudaFree(ctx->g_chain_agg2); cudaFree(ctx->g_chain_exact);
      ctx->g_tile_li = ctx->g_ticket_rec = ctx->g_long = ctx->g_chain_status = nullptr; ctx->g_chain_agg = ctx->g_chain_agg2 = ctx->g_chain_exact = nullptr;
      const size_t cap = tiles > ctx->g_chain_cap ? tiles : ctx->g_chain_cap;
      const int RR = ctx->R > ctx->g_chain_R ? ctx->R : ctx->g_chain_R;
      ctx->g_chain_cap = 0; ctx->g_chain_R = 0;
      CU(cudaMalloc(&ctx->g_tile_li, cap * 4));
      CU(cudaMalloc(&ctx->g_ticket_rec, cap * 16));
      CU(cudaMalloc(&ctx->g_long, cap * 2 * 4));  // namespaces and tile counts of the long segments (fewer than tiles)
      CU(cudaMalloc(&ctx->g_chain_status, (cap + 16) * 4));  // + the ticket counter and the statistics
      CU(cudaMalloc(&ctx->g_chain_agg, cap * RR * 8));
      CU(cudaMalloc(&ctx->g_chain_agg2, cap * RR * 8));
      CU(cudaMalloc(&ctx->g_chain_exact, cap * RR * 8));
      ctx->g_chain_cap = cap; ctx->g_chain_R = RR;
    }
  }
  size_t hist_need = (size_t)256 * ((n + kSortTile - 1) / kSortTile);
  hist_need += (hist_need + kScanBlock - 1) / kScanBlock + 1;  // + the block totals of the scan
  if (hist_need > ctx->g_hist_cap) {
    CU(cudaDeviceSynchronize());
    cudaFree(ctx->g_hist); ctx->g_hist = nullptr; ctx->g_hist_cap = 0;
    CU(cudaMalloc(&ctx->g_hist, hist_need * 4));
    ctx->g_hist_cap = hist_need;
  }
  return BGC_OK;
}

}  // namespace

namespace {
template <typename VT>
int admit_greedy_dev_impl(bgc_ctx* ctx, const uint32_t* d_ns_id, const int64_t* const* d_req,
                          const uint8_t* d_req_mask, size_t n, VT* d_verdict) {
  if (!ctx) return BGC_E_BADARG;
  if (!ctx->d_used) return fail(ctx, BGC_E_STATE, "quota table not set");
  NEED(n == 0 || (d_ns_id && d_req && d_verdict), "null column pointer");
  NEED(n < (size_t(1) << 31), "greedy admission handles fewer than 2^31 pods per call");
  NEED(sizeof(VT) == 2 || ctx->R <= kMaxR8, "u8 verdicts hold at most 4 resources: use bgc_admit_greedy16 for more");
  const long long* cols[kMaxR] = {};
  for (int r = 0; r < ctx->R && n; ++r) {
    NEED(d_req[r], "null request column");
    cols[r] = reinterpret_cast<const long long*>(d_req[r]);
  }
  if (n == 0) return BGC_OK;
  CU(cudaSetDevice(ctx->device));
  int end_bit = 1;
  while ((1ull << end_bit) <= (unsigned long long)ctx->NS) ++end_bit;  // keys are 0..NS
  int rc = ensure_greedy_buffers(ctx, n);
  if (rc) return rc;
  cudaStream_t st = ctx->stream;
  {
    // stable grouping by namespace: LSD radix passes of 8 bits over the end_bit significant key bits; buffers
    // ping-pong so that the last pass lands in (g_keys_sorted, g_order)
    const int passes = (end_bit + 7) / 8;
    const uint32_t ntile = (uint32_t)((n + kSortTile - 1) / kSortTile);
    const int grid = (int)(ntile < (uint32_t)ctx->sm_count * 8 ? ntile : (uint32_t)ctx->sm_count * 8);
    const uint32_t *kin = nullptr, *iin = nullptr;
    CU(cudaFuncSetAttribute(k_rsort_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, kSortTile * 10));
    for (int p = 0; p < passes; ++p) {
      const bool to_final = ((passes - 1 - p) % 2) == 0;
      uint32_t* kout = to_final ? ctx->g_keys_sorted : ctx->g_keys;
      uint32_t* iout = to_final ? ctx->g_order : ctx->g_idx;
      k_rsort_hist<<<grid, kSortThreads, 0, st>>>(kin, d_ns_id, ctx->NS, n, 8 * p, ctx->g_hist, ntile);
      {
        const size_t cnt = (size_t)256 * ntile;
        const unsigned blocks = (unsigned)((cnt + kScanBlock - 1) / kScanBlock);
        uint32_t* totals = ctx->g_hist + cnt;  // room reserved behind the histogram
        k_scan_blocks<<<blocks, 1024, 0, st>>>(ctx->g_hist, cnt, totals);
        k_scan_u32<<<1, 1024, 0, st>>>(totals, blocks);
        k_scan_add<<<blocks, 1024, 0, st>>>(ctx->g_hist, cnt, totals);
        ctx->tm.launches += 2;
      }
      SortPayload P{};
      P.R = ctx->R;
      for (int r = 0; r < ctx->R; ++r) {
        P.in[r] = p == 0 ? cols[r] : (to_final ? ctx->g_req2[r] : ctx->g_req[r]);   // the previous pass wrote the other set
        P.out[r] = to_final ? ctx->g_req[r] : ctx->g_req2[r];
      }
      if (d_req_mask) {
        P.mask_in = p == 0 ? d_req_mask : (to_final ? ctx->g_mask2 : ctx->g_mask);
        P.mask_out = to_final ? ctx->g_mask : ctx->g_mask2;
      }
      static const int sort_form = [] { const char* e = std::getenv("BGC_K3_SORT"); return e ? std::atoi(e) : 0; }();
      if (sort_form == 1) {  // the round-2 form before warp-private ranking and the staged payload (kept for comparison)
        k_rsort_scatter<<<grid, kSortThreads, (size_t)kSortTile * 10, st>>>(kin, iin, d_ns_id, ctx->NS, n, 8 * p, ctx->g_hist, ntile, kout, iout, P);
      } else {
        CU(cudaFuncSetAttribute(k_rsort_scatter_wp, cudaFuncAttributeMaxDynamicSharedMemorySize, kSortTile * 10));
        k_rsort_scatter_wp<<<grid, kSort2Threads, (size_t)kSortTile * 10, st>>>(kin, iin, d_ns_id, ctx->NS, n, 8 * p, ctx->g_hist, ntile, kout, iout, P);
      }
      ctx->tm.launches += 3;
      CU(cudaGetLastError());
      kin = kout;
      iin = iout;
    }
  }
  const int g2 = grid_for((size_t)ctx->NS + 2, 256, ctx->sm_count * 4);
  k_greedy_segments<<<g2, 256, 0, st>>>(ctx->g_keys_sorted, n, ctx->NS, ctx->g_seg);
  ctx->tm.launches++;
  CU(cudaGetLastError());
  k_greedy_bad_ns<VT><<<ctx->sm_count, 256, 0, st>>>(ctx->g_order, ctx->g_seg, ctx->NS, d_verdict);
  ctx->tm.launches++;
  CU(cudaGetLastError());
  ctx->head_stale = true;
  switch (ctx->R) {
    case 1: return launch_k3_r<1, VT>(ctx, st, n, d_req_mask, d_verdict);
    case 2: return launch_k3_r<2, VT>(ctx, st, n, d_req_mask, d_verdict);
    case 3: return launch_k3_r<3, VT>(ctx, st, n, d_req_mask, d_verdict);
    case 4: return launch_k3_r<4, VT>(ctx, st, n, d_req_mask, d_verdict);
    default: break;
  }
  if constexpr (sizeof(VT) == 2) {
    switch (ctx->R) {
      case 5: return launch_k3_r<5, VT>(ctx, st, n, d_req_mask, d_verdict);
      case 6: return launch_k3_r<6, VT>(ctx, st, n, d_req_mask, d_verdict);
      case 7: return launch_k3_r<7, VT>(ctx, st, n, d_req_mask, d_verdict);
      default: return launch_k3_r<8, VT>(ctx, st, n, d_req_mask, d_verdict);
    }
  }
  return fail(ctx, BGC_E_BADARG, "R out of range");
}

template <typename VT>
int admit_greedy_host_impl(bgc_ctx* ctx, const uint32_t* ns_id, const int64_t* const* req,
                           const uint8_t* req_mask, size_t n, VT* verdict) {
  if (!ctx) return BGC_E_BADARG;
  if (!ctx->d_used) return fail(ctx, BGC_E_STATE, "quota table not set");
  NEED(n == 0 || (ns_id && req && verdict), "null column pointer");
  const int R = ctx->R;
  for (int r = 0; r < R && n; ++r) NEED(req[r], "null request column");
  if (n == 0) return BGC_OK;
  CU(cudaSetDevice(ctx->device));
  if (ctx->stream != ctx->own_stream) CU(cudaStreamSynchronize(ctx->stream));
  // the whole batch must be resident for the grouping pass: device copies owned by the context, grown on
  // demand and kept (a call with a batch no larger than an earlier one allocates nothing)
  if (n > ctx->gh_cap || R > ctx->gh_R) {
    CU(cudaDeviceSynchronize());
    cudaFree(ctx->gh_ns); cudaFree(ctx->gh_mask); cudaFree(ctx->gh_v);
    for (auto& q : ctx->gh_req) { cudaFree(q); q = nullptr; }
    ctx->gh_ns = nullptr; ctx->gh_mask = ctx->gh_v = nullptr;
    const size_t cap = n > ctx->gh_cap ? n : ctx->gh_cap;
    const int RR = R > ctx->gh_R ? R : ctx->gh_R;
    ctx->gh_cap = 0; ctx->gh_R = 0;
    CU(cudaMalloc(&ctx->gh_ns, cap * 4));
    for (int r = 0; r < RR; ++r) CU(cudaMalloc(&ctx->gh_req[r], cap * 8));
    CU(cudaMalloc(&ctx->gh_mask, cap));
    CU(cudaMalloc(&ctx->gh_v, cap * 2));
    ctx->gh_cap = cap; ctx->gh_R = RR;
  }
  cudaStream_t keep = ctx->stream;
  ctx->stream = ctx->own_stream;
  int rc = BGC_OK;
  cudaError_t e = cudaMemcpyAsync(ctx->gh_ns, ns_id, n * 4, cudaMemcpyHostToDevice, ctx->own_stream);
  for (int r = 0; r < R && e == cudaSuccess; ++r)
    e = cudaMemcpyAsync(ctx->gh_req[r], req[r], n * 8, cudaMemcpyHostToDevice, ctx->own_stream);
  if (e == cudaSuccess && req_mask) e = cudaMemcpyAsync(ctx->gh_mask, req_mask, n, cudaMemcpyHostToDevice, ctx->own_stream);
  if (e != cudaSuccess) {
    rc = fail(ctx, BGC_E_CUDA, "greedy staging: %s", cudaGetErrorString(e));
  } else {
    const int64_t* cols[kMaxR];
    for (int r = 0; r < kMaxR; ++r) cols[r] = reinterpret_cast<const int64_t*>(ctx->gh_req[r]);
    rc = admit_greedy_dev_impl<VT>(ctx, ctx->gh_ns, cols, req_mask ? ctx->gh_mask : nullptr, n, reinterpret_cast<VT*>(ctx->gh_v));
    if (rc == BGC_OK) {
      e = cudaMemcpyAsync(verdict, ctx->gh_v, n * sizeof(VT), cudaMemcpyDeviceToHost, ctx->own_stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->own_stream);
      if (e != cudaSuccess) rc = fail(ctx, BGC_E_CUDA, "greedy copy-out: %s", cudaGetErrorString(e));
    }
  }
  cudaStreamSynchronize(ctx->own_stream);
  ctx->stream = keep;
  return rc;
}
}  // namespace

extern "C" {
int bgc_admit_greedy_dev(bgc_ctx* ctx, const uint32_t* d_ns_id, const int64_t* const* d_req, const uint8_t* d_req_mask, size_t n,
                         uint8_t* d_verdict) {
  return admit_greedy_dev_impl<uint8_t>(ctx, d_ns_id, d_req, d_req_mask, n, d_verdict);
}
int bgc_admit_greedy16_dev(bgc_ctx* ctx, const uint32_t* d_ns_id, const int64_t* const* d_req, const uint8_t* d_req_mask, size_t n,
                           uint16_t* d_verdict) {
  return admit_greedy_dev_impl<uint16_t>(ctx, d_ns_id, d_req, d_req_mask, n, d_verdict);
}
int bgc_admit_greedy(bgc_ctx* ctx, const uint32_t* ns_id, const int64_t* const* req, const uint8_t* req_mask, size_t n,
                     uint8_t* verdict) {
  return admit_greedy_host_impl<uint8_t>(ctx, ns_id, req, req_mask, n, verdict);
}
int bgc_admit_greedy16(bgc_ctx* ctx, const uint32_t* ns_id, const int64_t* const* req, const uint8_t* req_mask, size_t n,
                       uint16_t* verdict) {
  return admit_greedy_host_impl<uint16_t>(ctx, ns_id, req, req_mask, n, verdict);
}
}  // extern "C"

extern "C" {
// ---------------------------------------------------------------- synthetic inputs / checksums
int bgc_synth_pods_dev(bgc_ctx* ctx, const bgc_synth_params* p, uint64_t first, size_t count,
                       uint32_t* d_ns_id, int64_t* const* d_req, uint8_t* d_req_mask) {
  if (!ctx) return BGC_E_BADARG;
  NEED(p && d_ns_id && d_req, "null pointer");
  NEED(p->R >= 1 && p->R <= BGC_MAX_R, "R must be 1..BGC_MAX_R");
  NEED(p->ns_count > 0, "ns_count must be positive");
  NEED(p->dist != BGC_DIST_ZIPF_CDF || p->d_cdf, "zipf needs a device CDF table");
  NEED(!p->grouped || p->n_total > 0, "grouped mode needs n_total");
  if (count == 0) return BGC_OK;
  CU(cudaSetDevice(ctx->device));
  SynthArgs A{};
  const uint32_t sbase = p->set * 16u;
  A.key_ns = stream_key(p->seed, sbase + 0);
  A.key_mask = stream_key(p->seed, sbase + 8);
  for (uint32_t r = 0; r < p->R; ++r) {
    A.key_req[r] = stream_key(p->seed, sbase + 1 + r);
    A.kind[r] = p->kind[r];
    A.req[r] = reinterpret_cast<long long*>(d_req[r]);
    NEED(d_req[r], "null request column");
  }
  A.R = p->R; A.ns_count = p->ns_count; A.dist = p->dist; A.grouped = p->grouped;
  A.step = p->n_total ? UINT64_MAX / p->n_total : 0;
  A.a = p->scramble_a; A.b = p->scramble_b;
  A.cdf = p->d_cdf; A.ns_id = d_ns_id; A.req_mask = d_req_mask;
  const int grid = grid_for(count, 256, ctx->sm_count * 16);
  k_synth_pods<<<grid, 256, 0, ctx->stream>>>(A, first, count);
  ctx->tm.launches++;
  CU(cudaGetLastError());
  return BGC_OK;
}

int bgc_synth_pods_shard_dev(bgc_ctx* ctx, const bgc_synth_params* p, uint64_t first, size_t count, uint32_t world,
                             uint32_t rank, const uint32_t* d_local_of, uint64_t out_pos, size_t out_capacity,
                             uint32_t* d_ns_id, int64_t* const* d_req, uint8_t* d_req_mask, uint64_t* d_src_index,
                             uint64_t* kept_out) {
  if (!ctx) return BGC_E_BADARG;
  NEED(p && kept_out, "null pointer");
  NEED(world >= 1 && rank < world, "rank must be below world");
  NEED(p->R >= 1 && p->R <= BGC_MAX_R, "R must be 1..BGC_MAX_R");
  NEED(p->ns_count > 0, "ns_count must be positive");
  NEED(p->dist != BGC_DIST_ZIPF_CDF || p->d_cdf, "zipf needs a device CDF table");
  NEED(!p->grouped || p->n_total > 0, "grouped mode needs n_total");
  NEED(!d_ns_id || d_req, "null request column array");
  *kept_out = 0;
  if (count == 0) return BGC_OK;
  CU(cudaSetDevice(ctx->device));
  constexpr size_t kChunkTiles = 32768;                         // 2^26 pods per chunk
  constexpr size_t kChunk = kChunkTiles * (size_t)kShardTile;
  if (!ctx->sh_tile_count) {
    CU(cudaMalloc(&ctx->sh_tile_count, kChunkTiles * 4));
    CU(cudaMalloc(&ctx->sh_tile_off, (kChunkTiles + 1) * 8));
  }
  SynthArgs A{};
  const uint32_t sbase = p->set * 16u;
  A.key_ns = stream_key(p->seed, sbase + 0);
  A.key_mask = stream_key(p->seed, sbase + 8);
  for (uint32_t r = 0; r < p->R; ++r) {
    A.key_req[r] = stream_key(p->seed, sbase + 1 + r);
    A.kind[r] = p->kind[r];
    if (d_ns_id) {
      NEED(d_req[r], "null request column");
      A.req[r] = reinterpret_cast<long long*>(d_req[r]);
    }
  }
  A.R = p->R; A.ns_count = p->ns_count; A.dist = p->dist; A.grouped = p->grouped;
  A.step = p->n_total ? UINT64_MAX / p->n_total : 0;
  A.a = p->scramble_a; A.b = p->scramble_b;
  A.cdf = p->d_cdf; A.ns_id = d_ns_id; A.req_mask = d_req_mask;
  unsigned long long kept = 0;
  for (size_t off = 0; off < count; off += kChunk) {
    const size_t m = count - off < kChunk ? count - off : kChunk;
    const size_t tiles = (m + kShardTile - 1) / kShardTile;
    const int grid = (int)(tiles < (size_t)ctx->sm_count * 16 ? tiles : (size_t)ctx->sm_count * 16);
    k_synth_shard_count<<<grid, 256, 0, ctx->stream>>>(A, first + off, m, world, rank, ctx->sh_tile_count);
    k_scan_tiles<<<1, 1024, 0, ctx->stream>>>(ctx->sh_tile_count, tiles, out_pos + kept, ctx->sh_tile_off,
                                              ctx->sh_tile_off + kChunkTiles);
    ctx->tm.launches += 2;
    CU(cudaGetLastError());
    unsigned long long got = 0;
    CU(cudaMemcpyAsync(&got, ctx->sh_tile_off + kChunkTiles, 8, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    if (d_ns_id) {
      if (out_pos + kept + got > out_capacity)
        return fail(ctx, BGC_E_BADARG, "shard columns hold %zu pods, this shard needs at least %llu", out_capacity,
                    (unsigned long long)(out_pos + kept + got));
      k_synth_shard_emit<<<grid, 256, 0, ctx->stream>>>(A, first + off, m, world, rank, d_local_of, ctx->sh_tile_off,
                                                        reinterpret_cast<unsigned long long*>(d_src_index));
      ctx->tm.launches++;
      CU(cudaGetLastError());
    }
    kept += got;
  }
  *kept_out = kept;
  return BGC_OK;
}

int bgc_synth_policy_dev(bgc_ctx* ctx, uint64_t seed, uint64_t first, size_t count, uint8_t* d_op,
                         uint16_t* d_flags) {
  if (!ctx) return BGC_E_BADARG;
  NEED(d_op && d_flags, "null pointer");
  if (count == 0) return BGC_OK;
  CU(cudaSetDevice(ctx->device));
  const int grid = grid_for(count, 256, ctx->sm_count * 16);
  k_synth_policy<<<grid, 256, 0, ctx->stream>>>(stream_key(seed, 64), first, count, d_op, d_flags);
  ctx->tm.launches++;
  CU(cudaGetLastError());
  return BGC_OK;
}

int bgc_checksum_u8_dev(bgc_ctx* ctx, const uint8_t* d_v, uint64_t first, size_t count, uint64_t* out_host) {
  if (!ctx) return BGC_E_BADARG;
  NEED(out_host && (d_v || count == 0), "null pointer");
  CU(cudaSetDevice(ctx->device));
  CU(cudaMemsetAsync(ctx->d_scratch, 0, 8, ctx->stream));
  if (count) {
    const int grid = grid_for(count, 256, ctx->sm_count * 8);
    k_checksum_u8<<<grid, 256, 0, ctx->stream>>>(d_v, first, count, ctx->d_scratch);
    ctx->tm.launches++;
    CU(cudaGetLastError());
  }
  CU(cudaMemcpyAsync(out_host, ctx->d_scratch, 8, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return BGC_OK;
}

int bgc_histogram_u8_dev(bgc_ctx* ctx, const uint8_t* d_v, size_t count, uint64_t* out_host256) {
  if (!ctx) return BGC_E_BADARG;
  NEED(out_host256 && (d_v || count == 0), "null pointer");
  CU(cudaSetDevice(ctx->device));
  CU(cudaMemsetAsync(ctx->d_scratch, 0, 256 * 8, ctx->stream));
  if (count) {
    const int grid = grid_for(count, 256, ctx->sm_count * 4);
    k_histogram_u8<<<grid, 256, 0, ctx->stream>>>(d_v, count, ctx->d_scratch);
    ctx->tm.launches++;
    CU(cudaGetLastError());
  }
  CU(cudaMemcpyAsync(out_host256, ctx->d_scratch, 256 * 8, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return BGC_OK;
}

}  // extern "C"

#include "bgc_api_typed.inl"
