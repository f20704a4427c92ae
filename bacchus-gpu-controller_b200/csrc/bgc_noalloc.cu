// bgc_noalloc.cu — bgc_usage_reduce_static_dev: per-namespace usage without any allocation.
//
// Who this is for: a caller that must not (or cannot) allocate device memory at call time — a
// serving process that handed all of HBM to its own pool at start-up, so a cudaMalloc issued by a
// library would fail or eat into memory someone else has budgeted; or a request handler on a
// latency-critical path, where the implicit device-wide synchronisation of cudaMalloc/cudaFree
// would stall every other stream.  The regular entry points (bgc_ctx_create,
// bgc_quota_table_set, bgc_usage_reduce*) allocate their accumulation table, staging buffers and
// streams; this one allocates nothing, ever: no cudaMalloc, no cudaFree, no stream or event
// creation, no library routine that allocates.
//
// How: the accumulation table the reduction kernel scatters into lives in a file-scope
// `static __device__` array of fixed size (BGC_STATIC_MAX_NS namespaces x 4 columns x 8 B = 512 KiB), which
// every GPU that loads libbgc.so holds from module load until the process exits.  The caller
// passes the inputs (pod columns) and the output (used[R][NS]) as device pointers, plus the stream
// to run on.
//
// Things a user must know:
//   * It works on the CURRENT device (the one the pointers belong to).  The first call on each
//     device performs a one-time set-up — it asks the runtime for the device address of the
//     static array (cudaGetSymbolAddress), and loads and configures every kernel any later call
//     can launch (all R, both table kinds, aligned and unaligned builds), so that no later call
//     triggers a lazy module load.  Make that first call before starting any stream capture on
//     that device.  The address is cached for the life of the process: after cudaDeviceReset()
//     on that device it is stale, so do not reset a device this entry point has been used on.
//   * ns_count above BGC_STATIC_MAX_NS (or R outside 1..4) is rejected with BGC_E_BADARG
//     and a message in `errbuf`; nothing is launched.
//   * Every call on a device shares the one array, and the array keeps whatever the previous call
//     left in it.  Each call therefore first clears the slots it is about to use; but two calls
//     whose kernels overlap on the GPU (different streams, or different host threads) race on the
//     table and produce garbage.  Serialise calls per device.
#include <cstdarg>
#include <cstdio>
#include <mutex>

#include "../../include/bgc.h"
#include "bgc_kernels.cuh"

using namespace bgc;

namespace {

constexpr int kMaxDevices = 64;
constexpr int kStaticMaxR = 4;  // columns the static table is sized for
constexpr size_t kStaticSmemBytes = 227 * 1024;  // per-CTA shared-memory table of the scatter kernel

// The scratch itself: the [R][ns_count] usage table the kernel accumulates into.  In HOST code the
// name `g_static_used` is only a handle for the runtime (a host-side shadow symbol), not a device
// address; kernels must be given the pointer obtained from cudaGetSymbolAddress below.
static __device__ __align__(256) long long g_static_used[(size_t)BGC_STATIC_MAX_NS * kStaticMaxR];

struct PerDevice {
  bool ready = false;
  long long* used = nullptr;  // device address of g_static_used on this device
  int sm_count = 0;
  int seg_blocks_per_sm[kStaticMaxR + 1][2] = {};  // occupancy of the segmented kernel, per R and alignment
};
PerDevice g_dev[kMaxDevices];
std::mutex g_mu;

int say(char* errbuf, size_t cap, int code, const char* fmt, ...) {
  if (errbuf && cap) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(errbuf, cap, fmt, ap);
    va_end(ap);
  }
  return code;
}

#define SCU(call)                                                                          \
  do {                                                                                     \
    cudaError_t e_ = (call);                                                               \
    if (e_ != cudaSuccess) return say(errbuf, errbuf_len, BGC_E_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); \
  } while (0)

// Everything the entry point can launch for one R: the segmented kernel (grouped input) and the
// scatter kernel with its direct and hash tables, each in an aligned and an unaligned build.  With
// CUDA's lazy module loading a kernel is loaded on first use, and loading may allocate or
// synchronise — so every one of them is touched here, during the one-time set-up, and the
// per-call path only launches.
template <int R>
int prepare_r(PerDevice& d, char* errbuf, size_t errbuf_len) {
  SCU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.seg_blocks_per_sm[R][0], k1_usage_reduce<R, SINK_GLOBAL, false>, 512, 0));
  SCU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.seg_blocks_per_sm[R][1], k1_usage_reduce<R, SINK_GLOBAL, true>, 512, 0));
  SCU(cudaFuncSetAttribute(k1_scatter<R, SINK_DIRECT, false, 1024, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kStaticSmemBytes));
  SCU(cudaFuncSetAttribute(k1_scatter<R, SINK_DIRECT, true, 1024, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kStaticSmemBytes));
  SCU(cudaFuncSetAttribute(k1_scatter<R, SINK_HASH, false, 1024, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kStaticSmemBytes));
  SCU(cudaFuncSetAttribute(k1_scatter<R, SINK_HASH, true, 1024, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kStaticSmemBytes));
  return BGC_OK;
}

// one-time set-up for the current device (called under g_mu)
int setup_device(int dev, char* errbuf, size_t errbuf_len) {
  PerDevice& d = g_dev[dev];
  void* p = nullptr;
  SCU(cudaGetSymbolAddress(&p, g_static_used));
  if (!p) return say(errbuf, errbuf_len, BGC_E_CUDA, "cudaGetSymbolAddress returned a null address");
  SCU(cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, dev));
  int rc;
  if ((rc = prepare_r<1>(d, errbuf, errbuf_len)) || (rc = prepare_r<2>(d, errbuf, errbuf_len)) ||
      (rc = prepare_r<3>(d, errbuf, errbuf_len)) || (rc = prepare_r<4>(d, errbuf, errbuf_len)))
    return rc;
  d.used = static_cast<long long*>(p);
  d.ready = true;
  return BGC_OK;
}

// grouped input: segmented warp-shuffle reduction, one global red per run of equal namespaces
template <int R, bool VEC>
int launch_segmented(const PerDevice& d, cudaStream_t st, const uint32_t* ns, const long long* const* req, size_t n,
                     uint32_t NS, char* errbuf, size_t errbuf_len) {
  ReqCols<R> cols;
  for (int r = 0; r < R; ++r) cols.p[r] = req[r];
  const int threads = 512;
  int blocks_per_sm = d.seg_blocks_per_sm[R][VEC ? 1 : 0];
  if (blocks_per_sm < 1) blocks_per_sm = 1;
  const size_t tiles = (n + kK1Tile - 1) / kK1Tile;
  const size_t wpb = threads / 32;
  int grid = d.sm_count * blocks_per_sm;
  if ((size_t)grid * wpb > tiles) grid = (int)((tiles + wpb - 1) / wpb);
  if (grid < 1) grid = 1;
  size_t chunk_tiles = tiles / ((size_t)grid * wpb * 4);
  if (chunk_tiles < 1) chunk_tiles = 1;
  if (chunk_tiles > 32) chunk_tiles = 32;
  k1_usage_reduce<R, SINK_GLOBAL, VEC><<<grid, threads, 0, st>>>(ns, cols, n, d.used, NS, NS, (uint32_t)chunk_tiles, (1u << R) - 1u);
  SCU(cudaGetLastError());
  return BGC_OK;
}

// arrival-order input: the scatter kernel with a CTA-private shared-memory table (indexed by
// namespace when the table fits, claim-once hash slots otherwise), flushed into the static array.
// Every column keeps both 32-bit halves in shared memory (this entry point is not told the units).
template <int R, int SINK, bool VEC>
int launch_scatter(const PerDevice& d, cudaStream_t st, const uint32_t* ns, const long long* const* req, size_t n,
                   uint32_t NS, uint32_t cap, char* errbuf, size_t errbuf_len) {
  ReqCols<R> cols;
  for (int r = 0; r < R; ++r) cols.p[r] = req[r];
  const size_t smem = (size_t)cap * (8 * R + (SINK == SINK_HASH ? 4 : 0));
  const int threads = 1024;
  const size_t groups = (n + 1) / 2;
  int grid = d.sm_count;
  if ((size_t)grid * threads > groups) grid = (int)((groups + threads - 1) / threads);
  if (grid < 1) grid = 1;
  k1_scatter<R, SINK, VEC, 1024, 1><<<grid, threads, smem, st>>>(ns, cols, n, d.used, NS, cap, (1u << R) - 1u, 0u);
  SCU(cudaGetLastError());
  return BGC_OK;
}

template <int R>
int reduce_r(const PerDevice& d, cudaStream_t st, const uint32_t* ns, const long long* const* req, size_t n,
             uint32_t NS, unsigned flags, char* errbuf, size_t errbuf_len) {
  bool vec = (reinterpret_cast<uintptr_t>(ns) & 15) == 0;
  for (int r = 0; r < R; ++r) vec = vec && (reinterpret_cast<uintptr_t>(req[r]) & 31) == 0;
  if (flags & BGC_REDUCE_GROUPED)
    return vec ? launch_segmented<R, true>(d, st, ns, req, n, NS, errbuf, errbuf_len)
               : launch_segmented<R, false>(d, st, ns, req, n, NS, errbuf, errbuf_len);
  if ((size_t)NS * R * 8 <= kStaticSmemBytes)
    return vec ? launch_scatter<R, SINK_DIRECT, true>(d, st, ns, req, n, NS, NS, errbuf, errbuf_len)
               : launch_scatter<R, SINK_DIRECT, false>(d, st, ns, req, n, NS, NS, errbuf, errbuf_len);
  const uint32_t cap = (uint32_t)(kStaticSmemBytes / (8 * R + 4));
  return vec ? launch_scatter<R, SINK_HASH, true>(d, st, ns, req, n, NS, cap, errbuf, errbuf_len)
             : launch_scatter<R, SINK_HASH, false>(d, st, ns, req, n, NS, cap, errbuf, errbuf_len);
}

}  // namespace

extern "C" int bgc_usage_reduce_static_dev(const uint32_t* d_ns_id, const int64_t* const* d_req, uint32_t R,
                                           size_t n, uint32_t ns_count, unsigned flags, int64_t* d_used_out,
                                           void* cuda_stream, char* errbuf, size_t errbuf_len) {
  if (errbuf && errbuf_len) errbuf[0] = 0;
  if (R < 1 || R > (uint32_t)kStaticMaxR)
    return say(errbuf, errbuf_len, BGC_E_BADARG, "R = %u is outside 1..%d", R, kStaticMaxR);
  if (ns_count < 1 || ns_count > BGC_STATIC_MAX_NS)
    return say(errbuf, errbuf_len, BGC_E_BADARG,
               "ns_count = %u exceeds the static table (BGC_STATIC_MAX_NS = %d namespaces); use bgc_usage_reduce_dev",
               ns_count, BGC_STATIC_MAX_NS);
  if (!d_used_out || (n && (!d_ns_id || !d_req)))
    return say(errbuf, errbuf_len, BGC_E_BADARG, "null pointer argument");
  if (flags & BGC_REDUCE_ACCUMULATE)
    return say(errbuf, errbuf_len, BGC_E_BADARG, "BGC_REDUCE_ACCUMULATE is not supported: the table is cleared on every call");
  const long long* cols[kMaxR] = {};
  for (uint32_t r = 0; r < R && n; ++r) {
    if (!d_req[r]) return say(errbuf, errbuf_len, BGC_E_BADARG, "null request column %u", r);
    cols[r] = reinterpret_cast<const long long*>(d_req[r]);
  }
  int dev = 0;
  SCU(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return say(errbuf, errbuf_len, BGC_E_BADARG, "device index %d not supported", dev);
  {
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_dev[dev].ready) {
      const int rc = setup_device(dev, errbuf, errbuf_len);
      if (rc) return rc;
    }
  }
  const PerDevice& d = g_dev[dev];
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);

  // the table keeps the previous call's sums: zero the [R][ns_count] slots this call accumulates into
  SCU(cudaMemsetAsync(d.used, 0, (size_t)ns_count * R * 8, st));
  int rc = BGC_OK;
  if (n) {
    switch (R) {
      case 1: rc = reduce_r<1>(d, st, d_ns_id, cols, n, ns_count, flags, errbuf, errbuf_len); break;
      case 2: rc = reduce_r<2>(d, st, d_ns_id, cols, n, ns_count, flags, errbuf, errbuf_len); break;
      case 3: rc = reduce_r<3>(d, st, d_ns_id, cols, n, ns_count, flags, errbuf, errbuf_len); break;
      default: rc = reduce_r<4>(d, st, d_ns_id, cols, n, ns_count, flags, errbuf, errbuf_len); break;
    }
    if (rc) return rc;
  }
  // the table already has the [R][ns_count] layout of the output
  SCU(cudaMemcpyAsync(d_used_out, d.used, (size_t)ns_count * R * 8, cudaMemcpyDeviceToDevice, st));
  return BGC_OK;
}
