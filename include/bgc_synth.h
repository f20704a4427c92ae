/* bgc_synth.h — device-side synthetic input generator and stream checksums (libbgc.so).
 *
 * No counterpart in the reference (it has no benchmarks).  These entry points exist
 * so that the 10^8 / 10^9-pod configurations of BASELINE.json can be produced directly in HBM
 * instead of crossing PCIe, from the counter-based integer generator specified in SURVEY.md §8(d):
 * every column value is a pure function of (seed, set, column, pod index), so the CPU oracle can
 * regenerate any slice of the same data and check the GPU results on it.
 */
#ifndef BGC_SYNTH_H
#define BGC_SYNTH_H
#include "bgc.h"

#ifdef __cplusplus
extern "C" {
#endif

enum { BGC_GEN_GPU = 0, BGC_GEN_CPU_MILLI = 1, BGC_GEN_MEM_BYTES = 2, BGC_GEN_MIG = 3 };
enum { BGC_DIST_UNIFORM = 0, BGC_DIST_ZIPF_CDF = 1 };

typedef struct bgc_synth_params {
  uint64_t seed;
  uint32_t set;        /* 0 = running set, 1 = candidate set */
  uint32_t ns_count;
  uint32_t dist;       /* BGC_DIST_* */
  uint32_t grouped;    /* 1 = pods grouped by namespace */
  uint64_t n_total;    /* batch size (grouped mode spreads ranks over it) */
  uint64_t scramble_a; /* id = (rank * a + b) % ns_count */
  uint64_t scramble_b;
  const uint64_t *d_cdf; /* DEVICE pointer: ns_count nondecreasing u64 thresholds (ZIPF_CDF) */
  uint32_t R;
  uint32_t kind[BGC_MAX_R]; /* BGC_GEN_* per resource column */
} bgc_synth_params;

/* fill pods [first, first+count) into device columns; d_req_mask may be NULL */
int bgc_synth_pods_dev(bgc_ctx *ctx, const bgc_synth_params *p, uint64_t first, size_t count,
                       uint32_t *d_ns_id, int64_t *const *d_req, uint8_t *d_req_mask);
/* Row E: ONE shard of ONE global stream.  Of pods [first, first+count) of the
 * stream, keep those whose namespace hashes to this rank — mix32(ns_id) % world == rank, the same
 * murmur3 finaliser as shard.py — in stream order, and append them at position out_pos of the shard
 * columns (out_capacity = pods the columns can hold).  d_local_of (optional, u32[ns_count]) rewrites
 * each kept pod's global namespace id into the shard's dense id (every rank owns about ns_count/world
 * namespaces: the host flattener would hand each shard its own dense ids); d_src_index (optional)
 * receives each kept pod's position in the global stream — what the host-side gather of verdicts
 * needs.  d_ns_id == NULL: count only.  *kept_out = pods kept.  Deterministic: the shard is the
 * global stream filtered, whatever the grid. */
int bgc_synth_pods_shard_dev(bgc_ctx *ctx, const bgc_synth_params *p, uint64_t first, size_t count,
                             uint32_t world, uint32_t rank, const uint32_t *d_local_of,
                             uint64_t out_pos, size_t out_capacity, uint32_t *d_ns_id,
                             int64_t *const *d_req, uint8_t *d_req_mask, uint64_t *d_src_index,
                             uint64_t *kept_out);
/* policy inputs: op in 0..3, flags in 0..1023 */
int bgc_synth_policy_dev(bgc_ctx *ctx, uint64_t seed, uint64_t first, size_t count, uint8_t *d_op,
                         uint16_t *d_flags);
/* position-keyed checksum of a device u8 stream: Σ_k mix64((first+k) ^ (v[k] << 56)) mod 2^64;
 * additive, so checksums of consecutive slices sum to the checksum of the whole stream. */
int bgc_checksum_u8_dev(bgc_ctx *ctx, const uint8_t *d_v, uint64_t first, size_t count,
                        uint64_t *out_host);
/* histogram of a device u8 stream (256 bins, host output) — verdict statistics */
int bgc_histogram_u8_dev(bgc_ctx *ctx, const uint8_t *d_v, size_t count, uint64_t *out_host256);

#ifdef __cplusplus
}
#endif
#endif
