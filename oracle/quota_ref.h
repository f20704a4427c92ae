/* oracle/quota_ref.h — TEST INFRASTRUCTURE ONLY (CPU oracle; never linked into the product).
 *
 * PARITY UNPINNED: the reference has NO implementation of pod-vs-quota evaluation
 * (SURVEY.md §0, §8 rows A3/A4): it only *writes* the ResourceQuota object
 * (/root/reference/src/synchronizer.rs:249-281, /root/reference/src/controller.rs:89-110) and
 * kube-apiserver's built-in ResourceQuota plugin (Go, k8s.io/apiserver, not vendored, API level
 * v1_26 per /root/reference/Cargo.toml:31) enforces it.  This file is the literal CPU statement
 * of THIS repo's spec for rows A3 (usage_reduce) and A4 (quota_eval, snapshot semantics), whose
 * resource columns come from the key set at synchronizer.rs:249-281.
 *
 * Also holds the counter-based synthetic generator spec of SURVEY.md §8(d) (gen_ref_*), which
 * the CUDA library re-implements independently; tests compare the two bit for bit.
 */
#ifndef QUOTA_REF_H
#define QUOTA_REF_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define QUOTA_REF_MAX_R 8
#define QUOTA_REF_VERDICT_BAD_NS 0xFFu
/* element types of the typed (narrow-column) forms; value = widened element << shift, wrapping */
enum { QUOTA_REF_T_I64 = 0, QUOTA_REF_T_U32 = 1, QUOTA_REF_T_U16 = 2, QUOTA_REF_T_U8 = 3, QUOTA_REF_T_I32 = 4 };

/* A3: used[r][ns] = sum over pods i with ns_id[i]==ns of req[r][i], wrapping 64-bit.
 * `used` is [R][NS], zeroed here.  Pods with ns_id >= NS contribute nothing. */
void quota_ref_usage_reduce(const uint32_t *ns_id, const int64_t *const *req, size_t n,
                            uint32_t NS, uint32_t R, int64_t *used, int threads);

/* A4 (snapshot semantics): verdict[j] = (missing<<4)|exceeded, 0 == allow.
 *   for r < R with hard_mask[ns] bit r set:
 *     req_mask given and bit r clear:  compute resource (compute_mask bit r) → missing bit r;
 *                                      otherwise the request counts as 0 → nothing
 *     else if req > 0 and used + req > hard, evaluated WITHOUT overflow (in Z) → exceeded bit r
 *       (SURVEY.md §8 A4 wrote the sum as wrapping only so that CPU and GPU agree; the exact
 *        comparison is what Kubernetes' Quantity arithmetic decides and is identical whenever the
 *        64-bit sum does not overflow — see DESIGN.md §3)
 *   ns_id >= NS → QUOTA_REF_VERDICT_BAD_NS.  req_mask may be NULL (all present). */
void quota_ref_quota_eval(const uint32_t *ns_id, const int64_t *const *req,
                          const uint8_t *req_mask, size_t n, uint32_t NS, uint32_t R,
                          const int64_t *used, const int64_t *hard, const uint8_t *hard_mask,
                          uint8_t compute_mask, uint8_t *verdict, int threads);

/* The u16 form for up to 8 resources: verdict = (missing<<8)|exceeded, unknown namespace 0xFFFF. */
void quota_ref_quota_eval16(const uint32_t *ns_id, const int64_t *const *req,
                            const uint8_t *req_mask, size_t n, uint32_t NS, uint32_t R,
                            const int64_t *used, const int64_t *hard, const uint8_t *hard_mask,
                            uint8_t compute_mask, uint16_t *verdict, int threads);

/* N1: Kubernetes-faithful sequential admission inside one batch.  Pods are taken
 * in arrival order; pod j is admitted iff its verdict — computed exactly like quota_ref_quota_eval
 * but against the CURRENT usage of its namespace — is 0, and only admitted pods add their requests
 * (all R columns, wrapping i64 like A3) to `used`, which is updated in place.  The literal loop. */
void quota_ref_admit_greedy(const uint32_t *ns_id, const int64_t *const *req,
                            const uint8_t *req_mask, size_t n, uint32_t NS, uint32_t R,
                            int64_t *used, const int64_t *hard, const uint8_t *hard_mask,
                            uint8_t compute_mask, uint8_t *verdict);

/* ---- synthetic generator spec ---- */
enum { GEN_KIND_GPU = 0, GEN_KIND_CPU_MILLI = 1, GEN_KIND_MEM_BYTES = 2, GEN_KIND_MIG = 3 };
enum { GEN_DIST_UNIFORM = 0, GEN_DIST_ZIPF_CDF = 1 };

typedef struct {
  uint64_t seed;
  uint32_t set;        /* 0 = running set, 1 = candidate set (independent draws) */
  uint32_t ns_count;
  uint32_t dist;       /* GEN_DIST_* */
  uint32_t grouped;    /* 0 = arrival order is hashed, 1 = pods grouped by namespace */
  uint64_t n_total;    /* batch size (used by grouped mode) */
  uint64_t scramble_a; /* id = (rank*a + b) % ns_count; a coprime to ns_count */
  uint64_t scramble_b;
  const uint64_t *cdf; /* ns_count entries, nondecreasing, last = 2^64-1 (ZIPF_CDF) */
  uint32_t R;
  uint32_t kind[QUOTA_REF_MAX_R];
} gen_ref_params;

uint64_t gen_ref_hash(uint64_t seed, uint32_t stream, uint64_t index);
/* fill pods [first, first+count) */
void gen_ref_fill(const gen_ref_params *p, uint64_t first, size_t count, uint32_t *ns_id,
                  int64_t *const *req, uint8_t *req_mask /* may be NULL */, int threads);
/* policy inputs: op in 0..3, flags in 0..1023 drawn from the hash */
void gen_ref_policy_fill(uint64_t seed, uint64_t first, size_t count, uint8_t *op,
                         uint16_t *flags, int threads);
/* order-independent checksum of a u8 stream: sum_i SM64(i ^ v[i]<<56 ...) */
uint64_t gen_ref_checksum_u8(const uint8_t *v, uint64_t first, size_t count);

void quota_ref_admit_greedy16(const uint32_t *ns_id, const int64_t *const *req,
                              const uint8_t *req_mask, size_t n, uint32_t NS, uint32_t R,
                              int64_t *used, const int64_t *hard, const uint8_t *hard_mask,
                              uint8_t compute_mask, uint16_t *verdict);
/* typed forms: the same results as the plain functions on the widened columns (quota_ref.c) */
void quota_ref_usage_reduce_typed(const void *ns_id, uint32_t ns_type, const void *const *cols,
                                  const uint8_t *type, const uint8_t *shift, size_t n, uint32_t NS,
                                  uint32_t R, int64_t *used, int threads, uint64_t *workspace);
void quota_ref_quota_eval_typed(const void *ns_id, uint32_t ns_type, const void *const *cols,
                                const uint8_t *type, const uint8_t *shift, const uint8_t *req_mask,
                                size_t n, uint32_t NS, uint32_t R, const int64_t *used,
                                const int64_t *hard, const uint8_t *hard_mask, uint8_t compute_mask,
                                uint8_t *verdict, int threads);
void quota_ref_usage_reduce_ws(const uint32_t *ns_id, const int64_t *const *req, size_t n, uint32_t NS,
                               uint32_t R, int64_t *used, int threads, uint64_t *workspace);

#ifdef __cplusplus
}
#endif
#endif
