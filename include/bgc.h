/* bgc.h — C ABI of libbgc.so: batched admission / quota evaluation on B200 (sm_100a).
 *
 * This is the drop-in boundary a maintainer of bacchus-snu/bacchus-gpu-controller would bind
 * from Rust with an `extern "C"` block (INTEGRATION.md shows the stub).  The reference has no
 * FFI/plugin interface of its own; the seam these entry points sit behind is
 *
 *     fn mutate(req: &AdmissionRequest<DynamicObject>, state: &AppState)
 *         -> Result<AdmissionResponse, Error>          reference: src/admission.rs:241-244
 *
 * called from `mutate_handler` (src/admission.rs:184-204).  Everything here is plain-old-data:
 * caller-owned buffers, sizes, and an opaque context.  No exceptions cross the boundary.
 *
 * Conventions
 *   - every function returns BGC_OK (0) or a negative BGC_E_* code; bgc_last_error(ctx) gives a
 *     human-readable message owned by the context (valid until the next call on that ctx);
 *   - a bgc_ctx is bound to one CUDA device; calls on one ctx must not overlap (use one ctx per
 *     thread, like the reference's stateless per-request handler: src/admission.rs:145-147);
 *   - functions without the `_dev` suffix take HOST pointers (pageable or pinned) and move the
 *     data themselves (chunked, double-buffered H2D/D2H); `_dev` variants take DEVICE pointers
 *     and only enqueue kernels on the context's stream (see bgc_ctx_set_stream), so a caller can
 *     time or overlap transfers itself;
 *   - there is NO CPU fallback: without a CUDA device bgc_ctx_create fails with BGC_E_CUDA;
 *   - allocation: bgc_quota_table_set (when NS or R change), the first host-pointer call of a context
 *     and the first one with a larger batch or more columns than any before (staging buffers, up to
 *     4 Mi pods per stage), and bgc_admit_greedy* with a larger batch call cudaMalloc/cudaFree and
 *     cudaDeviceSynchronize; every other call only enqueues work.  A caller that cannot allocate at
 *     call time warms the context up once with its largest batch, or uses the `_dev` forms (which
 *     never allocate after bgc_quota_table_set), or bgc_usage_reduce_static_dev below.
 */
#ifndef BGC_H
#define BGC_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BGC_OK 0
#define BGC_E_BADARG (-1) /* NULL pointer, R out of range, size mismatch, misaligned device pointer */
#define BGC_E_CUDA (-2)   /* a CUDA runtime call failed; message holds cudaGetErrorString */
#define BGC_E_OOM (-3)    /* device or pinned-host allocation failed */
#define BGC_E_STATE (-4)  /* call order: quota table not set before reduce/eval */

#define BGC_MAX_R 8  /* resources per pod: up to 4 with u8 verdicts, up to 8 with u16 verdicts */
#define BGC_MAX_R8 4 /* widest R of the u8 verdict encoding */

/* ---- row A1/A2: decision codes of the batched form of mutate() --------------------------- */
/* op: kube-core `Operation` as matched at src/admission.rs:272-310 */
enum { BGC_OP_CREATE = 0, BGC_OP_UPDATE = 1, BGC_OP_DELETE = 2, BGC_OP_OTHER = 3 };
/* code = class << 4 | reason */
enum { BGC_ALLOW = 0, BGC_ALLOW_PATCH = 1, BGC_DENY = 2, BGC_INVALID = 3 };
enum {
  BGC_R_OK = 0,                   /* src/admission.rs:427,429 */
  BGC_R_NO_USERNAME = 1,          /* :251 */
  BGC_R_NOT_IN_GROUP = 2,         /* :277 */
  BGC_R_NORMAL_DELETE = 3,        /* :287 */
  BGC_R_NORMAL_UPDATE = 4,        /* :298 */
  BGC_R_INVALID_OPERATION = 5,    /* :306 */
  BGC_R_NO_RESOURCE_NAME = 6,     /* :324 */
  BGC_R_NAME_MISMATCH = 7,        /* :333 */
  BGC_R_NOT_USERBOOTSTRAP = 8,    /* :344-345 */
  BGC_R_ADMIN_EMPTY_KUBE_USERNAME = 9,   /* :369 */
  BGC_R_NORMAL_SETS_QUOTA = 10,          /* :379 */
  BGC_R_NORMAL_SETS_ROLEBINDING = 11,    /* :420 */
  BGC_R_ADMIN_DELETE_EARLY = 12,  /* :293 */
  BGC_R_NO_OBJECT = 13            /* :317 */
};
/* patch_mask bits, in the reference's push order */
enum {
  BGC_P_USER = 1,     /* add /spec/kube_username            src/admission.rs:354-357 */
  BGC_P_RB = 2,       /* add /spec/rolebinding {} + default src/admission.rs:387-416 */
  BGC_P_SUBJ_ORIG = 4 /* default subject = original name    src/admission.rs:394     */
};
/* predicate bits of one request (flags column), SURVEY.md §8 row A2 */
enum {
  BGC_F_HAS_USERNAME = 1 << 0,           /* src/admission.rs:248 */
  BGC_F_IS_NORMAL = 1 << 1,              /* :219 */
  BGC_F_IN_GROUP = 1 << 2,               /* :263-270 */
  BGC_F_HAS_OBJECT = 1 << 3,             /* :312 */
  BGC_F_HAS_NAME = 1 << 4,               /* :321 */
  BGC_F_NAME_EQ = 1 << 5,                /* :332 */
  BGC_F_PARSE_OK = 1 << 6,               /* :341 */
  BGC_F_KUBE_USERNAME_NONEMPTY = 1 << 7, /* :362-367 */
  BGC_F_HAS_QUOTA = 1 << 8,              /* :376 */
  BGC_F_HAS_ROLEBINDING = 1 << 9         /* :386 */
};

/* ---- rows A3/A4: resource columns ---------------------------------------------------------
 * Canonical integer units for the key set the reference writes (src/synchronizer.rs:249-281):
 * cpu → millicores, memory → bytes, extended resources (nvidia.com/gpu, mig-1g.10gb) → count. */
enum { BGC_UNIT_COUNT = 0, BGC_UNIT_MILLI = 1, BGC_UNIT_BYTES = 2 };

#define BGC_VERDICT_ALLOW 0x00u
#define BGC_VERDICT_BAD_NS 0xFFu     /* ns_id >= NS (u8 form) */
#define BGC_VERDICT16_BAD_NS 0xFFFFu /* ns_id >= NS (u16 form) */
/* verdict = missing << 4 | exceeded, one bit per resource r < R.
 * exceeded bit r: namespace constrains r, the pod states r, req > 0 and used + req > hard, decided
 * without 64-bit overflow (as Kubernetes' Quantity arithmetic would);
 * missing bit r : namespace constrains r, r is a compute resource (MILLI/BYTES) and the pod does
 * not state it (req_mask bit clear).  0 = allow. */

/* flags of bgc_usage_reduce* */
#define BGC_REDUCE_GROUPED 1u    /* hint: pods arrive grouped by namespace (any order of groups); the host-pointer
                                    entry point infers it from a sample of the batch when not given */
#define BGC_REDUCE_ACCUMULATE 2u /* add to the current usage table instead of zeroing it first */

typedef struct bgc_ctx bgc_ctx;

typedef struct bgc_timings {
  uint64_t h2d_ns;    /* device time of host→device copies of the last host-pointer call */
  uint64_t kernel_ns; /* device time of the kernels of the last call */
  uint64_t d2h_ns;    /* device time of device→host copies of the last host-pointer call */
  uint64_t launches;  /* kernels launched by this ctx since creation */
} bgc_timings;

/* ---- context (replaces the reference's Arc<AppState>, src/admission.rs:128-131,145-147) ---- */
int bgc_ctx_create(int device, bgc_ctx **out);
int bgc_ctx_destroy(bgc_ctx *ctx);
const char *bgc_last_error(const bgc_ctx *ctx);
/* stream the `_dev` entry points enqueue on: any cudaStream_t, NULL being the legacy default
 * stream (what PyTorch uses unless told otherwise); a new ctx starts on its own stream */
int bgc_ctx_set_stream(bgc_ctx *ctx, void *cuda_stream);
/* back to the context's own non-blocking stream */
int bgc_ctx_reset_stream(bgc_ctx *ctx);
int bgc_ctx_sync(bgc_ctx *ctx);
int bgc_timings_get(bgc_ctx *ctx, bgc_timings *out);

/* ---- row A1: K0 policy_eval — batched mutate() decision, src/admission.rs:241-431 ----------
 * in : op[n] (BGC_OP_*), flags[n] (BGC_F_*);  out: code[n], patch_mask[n]. */
int bgc_policy_eval(bgc_ctx *ctx, const uint8_t *op, const uint16_t *flags, size_t n,
                    uint8_t *code, uint8_t *patch_mask);
int bgc_policy_eval_dev(bgc_ctx *ctx, const uint8_t *d_op, const uint16_t *d_flags, size_t n,
                        uint8_t *d_code, uint8_t *d_patch_mask);

/* ---- quota table: the `hard` map of the ResourceQuota the reference writes
 * (src/synchronizer.rs:249-281 → src/controller.rs:89-110).  hard is [R][NS] (host), hard_mask[ns]
 * bit r = resource r is constrained in namespace ns (0 = namespace has no quota object, cf.
 * src/controller.rs:90); unit[r] is BGC_UNIT_* (MILLI/BYTES columns are "compute resources": a
 * pod that omits them in a namespace that constrains them is rejected as `missing`).
 * Resets usage to zero. */
int bgc_quota_table_set(bgc_ctx *ctx, uint32_t ns_count, uint32_t R, const int64_t *hard,
                        const uint8_t *hard_mask, const uint8_t *unit);

/* ---- row A3: K1 usage_reduce — used[r][ns] = Σ req[r][i] over pods of ns (wrapping i64) ------
 * No reference function; the Kubernetes ResourceQuota controller computes this
 * for the object written at src/controller.rs:89-110.  req is an array of R column pointers.
 * used_out: [R][NS] or NULL to leave the result on the device only. */
int bgc_usage_reduce(bgc_ctx *ctx, const uint32_t *ns_id, const int64_t *const *req, size_t n,
                     unsigned flags, int64_t *used_out);
int bgc_usage_reduce_dev(bgc_ctx *ctx, const uint32_t *d_ns_id, const int64_t *const *d_req,
                         size_t n, unsigned flags, int64_t *d_used_out);
int bgc_usage_get(bgc_ctx *ctx, int64_t *used_out /* host [R][NS] */);
int bgc_usage_set(bgc_ctx *ctx, const int64_t *used /* host [R][NS] */);
/* device [R][NS] forms on the context's stream — what the index-sharded multi-GPU variant
 *  all-reduces between usage_reduce and quota_eval */
int bgc_usage_get_dev(bgc_ctx *ctx, int64_t *d_used_out);
int bgc_usage_set_dev(bgc_ctx *ctx, const int64_t *d_used);

/* ---- row A4: K2 quota_eval — snapshot verdict per candidate pod ------------------------------
 * No reference function; semantic model = kube-apiserver ResourceQuota
 * admission over the object from src/controller.rs:89-110.  req_mask may be NULL (every pod
 * states every resource). */
int bgc_quota_eval(bgc_ctx *ctx, const uint32_t *ns_id, const int64_t *const *req,
                   const uint8_t *req_mask, size_t n, uint8_t *verdict);
int bgc_quota_eval_dev(bgc_ctx *ctx, const uint32_t *d_ns_id, const int64_t *const *d_req,
                       const uint8_t *d_req_mask, size_t n, uint8_t *d_verdict);
/* u16 form for up to 8 resources — e.g. the six pod-relevant keys of src/synchronizer.rs:249-281 in
 * one pass: verdict = missing << 8 | exceeded.  Works for any R <= BGC_MAX_R; the
 * u8 entry points above refuse R > 4. */
int bgc_quota_eval16(bgc_ctx *ctx, const uint32_t *ns_id, const int64_t *const *req,
                     const uint8_t *req_mask, size_t n, uint16_t *verdict);
int bgc_quota_eval16_dev(bgc_ctx *ctx, const uint32_t *d_ns_id, const int64_t *const *d_req,
                         const uint8_t *d_req_mask, size_t n, uint16_t *d_verdict);

/* ---- rows A3/A4 over NARROW columns ------------------------------------------------------------
 * Same results as bgc_usage_reduce / bgc_quota_eval, for callers whose flattener (row A5) can emit compact
 * columns: end to end the path is PCIe-bound (57 bytes per pod-step with int64 columns), and a GPU count
 * fits u8, cpu millicores u16/u32, memory is a multiple of Mi, a namespace id fits u16 up to 65,536
 * namespaces — the key set of src/synchronizer.rs:249-281 in 9 bytes per pod instead of 28.
 *   value of element x of column r = (int64) x << req[r].shift, wrapping; x is zero- (U8/U16/U32) or
 *   sign-extended (I32); ns_type is BGC_T_U32 or BGC_T_U16 (ids >= NS are unknown namespaces as usual).
 * The columns are widened inside the kernels' loads (nothing is expanded in memory), so the narrow bytes
 * are what crosses PCIe and what is read from HBM.  Input order does not matter (arrival-order kernel;
 * flags: BGC_REDUCE_ACCUMULATE only).  R and the table come from bgc_quota_table_set as for the int64
 * forms; verdict encoding as bgc_quota_eval (R <= 4). */
enum { BGC_T_I64 = 0, BGC_T_U32 = 1, BGC_T_U16 = 2, BGC_T_U8 = 3, BGC_T_I32 = 4 };
typedef struct bgc_col {
  const void *data; /* n elements of `type` */
  uint8_t type;     /* BGC_T_* */
  uint8_t shift;    /* 0..63 */
  uint8_t reserved[6];
} bgc_col;
int bgc_usage_reduce_typed(bgc_ctx *ctx, const void *ns_id, int ns_type, const bgc_col *req, size_t n,
                           unsigned flags, int64_t *used_out);
int bgc_usage_reduce_typed_dev(bgc_ctx *ctx, const void *d_ns_id, int ns_type, const bgc_col *d_req,
                               size_t n, unsigned flags, int64_t *d_used_out);
int bgc_quota_eval_typed(bgc_ctx *ctx, const void *ns_id, int ns_type, const bgc_col *req,
                         const uint8_t *req_mask, size_t n, uint8_t *verdict);
int bgc_quota_eval_typed_dev(bgc_ctx *ctx, const void *d_ns_id, int ns_type, const bgc_col *d_req,
                             const uint8_t *d_req_mask, size_t n, uint8_t *d_verdict);

/* ---- row N1: K3 admit_greedy — sequential admission inside one batch -----------
 * No reference function; this is what kube-apiserver does one pod at a time against the
 * ResourceQuota the reference writes (src/controller.rs:89-110): per namespace, in arrival order, a
 * pod is admitted iff its verdict (same encoding and rules as bgc_quota_eval) against the CURRENT
 * usage is 0, and only admitted pods add their requests to the usage table, which is updated in
 * place.  Order-dependent, so the checker is the literal loop.  n < 2^31, R <= 4 (u8 verdicts; see the u16
 * form below).  Pods are grouped by namespace with the library's own stable radix sort (no CUB). */
int bgc_admit_greedy(bgc_ctx *ctx, const uint32_t *ns_id, const int64_t *const *req,
                     const uint8_t *req_mask, size_t n, uint8_t *verdict);
int bgc_admit_greedy_dev(bgc_ctx *ctx, const uint32_t *d_ns_id, const int64_t *const *d_req,
                         const uint8_t *d_req_mask, size_t n, uint8_t *d_verdict);
/* u16 form (verdict = missing << 8 | exceeded) for up to BGC_MAX_R resources — the six pod-relevant keys of
 * src/synchronizer.rs:249-281 in one pass; works for any R <= 8, the u8 entry points refuse R > 4 */
int bgc_admit_greedy16(bgc_ctx *ctx, const uint32_t *ns_id, const int64_t *const *req,
                       const uint8_t *req_mask, size_t n, uint16_t *verdict);
int bgc_admit_greedy16_dev(bgc_ctx *ctx, const uint32_t *d_ns_id, const int64_t *const *d_req,
                           const uint8_t *d_req_mask, size_t n, uint16_t *d_verdict);

/* ---- allocation-free usage reduction ---------------------------------------------------------
 * No counterpart in the reference.  For callers that cannot allocate device memory when they
 * call (all of HBM already handed to their own pool, or a latency-critical handler that must not
 * trigger cudaMalloc's device-wide synchronisation): the same result as bgc_usage_reduce_dev, but
 * without a bgc_ctx and without a single allocation — the accumulation table is a static device
 * array of BGC_STATIC_MAX_NS namespaces (512 KiB per GPU, held from library load to process exit).
 *   - runs on the CURRENT device, on `cuda_stream`; inputs and d_used_out ([R][ns_count]) are
 *     device pointers owned by the caller;
 *   - ns_count > BGC_STATIC_MAX_NS or R outside 1..4 → BGC_E_BADARG, message in errbuf
 *     (errbuf may be NULL);
 *   - the first call on a device does a one-time set-up (resolves the array's device address and
 *     loads every kernel a later call can launch); make that call before any stream capture on the
 *     device, and do not cudaDeviceReset() a device afterwards (the cached address would be stale);
 *   - all calls on a device share the one table, which keeps the last call's contents: each call
 *     clears what it uses first, and calls that overlap on the GPU race — serialise them;
 *   - flags: BGC_REDUCE_GROUPED only. */
#define BGC_STATIC_MAX_NS 16384
int bgc_usage_reduce_static_dev(const uint32_t *d_ns_id, const int64_t *const *d_req, uint32_t R,
                                size_t n, uint32_t ns_count, unsigned flags, int64_t *d_used_out,
                                void *cuda_stream, char *errbuf, size_t errbuf_len);

/* ---- tuning / introspection ---------------------------------------------------------------- */
/* K1 kernel choice: 0 = automatic; otherwise force BGC_K1_* (tests exercise every variant) */
enum { BGC_K1_AUTO = 0, BGC_K1_GLOBAL = 1, BGC_K1_SMEM_DIRECT = 2, BGC_K1_SMEM_HASH = 3, BGC_K1_CLUSTER8 = 4, BGC_K1_CLUSTER16 = 5 };
int bgc_ctx_set_k1_variant(bgc_ctx *ctx, int variant);
/* name of the K1 variant the last bgc_usage_reduce* launched */
const char *bgc_last_k1_variant(const bgc_ctx *ctx);
const char *bgc_version(void);

#ifdef __cplusplus
}
#endif
#endif /* BGC_H */
