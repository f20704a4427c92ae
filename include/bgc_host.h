/* bgc_host.h — host-side half of the boundary (libbgc.so, plain C++ behind a C ABI).
 *
 * Mirrors the reference's handler interface for the admission path: the same inputs the
 * reference's `mutate(req, state)` reads (src/admission.rs:241-431), flattened on the host into
 * the integer columns the kernels consume, and the response fields
 * it produces.  The reference is Rust; with no Rust toolchain in this image the host layer is
 * C++ exporting C symbols, which is what a Rust `extern "C"` block binds to.
 */
#ifndef BGC_HOST_H
#define BGC_HOST_H
#include "bgc.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- row A2: the fields mutate() reads from Config / AdmissionRequest ----------------------- */
/* reference: struct Config, src/admission.rs:22-39 (only what mutate() uses) */
typedef struct bgc_admission_config {
  const char *oidc_username_prefix;          /* src/admission.rs:33  */
  const char *default_role_name;             /* src/admission.rs:35  */
  const char *const *authorized_group_names; /* src/admission.rs:38  */
  size_t n_authorized;
} bgc_admission_config;

/* reference: AdmissionRequest<DynamicObject> as consumed by mutate() */
typedef struct bgc_admission_request {
  const char *username;           /* req.user_info.username; NULL = None   src/admission.rs:248 */
  const char *const *groups;      /* req.user_info.groups;   NULL = None   src/admission.rs:263-267 */
  size_t n_groups;
  int operation;                  /* BGC_OP_*                               src/admission.rs:272 */
  int has_object;                 /* req.object.is_some()                   src/admission.rs:312 */
  const char *name;               /* obj.metadata.name; NULL = None         src/admission.rs:321 */
  int parse_ok;                   /* obj parses as UserBootstrap            src/admission.rs:341 */
  const char *spec_kube_username; /* ub.spec.kube_username; NULL = None     src/admission.rs:362-367 */
  int has_quota;                  /* ub.spec.quota.is_some()                src/admission.rs:376 */
  int has_rolebinding;            /* ub.spec.rolebinding.is_some()          src/admission.rs:386 */
} bgc_admission_request;

/* reference: AdmissionResponse fields set by mutate() (allowed, result.message, patch) */
typedef struct bgc_admission_response {
  uint8_t allowed;     /* 1 for ALLOW / ALLOW_PATCH */
  uint8_t klass;       /* BGC_ALLOW .. BGC_INVALID */
  uint8_t reason;      /* BGC_R_* */
  uint8_t patch_mask;  /* BGC_P_* */
  const char *message; /* static literal from src/admission.rs, NULL when allowed */
} bgc_admission_response;

/* comma_separated_deserialize, src/admission.rs:41-50: split on ',' without trimming, "" → [""].
 * `scratch` must hold strlen(value)+1 bytes; returns the number of items (may exceed cap). */
size_t bgc_split_groups(const char *value, char *scratch, const char **out, size_t cap);

/* strings → (op, flags) columns; the string predicates of src/admission.rs:217-239,263-270,
 * 312,321,332,341,362-367,376,386 */
int bgc_flatten_requests(const bgc_admission_config *cfg, const bgc_admission_request *reqs, size_t n,
                         uint8_t *op, uint16_t *flags);

/* exact message literal of a reason (src/admission.rs:251,277,287,298,306,324,333,369,379,420);
 * NULL for allow reasons; a placeholder for BGC_R_NOT_USERBOOTSTRAP (serde text, unpinned) */
const char *bgc_reason_message(int reason);

/* batched mutate(): flatten on the host, decide on the GPU (bgc_policy_eval), fill responses.
 * Error convention follows the reference: a decision is never an error (deny/invalid are
 * ordinary responses, src/admission.rs:253,279,...); only library failures return non-zero. */
int bgc_mutate_batch(bgc_ctx *ctx, const bgc_admission_config *cfg, const bgc_admission_request *reqs,
                     size_t n, bgc_admission_response *out);

/* JSON-patch text the reference would attach (src/admission.rs:354-357,387-416,429) for one
 * request, in push order.  Returns the length written (excluding NUL), 0 when there is no
 * patch, or a negative BGC_E_* (buffer too small → BGC_E_BADARG).  Byte-level parity with
 * json-patch 1.0 / serde_json is unpinned; field order follows k8s-openapi
 * 0.18 RoleRef/Subject declaration order. */
long bgc_render_patch(const bgc_admission_config *cfg, const bgc_admission_request *req,
                      uint8_t patch_mask, char *buf, size_t cap);

/* ---- row A5: pod specs → resource columns ----------------------------------------------------
 * No reference function (the reference never reads pods).  Rules = public Kubernetes pod-usage
 * semantics for the quota keys written at src/synchronizer.rs:249-281. */
typedef struct bgc_resource_list {
  const char *const *names;      /* e.g. "cpu", "memory", "nvidia.com/gpu" */
  const char *const *quantities; /* Kubernetes quantity strings: "500m", "2", "4Gi", "1e3" */
  size_t n;
} bgc_resource_list;

typedef struct bgc_container_spec {
  bgc_resource_list requests;
  bgc_resource_list limits;
} bgc_container_spec;

enum { BGC_PHASE_PENDING = 0, BGC_PHASE_RUNNING = 1, BGC_PHASE_SUCCEEDED = 2, BGC_PHASE_FAILED = 3, BGC_PHASE_UNKNOWN = 4 };

typedef struct bgc_pod_spec {
  const char *ns; /* namespace name; the reference names it after the lowercased UserBootstrap, src/controller.rs:55-63 */
  int phase;
  const bgc_container_spec *containers;
  size_t n_containers;
  const bgc_container_spec *init_containers;
  size_t n_init_containers;
  const bgc_resource_list *overhead; /* may be NULL */
} bgc_pod_spec;

/* quantity string → canonical integer (BGC_UNIT_MILLI: millicores; otherwise whole units), rounding
 * up like Kubernetes' Quantity.MilliValue()/Value().  Returns BGC_E_BADARG on syntax error or
 * overflow. */
int bgc_parse_quantity(const char *s, int unit, int64_t *out);

/* canonical integer → Kubernetes quantity text, the inverse of bgc_parse_quantity for the forms the
 * reference writes (src/synchronizer.rs:253-277) and Kubernetes prints: MILLI → "<n>" when a whole
 * number of cores else "<n>m"; BYTES → the largest exact binary suffix ("64Gi", "1536Mi") else the
 * plain integer; COUNT → the plain integer.  Returns the length written or BGC_E_BADARG. */
long bgc_format_quantity(int64_t value, int unit, char *buf, size_t cap);

/* The message kube-apiserver's ResourceQuota admission attaches to a rejected pod, for one verdict of
 * bgc_quota_eval / bgc_quota_eval16 (wide != 0: the u16 layout):
 *   missing  → "failed quota: <quota>: must specify k1,k2"
 *   exceeded → "exceeded quota: <quota>, requested: k=v,..., used: k=v,..., limited: k=v,..."
 * (missing wins, as the plugin checks it first; keys in the order given; every failing resource is
 * listed).  <quota> is the ResourceQuota's name — the reference names it after the namespace
 * (src/controller.rs:92).  req / used / hard are the R values of this pod / its namespace.
 * Message wording restated from Kubernetes' plugin from memory — unpinned.  Returns the length
 * written, 0 for an allowed pod, or BGC_E_BADARG. */
long bgc_render_quota_denial(unsigned verdict, int wide, const char *quota_name, const char *const *quota_keys,
                             const uint8_t *unit, uint32_t R, const int64_t *req, const int64_t *used,
                             const int64_t *hard, char *buf, size_t cap);

/* namespace name ↔ dense u32 id */
typedef struct bgc_nsmap bgc_nsmap;
bgc_nsmap *bgc_nsmap_create(void);
void bgc_nsmap_destroy(bgc_nsmap *m);
uint32_t bgc_nsmap_intern(bgc_nsmap *m, const char *name);        /* id, assigning the next one if new */
uint32_t bgc_nsmap_lookup(const bgc_nsmap *m, const char *name);  /* id or 0xFFFFFFFF */
uint32_t bgc_nsmap_size(const bgc_nsmap *m);
const char *bgc_nsmap_name(const bgc_nsmap *m, uint32_t id);

/* Flatten n pod specs into columns.  quota_keys[r] is a ResourceQuota key as written by the
 * reference ("requests.cpu", "limits.memory", "requests.nvidia.com/gpu", ...,
 * src/synchronizer.rs:251-279): the "limits." prefix selects container limits, anything else
 * container requests.  Per resource: max(Σ containers, max over init containers) + overhead.
 * Terminal pods (Succeeded/Failed) are skipped.  req_mask bit r is set when every container states
 * the resource.  Unknown namespaces are interned when `intern` != 0, else get id 0xFFFFFFFF.
 * Returns the number of rows written through *n_out. */
int bgc_flatten_pods(const bgc_pod_spec *pods, size_t n, const char *const *quota_keys,
                     const uint8_t *unit, uint32_t R, bgc_nsmap *map, int intern, uint32_t *ns_id,
                     int64_t *const *req, uint8_t *req_mask, size_t *n_out);

/* ---- compact columns for the narrow-column entry points (bgc_*_typed, bgc.h) -----------------------
 * No counterpart in the reference.  The columns bgc_flatten_pods emits are int64; what crosses PCIe should not
 * be.  bgc_compact_plan picks, for one column, the largest shift (trailing zero bits common to every value, so
 * memory quantities written as <n>Mi / <n>Gi — src/synchronizer.rs:257,265 — lose 20-30 bits) and then the
 * narrowest element type (BGC_T_U8 / U16 / U32 for non-negative values, BGC_T_I32, else BGC_T_I64 with shift 0)
 * that represents every value exactly; bgc_compact_column writes the column in a given representation and
 * returns BGC_E_BADARG if some value would not survive (int64)element << shift.  An all-zero column is U8 / 0. */
int bgc_compact_plan(const int64_t *col, size_t n, uint8_t *type_out, uint8_t *shift_out);
int bgc_compact_column(const int64_t *col, size_t n, uint8_t type, uint8_t shift, void *out);
/* the same for namespace ids: u16 when every id (including the unknown-namespace marker 0xFFFFFFFF, which becomes
 * 0xFFFF) stays distinguishable, i.e. ns_count <= 65535; returns the element type chosen (BGC_T_U16 / BGC_T_U32) */
int bgc_compact_ns_ids(const uint32_t *ns_id, size_t n, uint32_t ns_count, void *out);

/* the `hard` values the reference's synchronizer writes for one sheet row
 * (src/synchronizer.rs:249-281), in canonical units, for the given quota keys:
 * cpu → cpu_request*1000 millicores, memory/storage → <n>Gi bytes, gpu/mig → count.
 * Unknown keys yield BGC_E_BADARG. */
int bgc_hard_from_sheet_row(int64_t cpu_request, int64_t memory_request_gi, int64_t gpu_request,
                            int64_t storage_request_gi, int64_t mig_request,
                            const char *const *quota_keys, uint32_t R, int64_t *hard_out,
                            uint8_t *unit_out);

/* ---- row N4: AdmissionReview v1 JSON in and out ---------------------------------
 * What `mutate_handler` does around `mutate` (src/admission.rs:184-204): the axum `Json` extractor
 * turns the body into AdmissionReview<DynamicObject>, `try_into()` yields the AdmissionRequest (a
 * review without `request` becomes an `invalid` response, :189-197), and `into_review()` wraps the
 * response.  Byte-level parity with kube-core 0.84 / serde_json is unpinned:
 * field order and the Status shape below are restated from memory. */
typedef struct bgc_review bgc_review;

/* NULL + message on malformed JSON (the reference answers 4xx from the extractor). */
bgc_review *bgc_review_parse(const char *json, size_t len, char *err, size_t err_len);
void bgc_review_destroy(bgc_review *r);
/* 1 when the document carries a `request` object (AdmissionReview::try_into succeeds) */
int bgc_review_has_request(const bgc_review *r);
const char *bgc_review_uid(const bgc_review *r);       /* request.uid or "" */
/* request.kind.kind, e.g. "UserBootstrap" or "Pod" ("" when absent) */
const char *bgc_review_kind(const bgc_review *r);
/* the fields mutate() reads; string pointers stay valid until bgc_review_destroy */
int bgc_review_request(const bgc_review *r, bgc_admission_request *out);
/* request.object as a pod spec for bgc_flatten_pods (namespace from request.namespace or
 * object.metadata.namespace); pointers valid until bgc_review_destroy */
int bgc_review_pod(const bgc_review *r, bgc_pod_spec *out);

/* The batched mutate_handler (src/admission.rs:184-204) in ONE call, wire bytes to wire bytes.  One pass per body
 * (OpenMP over the batch, everything in the thread that parsed it): parse, the fields mutate() reads, their predicate
 * bits (row A2), release of the parsed tree; then one K0 launch (bgc_policy_eval) over the batch; then the responses,
 * patches included.  A body that is not a well-formed AdmissionReview, or carries no `request`, is answered `invalid`
 * on its own, as the reference's handler rejects only the offending request (:189-197).  The responses are written
 * back to back into `out` (NUL-terminated each); out_off[i] = offset of response i, out_off[n] = bytes used.
 * Returns BGC_E_BADARG when `out` is too small (out_off[n] then holds the size needed).  stage_ns (optional, 4
 * values): host time of the per-body pass, the decision (gather + K0), rendering, copy-out. */
int bgc_handle_reviews(bgc_ctx *ctx, const bgc_admission_config *cfg, const char *const *bodies,
                       const size_t *body_len, size_t n, char *out, size_t out_cap, size_t *out_off,
                       uint64_t *stage_ns);

/* Pod-create AdmissionReviews wire to wire in ONE call: parse (OpenMP) -> bgc_review_pod -> bgc_flatten_pods (row A5)
 * -> snapshot verdicts on the GPU (bgc_quota_eval, or bgc_quota_eval16 when wide != 0) -> responses, a denied pod
 * carrying bgc_render_quota_denial's message when used / hard ([R][ns_count], as given to bgc_quota_table_set /
 * read by bgc_usage_get) are passed, else "exceeded quota".  Not a reference path: the reference's webhook never sees
 * pods; the seam is the same handler, src/admission.rs:184-204.  A body that does not parse, or
 * carries no pod object, is answered `invalid` on its own; a pod of a namespace the map does not know is denied.
 * Output layout as bgc_handle_reviews; stage_ns (optional, 5 values): the per-body pass (parse, pod spec, flatten into
 * the body's row, release of the parsed tree — all in the thread that parsed it), gathering the rows, verdicts,
 * rendering, copy-out.
 * The quota table of `ctx` must be set for `quota_keys`' columns.  A terminal pod in a create review -> BGC_E_BADARG. */
int bgc_handle_pod_reviews(bgc_ctx *ctx, const char *const *bodies, const size_t *body_len, size_t n,
                           const char *const *quota_keys, const uint8_t *unit, uint32_t R, bgc_nsmap *map, int wide,
                           const int64_t *used, const int64_t *hard, uint32_t ns_count, char *out, size_t out_cap,
                           size_t *out_off, uint64_t *stage_ns);

/* AdmissionReview response text for one decision.  patch_json = bgc_render_patch output or NULL.
 * allow   → {"uid","allowed":true,"status":{}} (+ "patch": base64, "patchType":"JSONPatch")
 * deny    → allowed:false, status.message = the reference's literal
 * invalid → uid "", allowed:false, status {"status":"Failure","code":400,"message",...,"reason":"InvalidRequest"}
 * Returns the length written or BGC_E_BADARG when the buffer is too small. */
long bgc_render_review_response(const char *uid, const bgc_admission_response *resp, const char *message_override,
                                const char *patch_json, char *buf, size_t cap);

/* ---- row N3: reconcile plan ------------------------------------------------------
 * The objects `reconcile()` server-side-applies for one UserBootstrap (src/controller.rs:50-155),
 * as request path + body pairs in the reference's order: Namespace (name lower-cased, :55-63),
 * ResourceQuota only if spec.quota is set (:89-110), Role if set (:112-124, owner reference
 * overwritten), RoleBinding only if set AND status.synchronized_with_sheet (:126-152); every object
 * carries the controller owner reference (:52); paths carry PatchParams::apply(PATCH_MANAGER).force()
 * (:22,:67).  Wire format only — nothing is sent.  Field order of the bodies is unpinned. */
typedef struct bgc_plan bgc_plan;
/* NULL + message for malformed JSON or a missing metadata.name / metadata.uid */
bgc_plan *bgc_reconcile_plan(const char *userbootstrap_json, size_t len, char *err, size_t err_len);
void bgc_plan_destroy(bgc_plan *p);
size_t bgc_plan_count(const bgc_plan *p);
/* ns is NULL for the cluster-scoped Namespace object; pointers valid until bgc_plan_destroy */
int bgc_plan_item(const bgc_plan *p, size_t i, const char **kind, const char **ns, const char **name,
                  const char **path, const char **body);
/* Action::requeue: 30 s after success (:154), 3 s after an error (:174) */
int bgc_plan_requeue_seconds(int failed);

/* ---- row N2: the quota-table producer ------------------------------------------
 * Batch form of the synchronizer's sheet logic, reference src/synchronizer.rs:63-169,208-281:
 * the Google-Form CSV export is parsed with the reference's header heuristics, filtered by GPU
 * server, and for every UserBootstrap name the LAST authorised matching row yields the seven `hard`
 * quantities.  The result feeds bgc_quota_table_set. */
typedef struct bgc_sheet bgc_sheet;

typedef struct bgc_sheet_row { /* struct Row, src/synchronizer.rs:63-93 */
  const char *name, *department, *id_username, *gpu_server, *authorized;
  int64_t gpu_request, cpu_request, memory_request, storage_request, mig_request;
} bgc_sheet_row;

/* try_infer_header, src/synchronizer.rs:97-143: canonical field name for one header cell, or NULL
 * for an unknown header (the reference fails the whole parse with CsvHeaderError). */
const char *bgc_sheet_infer_header(const char *header);

/* parse_csv, src/synchronizer.rs:145-169: returns NULL and a message in err on a header error;
 * rows that fail to deserialise (missing column, non-integer request, wrong field count) are
 * skipped, as the reference does with a warning. */
bgc_sheet *bgc_sheet_parse_csv(const char *content, size_t len, char *err, size_t err_len);
void bgc_sheet_destroy(bgc_sheet *s);
size_t bgc_sheet_row_count(const bgc_sheet *s);
size_t bgc_sheet_skipped_rows(const bgc_sheet *s);
int bgc_sheet_get_row(const bgc_sheet *s, size_t i, bgc_sheet_row *out);

/* src/synchronizer.rs:208-236: keep rows whose gpu_server CONTAINS server_name (not exact), then
 * for each name take the last row with authorized.trim().to_lowercase() == "o" and
 * id_username == name.  row_index[i] = index into the sheet, or -1 when there is no such row
 * (the reference `continue`s: quota untouched). */
int bgc_sheet_match(const bgc_sheet *s, const char *server_name, const char *const *names, size_t n,
                    long *row_index);

/* The seven (key, quantity string) pairs of src/synchronizer.rs:249-281 for one row, in BTreeMap
 * (sorted-key) order as serde_json would emit them, and the same values in canonical units.
 * keys[k], quantities[k] point into buffers owned by `scratch` (>= 512 bytes). */
#define BGC_SHEET_KEYS 7
int bgc_sheet_row_quota(const bgc_sheet_row *row, const char **keys, const char **quantities,
                        int64_t *canonical, char *scratch, size_t scratch_len);

#ifdef __cplusplus
}
#endif
#endif /* BGC_HOST_H */
