/* examples/mutate_batch.c — the C ABI from plain C: a batch of UserBootstrap admission requests is
 * flattened on the host, decided on the GPU and printed; then a handful of pods are checked against a
 * quota table, and three raw AdmissionReview bodies go through the one-call handler.  Build (after `make -C bacchus-gpu-controller_b200/csrc`):
 *   gcc -std=c99 -Iinclude examples/mutate_batch.c -Lbacchus-gpu-controller_b200 -lbgc \
 *       -Wl,-rpath,$PWD/bacchus-gpu-controller_b200 -o /tmp/mutate_batch && /tmp/mutate_batch
 * Needs a CUDA device: there is no CPU fallback (bgc_ctx_create fails with BGC_E_CUDA otherwise). */
#include <stdio.h>

#include "bgc.h"
#include "bgc_host.h"

int main(void) {
  bgc_ctx *ctx = NULL;
  int rc = bgc_ctx_create(0, &ctx);
  if (rc != BGC_OK) {
    fprintf(stderr, "bgc_ctx_create: %d (no CUDA device?)\n", rc);
    return 2;
  }
  /* Config as the chart sets it (charts/bacchus-gpu-controller/values.yaml:84-89) */
  const char *authorized[] = {"gpu", "admin"};
  bgc_admission_config cfg = {"oidc:", "edit", authorized, 2};
  const char *groups_gpu[] = {"system:authenticated", "gpu"};
  const char *groups_none[] = {"system:authenticated"};
  bgc_admission_request reqs[3] = {
      {"oidc:alice", groups_gpu, 2, BGC_OP_CREATE, 1, "alice", 1, NULL, 0, 0},  /* allowed, patched */
      {"oidc:bob", groups_none, 1, BGC_OP_CREATE, 1, "bob", 1, NULL, 0, 0},     /* not in an authorised group */
      {"admin", NULL, 0, BGC_OP_DELETE, 0, NULL, 0, NULL, 0, 0},                /* admin delete */
  };
  bgc_admission_response out[3];
  rc = bgc_mutate_batch(ctx, &cfg, reqs, 3, out);
  if (rc != BGC_OK) {
    fprintf(stderr, "bgc_mutate_batch: %s\n", bgc_last_error(ctx));
    return 1;
  }
  for (int i = 0; i < 3; ++i) {
    char patch[1024];
    long n = bgc_render_patch(&cfg, &reqs[i], out[i].patch_mask, patch, sizeof patch);
    printf("request %d: allowed=%d class=%d reason=%d message=%s patch=%s\n", i, out[i].allowed, out[i].klass,
           out[i].reason, out[i].message ? out[i].message : "-", n > 0 ? patch : "-");
  }

  /* quota: 2 namespaces, 2 resources (gpu count, cpu millicores) */
  const int64_t hard[2 * 2] = {/* gpu */ 2, 1, /* cpu */ 8000, 4000};
  const uint8_t hard_mask[2] = {3, 3}, unit[2] = {BGC_UNIT_COUNT, BGC_UNIT_MILLI};
  rc = bgc_quota_table_set(ctx, 2, 2, hard, hard_mask, unit);
  const uint32_t run_ns[3] = {0, 0, 1};
  const int64_t run_gpu[3] = {1, 0, 1}, run_cpu[3] = {2000, 4000, 1000};
  const int64_t *run_req[2] = {run_gpu, run_cpu};
  int64_t used[4];
  if (rc == BGC_OK) rc = bgc_usage_reduce(ctx, run_ns, run_req, 3, 0, used);
  const uint32_t cand_ns[3] = {0, 0, 1};
  const int64_t cand_gpu[3] = {1, 2, 1}, cand_cpu[3] = {2000, 1000, 500};
  const int64_t *cand_req[2] = {cand_gpu, cand_cpu};
  uint8_t verdict[3];
  if (rc == BGC_OK) rc = bgc_quota_eval(ctx, cand_ns, cand_req, NULL, 3, verdict);
  if (rc != BGC_OK) {
    fprintf(stderr, "quota path: %s\n", bgc_last_error(ctx));
    return 1;
  }
  printf("used gpu=[%lld,%lld] cpu=[%lld,%lld]; verdicts = %u %u %u (0 = allow, bit r = resource r exceeded)\n",
         (long long)used[0], (long long)used[1], (long long)used[2], (long long)used[3], verdict[0], verdict[1], verdict[2]);

  /* the same pods as narrow columns (bgc_*_typed): u16 namespace ids, u8 gpu counts, u16 cpu millicores —
   * 5 bytes per pod instead of 20 across PCIe; values are (int64) element << shift */
  const uint16_t run_ns16[3] = {0, 0, 1}, cand_ns16[3] = {0, 0, 1};
  const uint8_t run_gpu8[3] = {1, 0, 1}, cand_gpu8[3] = {1, 2, 1};
  const uint16_t run_cpu16[3] = {2000, 4000, 1000}, cand_cpu16[3] = {2000, 1000, 500};
  const bgc_col run_cols[2] = {{run_gpu8, BGC_T_U8, 0, {0}}, {run_cpu16, BGC_T_U16, 0, {0}}};
  const bgc_col cand_cols[2] = {{cand_gpu8, BGC_T_U8, 0, {0}}, {cand_cpu16, BGC_T_U16, 0, {0}}};
  int64_t used_t[4];
  uint8_t verdict_t[3];
  rc = bgc_usage_reduce_typed(ctx, run_ns16, BGC_T_U16, run_cols, 3, 0, used_t);
  if (rc == BGC_OK) rc = bgc_quota_eval_typed(ctx, cand_ns16, BGC_T_U16, cand_cols, NULL, 3, verdict_t);
  if (rc != BGC_OK) {
    fprintf(stderr, "narrow-column path: %s\n", bgc_last_error(ctx));
    return 1;
  }
  printf("narrow columns: used gpu=[%lld,%lld] cpu=[%lld,%lld]; verdicts = %u %u %u\n", (long long)used_t[0], (long long)used_t[1],
         (long long)used_t[2], (long long)used_t[3], verdict_t[0], verdict_t[1], verdict_t[2]);
  /* The handler itself, wire bytes to wire bytes (src/admission.rs:184-204): two review bodies and a broken one. */
  const char *bodies[3] = {
      "{\"apiVersion\":\"admission.k8s.io/v1\",\"kind\":\"AdmissionReview\",\"request\":{\"uid\":\"u-1\",\"operation\":\"CREATE\","
      "\"userInfo\":{\"username\":\"oidc:alice\",\"groups\":[\"gpu\"]},\"object\":{\"metadata\":{\"name\":\"alice\"},\"spec\":{}}}}",
      "{\"apiVersion\":\"admission.k8s.io/v1\",\"kind\":\"AdmissionReview\",\"request\":{\"uid\":\"u-2\",\"operation\":\"CREATE\","
      "\"userInfo\":{\"username\":\"oidc:bob\",\"groups\":[]},\"object\":{\"metadata\":{\"name\":\"bob\"},\"spec\":{}}}}",
      "{\"apiVersion\":"};
  size_t lens[3], off[4];
  char text[4096];
  for (int i = 0; i < 3; ++i) {
    lens[i] = 0;
    while (bodies[i][lens[i]]) ++lens[i];
  }
  rc = bgc_handle_reviews(ctx, &cfg, bodies, lens, 3, text, sizeof text, off, NULL);
  if (rc != BGC_OK) {
    fprintf(stderr, "bgc_handle_reviews: %s\n", bgc_last_error(ctx));
    return 1;
  }
  for (int i = 0; i < 3; ++i) printf("review %d -> %s\n", i, text + off[i]);
  bgc_ctx_destroy(ctx);
  return 0;
}
