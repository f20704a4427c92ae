/* oracle/policy_ref.c — TEST INFRASTRUCTURE ONLY.  See policy_ref.h for the rules.
 *
 * Restates /root/reference/src/admission.rs:206-431 on the CPU.  PARITY UNPINNED (the
 * reference holds no golden vectors and cannot be built here); the restatement follows the
 * source text and cites the line behind every branch.
 */
#include "policy_ref.h"
#include <string.h>
#include <stdio.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static const char *const k_messages[] = {
    /* 0  */ NULL,
    /* 1  admission.rs:251 */ "cannot get requester's username from request",
    /* 2  admission.rs:277 */ "user is not in authorized group",
    /* 3  admission.rs:287 */ "normal user is not allowed to delete resource",
    /* 4  admission.rs:298 */ "normal user is not allowed to update resource",
    /* 5  admission.rs:306 */ "invalid operation",
    /* 6  admission.rs:324 */ "cannot get resource name from request",
    /* 7  admission.rs:333 */ "username not match with resource name",
    /* 8  admission.rs:345 */ "<serde error text: object is not a UserBootstrap>",
    /* 9  admission.rs:369 */ "kube_username field is empty. you are an admin, so fill it",
    /* 10 admission.rs:379 */ "quota field is not empty. you are a normal user, so leave it empty",
    /* 11 admission.rs:420 */ "rolebinding field is not empty. you are a normal user, so leave it empty",
    /* 12 */ NULL,
    /* 13 */ NULL,
};

const char *policy_ref_message(int reason) {
  if (reason < 0 || reason > 13) return NULL;
  return k_messages[reason];
}

static int starts_with(const char *s, const char *prefix) {
  size_t lp = strlen(prefix);
  return strncmp(s, prefix, lp) == 0; /* str::starts_with: "" is a prefix of everything */
}

/* admission.rs:263-270 — any request group is contained (exact, case-sensitive) in
 * authorized_group_names; groups None → empty list → false. */
static int in_group(const policy_ref_request *req, const policy_ref_config *cfg) {
  if (!req->groups) return 0;
  for (size_t g = 0; g < req->n_groups; ++g)
    for (size_t a = 0; a < cfg->n_authorized; ++a)
      if (strcmp(req->groups[g], cfg->authorized_group_names[a]) == 0) return 1;
  return 0;
}

static void finish(policy_ref_decision *out, int klass, int reason) {
  out->klass = (uint8_t)klass;
  out->reason = (uint8_t)reason;
  out->message = k_messages[reason];
  if (klass == POLICY_DENY || klass == POLICY_INVALID) {
    /* deny()/invalid() responses carry no patch (admission.rs returns before :429) */
    out->patch_mask = 0;
    out->kube_username_patch[0] = 0;
    out->subject_name[0] = 0;
  }
}

void policy_ref_mutate(const policy_ref_request *req, const policy_ref_config *cfg,
                       policy_ref_decision *out) {
  memset(out, 0, sizeof *out);

  /* :246-255 username None → invalid */
  if (!req->username) { finish(out, POLICY_INVALID, POLICY_R_NO_USERNAME); return; }

  /* :217-239 Username::from */
  const char *original_username = req->username;
  const char *kube_username;
  int normal;
  if (starts_with(req->username, cfg->oidc_username_prefix)) { /* :219 */
    kube_username = req->username + strlen(cfg->oidc_username_prefix); /* strip_prefix :225 */
    normal = 1;                                                 /* :228 */
  } else {
    kube_username = req->username;                              /* :234 */
    normal = 0;                                                 /* :235 */
  }

  int is_in_group = in_group(req, cfg); /* :263-270 */

  switch (req->operation) { /* :272 */
    case POLICY_OP_CREATE: /* :273-283 */
      if (normal && !is_in_group) { finish(out, POLICY_DENY, POLICY_R_NOT_IN_GROUP); return; }
      break;
    case POLICY_OP_DELETE: /* :284-294 */
      if (normal) { finish(out, POLICY_DENY, POLICY_R_NORMAL_DELETE); return; }
      finish(out, POLICY_ALLOW, POLICY_R_ADMIN_DELETE_EARLY); /* :293 */
      return;
    case POLICY_OP_UPDATE: /* :295-304 */
      if (normal) { finish(out, POLICY_DENY, POLICY_R_NORMAL_UPDATE); return; }
      break;
    default: /* :305-309 */
      finish(out, POLICY_INVALID, POLICY_R_INVALID_OPERATION);
      return;
  }

  /* :312-318 object None → untouched allow */
  if (!req->has_object) { finish(out, POLICY_ALLOW, POLICY_R_NO_OBJECT); return; }

  /* :321-328 */
  if (!req->name) { finish(out, POLICY_INVALID, POLICY_R_NO_RESOURCE_NAME); return; }

  /* :331-338 case-sensitive comparison of the stripped username with metadata.name */
  if (normal && strcmp(kube_username, req->name) != 0) {
    finish(out, POLICY_DENY, POLICY_R_NAME_MISMATCH);
    return;
  }

  /* :341-347 */
  if (!req->parse_ok) { finish(out, POLICY_INVALID, POLICY_R_NOT_USERBOOTSTRAP); return; }

  unsigned patches = 0;
  if (normal) { /* :352-358 */
    patches |= POLICY_P_USER;
    snprintf(out->kube_username_patch, sizeof out->kube_username_patch, "%s", kube_username);
  } else {      /* :359-373 unwrap_or("").is_empty() */
    if (!req->spec_kube_username || req->spec_kube_username[0] == 0) {
      finish(out, POLICY_DENY, POLICY_R_ADMIN_EMPTY_KUBE_USERNAME);
      return;
    }
  }

  /* :376-383 */
  if (req->has_quota && normal) { finish(out, POLICY_DENY, POLICY_R_NORMAL_SETS_QUOTA); return; }

  if (!req->has_rolebinding) { /* :386-416 */
    patches |= POLICY_P_RB;
    const char *subject;
    if (normal) { subject = original_username; patches |= POLICY_P_SUBJ_ORIG; } /* :394 */
    else subject = req->spec_kube_username;                                      /* :396 */
    snprintf(out->subject_name, sizeof out->subject_name, "%s", subject);
  } else if (normal) {         /* :417-424 */
    finish(out, POLICY_DENY, POLICY_R_NORMAL_SETS_ROLEBINDING);
    return;
  }

  out->patch_mask = (uint8_t)patches;
  finish(out, patches ? POLICY_ALLOW_PATCH : POLICY_ALLOW, POLICY_R_OK); /* :426-430 */
}

void policy_ref_flatten(const policy_ref_request *req, const policy_ref_config *cfg,
                        uint8_t *op, uint16_t *flags) {
  unsigned f = 0;
  const char *kube_username = NULL;
  if (req->username) {
    f |= POLICY_F_HAS_USERNAME;
    if (starts_with(req->username, cfg->oidc_username_prefix)) {
      f |= POLICY_F_IS_NORMAL;
      kube_username = req->username + strlen(cfg->oidc_username_prefix);
    } else {
      kube_username = req->username;
    }
  }
  if (in_group(req, cfg)) f |= POLICY_F_IN_GROUP;
  if (req->has_object) {
    f |= POLICY_F_HAS_OBJECT;
    if (req->name) {
      f |= POLICY_F_HAS_NAME;
      if (kube_username && strcmp(kube_username, req->name) == 0) f |= POLICY_F_NAME_EQ;
    }
    if (req->parse_ok) {
      f |= POLICY_F_PARSE_OK;
      if (req->spec_kube_username && req->spec_kube_username[0]) f |= POLICY_F_KUBE_USERNAME_NONEMPTY;
      if (req->has_quota) f |= POLICY_F_HAS_QUOTA;
      if (req->has_rolebinding) f |= POLICY_F_HAS_ROLEBINDING;
    }
  }
  *op = (uint8_t)(req->operation >= 0 && req->operation <= 2 ? req->operation : POLICY_OP_OTHER);
  *flags = (uint16_t)f;
}

void policy_ref_bits(uint8_t op, uint16_t flags, uint8_t *code, uint8_t *patch_mask) {
#define RET(k, r, p) do { *code = (uint8_t)(((k) << 4) | (r)); *patch_mask = (uint8_t)(p); return; } while (0)
  const int normal = (flags & POLICY_F_IS_NORMAL) != 0;
  if (!(flags & POLICY_F_HAS_USERNAME)) RET(POLICY_INVALID, POLICY_R_NO_USERNAME, 0);      /* :248-255 */
  switch (op) {
    case POLICY_OP_CREATE:
      if (normal && !(flags & POLICY_F_IN_GROUP)) RET(POLICY_DENY, POLICY_R_NOT_IN_GROUP, 0); /* :276-280 */
      break;
    case POLICY_OP_DELETE:
      if (normal) RET(POLICY_DENY, POLICY_R_NORMAL_DELETE, 0);                              /* :286-290 */
      RET(POLICY_ALLOW, POLICY_R_ADMIN_DELETE_EARLY, 0);                                    /* :293 */
    case POLICY_OP_UPDATE:
      if (normal) RET(POLICY_DENY, POLICY_R_NORMAL_UPDATE, 0);                              /* :297-301 */
      break;
    default:
      RET(POLICY_INVALID, POLICY_R_INVALID_OPERATION, 0);                                   /* :305-309 */
  }
  if (!(flags & POLICY_F_HAS_OBJECT)) RET(POLICY_ALLOW, POLICY_R_NO_OBJECT, 0);             /* :312-318 */
  if (!(flags & POLICY_F_HAS_NAME)) RET(POLICY_INVALID, POLICY_R_NO_RESOURCE_NAME, 0);      /* :321-328 */
  if (normal && !(flags & POLICY_F_NAME_EQ)) RET(POLICY_DENY, POLICY_R_NAME_MISMATCH, 0);   /* :331-338 */
  if (!(flags & POLICY_F_PARSE_OK)) RET(POLICY_INVALID, POLICY_R_NOT_USERBOOTSTRAP, 0);     /* :341-347 */
  unsigned p = 0;
  if (normal) p |= POLICY_P_USER;                                                           /* :352-358 */
  else if (!(flags & POLICY_F_KUBE_USERNAME_NONEMPTY))
    RET(POLICY_DENY, POLICY_R_ADMIN_EMPTY_KUBE_USERNAME, 0);                                /* :359-373 */
  if ((flags & POLICY_F_HAS_QUOTA) && normal) RET(POLICY_DENY, POLICY_R_NORMAL_SETS_QUOTA, 0); /* :376-383 */
  if (!(flags & POLICY_F_HAS_ROLEBINDING)) {                                                /* :386-416 */
    p |= POLICY_P_RB;
    if (normal) p |= POLICY_P_SUBJ_ORIG;
  } else if (normal) {
    RET(POLICY_DENY, POLICY_R_NORMAL_SETS_ROLEBINDING, 0);                                  /* :417-424 */
  }
  RET(p ? POLICY_ALLOW_PATCH : POLICY_ALLOW, POLICY_R_OK, p);                               /* :426-430 */
#undef RET
}

void policy_ref_bits_batch(const uint8_t *op, const uint16_t *flags, size_t n, uint8_t *code,
                           uint8_t *patch_mask, int threads) {
  (void)threads;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(threads > 0 ? threads : 1)
#endif
  for (long long i = 0; i < (long long)n; ++i) policy_ref_bits(op[i], flags[i], &code[i], &patch_mask[i]);
}

size_t policy_ref_split_groups(const char *s, char *scratch, const char **out, size_t cap) {
  /* admission.rs:45-49: String::split(',') keeps empty items and does not trim */
  size_t n = 0;
  strcpy(scratch, s);
  char *p = scratch;
  for (;;) {
    if (n < cap) out[n] = p;
    ++n;
    char *c = strchr(p, ',');
    if (!c) break;
    *c = 0;
    p = c + 1;
  }
  return n;
}
