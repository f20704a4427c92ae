/* oracle/policy_ref.h — TEST INFRASTRUCTURE ONLY (CPU oracle; never linked into the product).
 *
 * CPU restatement of the reference's admission decision function
 *   fn mutate(req, state) -> AdmissionResponse      /root/reference/src/admission.rs:241-431
 *   Username::from / UsernameKind                   /root/reference/src/admission.rs:206-239
 *
 * PARITY UNPINNED: the reference has no tests, golden vectors or fixtures for this path
 *  and cannot be compiled here (no cargo/rustc).  This file follows the
 * reference SOURCE TEXT line by line; every branch cites the line it restates.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg
 * may build, load or call anything in oracle/.
 */
#ifndef POLICY_REF_H
#define POLICY_REF_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* kube-core `Operation` (imported at admission.rs:15); anything that is not
 * Create/Update/Delete falls in the `_` arm at admission.rs:305. */
enum { POLICY_OP_CREATE = 0, POLICY_OP_UPDATE = 1, POLICY_OP_DELETE = 2, POLICY_OP_OTHER = 3 };

/* decision class */
enum { POLICY_ALLOW = 0, POLICY_ALLOW_PATCH = 1, POLICY_DENY = 2, POLICY_INVALID = 3 };

/* reason = which return statement of mutate() produced the response */
enum {
  POLICY_R_OK                = 0,  /* admission.rs:427 / :429 */
  POLICY_R_NO_USERNAME       = 1,  /* :251-253 invalid */
  POLICY_R_NOT_IN_GROUP      = 2,  /* :277-279 deny    */
  POLICY_R_NORMAL_DELETE     = 3,  /* :287-289 deny    */
  POLICY_R_NORMAL_UPDATE     = 4,  /* :298-300 deny    */
  POLICY_R_INVALID_OPERATION = 5,  /* :306-308 invalid */
  POLICY_R_NO_RESOURCE_NAME  = 6,  /* :324-326 invalid */
  POLICY_R_NAME_MISMATCH     = 7,  /* :333-335 deny    */
  POLICY_R_NOT_USERBOOTSTRAP = 8,  /* :344-345 invalid (message = serde error text) */
  POLICY_R_ADMIN_EMPTY_KUBE_USERNAME = 9, /* :369-371 deny */
  POLICY_R_NORMAL_SETS_QUOTA = 10, /* :379-381 deny    */
  POLICY_R_NORMAL_SETS_ROLEBINDING = 11, /* :420-422 deny */
  POLICY_R_ADMIN_DELETE_EARLY = 12, /* :293 allow, untouched response */
  POLICY_R_NO_OBJECT         = 13  /* :317 allow, untouched response */
};

/* patch_mask bits, in push order (admission.rs:354, :387+:413) */
enum {
  POLICY_P_USER      = 1, /* add /spec/kube_username = stripped username   :354-357 */
  POLICY_P_RB        = 2, /* add /spec/rolebinding {} then the default one :387-416 */
  POLICY_P_SUBJ_ORIG = 4  /* subject name = original (prefixed) username   :394     */
};

/* flattened predicate bits */
enum {
  POLICY_F_HAS_USERNAME = 1 << 0, /* req.user_info.username is Some           :248 */
  POLICY_F_IS_NORMAL    = 1 << 1, /* username.starts_with(prefix)             :219 */
  POLICY_F_IN_GROUP     = 1 << 2, /* any group in authorized_group_names      :263-270 */
  POLICY_F_HAS_OBJECT   = 1 << 3, /* req.object is Some                       :312 */
  POLICY_F_HAS_NAME     = 1 << 4, /* obj.metadata.name is Some                :321 */
  POLICY_F_NAME_EQ      = 1 << 5, /* kube_username == resource_name           :332 */
  POLICY_F_PARSE_OK     = 1 << 6, /* obj.try_parse::<UserBootstrap>() is Ok   :341 */
  POLICY_F_KUBE_USERNAME_NONEMPTY = 1 << 7, /* spec.kube_username Some && != "" :362-367 */
  POLICY_F_HAS_QUOTA    = 1 << 8, /* spec.quota.is_some()                     :376 */
  POLICY_F_HAS_ROLEBINDING = 1 << 9 /* !spec.rolebinding.is_none()            :386 */
};

typedef struct {
  const char *username;            /* NULL = None */
  const char *const *groups;       /* NULL = None (unwrap_or_default → empty) :267 */
  size_t n_groups;
  int operation;                   /* POLICY_OP_* */
  int has_object;
  const char *name;                /* metadata.name; NULL = None */
  int parse_ok;                    /* whether the object deserialises as UserBootstrap */
  const char *spec_kube_username;  /* NULL = None */
  int has_quota;
  int has_rolebinding;
} policy_ref_request;

typedef struct {
  const char *oidc_username_prefix;
  const char *default_role_name;
  const char *const *authorized_group_names;
  size_t n_authorized;
} policy_ref_config;

typedef struct {
  uint8_t klass, reason, patch_mask;
  const char *message;             /* static string; NULL when allowed; for
                                      NOT_USERBOOTSTRAP a placeholder (serde text unpinned) */
  char kube_username_patch[256];   /* value of P_USER, if pushed */
  char subject_name[256];          /* subject of the default rolebinding, if pushed */
} policy_ref_decision;

/* string-level restatement of mutate() */
void policy_ref_mutate(const policy_ref_request *req, const policy_ref_config *cfg,
                       policy_ref_decision *out);
/* string-level restatement of the predicates (row A2) */
void policy_ref_flatten(const policy_ref_request *req, const policy_ref_config *cfg,
                        uint8_t *op, uint16_t *flags);
/* bit-level restatement (row A1 batched form); code = klass<<4 | reason */
void policy_ref_bits(uint8_t op, uint16_t flags, uint8_t *code, uint8_t *patch_mask);
/* batched, single thread and OpenMP */
void policy_ref_bits_batch(const uint8_t *op, const uint16_t *flags, size_t n,
                           uint8_t *code, uint8_t *patch_mask, int threads);
/* exact message text for a reason (admission.rs string literals) */
const char *policy_ref_message(int reason);
/* comma_separated_deserialize, admission.rs:41-50: split on ',' with no trimming;
 * "" → [""] .  Writes up to cap pointers into `out` (pointing into `scratch`, which is a
 * mutated copy of s); returns the number of items. */
size_t policy_ref_split_groups(const char *s, char *scratch, const char **out, size_t cap);

#ifdef __cplusplus
}
#endif
#endif
