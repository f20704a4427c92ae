// bgc_host.cpp — host-side flattening and response rendering (include/bgc_host.h).
//
// Row A2: the string predicates mutate() evaluates (reference src/admission.rs:217-239, 263-270,
// 312-386) become the (op, flags) columns K0 consumes.  Row A5: pod specs become the
// (ns_id, req[r], req_mask) columns K1/K2 consume.  Plain C++17, no CUDA in this file; the only
// device work is reached through the C ABI (bgc_policy_eval).
#include <chrono>
#include <deque>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <string_view>
#include <unordered_map>
#include <vector>

#include "../../include/bgc_host.h"

namespace {

constexpr const char* kMessages[] = {
    nullptr,
    "cannot get requester's username from request",                                   // src/admission.rs:251
    "user is not in authorized group",                                                // :277
    "normal user is not allowed to delete resource",                                  // :287
    "normal user is not allowed to update resource",                                  // :298
    "invalid operation",                                                              // :306
    "cannot get resource name from request",                                          // :324
    "username not match with resource name",                                          // :333
    "<serde error text: object is not a UserBootstrap>",                              // :345 (unpinned)
    "kube_username field is empty. you are an admin, so fill it",                     // :369
    "quota field is not empty. you are a normal user, so leave it empty",             // :379
    "rolebinding field is not empty. you are a normal user, so leave it empty",       // :420
    nullptr,
    nullptr,
};

struct NameParts {
  bool normal;
  std::string_view kube_username;  // stripped for Normal, whole for Admin
};

// Username::from, src/admission.rs:217-239
NameParts classify(std::string_view username, std::string_view prefix) {
  if (username.substr(0, prefix.size()) == prefix) return {true, username.substr(prefix.size())};
  return {false, username};
}

void json_escape(std::string& out, std::string_view s) {
  out.push_back('"');
  for (unsigned char c : s) {
    switch (c) {
      case '"': out += "\\\""; break;
      case '\\': out += "\\\\"; break;
      case '\n': out += "\\n"; break;
      case '\r': out += "\\r"; break;
      case '\t': out += "\\t"; break;
      case '\b': out += "\\b"; break;
      case '\f': out += "\\f"; break;
      default:
        if (c < 0x20) {
          char b[8];
          snprintf(b, sizeof b, "\\u%04x", c);
          out += b;
        } else {
          out.push_back((char)c);
        }
    }
  }
  out.push_back('"');
}

}  // namespace

// ------------------------------------------------------------------------------ row N2 helpers
struct SheetRowOwned {
  std::string name, department, id_username, gpu_server, authorized;
  int64_t gpu_request = 0, cpu_request = 0, memory_request = 0, storage_request = 0, mig_request = 0;
};

struct bgc_sheet {
  std::vector<SheetRowOwned> rows;
  size_t skipped = 0;
};

namespace {

// CSV records as the rust `csv` crate reads them with default settings (what the reference uses,
// src/synchronizer.rs:146): ',' delimiter, '"' quoting with "" as an escaped quote, \n / \r\n / \r
// record terminators, no trimming, empty lines skipped.  (csv 1.2 behaviour restated from its
// documentation — unpinned; a leading UTF-8 BOM is dropped.)
std::vector<std::vector<std::string>> csv_records(std::string_view text) {
  std::vector<std::vector<std::string>> out;
  if (text.size() >= 3 && (unsigned char)text[0] == 0xEF && (unsigned char)text[1] == 0xBB && (unsigned char)text[2] == 0xBF)
    text.remove_prefix(3);
  std::vector<std::string> rec;
  std::string field;
  bool in_quotes = false, field_started = false, any = false;
  auto end_field = [&] { rec.push_back(field); field.clear(); field_started = false; };
  auto end_record = [&] {
    if (any || !rec.empty() || field_started) { end_field(); out.push_back(rec); }
    rec.clear(); any = false;
  };
  for (size_t i = 0; i < text.size(); ++i) {
    const char c = text[i];
    if (in_quotes) {
      if (c == '"') {
        if (i + 1 < text.size() && text[i + 1] == '"') { field.push_back('"'); ++i; }
        else in_quotes = false;
      } else field.push_back(c);
      continue;
    }
    if (c == '"' && !field_started) { in_quotes = true; field_started = true; any = true; }
    else if (c == ',') { end_field(); any = true; }
    else if (c == '\n' || c == '\r') {
      if (c == '\r' && i + 1 < text.size() && text[i + 1] == '\n') ++i;
      end_record();
    } else { field.push_back(c); field_started = true; any = true; }
  }
  end_record();
  return out;
}

// i64 from a CSV field the way serde's integer visitor over csv does: str::parse::<i64>()
bool parse_i64_strict(const std::string& s, int64_t* out) {
  if (s.empty()) return false;
  size_t i = 0;
  bool neg = false;
  if (s[0] == '+' || s[0] == '-') { neg = s[0] == '-'; i = 1; }
  if (i == s.size()) return false;
  unsigned long long v = 0;
  const unsigned long long lim = neg ? 9223372036854775808ull : 9223372036854775807ull;
  for (; i < s.size(); ++i) {
    if (s[i] < '0' || s[i] > '9') return false;
    const unsigned d = (unsigned)(s[i] - '0');
    if (v > (lim - d) / 10) return false;
    v = v * 10 + d;
  }
  *out = neg ? (int64_t)(0 - v) : (int64_t)v;
  return true;
}

bool contains(std::string_view hay, std::string_view needle) { return hay.find(needle) != std::string_view::npos; }

// str::trim(): Unicode White_Space at both ends
std::string_view trim_unicode(std::string_view s) {
  auto ws_len_front = [](std::string_view t) -> size_t {
    if (t.empty()) return 0;
    const unsigned char c = (unsigned char)t[0];
    if (c == ' ' || (c >= 0x09 && c <= 0x0D)) return 1;
    if (t.size() >= 2 && c == 0xC2 && ((unsigned char)t[1] == 0x85 || (unsigned char)t[1] == 0xA0)) return 2;
    if (t.size() >= 3) {
      const unsigned char d = (unsigned char)t[1], e = (unsigned char)t[2];
      if (c == 0xE1 && d == 0x9A && e == 0x80) return 3;                                  // U+1680
      if (c == 0xE2 && d == 0x80 && ((e >= 0x80 && e <= 0x8A) || e == 0xA8 || e == 0xA9 || e == 0xAF)) return 3;
      if (c == 0xE2 && d == 0x81 && e == 0x9F) return 3;                                  // U+205F
      if (c == 0xE3 && d == 0x80 && e == 0x80) return 3;                                  // U+3000
    }
    return 0;
  };
  for (size_t k; (k = ws_len_front(s)) != 0;) s.remove_prefix(k);
  for (bool again = true; again && !s.empty();) {
    again = false;
    for (size_t k = 1; k <= 3 && k <= s.size(); ++k) {
      const std::string_view tail = s.substr(s.size() - k);
      if (ws_len_front(tail) == k) { s.remove_suffix(k); again = true; break; }
    }
  }
  return s;
}

}  // namespace

struct bgc_nsmap {
  std::unordered_map<std::string, uint32_t> ids;
  std::vector<std::string> names;
};

extern "C" {

size_t bgc_split_groups(const char* value, char* scratch, const char** out, size_t cap) {
  size_t n = 0;
  std::strcpy(scratch, value);
  char* p = scratch;
  for (;;) {
    if (n < cap) out[n] = p;
    ++n;
    char* comma = std::strchr(p, ',');
    if (!comma) break;
    *comma = 0;
    p = comma + 1;
  }
  return n;
}

const char* bgc_reason_message(int reason) {
  if (reason < 0 || reason > 13) return nullptr;
  return kMessages[reason];
}

int bgc_flatten_requests(const bgc_admission_config* cfg, const bgc_admission_request* reqs, size_t n,
                         uint8_t* op, uint16_t* flags) {
  if (!cfg || !cfg->oidc_username_prefix || (n && (!reqs || !op || !flags))) return BGC_E_BADARG;
  const std::string_view prefix(cfg->oidc_username_prefix);
#pragma omp parallel for schedule(static) if (n >= 4096)
  for (long long i = 0; i < (long long)n; ++i) {
    const bgc_admission_request& q = reqs[i];
    unsigned f = 0;
    NameParts who{false, {}};
    if (q.username) {
      f |= BGC_F_HAS_USERNAME;
      who = classify(q.username, prefix);
      if (who.normal) f |= BGC_F_IS_NORMAL;
    }
    // groups.iter().any(|g| authorized_group_names.contains(g)), src/admission.rs:263-270
    bool in_group = false;
    for (size_t g = 0; q.groups && g < q.n_groups && !in_group; ++g)
      for (size_t a = 0; a < cfg->n_authorized; ++a)
        if (std::strcmp(q.groups[g], cfg->authorized_group_names[a]) == 0) { in_group = true; break; }
    if (in_group) f |= BGC_F_IN_GROUP;
    if (q.has_object) {
      f |= BGC_F_HAS_OBJECT;
      if (q.name) {
        f |= BGC_F_HAS_NAME;
        if (q.username && who.kube_username == std::string_view(q.name)) f |= BGC_F_NAME_EQ;
      }
      if (q.parse_ok) {
        f |= BGC_F_PARSE_OK;
        if (q.spec_kube_username && q.spec_kube_username[0]) f |= BGC_F_KUBE_USERNAME_NONEMPTY;
        if (q.has_quota) f |= BGC_F_HAS_QUOTA;
        if (q.has_rolebinding) f |= BGC_F_HAS_ROLEBINDING;
      }
    }
    op[i] = (uint8_t)((q.operation >= 0 && q.operation <= 2) ? q.operation : BGC_OP_OTHER);
    flags[i] = (uint16_t)f;
  }
  return BGC_OK;
}

int bgc_mutate_batch(bgc_ctx* ctx, const bgc_admission_config* cfg, const bgc_admission_request* reqs,
                     size_t n, bgc_admission_response* out) {
  if (!ctx || (n && !out)) return BGC_E_BADARG;
  std::vector<uint8_t> op(n), code(n), pm(n);
  std::vector<uint16_t> flags(n);
  int rc = bgc_flatten_requests(cfg, reqs, n, op.data(), flags.data());
  if (rc) return rc;
  rc = bgc_policy_eval(ctx, op.data(), flags.data(), n, code.data(), pm.data());
  if (rc) return rc;
  for (size_t i = 0; i < n; ++i) {
    out[i].klass = code[i] >> 4;
    out[i].reason = code[i] & 15;
    out[i].patch_mask = pm[i];
    out[i].allowed = out[i].klass <= BGC_ALLOW_PATCH;
    out[i].message = bgc_reason_message(out[i].reason);
  }
  return BGC_OK;
}

long bgc_render_patch(const bgc_admission_config* cfg, const bgc_admission_request* req,
                      uint8_t patch_mask, char* buf, size_t cap) {
  if (!cfg || !req || !buf) return BGC_E_BADARG;
  if (!(patch_mask & (BGC_P_USER | BGC_P_RB))) {
    if (cap) buf[0] = 0;
    return 0;
  }
  if (!req->username) return BGC_E_BADARG;
  const NameParts who = classify(req->username, cfg->oidc_username_prefix);
  std::string s = "[";
  bool first = true;
  if (patch_mask & BGC_P_USER) {  // src/admission.rs:354-357
    s += "{\"op\":\"add\",\"path\":\"/spec/kube_username\",\"value\":";
    json_escape(s, who.kube_username);
    s += "}";
    first = false;
  }
  if (patch_mask & BGC_P_RB) {  // src/admission.rs:387-416
    if (!first) s += ",";
    s += "{\"op\":\"add\",\"path\":\"/spec/rolebinding\",\"value\":{}},";
    s += "{\"op\":\"add\",\"path\":\"/spec/rolebinding\",\"value\":{\"role_ref\":{\"apiGroup\":"
         "\"rbac.authorization.k8s.io\",\"kind\":\"ClusterRole\",\"name\":";
    json_escape(s, cfg->default_role_name ? cfg->default_role_name : "");
    s += "},\"subjects\":[{\"apiGroup\":\"rbac.authorization.k8s.io\",\"kind\":\"User\",\"name\":";
    const char* subject = (patch_mask & BGC_P_SUBJ_ORIG) ? req->username                      // :394
                                                         : (req->spec_kube_username ? req->spec_kube_username : "");  // :396
    json_escape(s, subject);
    s += "}]}}";
  }
  s += "]";
  if (s.size() + 1 > cap) return BGC_E_BADARG;
  std::memcpy(buf, s.c_str(), s.size() + 1);
  return (long)s.size();
}

// ------------------------------------------------------------------------------ row A5
// ---- compact columns (what the narrow-column entry points take) ------------------------------------
int bgc_compact_plan(const int64_t* col, size_t n, uint8_t* type_out, uint8_t* shift_out) {
  if ((!col && n) || !type_out || !shift_out) return BGC_E_BADARG;
  uint64_t any = 0;
  int64_t lo = 0, hi = 0;
#pragma omp parallel for schedule(static) reduction(| : any) reduction(min : lo) reduction(max : hi) if (n >= (1u << 16))
  for (long long i = 0; i < (long long)n; ++i) {
    any |= (uint64_t)col[i];
    lo = col[i] < lo ? col[i] : lo;
    hi = col[i] > hi ? col[i] : hi;
  }
  int shift = any ? __builtin_ctzll(any) : 0;
  // (arithmetic shift keeps the sign: value >> shift is exact because the low bits are zero)
  const int64_t slo = lo >> shift, shi = hi >> shift;
  uint8_t type;
  if (slo >= 0 && shi <= 0xFF) type = BGC_T_U8;
  else if (slo >= 0 && shi <= 0xFFFF) type = BGC_T_U16;
  else if (slo >= 0 && shi <= 0xFFFFFFFFll) type = BGC_T_U32;
  else if (slo >= INT32_MIN && shi <= INT32_MAX) type = BGC_T_I32;
  else { type = BGC_T_I64; shift = 0; }
  *type_out = type;
  *shift_out = (uint8_t)shift;
  return BGC_OK;
}

int bgc_compact_column(const int64_t* col, size_t n, uint8_t type, uint8_t shift, void* out) {
  if ((n && (!col || !out)) || type > BGC_T_I32 || shift > 63) return BGC_E_BADARG;
  int bad = 0;
#pragma omp parallel for schedule(static) reduction(| : bad) if (n >= (1u << 16))
  for (long long i = 0; i < (long long)n; ++i) {
    const int64_t v = col[i], x = v >> shift;
    int64_t back;
    switch (type) {
      case BGC_T_U8: static_cast<uint8_t*>(out)[i] = (uint8_t)x; back = (int64_t)(uint8_t)x; break;
      case BGC_T_U16: static_cast<uint16_t*>(out)[i] = (uint16_t)x; back = (int64_t)(uint16_t)x; break;
      case BGC_T_U32: static_cast<uint32_t*>(out)[i] = (uint32_t)x; back = (int64_t)(uint32_t)x; break;
      case BGC_T_I32: static_cast<int32_t*>(out)[i] = (int32_t)x; back = (int64_t)(int32_t)x; break;
      default: static_cast<int64_t*>(out)[i] = x; back = x; break;
    }
    bad |= (int64_t)((uint64_t)back << shift) != v;
  }
  return bad ? BGC_E_BADARG : BGC_OK;
}

int bgc_compact_ns_ids(const uint32_t* ns_id, size_t n, uint32_t ns_count, void* out) {
  if (n && (!ns_id || !out)) return BGC_E_BADARG;
  if (ns_count > 65535) {
    if (out != ns_id) std::memcpy(out, ns_id, n * 4);
    return BGC_T_U32;
  }
#pragma omp parallel for schedule(static) if (n >= (1u << 16))
  for (long long i = 0; i < (long long)n; ++i)
    static_cast<uint16_t*>(out)[i] = ns_id[i] < ns_count ? (uint16_t)ns_id[i] : (uint16_t)0xFFFF;  // any id >= ns_count is "unknown"
  return BGC_T_U16;
}

int bgc_parse_quantity(const char* s, int unit, int64_t* out) {
  if (!s || !out) return BGC_E_BADARG;
  const char* p = s;
  bool neg = false;
  if (*p == '+' || *p == '-') { neg = (*p == '-'); ++p; }
  // mantissa digits → 128-bit integer and a decimal exponent
  unsigned __int128 mant = 0;
  int dec_exp = 0;
  bool any = false, seen_dot = false;
  const unsigned __int128 kLimit = (unsigned __int128)1 << 100;
  for (;; ++p) {
    if (*p >= '0' && *p <= '9') {
      any = true;
      if (mant >= kLimit) return BGC_E_BADARG;
      mant = mant * 10 + (unsigned)(*p - '0');
      if (seen_dot) --dec_exp;
    } else if (*p == '.' && !seen_dot) {
      seen_dot = true;
    } else {
      break;
    }
  }
  if (!any) return BGC_E_BADARG;
  int bin_exp = 0;
  // suffix: binary SI, decimal SI, or decimal exponent
  if (p[0] && p[1] == 'i' && !p[2]) {
    switch (p[0]) {
      case 'K': bin_exp = 10; break;
      case 'M': bin_exp = 20; break;
      case 'G': bin_exp = 30; break;
      case 'T': bin_exp = 40; break;
      case 'P': bin_exp = 50; break;
      case 'E': bin_exp = 60; break;
      default: return BGC_E_BADARG;
    }
  } else if ((p[0] == 'e' || p[0] == 'E') && (p[1] == '+' || p[1] == '-' || (p[1] >= '0' && p[1] <= '9'))) {
    const char* e = p + 1;
    bool eneg = false;
    if (*e == '+' || *e == '-') { eneg = (*e == '-'); ++e; }
    if (!(*e >= '0' && *e <= '9')) return BGC_E_BADARG;
    int ev = 0;
    for (; *e >= '0' && *e <= '9'; ++e) {
      ev = ev * 10 + (*e - '0');
      if (ev > 1000) return BGC_E_BADARG;
    }
    if (*e) return BGC_E_BADARG;
    dec_exp += eneg ? -ev : ev;
  } else if (p[0] && !p[1]) {
    switch (p[0]) {
      case 'n': dec_exp -= 9; break;
      case 'u': dec_exp -= 6; break;
      case 'm': dec_exp -= 3; break;
      case 'k': dec_exp += 3; break;
      case 'M': dec_exp += 6; break;
      case 'G': dec_exp += 9; break;
      case 'T': dec_exp += 12; break;
      case 'P': dec_exp += 15; break;
      case 'E': dec_exp += 18; break;
      default: return BGC_E_BADARG;
    }
  } else if (p[0]) {
    return BGC_E_BADARG;
  }
  if (unit == BGC_UNIT_MILLI) dec_exp += 3;
  // value = mant * 2^bin_exp * 10^dec_exp, rounded away from zero for positives (ceil of |v|, as
  // Kubernetes' Value()/MilliValue() round up)
  if (mant == 0) { *out = 0; return BGC_OK; }
  const unsigned __int128 kMax = (unsigned __int128)INT64_MAX;
  for (; dec_exp > 0; --dec_exp) {
    if (mant > kMax) return BGC_E_BADARG;
    mant *= 10;
  }
  if (bin_exp) {
    if (mant > (kMax >> 2)) {
      // shifting would exceed 128 bits only for absurd inputs; reject instead
      if (mant >> (127 - bin_exp)) return BGC_E_BADARG;
    }
    mant <<= bin_exp;
  }
  bool inexact = false;
  for (; dec_exp < 0; ++dec_exp) {
    if (mant % 10) inexact = true;
    mant /= 10;
    if (mant == 0 && dec_exp < -1) { dec_exp = -1; }
  }
  if (inexact && !neg) mant += 1;  // round up; negative values round toward zero (= up)
  if (mant > kMax) return BGC_E_BADARG;
  *out = neg ? -(int64_t)mant : (int64_t)mant;
  return BGC_OK;
}

long bgc_format_quantity(int64_t value, int unit, char* buf, size_t cap) {
  if (!buf || unit < BGC_UNIT_COUNT || unit > BGC_UNIT_BYTES) return BGC_E_BADARG;
  char tmp[48];
  int w;
  if (unit == BGC_UNIT_MILLI) {
    if (value % 1000 == 0) w = snprintf(tmp, sizeof tmp, "%lld", (long long)(value / 1000));
    else w = snprintf(tmp, sizeof tmp, "%lldm", (long long)value);
  } else if (unit == BGC_UNIT_BYTES && value != 0) {
    static const char* const suf[] = {"", "Ki", "Mi", "Gi", "Ti", "Pi", "Ei"};
    int k = 0;
    int64_t v = value;
    while (k < 6 && v % 1024 == 0) { v /= 1024; ++k; }
    w = snprintf(tmp, sizeof tmp, "%lld%s", (long long)v, suf[k]);
  } else {
    w = snprintf(tmp, sizeof tmp, "%lld", (long long)value);
  }
  if (w < 0 || (size_t)w + 1 > cap) return BGC_E_BADARG;
  std::memcpy(buf, tmp, (size_t)w + 1);
  return w;
}

long bgc_render_quota_denial(unsigned verdict, int wide, const char* quota_name, const char* const* quota_keys,
                             const uint8_t* unit, uint32_t R, const int64_t* req, const int64_t* used,
                             const int64_t* hard, char* buf, size_t cap) {
  if (!buf || !quota_name || !quota_keys || !unit || !req || !used || !hard || R < 1 || R > BGC_MAX_R) return BGC_E_BADARG;
  if ((!wide && R > BGC_MAX_R8) || verdict == (wide ? BGC_VERDICT16_BAD_NS : BGC_VERDICT_BAD_NS)) return BGC_E_BADARG;
  const unsigned shift = wide ? 8 : 4;
  const unsigned missing = (verdict >> shift) & ((1u << R) - 1u), exceeded = verdict & ((1u << R) - 1u);
  std::string s;
  if (missing) {
    s = "failed quota: ";
    s += quota_name;
    s += ": must specify ";
    bool first = true;
    for (uint32_t r = 0; r < R; ++r)
      if ((missing >> r) & 1u) { if (!first) s += ","; s += quota_keys[r]; first = false; }
  } else if (exceeded) {
    s = "exceeded quota: ";
    s += quota_name;
    const char* const heads[3] = {", requested: ", ", used: ", ", limited: "};
    const int64_t* const vals[3] = {req, used, hard};
    for (int part = 0; part < 3; ++part) {
      s += heads[part];
      bool first = true;
      for (uint32_t r = 0; r < R; ++r) {
        if (!((exceeded >> r) & 1u)) continue;
        char q[48];
        if (bgc_format_quantity(vals[part][r], unit[r], q, sizeof q) < 0) return BGC_E_BADARG;
        if (!first) s += ",";
        s += quota_keys[r];
        s += "=";
        s += q;
        first = false;
      }
    }
  }
  if (s.size() + 1 > cap) return BGC_E_BADARG;
  std::memcpy(buf, s.c_str(), s.size() + 1);
  return (long)s.size();
}

bgc_nsmap* bgc_nsmap_create(void) { return new (std::nothrow) bgc_nsmap; }
void bgc_nsmap_destroy(bgc_nsmap* m) { delete m; }
uint32_t bgc_nsmap_intern(bgc_nsmap* m, const char* name) {
  auto it = m->ids.find(name);
  if (it != m->ids.end()) return it->second;
  const uint32_t id = (uint32_t)m->names.size();
  m->names.emplace_back(name);
  m->ids.emplace(m->names.back(), id);
  return id;
}
uint32_t bgc_nsmap_lookup(const bgc_nsmap* m, const char* name) {
  auto it = m->ids.find(name);
  return it == m->ids.end() ? 0xFFFFFFFFu : it->second;
}
uint32_t bgc_nsmap_size(const bgc_nsmap* m) { return (uint32_t)m->names.size(); }
const char* bgc_nsmap_name(const bgc_nsmap* m, uint32_t id) {
  return id < m->names.size() ? m->names[id].c_str() : nullptr;
}

static int sum_list(const bgc_resource_list& l, const char* name, int unit, int64_t* v, bool* found) {
  *found = false;
  *v = 0;
  for (size_t k = 0; k < l.n; ++k)
    if (std::strcmp(l.names[k], name) == 0) {
      *found = true;
      return bgc_parse_quantity(l.quantities[k], unit, v);
    }
  return BGC_OK;
}

int bgc_flatten_pods(const bgc_pod_spec* pods, size_t n, const char* const* quota_keys,
                     const uint8_t* unit, uint32_t R, bgc_nsmap* map, int intern, uint32_t* ns_id,
                     int64_t* const* req, uint8_t* req_mask, size_t* n_out) {
  if (!quota_keys || !unit || !map || !n_out || R < 1 || R > BGC_MAX_R) return BGC_E_BADARG;
  if (n && (!pods || !ns_id || !req)) return BGC_E_BADARG;
  struct Key { bool limits; const char* name; };
  Key keys[BGC_MAX_R];
  for (uint32_t r = 0; r < R; ++r) {
    const char* k = quota_keys[r];
    if (std::strncmp(k, "limits.", 7) == 0) keys[r] = {true, k + 7};
    else if (std::strncmp(k, "requests.", 9) == 0) keys[r] = {false, k + 9};
    else keys[r] = {false, k};
  }
  // Phase 1 (parallel when the namespace map is read-only): every pod's row is computed in place at
  // its own index; phase 2 compacts away the terminal pods sequentially.
  std::vector<uint8_t> keep(n, 0);
  int first_error = BGC_OK;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) if (n >= 2048 && !intern)
#endif
  for (long long ii = 0; ii < (long long)n; ++ii) {
    const size_t i = (size_t)ii;
    const bgc_pod_spec& pod = pods[i];
    if (pod.phase == BGC_PHASE_SUCCEEDED || pod.phase == BGC_PHASE_FAILED) continue;  // terminal pods hold no quota
    unsigned mask = 0;
    int err = BGC_OK;
    for (uint32_t r = 0; r < R && !err; ++r) {
      bool all_state = true;
      uint64_t app = 0;
      int64_t init_max = 0;
      for (size_t c = 0; c < pod.n_containers && !err; ++c) {
        int64_t v; bool found;
        const bgc_resource_list& l = keys[r].limits ? pod.containers[c].limits : pod.containers[c].requests;
        err = sum_list(l, keys[r].name, unit[r], &v, &found);
        all_state = all_state && found;
        app += (uint64_t)v;
      }
      for (size_t c = 0; c < pod.n_init_containers && !err; ++c) {
        int64_t v; bool found;
        const bgc_resource_list& l = keys[r].limits ? pod.init_containers[c].limits : pod.init_containers[c].requests;
        err = sum_list(l, keys[r].name, unit[r], &v, &found);
        all_state = all_state && found;
        if (v > init_max) init_max = v;
      }
      int64_t total = (int64_t)app > init_max ? (int64_t)app : init_max;
      if (pod.overhead && !err) {
        int64_t v; bool found;
        err = sum_list(*pod.overhead, keys[r].name, unit[r], &v, &found);
        total = (int64_t)((uint64_t)total + (uint64_t)v);
      }
      req[r][i] = total;
      if (all_state && pod.n_containers + pod.n_init_containers > 0) mask |= 1u << r;
    }
    if (err) {
#ifdef _OPENMP
#pragma omp critical
#endif
      if (!first_error) first_error = err;
      continue;
    }
    ns_id[i] = intern ? bgc_nsmap_intern(map, pod.ns) : bgc_nsmap_lookup(map, pod.ns);
    if (req_mask) req_mask[i] = (uint8_t)mask;
    keep[i] = 1;
  }
  if (first_error) return first_error;
  size_t w = 0;
  for (size_t i = 0; i < n; ++i) {
    if (!keep[i]) continue;
    if (w != i) {
      ns_id[w] = ns_id[i];
      for (uint32_t r = 0; r < R; ++r) req[r][w] = req[r][i];
      if (req_mask) req_mask[w] = req_mask[i];
    }
    ++w;
  }
  *n_out = w;
  return BGC_OK;
}

// ------------------------------------------------------------------------------ row N2
const char* bgc_sheet_infer_header(const char* header) {
  if (!header) return nullptr;
  const std::string_view h(header);
  // first match wins, in the reference's order (src/synchronizer.rs:99-137)
  if (h == "타임스탬프") return "timestamp";
  if (h == "이름") return "name";
  if (h == "소속") return "department";
  if (contains(h, "SNUCSE ID")) return "id_username";
  if (contains(h, "사용할 서버")) return "gpu_server";
  if (contains(h, "GPU 개수")) return "gpu_request";
  if (contains(h, "vCPU 개수")) return "cpu_request";
  if (contains(h, "메모리")) return "memory_request";
  if (contains(h, "스토리지")) return "storage_request";
  if (contains(h, "MiG 개수")) return "mig_request";
  if (contains(h, "요청 사유")) return "description";
  if (contains(h, "승인")) return "authorized";
  if (contains(h, "이메일")) return "email";
  return nullptr;
}

bgc_sheet* bgc_sheet_parse_csv(const char* content, size_t len, char* err, size_t err_len) {
  if (err && err_len) err[0] = 0;
  if (!content) return nullptr;
  auto recs = csv_records(std::string_view(content, len));
  auto sheet = new (std::nothrow) bgc_sheet;
  if (!sheet) return nullptr;
  if (recs.empty()) return sheet;
  // header row → canonical names (src/synchronizer.rs:149-154)
  std::vector<std::string> names;
  for (const auto& h : recs[0]) {
    const char* n = bgc_sheet_infer_header(h.c_str());
    if (!n) {
      if (err && err_len) snprintf(err, err_len, "unknown header: \"%s\"", h.c_str());  // :139-142
      delete sheet;
      return nullptr;
    }
    names.emplace_back(n);
  }
  auto col = [&](const char* want) -> long {
    long found = -1;
    for (size_t i = 0; i < names.size(); ++i)
      if (names[i] == want) { if (found >= 0) return -2; found = (long)i; }  // duplicate field → serde error per row
    return found;
  };
  const long c_name = col("name"), c_dep = col("department"), c_id = col("id_username"), c_srv = col("gpu_server"),
             c_gpu = col("gpu_request"), c_cpu = col("cpu_request"), c_mem = col("memory_request"),
             c_sto = col("storage_request"), c_mig = col("mig_request"), c_auth = col("authorized");
  const long need[] = {c_name, c_dep, c_id, c_srv, c_gpu, c_cpu, c_mem, c_sto, c_mig, c_auth};
  bool schema_ok = true;
  for (long c : need) schema_ok = schema_ok && c >= 0;
  for (size_t r = 1; r < recs.size(); ++r) {
    const auto& f = recs[r];
    SheetRowOwned row;
    // the csv reader is not flexible: a record whose length differs from the header is an error
    bool ok = schema_ok && f.size() == names.size();
    ok = ok && parse_i64_strict(f[c_gpu], &row.gpu_request) && parse_i64_strict(f[c_cpu], &row.cpu_request) &&
         parse_i64_strict(f[c_mem], &row.memory_request) && parse_i64_strict(f[c_sto], &row.storage_request) &&
         parse_i64_strict(f[c_mig], &row.mig_request);
    if (!ok) { ++sheet->skipped; continue; }  // "row parsing error. skipping" :160-163
    row.name = f[c_name]; row.department = f[c_dep]; row.id_username = f[c_id];
    row.gpu_server = f[c_srv]; row.authorized = f[c_auth];
    sheet->rows.push_back(std::move(row));
  }
  return sheet;
}

void bgc_sheet_destroy(bgc_sheet* s) { delete s; }
size_t bgc_sheet_row_count(const bgc_sheet* s) { return s ? s->rows.size() : 0; }
size_t bgc_sheet_skipped_rows(const bgc_sheet* s) { return s ? s->skipped : 0; }

int bgc_sheet_get_row(const bgc_sheet* s, size_t i, bgc_sheet_row* out) {
  if (!s || !out || i >= s->rows.size()) return BGC_E_BADARG;
  const SheetRowOwned& r = s->rows[i];
  out->name = r.name.c_str(); out->department = r.department.c_str(); out->id_username = r.id_username.c_str();
  out->gpu_server = r.gpu_server.c_str(); out->authorized = r.authorized.c_str();
  out->gpu_request = r.gpu_request; out->cpu_request = r.cpu_request; out->memory_request = r.memory_request;
  out->storage_request = r.storage_request; out->mig_request = r.mig_request;
  return BGC_OK;
}

int bgc_sheet_match(const bgc_sheet* s, const char* server_name, const char* const* names, size_t n, long* row_index) {
  if (!s || !server_name || (n && (!names || !row_index))) return BGC_E_BADARG;
  // rows of this server, in sheet order (src/synchronizer.rs:208-213); "o" test at :227-231
  std::vector<size_t> mine;
  std::vector<char> authorised;
  for (size_t i = 0; i < s->rows.size(); ++i) {
    if (!contains(s->rows[i].gpu_server, server_name)) continue;
    mine.push_back(i);
    const std::string_view a = trim_unicode(s->rows[i].authorized);
    // to_lowercase() == "o": the only strings that lower-case to "o" are "o" and "O"
    authorised.push_back(a == "o" || a == "O");
  }
  for (size_t k = 0; k < n; ++k) {
    row_index[k] = -1;
    for (size_t j = mine.size(); j-- > 0;) {  // rows.iter().rev().find(..) :225
      if (authorised[j] && s->rows[mine[j]].id_username == names[k]) { row_index[k] = (long)mine[j]; break; }
    }
  }
  return BGC_OK;
}

int bgc_sheet_row_quota(const bgc_sheet_row* row, const char** keys, const char** quantities, int64_t* canonical,
                        char* scratch, size_t scratch_len) {
  if (!row || !keys || !quantities || !canonical || !scratch || scratch_len < 512) return BGC_E_BADARG;
  // BTreeMap iteration order = byte order of the keys (src/synchronizer.rs:250)
  struct Item { const char* key; int64_t v; bool gi; int64_t canon; };
  const Item items[BGC_SHEET_KEYS] = {
      {"limits.cpu", row->cpu_request, false, (int64_t)((uint64_t)row->cpu_request * 1000u)},             // :259-261
      {"limits.memory", row->memory_request, true, (int64_t)((uint64_t)row->memory_request << 30)},       // :263-265
      {"requests.cpu", row->cpu_request, false, (int64_t)((uint64_t)row->cpu_request * 1000u)},           // :251-253
      {"requests.memory", row->memory_request, true, (int64_t)((uint64_t)row->memory_request << 30)},     // :255-257
      {"requests.nvidia.com/gpu", row->gpu_request, false, row->gpu_request},                             // :267-269
      {"requests.nvidia.com/mig-1g.10gb", row->mig_request, false, row->mig_request},                     // :275-277
      {"requests.storage", row->storage_request, true, (int64_t)((uint64_t)row->storage_request << 30)},  // :271-273
  };
  size_t off = 0;
  for (int k = 0; k < BGC_SHEET_KEYS; ++k) {
    const int w = snprintf(scratch + off, scratch_len - off, items[k].gi ? "%lldGi" : "%lld", (long long)items[k].v);
    if (w < 0 || off + (size_t)w + 1 > scratch_len) return BGC_E_BADARG;
    keys[k] = items[k].key;
    quantities[k] = scratch + off;
    canonical[k] = items[k].canon;
    off += (size_t)w + 1;
  }
  return BGC_OK;
}

int bgc_hard_from_sheet_row(int64_t cpu_request, int64_t memory_request_gi, int64_t gpu_request,
                            int64_t storage_request_gi, int64_t mig_request,
                            const char* const* quota_keys, uint32_t R, int64_t* hard_out,
                            uint8_t* unit_out) {
  if (!quota_keys || !hard_out || !unit_out) return BGC_E_BADARG;
  for (uint32_t r = 0; r < R; ++r) {
    const std::string_view k(quota_keys[r]);
    if (k == "requests.cpu" || k == "limits.cpu") {              // src/synchronizer.rs:251-253,259-261
      hard_out[r] = cpu_request * 1000; unit_out[r] = BGC_UNIT_MILLI;
    } else if (k == "requests.memory" || k == "limits.memory") { // :255-257,263-265 "<n>Gi"
      hard_out[r] = (int64_t)((uint64_t)memory_request_gi << 30); unit_out[r] = BGC_UNIT_BYTES;
    } else if (k == "requests.nvidia.com/gpu") {                 // :267-269
      hard_out[r] = gpu_request; unit_out[r] = BGC_UNIT_COUNT;
    } else if (k == "requests.storage") {                        // :271-273 "<n>Gi" (PVC-side key)
      hard_out[r] = (int64_t)((uint64_t)storage_request_gi << 30); unit_out[r] = BGC_UNIT_COUNT;
    } else if (k == "requests.nvidia.com/mig-1g.10gb") {         // :275-277
      hard_out[r] = mig_request; unit_out[r] = BGC_UNIT_COUNT;
    } else {
      return BGC_E_BADARG;
    }
  }
  return BGC_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------ row N4
// A small order-preserving JSON DOM: enough for AdmissionReview documents.
namespace {

struct JVal {
  enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
  bool b = false;
  std::string s;  // string value, or the raw text of a number
  std::vector<JVal> arr;
  std::vector<std::pair<std::string, JVal>> obj;
  const JVal* get(std::string_view key) const {
    if (kind != Obj) return nullptr;
    const JVal* hit = nullptr;
    for (const auto& kv : obj)
      if (kv.first == key) hit = &kv.second;  // serde_json keeps the last duplicate
    return hit;
  }
  const JVal* get_obj(std::string_view key) const {
    const JVal* v = get(key);
    return v && v->kind == Obj ? v : nullptr;
  }
  const char* get_str(std::string_view key) const {
    const JVal* v = get(key);
    return v && v->kind == Str ? v->s.c_str() : nullptr;
  }
};

struct JParser {
  const char* p;
  const char* end;
  std::string err;
  int depth = 0;
  bool fail(const char* m) { if (err.empty()) err = m; return false; }
  void ws() { while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) ++p; }
  static void utf8(std::string& out, unsigned cp) {
    if (cp < 0x80) out.push_back((char)cp);
    else if (cp < 0x800) { out.push_back((char)(0xC0 | (cp >> 6))); out.push_back((char)(0x80 | (cp & 0x3F))); }
    else if (cp < 0x10000) { out.push_back((char)(0xE0 | (cp >> 12))); out.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); out.push_back((char)(0x80 | (cp & 0x3F))); }
    else { out.push_back((char)(0xF0 | (cp >> 18))); out.push_back((char)(0x80 | ((cp >> 12) & 0x3F))); out.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); out.push_back((char)(0x80 | (cp & 0x3F))); }
  }
  bool hex4(unsigned* v) {
    if (end - p < 4) return fail("truncated \\u escape");
    unsigned x = 0;
    for (int i = 0; i < 4; ++i) {
      const char c = p[i];
      x <<= 4;
      if (c >= '0' && c <= '9') x |= (unsigned)(c - '0');
      else if (c >= 'a' && c <= 'f') x |= (unsigned)(c - 'a' + 10);
      else if (c >= 'A' && c <= 'F') x |= (unsigned)(c - 'A' + 10);
      else return fail("bad \\u escape");
    }
    p += 4;
    *v = x;
    return true;
  }
  bool str(std::string& out) {
    if (p >= end || *p != '"') return fail("expected string");
    ++p;
    while (p < end && *p != '"') {
      if ((unsigned char)*p < 0x20) return fail("control character in string");
      if (*p == '\\') {
        if (++p >= end) return fail("truncated escape");
        switch (*p++) {
          case '"': out.push_back('"'); break;
          case '\\': out.push_back('\\'); break;
          case '/': out.push_back('/'); break;
          case 'b': out.push_back('\b'); break;
          case 'f': out.push_back('\f'); break;
          case 'n': out.push_back('\n'); break;
          case 'r': out.push_back('\r'); break;
          case 't': out.push_back('\t'); break;
          case 'u': {
            unsigned cp;
            if (!hex4(&cp)) return false;
            if (cp >= 0xD800 && cp < 0xDC00) {  // surrogate pair
              unsigned lo;
              if (end - p < 6 || p[0] != '\\' || p[1] != 'u') return fail("lone surrogate");
              p += 2;
              if (!hex4(&lo) || lo < 0xDC00 || lo > 0xDFFF) return fail("bad surrogate pair");
              cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
            } else if (cp >= 0xDC00 && cp < 0xE000) return fail("lone surrogate");
            utf8(out, cp);
            break;
          }
          default: return fail("bad escape");
        }
      } else if ((unsigned char)*p < 0x80) {
        out.push_back(*p++);
      } else {  // raw multi-byte sequence: must be well-formed UTF-8, as serde_json requires
        const unsigned char c = (unsigned char)*p;
        const int len = c >= 0xF0 ? 4 : c >= 0xE0 ? 3 : c >= 0xC2 ? 2 : 0;
        if (!len || c > 0xF4 || end - p < len) return fail("invalid UTF-8 in string");
        unsigned cp = c & (0xFF >> (len + 1));
        for (int i = 1; i < len; ++i) {
          const unsigned char d = (unsigned char)p[i];
          if ((d & 0xC0) != 0x80) return fail("invalid UTF-8 in string");
          cp = (cp << 6) | (d & 0x3F);
        }
        if ((len == 3 && (cp < 0x800 || (cp >= 0xD800 && cp < 0xE000))) || (len == 4 && (cp < 0x10000 || cp > 0x10FFFF)))
          return fail("invalid UTF-8 in string");
        out.append(p, (size_t)len);
        p += len;
      }
    }
    if (p >= end) return fail("unterminated string");
    ++p;
    return true;
  }
  bool value(JVal& v) {
    if (++depth > 128) return fail("nesting too deep");
    ws();
    if (p >= end) return fail("unexpected end");
    bool ok = true;
    if (*p == '{') {
      v.kind = JVal::Obj;
      ++p; ws();
      if (p < end && *p == '}') ++p;
      else for (;;) {
        ws();
        std::string k;
        if (!str(k)) { ok = false; break; }
        ws();
        if (p >= end || *p != ':') { ok = fail("expected ':'"); break; }
        ++p;
        v.obj.emplace_back(std::move(k), JVal());
        if (!value(v.obj.back().second)) { ok = false; break; }
        ws();
        if (p < end && *p == ',') { ++p; continue; }
        if (p < end && *p == '}') { ++p; break; }
        ok = fail("expected ',' or '}'"); break;
      }
    } else if (*p == '[') {
      v.kind = JVal::Arr;
      ++p; ws();
      if (p < end && *p == ']') ++p;
      else for (;;) {
        v.arr.emplace_back();
        if (!value(v.arr.back())) { ok = false; break; }
        ws();
        if (p < end && *p == ',') { ++p; continue; }
        if (p < end && *p == ']') { ++p; break; }
        ok = fail("expected ',' or ']'"); break;
      }
    } else if (*p == '"') {
      v.kind = JVal::Str;
      ok = str(v.s);
    } else if (end - p >= 4 && !std::strncmp(p, "true", 4)) { v.kind = JVal::Bool; v.b = true; p += 4; }
    else if (end - p >= 5 && !std::strncmp(p, "false", 5)) { v.kind = JVal::Bool; v.b = false; p += 5; }
    else if (end - p >= 4 && !std::strncmp(p, "null", 4)) { v.kind = JVal::Null; p += 4; }
    else if (*p == '-' || (*p >= '0' && *p <= '9')) {
      v.kind = JVal::Num;
      const char* q = p;
      if (*p == '-') ++p;
      if (p >= end || *p < '0' || *p > '9') return fail("bad number");
      while (p < end && ((*p >= '0' && *p <= '9') || *p == '.' || *p == 'e' || *p == 'E' || *p == '+' || *p == '-')) ++p;
      v.s.assign(q, p);
    } else ok = fail("unexpected character");
    --depth;
    return ok;
  }
};

// DynamicObject::try_parse::<UserBootstrap>() approximated as in the Python mirror
// (bacchus-gpu-controller_b200/admission.py::_parses_as_userbootstrap; crd.rs:19-42).
bool parses_as_userbootstrap(const JVal& obj) {
  const JVal* spec = obj.get_obj("spec");
  if (!spec) return false;
  auto opt_is = [&](const char* k, JVal::Kind kind) {
    const JVal* v = spec->get(k);
    return !v || v->kind == JVal::Null || v->kind == kind;
  };
  if (!opt_is("kube_username", JVal::Str) || !opt_is("quota", JVal::Obj) || !opt_is("role", JVal::Obj)) return false;
  const JVal* rb = spec->get("rolebinding");
  if (rb && rb->kind != JVal::Null) {
    if (rb->kind != JVal::Obj) return false;
    const JVal* rr = rb->get_obj("role_ref");
    if (!rr || !rr->get_str("apiGroup") || !rr->get_str("kind") || !rr->get_str("name")) return false;
    const JVal* subs = rb->get("subjects");
    if (subs && subs->kind != JVal::Null) {
      if (subs->kind != JVal::Arr) return false;
      for (const auto& sub : subs->arr)  // k8s-openapi Subject: `kind` and `name` are required strings
        if (sub.kind != JVal::Obj || !sub.get_str("kind") || !sub.get_str("name")) return false;
    }
  }
  // ResourceQuotaSpec.hard: a map of Quantity, which deserialises from a string or a number
  if (const JVal* quota = spec->get("quota"); quota && quota->kind == JVal::Obj) {
    const JVal* hard = quota->get("hard");
    if (hard && hard->kind != JVal::Null) {
      if (hard->kind != JVal::Obj) return false;
      for (const auto& kv : hard->obj)
        if (kv.second.kind != JVal::Str && kv.second.kind != JVal::Num) return false;
    }
  }
  return true;
}

const char kB64[] = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/";
void base64(std::string& out, std::string_view in) {
  size_t i = 0;
  for (; i + 2 < in.size(); i += 3) {
    const unsigned v = ((unsigned char)in[i] << 16) | ((unsigned char)in[i + 1] << 8) | (unsigned char)in[i + 2];
    out.push_back(kB64[v >> 18]); out.push_back(kB64[(v >> 12) & 63]); out.push_back(kB64[(v >> 6) & 63]); out.push_back(kB64[v & 63]);
  }
  if (i + 1 == in.size()) {
    const unsigned v = (unsigned char)in[i] << 16;
    out.push_back(kB64[v >> 18]); out.push_back(kB64[(v >> 12) & 63]); out += "==";
  } else if (i + 2 == in.size()) {
    const unsigned v = ((unsigned char)in[i] << 16) | ((unsigned char)in[i + 1] << 8);
    out.push_back(kB64[v >> 18]); out.push_back(kB64[(v >> 12) & 63]); out.push_back(kB64[(v >> 6) & 63]); out.push_back('=');
  }
}

}  // namespace

struct bgc_review {
  JVal doc;
  const JVal* request = nullptr;
  std::vector<const char*> groups;
  // pod view
  std::deque<std::vector<const char*>> name_store, quant_store;  // a deque never moves its elements: lists handed out stay put
  std::vector<bgc_container_spec> containers, init_containers;
  bgc_resource_list overhead{};
  bool has_overhead = false;
  std::string ns;
};

extern "C" {

bgc_review* bgc_review_parse(const char* json, size_t len, char* err, size_t err_len) {
  if (err && err_len) err[0] = 0;
  if (!json) return nullptr;
  auto r = new (std::nothrow) bgc_review;
  if (!r) return nullptr;
  JParser ps{json, json + len, {}, 0};
  bool ok = ps.value(r->doc);
  if (ok) { ps.ws(); if (ps.p != ps.end) ok = ps.fail("trailing characters"); }
  if (!ok || r->doc.kind != JVal::Obj) {
    if (err && err_len) snprintf(err, err_len, "%s", ok ? "AdmissionReview must be a JSON object" : ps.err.c_str());
    delete r;
    return nullptr;
  }
  r->request = r->doc.get_obj("request");
  // serde would reject the whole AdmissionReview for a wrong-typed userInfo.groups (Option<Vec<String>>): so does
  // this parser, instead of quietly dropping the odd entries
  if (r->request) {
    const JVal* ui = r->request->get_obj("userInfo");
    const JVal* groups = ui ? ui->get("groups") : nullptr;
    bool bad = groups && groups->kind != JVal::Null && groups->kind != JVal::Arr;
    if (groups && groups->kind == JVal::Arr)
      for (const auto& g : groups->arr) bad = bad || g.kind != JVal::Str;
    if (bad) {
      if (err && err_len) snprintf(err, err_len, "request.userInfo.groups must be an array of strings");
      delete r;
      return nullptr;
    }
  }
  return r;
}

void bgc_review_destroy(bgc_review* r) { delete r; }
int bgc_review_has_request(const bgc_review* r) { return r && r->request ? 1 : 0; }

const char* bgc_review_uid(const bgc_review* r) {
  const char* u = r && r->request ? r->request->get_str("uid") : nullptr;
  return u ? u : "";
}

const char* bgc_review_kind(const bgc_review* r) {
  const JVal* k = r && r->request ? r->request->get_obj("kind") : nullptr;
  const char* s = k ? k->get_str("kind") : nullptr;
  return s ? s : "";
}

int bgc_review_request(const bgc_review* rc, bgc_admission_request* out) {
  if (!rc || !out || !rc->request) return BGC_E_BADARG;
  bgc_review* r = const_cast<bgc_review*>(rc);
  const JVal& q = *r->request;
  std::memset(out, 0, sizeof *out);
  const JVal* ui = q.get_obj("userInfo");
  out->username = ui ? ui->get_str("username") : nullptr;                     // src/admission.rs:248
  const JVal* groups = ui ? ui->get("groups") : nullptr;
  r->groups.clear();
  if (groups && groups->kind == JVal::Arr) {                                  // :263-267
    for (const auto& g : groups->arr)
      if (g.kind == JVal::Str) r->groups.push_back(g.s.c_str());
    static const char* const kNoGroups[1] = {nullptr};  // Some([]) is distinct from None
    out->groups = r->groups.empty() ? kNoGroups : r->groups.data();
    out->n_groups = r->groups.size();
  }
  const char* op = q.get_str("operation");                                    // :272
  out->operation = !op ? BGC_OP_OTHER : !std::strcmp(op, "CREATE") ? BGC_OP_CREATE : !std::strcmp(op, "UPDATE") ? BGC_OP_UPDATE
                   : !std::strcmp(op, "DELETE") ? BGC_OP_DELETE : BGC_OP_OTHER;
  const JVal* obj = q.get("object");
  if (obj && obj->kind == JVal::Obj) {                                        // :312
    out->has_object = 1;
    const JVal* meta = obj->get_obj("metadata");
    out->name = meta ? meta->get_str("name") : nullptr;                       // :321
    if (parses_as_userbootstrap(*obj)) {                                      // :341
      out->parse_ok = 1;
      const JVal* spec = obj->get_obj("spec");
      out->spec_kube_username = spec->get_str("kube_username");               // :362-367
      const JVal* quota = spec->get("quota");
      out->has_quota = quota && quota->kind != JVal::Null;                    // :376
      const JVal* rb = spec->get("rolebinding");
      out->has_rolebinding = rb && rb->kind != JVal::Null;                    // :386
    }
  }
  return BGC_OK;
}

static void fill_list(bgc_review* r, const JVal* m, bgc_resource_list* out) {
  r->name_store.emplace_back();
  r->quant_store.emplace_back();
  auto& names = r->name_store.back();
  auto& quants = r->quant_store.back();
  if (m && m->kind == JVal::Obj)
    for (const auto& kv : m->obj)
      if (kv.second.kind == JVal::Str || kv.second.kind == JVal::Num) {  // quantities arrive as strings or bare numbers
        names.push_back(kv.first.c_str());
        quants.push_back(kv.second.s.c_str());
      }
  out->names = names.data();
  out->quantities = quants.data();
  out->n = names.size();
}

static void fill_containers(bgc_review* r, const JVal* list, std::vector<bgc_container_spec>& out) {
  out.clear();
  if (!list || list->kind != JVal::Arr) return;
  for (const auto& c : list->arr) {
    bgc_container_spec cs{};
    const JVal* res = c.get_obj("resources");
    fill_list(r, res ? res->get("requests") : nullptr, &cs.requests);
    fill_list(r, res ? res->get("limits") : nullptr, &cs.limits);
    out.push_back(cs);
  }
}

int bgc_review_pod(const bgc_review* rc, bgc_pod_spec* out) {
  if (!rc || !out || !rc->request) return BGC_E_BADARG;
  bgc_review* r = const_cast<bgc_review*>(rc);
  const JVal* obj = r->request->get_obj("object");
  if (!obj) return BGC_E_BADARG;
  r->name_store.clear(); r->quant_store.clear();
  const JVal* spec = obj->get_obj("spec");
  const JVal* meta = obj->get_obj("metadata");
  const char* ns = r->request->get_str("namespace");
  if (!ns && meta) ns = meta->get_str("namespace");
  r->ns = ns ? ns : "default";
  fill_containers(r, spec ? spec->get("containers") : nullptr, r->containers);
  fill_containers(r, spec ? spec->get("initContainers") : nullptr, r->init_containers);
  const JVal* oh = spec ? spec->get("overhead") : nullptr;
  r->has_overhead = oh && oh->kind == JVal::Obj;
  if (r->has_overhead) fill_list(r, oh, &r->overhead);
  const JVal* status = obj->get_obj("status");
  const char* phase = status ? status->get_str("phase") : nullptr;
  out->ns = r->ns.c_str();
  out->phase = !phase ? BGC_PHASE_PENDING : !std::strcmp(phase, "Running") ? BGC_PHASE_RUNNING : !std::strcmp(phase, "Succeeded") ? BGC_PHASE_SUCCEEDED
               : !std::strcmp(phase, "Failed") ? BGC_PHASE_FAILED : !std::strcmp(phase, "Pending") ? BGC_PHASE_PENDING : BGC_PHASE_UNKNOWN;
  out->containers = r->containers.data(); out->n_containers = r->containers.size();
  out->init_containers = r->init_containers.data(); out->n_init_containers = r->init_containers.size();
  out->overhead = r->has_overhead ? &r->overhead : nullptr;
  return BGC_OK;
}

long bgc_render_review_response(const char* uid, const bgc_admission_response* resp, const char* message_override,
                                const char* patch_json, char* buf, size_t cap) {
  if (!resp || !buf) return BGC_E_BADARG;
  std::string s = "{\"apiVersion\":\"admission.k8s.io/v1\",\"kind\":\"AdmissionReview\",\"response\":{\"uid\":";
  const char* msg = message_override ? message_override : resp->message;
  if (resp->klass == BGC_INVALID) {  // AdmissionResponse::invalid: fresh response without the request's uid
    json_escape(s, "");
    s += ",\"allowed\":false,\"status\":{\"status\":\"Failure\",\"code\":400,\"message\":";
    json_escape(s, msg ? msg : "");
    s += ",\"reason\":\"InvalidRequest\"}";
  } else {
    json_escape(s, uid ? uid : "");
    s += resp->allowed ? ",\"allowed\":true,\"status\":{" : ",\"allowed\":false,\"status\":{";
    if (!resp->allowed) { s += "\"message\":"; json_escape(s, msg ? msg : ""); }
    s += "}";
    if (resp->allowed && patch_json && patch_json[0]) {  // with_patch, src/admission.rs:429
      s += ",\"patch\":\"";
      base64(s, patch_json);
      s += "\",\"patchType\":\"JSONPatch\"";
    }
  }
  s += "}}";
  if (s.size() + 1 > cap) return BGC_E_BADARG;
  std::memcpy(buf, s.c_str(), s.size() + 1);
  return (long)s.size();
}

}  // extern "C"

// ------------------------------------------------------------------------------ row N3
// Reconcile plan: the objects `reconcile()` server-side-applies for one UserBootstrap
// (reference src/controller.rs:50-155), as (path, body) pairs in the reference's order.
namespace {

void jdump(std::string& out, const JVal& v) {
  switch (v.kind) {
    case JVal::Null: out += "null"; break;
    case JVal::Bool: out += v.b ? "true" : "false"; break;
    case JVal::Num: out += v.s; break;
    case JVal::Str: json_escape(out, v.s); break;
    case JVal::Arr:
      out.push_back('[');
      for (size_t i = 0; i < v.arr.size(); ++i) { if (i) out.push_back(','); jdump(out, v.arr[i]); }
      out.push_back(']');
      break;
    case JVal::Obj:
      out.push_back('{');
      for (size_t i = 0; i < v.obj.size(); ++i) {
        if (i) out.push_back(',');
        json_escape(out, v.obj[i].first);
        out.push_back(':');
        jdump(out, v.obj[i].second);
      }
      out.push_back('}');
      break;
  }
}

}  // namespace

struct bgc_plan_item_s { std::string kind, ns, name, path, body; bool namespaced; };
struct bgc_plan { std::vector<bgc_plan_item_s> items; };

extern "C" {

bgc_plan* bgc_reconcile_plan(const char* ub_json, size_t len, char* err, size_t err_len) {
  if (err && err_len) err[0] = 0;
  auto say = [&](const char* m) -> bgc_plan* { if (err && err_len) snprintf(err, err_len, "%s", m); return nullptr; };
  if (!ub_json) return say("null document");
  JVal doc;
  JParser ps{ub_json, ub_json + len, {}, 0};
  bool ok = ps.value(doc);
  if (ok) { ps.ws(); ok = ps.p == ps.end; }
  if (!ok || doc.kind != JVal::Obj) return say(ps.err.empty() ? "UserBootstrap must be a JSON object" : ps.err.c_str());
  const JVal* meta = doc.get_obj("metadata");
  const char* ub_name = meta ? meta->get_str("name") : nullptr;
  const char* ub_uid = meta ? meta->get_str("uid") : nullptr;
  // controller_owner_ref(&()).unwrap() at src/controller.rs:52 needs both name and uid
  if (!ub_name || !ub_uid) return say("missing object key: .metadata.name / .metadata.uid (no controller owner reference)");
  std::string name(ub_name);
  for (auto& c : name)
    if (c >= 'A' && c <= 'Z') c = (char)(c - 'A' + 'a');  // .to_lowercase() :63 (object names are ASCII)
  std::string oref = "{\"apiVersion\":\"bacchus.io/v1\",\"blockOwnerDeletion\":true,\"controller\":true,\"kind\":\"UserBootstrap\",\"name\":";
  json_escape(oref, ub_name);
  oref += ",\"uid\":";
  json_escape(oref, ub_uid);
  oref += "}";
  std::string qname;
  json_escape(qname, name);
  const std::string metadata = "{\"name\":" + qname + ",\"ownerReferences\":[" + oref + "]}";
  const std::string params = "?fieldManager=bacchus-gpu-controller.bacchus.io&force=true";  // PatchParams::apply(PATCH_MANAGER).force() :67
  auto plan = new (std::nothrow) bgc_plan;
  if (!plan) return say("out of memory");
  // Namespace :69-87
  plan->items.push_back({"Namespace", "", name, "/api/v1/namespaces/" + name + params,
                         "{\"apiVersion\":\"v1\",\"kind\":\"Namespace\",\"metadata\":" + metadata + "}", false});
  const JVal* spec = doc.get_obj("spec");
  const JVal* quota = spec ? spec->get("quota") : nullptr;
  if (quota && quota->kind != JVal::Null) {  // ResourceQuota only if spec.quota is Some :89-110
    std::string body = "{\"apiVersion\":\"v1\",\"kind\":\"ResourceQuota\",\"metadata\":" + metadata + ",\"spec\":";
    jdump(body, *quota);
    body += "}";
    plan->items.push_back({"ResourceQuota", name, name, "/api/v1/namespaces/" + name + "/resourcequotas/" + name + params, body, true});
  }
  const JVal* role = spec ? spec->get("role") : nullptr;
  if (role && role->kind == JVal::Obj) {  // Role if Some, owner reference overwritten :112-124
    std::string body = "{\"apiVersion\":\"rbac.authorization.k8s.io/v1\",\"kind\":\"Role\",\"metadata\":{";
    bool first = true;
    if (const JVal* rmeta = role->get_obj("metadata"))
      for (const auto& kv : rmeta->obj) {
        if (kv.first == "ownerReferences") continue;
        if (!first) body.push_back(',');
        json_escape(body, kv.first); body.push_back(':'); jdump(body, kv.second);
        first = false;
      }
    if (!first) body.push_back(',');
    body += "\"ownerReferences\":[" + oref + "]}";
    if (const JVal* rules = role->get("rules")) { body += ",\"rules\":"; jdump(body, *rules); }
    body += "}";
    plan->items.push_back({"Role", name, name, "/apis/rbac.authorization.k8s.io/v1/namespaces/" + name + "/roles/" + name + params, body, true});
  }
  const JVal* rb = spec ? spec->get("rolebinding") : nullptr;
  const JVal* status = doc.get_obj("status");
  const JVal* synced = status ? status->get("synchronized_with_sheet") : nullptr;
  // RoleBinding only once the sheet synchroniser has vouched for the user :126-152
  if (rb && rb->kind == JVal::Obj && synced && synced->kind == JVal::Bool && synced->b) {
    std::string body = "{\"apiVersion\":\"rbac.authorization.k8s.io/v1\",\"kind\":\"RoleBinding\",\"metadata\":" + metadata + ",\"roleRef\":";
    if (const JVal* rr = rb->get("role_ref")) jdump(body, *rr); else body += "null";
    const JVal* subs = rb->get("subjects");
    if (subs && subs->kind != JVal::Null) { body += ",\"subjects\":"; jdump(body, *subs); }
    body += "}";
    plan->items.push_back({"RoleBinding", name, name, "/apis/rbac.authorization.k8s.io/v1/namespaces/" + name + "/rolebindings/" + name + params, body, true});
  }
  return plan;
}

void bgc_plan_destroy(bgc_plan* p) { delete p; }
size_t bgc_plan_count(const bgc_plan* p) { return p ? p->items.size() : 0; }

int bgc_plan_item(const bgc_plan* p, size_t i, const char** kind, const char** ns, const char** name, const char** path,
                  const char** body) {
  if (!p || i >= p->items.size()) return BGC_E_BADARG;
  const auto& it = p->items[i];
  if (kind) *kind = it.kind.c_str();
  if (ns) *ns = it.namespaced ? it.ns.c_str() : nullptr;
  if (name) *name = it.name.c_str();
  if (path) *path = it.path.c_str();
  if (body) *body = it.body.c_str();
  return BGC_OK;
}

int bgc_plan_requeue_seconds(int failed) { return failed ? 3 : 30; }  // src/controller.rs:154,174

}  // extern "C"

// ---- the batched mutate_handler, wire bytes to wire bytes (reference src/admission.rs:184-204) ----------------
static int handle_reviews_impl(bgc_ctx* ctx, const bgc_admission_config* cfg, const char* const* bodies, const size_t* body_len,
                       size_t n, char* out, size_t out_cap, size_t* out_off, uint64_t* stage_ns) {
  if (!ctx || !cfg || (n && (!bodies || !body_len)) || !out_off || (!out && out_cap)) return BGC_E_BADARG;
  using clk = std::chrono::steady_clock;
  auto ns_since = [](clk::time_point t0) { return (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(clk::now() - t0).count(); };
  auto t0 = clk::now();
  // One pass per body, all in the thread that parsed it: parse, the fields mutate() reads, their predicate bits
  // (row A2), copies of the three strings a response can need, release of the tree.  (Freeing 10k trees after the
  // batch cost more than parsing them, even spread over the cores.)
  struct Kept {
    std::string uid, username, spec_user;
    bool has_username = false, has_spec_user = false;
  };
  std::vector<uint8_t> state(n, 0);  // 0 malformed, 1 no request, 2 decided below
  std::vector<Kept> kept(n);
  std::vector<uint8_t> op_all(n ? n : 1);
  std::vector<uint16_t> flags_all(n ? n : 1);
  int rc = BGC_OK;
#pragma omp parallel for schedule(dynamic, 64) if (n >= 256)
  for (long long i = 0; i < (long long)n; ++i) {
    bgc_review* r = bodies[i] ? bgc_review_parse(bodies[i], body_len[i], nullptr, 0) : nullptr;
    if (!r) continue;
    state[i] = 1;
    bgc_admission_request q;
    if (bgc_review_has_request(r) && bgc_review_request(r, &q) == BGC_OK) {
      const int e = bgc_flatten_requests(cfg, &q, 1, &op_all[i], &flags_all[i]);
      if (e != BGC_OK) {
#pragma omp atomic write
        rc = e;
      } else {
        state[i] = 2;
        Kept& k = kept[i];
        k.uid = bgc_review_uid(r);
        if (q.username) { k.username = q.username; k.has_username = true; }
        if (q.spec_kube_username) { k.spec_user = q.spec_kube_username; k.has_spec_user = true; }
      }
    }
    bgc_review_destroy(r);
  }
  if (rc) return rc;
  const uint64_t t_parse = ns_since(t0);
  t0 = clk::now();
  std::vector<size_t> live;
  live.reserve(n);
  for (size_t i = 0; i < n; ++i)
    if (state[i] == 2) live.push_back(i);
  const size_t m = live.size();
  std::vector<uint8_t> op(m ? m : 1), code(m ? m : 1), pm(m ? m : 1);
  std::vector<uint16_t> flags(m ? m : 1);
  for (size_t k = 0; k < m; ++k) { op[k] = op_all[live[k]]; flags[k] = flags_all[live[k]]; }
  if (m) rc = bgc_policy_eval(ctx, op.data(), flags.data(), m, code.data(), pm.data());  // one K0 launch
  if (rc) return rc;
  const uint64_t t_decide = ns_since(t0);
  t0 = clk::now();
  std::vector<long long> slot(n, -1);
  for (size_t k = 0; k < m; ++k) slot[live[k]] = (long long)k;
  std::vector<std::string> text(n);
  int bad_render = 0;
#pragma omp parallel for schedule(static) if (n >= 256)
  for (long long i = 0; i < (long long)n; ++i) {
    bgc_admission_response resp = {0, BGC_INVALID, 0, 0, nullptr};
    const char* uid = "";
    const char* msg = nullptr;
    std::string patch;
    bool fail = false;
    if (slot[i] < 0) {
      msg = state[i] ? "invalid request: AdmissionReview carries no request" : "malformed AdmissionReview";
    } else {
      const size_t k = (size_t)slot[i];
      resp.klass = code[k] >> 4;
      resp.reason = code[k] & 15;
      resp.patch_mask = pm[k];
      resp.allowed = resp.klass <= BGC_ALLOW_PATCH;
      resp.message = bgc_reason_message(resp.reason);
      uid = kept[i].uid.c_str();
      if (resp.patch_mask) {  // the patch reads two strings of the request only (src/admission.rs:354-357,387-416)
        bgc_admission_request q;
        std::memset(&q, 0, sizeof q);
        q.username = kept[i].has_username ? kept[i].username.c_str() : nullptr;
        q.spec_kube_username = kept[i].has_spec_user ? kept[i].spec_user.c_str() : nullptr;
        std::string buf(1024, '\0');
        long len;
        while ((len = bgc_render_patch(cfg, &q, resp.patch_mask, &buf[0], buf.size())) == BGC_E_BADARG && q.username && buf.size() < (1u << 24))
          buf.resize(buf.size() * 4);
        if (len > 0) patch.assign(buf.data(), (size_t)len);
        else fail = true;  // a patch that was asked for and did not come out fails the call, never a response without it
      }
    }
    std::string buf(1024, '\0');
    long len = -1;
    while (!fail && (len = bgc_render_review_response(uid, &resp, msg, patch.empty() ? nullptr : patch.c_str(), &buf[0], buf.size())) == BGC_E_BADARG &&
           buf.size() < (1u << 24))
      buf.resize(buf.size() * 4);
    if (len > 0) text[i].assign(buf.data(), (size_t)len);
    else
#pragma omp atomic write
      bad_render = 1;
  }
  if (bad_render) return BGC_E_BADARG;
  const uint64_t t_render = ns_since(t0);
  t0 = clk::now();
  size_t need = 0;
  for (size_t i = 0; i < n; ++i) {
    out_off[i] = need;
    need += text[i].size() + 1;
  }
  out_off[n] = need;
  if (need > out_cap) return BGC_E_BADARG;
  for (size_t i = 0; i < n; ++i) std::memcpy(out + out_off[i], text[i].c_str(), text[i].size() + 1);
  if (stage_ns) { stage_ns[0] = t_parse; stage_ns[1] = t_decide; stage_ns[2] = t_render; stage_ns[3] = ns_since(t0); }
  return BGC_OK;
}

// ---- Pod-create reviews, wire bytes to wire bytes (no reference counterpart; same seam as above) ------------------
static int handle_pod_reviews_impl(bgc_ctx* ctx, const char* const* bodies, const size_t* body_len, size_t n,
                                      const char* const* quota_keys, const uint8_t* unit, uint32_t R, bgc_nsmap* map, int wide,
                                      const int64_t* used, const int64_t* hard, uint32_t ns_count, char* out, size_t out_cap,
                                      size_t* out_off, uint64_t* stage_ns) {
  if (!ctx || (n && (!bodies || !body_len)) || !quota_keys || !unit || !map || R < 1 || R > BGC_MAX_R || !out_off || (!out && out_cap))
    return BGC_E_BADARG;
  if (!wide && R > 4) return BGC_E_BADARG;
  using clk = std::chrono::steady_clock;
  auto ns_since = [](clk::time_point t0) { return (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(clk::now() - t0).count(); };
  auto t0 = clk::now();
  // one pass per body, all in the thread that parsed it: parse, pod spec, flatten (row A5) into row i, keep the uid,
  // release the tree (10k trees freed after the batch, even in parallel, cost more than parsing them)
  std::vector<uint8_t> state(n, 0);  // 0 malformed, 1 no pod object, 2 has a row
  std::vector<std::string> uid_of(n);
  std::vector<uint32_t> ns_all(n ? n : 1);
  std::vector<std::vector<int64_t>> req_all(R, std::vector<int64_t>(n ? n : 1));
  std::vector<uint8_t> mask_all(n ? n : 1);
  int rc = BGC_OK;
#pragma omp parallel for schedule(dynamic, 64) if (n >= 256)
  for (long long i = 0; i < (long long)n; ++i) {
    bgc_review* r = bodies[i] ? bgc_review_parse(bodies[i], body_len[i], nullptr, 0) : nullptr;
    if (!r) continue;
    bgc_pod_spec spec;
    state[i] = 1;
    if (bgc_review_pod(r, &spec) == BGC_OK) {
      int64_t* row[BGC_MAX_R];
      for (uint32_t q = 0; q < R; ++q) row[q] = req_all[q].data() + i;
      size_t wrote = 0;
      int e = bgc_flatten_pods(&spec, 1, quota_keys, unit, R, map, 0, ns_all.data() + i, row, mask_all.data() + i, &wrote);
      if (e == BGC_OK && wrote != 1) e = BGC_E_BADARG;  // terminal pods cannot appear in create reviews
      if (e != BGC_OK) {
#pragma omp atomic write
        rc = e;
      } else {
        state[i] = 2;
        uid_of[i] = bgc_review_uid(r);
      }
    }
    bgc_review_destroy(r);
  }
  const uint64_t t_parse = ns_since(t0);
  t0 = clk::now();
  std::vector<size_t> live;
  live.reserve(n);
  for (size_t i = 0; i < n; ++i)
    if (state[i] == 2) live.push_back(i);
  const size_t m = live.size();
  std::vector<uint32_t> ns_id(m ? m : 1);
  std::vector<std::vector<int64_t>> req(R, std::vector<int64_t>(m ? m : 1));
  std::vector<const int64_t*> reqc(R);
  std::vector<uint8_t> mask(m ? m : 1);
  for (size_t k = 0; k < m; ++k) {
    ns_id[k] = ns_all[live[k]];
    mask[k] = mask_all[live[k]];
    for (uint32_t q = 0; q < R; ++q) req[q][k] = req_all[q][live[k]];
  }
  for (uint32_t q = 0; q < R; ++q) reqc[q] = req[q].data();
  const uint64_t t_flat = ns_since(t0);
  t0 = clk::now();
  std::vector<uint16_t> v16(wide && m ? m : 1);
  std::vector<uint8_t> v8(!wide && m ? m : 1);
  if (rc == BGC_OK && m)
    rc = wide ? bgc_quota_eval16(ctx, ns_id.data(), reqc.data(), mask.data(), m, v16.data())
              : bgc_quota_eval(ctx, ns_id.data(), reqc.data(), mask.data(), m, v8.data());
  const uint64_t t_eval = ns_since(t0);
  t0 = clk::now();
  std::vector<std::string> text(n);
  int bad_render = 0;
  if (rc == BGC_OK) {
    std::vector<long long> slot(n, -1);
    for (size_t k = 0; k < m; ++k) slot[live[k]] = (long long)k;
    const unsigned unknown = wide ? 0xFFFFu : 0xFFu;
#pragma omp parallel for schedule(static) if (n >= 256)
    for (long long i = 0; i < (long long)n; ++i) {
      bgc_admission_response resp = {0, BGC_DENY, 0, 0, nullptr};
      std::string msg;
      const char* uid = "";
      if (slot[i] < 0) {
        resp.klass = BGC_INVALID;
        msg = state[i] ? "invalid request: AdmissionReview carries no pod object" : "malformed AdmissionReview";
      } else {
        const size_t k = (size_t)slot[i];
        const unsigned v = wide ? v16[k] : v8[k];
        uid = uid_of[i].c_str();
        if (v == 0) {
          resp.allowed = 1; resp.klass = BGC_ALLOW;
        } else if (v == unknown) {
          msg = "namespace is not known to the quota engine";
        } else if (!used || !hard || ns_id[k] >= ns_count) {
          msg = "exceeded quota";
        } else {
          int64_t q[BGC_MAX_R], u[BGC_MAX_R], h[BGC_MAX_R];
          for (uint32_t r = 0; r < R; ++r) {
            q[r] = req[r][k];
            u[r] = used[(size_t)r * ns_count + ns_id[k]];
            h[r] = hard[(size_t)r * ns_count + ns_id[k]];
          }
          std::string buf(1024, '\0');
          long len;
          while ((len = bgc_render_quota_denial(v, wide, bgc_nsmap_name(map, ns_id[k]), quota_keys, unit, R, q, u, h, &buf[0], buf.size())) == BGC_E_BADARG &&
                 buf.size() < (1u << 20))
            buf.resize(buf.size() * 4);
          if (len > 0) msg.assign(buf.data(), (size_t)len);
          else msg = "exceeded quota";
        }
      }
      std::string buf(1024, '\0');
      long len;
      while ((len = bgc_render_review_response(uid, &resp, resp.allowed ? nullptr : msg.c_str(), nullptr, &buf[0], buf.size())) == BGC_E_BADARG &&
             buf.size() < (1u << 24))
        buf.resize(buf.size() * 4);
      if (len > 0) text[i].assign(buf.data(), (size_t)len);
      else
#pragma omp atomic write
        bad_render = 1;
    }
  }
  const uint64_t t_render = ns_since(t0);
  t0 = clk::now();
  if (rc) return rc;
  if (bad_render) return BGC_E_BADARG;
  size_t need = 0;
  for (size_t i = 0; i < n; ++i) {
    out_off[i] = need;
    need += text[i].size() + 1;
  }
  out_off[n] = need;
  if (need > out_cap) return BGC_E_BADARG;
  for (size_t i = 0; i < n; ++i) std::memcpy(out + out_off[i], text[i].c_str(), text[i].size() + 1);
  if (stage_ns) { stage_ns[0] = t_parse; stage_ns[1] = t_flat; stage_ns[2] = t_eval; stage_ns[3] = t_render; stage_ns[4] = ns_since(t0); }
  return BGC_OK;
}

// the C boundary: nothing is thrown across it (the batch handlers allocate in proportion to the batch)
extern "C" int bgc_handle_reviews(bgc_ctx* ctx, const bgc_admission_config* cfg, const char* const* bodies, const size_t* body_len,
                                  size_t n, char* out, size_t out_cap, size_t* out_off, uint64_t* stage_ns) {
  try {
    return handle_reviews_impl(ctx, cfg, bodies, body_len, n, out, out_cap, out_off, stage_ns);
  } catch (const std::bad_alloc&) {
    return BGC_E_OOM;
  } catch (...) {
    return BGC_E_STATE;
  }
}
extern "C" int bgc_handle_pod_reviews(bgc_ctx* ctx, const char* const* bodies, const size_t* body_len, size_t n,
                                      const char* const* quota_keys, const uint8_t* unit, uint32_t R, bgc_nsmap* map, int wide,
                                      const int64_t* used, const int64_t* hard, uint32_t ns_count, char* out, size_t out_cap,
                                      size_t* out_off, uint64_t* stage_ns) {
  try {
    return handle_pod_reviews_impl(ctx, bodies, body_len, n, quota_keys, unit, R, map, wide, used, hard, ns_count, out, out_cap, out_off, stage_ns);
  } catch (const std::bad_alloc&) {
    return BGC_E_OOM;
  } catch (...) {
    return BGC_E_STATE;
  }
}
