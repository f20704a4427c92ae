/* oracle/quota_ref.c — TEST INFRASTRUCTURE ONLY.  See quota_ref.h: PARITY UNPINNED (the
 * reference implements none of this; spec source = synchronizer.rs:249-281 key set + public
 * Kubernetes ResourceQuota semantics, restated in SURVEY.md §8 rows A3/A4). */
#include "quota_ref.h"
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static inline int nthreads(int threads) { return threads > 0 ? threads : 1; }

void quota_ref_usage_reduce(const uint32_t *ns_id, const int64_t *const *req, size_t n,
                            uint32_t NS, uint32_t R, int64_t *used, int threads) {
  const size_t tab = (size_t)R * NS;
  memset(used, 0, tab * sizeof(int64_t));
  const int T = nthreads(threads);
  if (T == 1) {
    /* the literal loop */
    for (size_t i = 0; i < n; ++i) {
      uint32_t ns = ns_id[i];
      if (ns >= NS) continue;
      for (uint32_t r = 0; r < R; ++r)
        used[(size_t)r * NS + ns] = (int64_t)((uint64_t)used[(size_t)r * NS + ns] + (uint64_t)req[r][i]);
    }
    return;
  }
#ifdef _OPENMP
  /* all-cores baseline: per-thread private tables merged at the end */
  uint64_t *priv = (uint64_t *)calloc((size_t)T * tab, sizeof(uint64_t));
#pragma omp parallel num_threads(T)
  {
    uint64_t *mine = priv + (size_t)omp_get_thread_num() * tab;
#pragma omp for schedule(static)
    for (long long i = 0; i < (long long)n; ++i) {
      uint32_t ns = ns_id[i];
      if (ns >= NS) continue;
      for (uint32_t r = 0; r < R; ++r) mine[(size_t)r * NS + ns] += (uint64_t)req[r][i];
    }
#pragma omp for schedule(static)
    for (long long k = 0; k < (long long)tab; ++k) {
      uint64_t s = 0;
      for (int t = 0; t < T; ++t) s += priv[(size_t)t * tab + k];
      used[k] = (int64_t)s;
    }
  }
  free(priv);
#else
  quota_ref_usage_reduce(ns_id, req, n, NS, R, used, 1);
#endif
}

/* returns missing | exceeded << 8 (or 0xFFFFFFFF for an unknown namespace); the callers pack it */
static inline uint32_t eval_bits(uint32_t ns, const int64_t *const *req, uint8_t rmask, size_t j,
                                 uint32_t NS, uint32_t R, const int64_t *used, const int64_t *hard,
                                 const uint8_t *hard_mask, uint8_t compute_mask) {
  if (ns >= NS) return 0xFFFFFFFFu;
  const uint8_t hm = hard_mask[ns];
  unsigned missing = 0, exceeded = 0;
  for (uint32_t r = 0; r < R; ++r) {
    if (!((hm >> r) & 1)) continue;               /* resource not constrained in this namespace */
    if (!((rmask >> r) & 1)) {                    /* pod does not state this resource */
      if ((compute_mask >> r) & 1) missing |= 1u << r;
      continue;
    }
    const int64_t q = req[r][j];
    if (q <= 0) continue;
    /* overflow-free: decided in 128 bits, as Kubernetes' arbitrary-precision Quantity would */
    const __int128 total = (__int128)used[(size_t)r * NS + ns] + (__int128)q;
    if (total > (__int128)hard[(size_t)r * NS + ns]) exceeded |= 1u << r;
  }
  return missing | (exceeded << 8);
}

static inline uint8_t eval_one(uint32_t ns, const int64_t *const *req, uint8_t rmask, size_t j,
                               uint32_t NS, uint32_t R, const int64_t *used, const int64_t *hard,
                               const uint8_t *hard_mask, uint8_t compute_mask) {
  const uint32_t b = eval_bits(ns, req, rmask, j, NS, R, used, hard, hard_mask, compute_mask);
  if (b == 0xFFFFFFFFu) return QUOTA_REF_VERDICT_BAD_NS;
  return (uint8_t)(((b & 0xFF) << 4) | (b >> 8));
}

void quota_ref_quota_eval(const uint32_t *ns_id, const int64_t *const *req,
                          const uint8_t *req_mask, size_t n, uint32_t NS, uint32_t R,
                          const int64_t *used, const int64_t *hard, const uint8_t *hard_mask,
                          uint8_t compute_mask, uint8_t *verdict, int threads) {
  (void)threads;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(nthreads(threads))
#endif
  for (long long j = 0; j < (long long)n; ++j)
    verdict[j] = eval_one(ns_id[j], req, req_mask ? req_mask[j] : 0xFF, (size_t)j, NS, R, used,
                          hard, hard_mask, compute_mask);
}

void quota_ref_quota_eval16(const uint32_t *ns_id, const int64_t *const *req,
                            const uint8_t *req_mask, size_t n, uint32_t NS, uint32_t R,
                            const int64_t *used, const int64_t *hard, const uint8_t *hard_mask,
                            uint8_t compute_mask, uint16_t *verdict, int threads) {
  (void)threads;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(nthreads(threads))
#endif
  for (long long j = 0; j < (long long)n; ++j) {
    const uint32_t b = eval_bits(ns_id[j], req, req_mask ? req_mask[j] : 0xFF, (size_t)j, NS, R, used, hard, hard_mask, compute_mask);
    verdict[j] = b == 0xFFFFFFFFu ? 0xFFFF : (uint16_t)(((b & 0xFF) << 8) | (b >> 8));
  }
}

void quota_ref_admit_greedy(const uint32_t *ns_id, const int64_t *const *req,
                            const uint8_t *req_mask, size_t n, uint32_t NS, uint32_t R,
                            int64_t *used, const int64_t *hard, const uint8_t *hard_mask,
                            uint8_t compute_mask, uint8_t *verdict) {
  for (size_t j = 0; j < n; ++j) {
    const uint32_t ns = ns_id[j];
    const uint8_t v = eval_one(ns, req, req_mask ? req_mask[j] : 0xFF, j, NS, R, used, hard, hard_mask, compute_mask);
    verdict[j] = v;
    if (v == 0) /* admitted: it now holds quota */
      for (uint32_t r = 0; r < R; ++r)
        used[(size_t)r * NS + ns] = (int64_t)((uint64_t)used[(size_t)r * NS + ns] + (uint64_t)req[r][j]);
  }
}

/* the same literal loop with u16 verdicts (missing << 8 | exceeded), any R <= 8 */
void quota_ref_admit_greedy16(const uint32_t *ns_id, const int64_t *const *req,
                              const uint8_t *req_mask, size_t n, uint32_t NS, uint32_t R,
                              int64_t *used, const int64_t *hard, const uint8_t *hard_mask,
                              uint8_t compute_mask, uint16_t *verdict) {
  for (size_t j = 0; j < n; ++j) {
    const uint32_t ns = ns_id[j];
    const uint32_t b = eval_bits(ns, req, req_mask ? req_mask[j] : 0xFF, j, NS, R, used, hard, hard_mask, compute_mask);
    const uint16_t v = b == 0xFFFFFFFFu ? 0xFFFF : (uint16_t)(((b & 0xFF) << 8) | (b >> 8));
    verdict[j] = v;
    if (v == 0)
      for (uint32_t r = 0; r < R; ++r)
        used[(size_t)r * NS + ns] = (int64_t)((uint64_t)used[(size_t)r * NS + ns] + (uint64_t)req[r][j]);
  }
}

/* ------------------------------------------------------------------ generator spec */
static inline uint64_t sm64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

uint64_t gen_ref_hash(uint64_t seed, uint32_t stream, uint64_t index) {
  const uint64_t key = sm64(seed ^ ((uint64_t)(stream + 1) * 0xD1B54A32D192ED03ull));
  return sm64(key ^ (index * 0x9E3779B97F4A7C15ull));
}

static inline uint32_t upper_bound_u64(const uint64_t *a, uint32_t n, uint64_t u) {
  uint32_t lo = 0, hi = n; /* first k with a[k] > u */
  while (lo < hi) {
    uint32_t mid = lo + ((hi - lo) >> 1);
    if (a[mid] > u) hi = mid; else lo = mid + 1;
  }
  return lo < n ? lo : n - 1;
}

static inline int64_t gen_value(uint32_t kind, uint64_t h) {
  const uint32_t t = (uint32_t)(h >> 32);
  switch (kind) {
    case GEN_KIND_GPU: { /* {0,1,2,4,8} w.p. {.5,.3,.1,.07,.03} */
      uint32_t c = t % 100u;
      return c < 50 ? 0 : c < 80 ? 1 : c < 90 ? 2 : c < 97 ? 4 : 8;
    }
    case GEN_KIND_CPU_MILLI: return (int64_t)250 << (t % 7u);          /* 250..16000 millicores */
    case GEN_KIND_MEM_BYTES: return (int64_t)1 << (28 + (t % 9u));     /* 2^28..2^36 bytes */
    default: { uint32_t c = t % 100u; return c < 90 ? 0 : c < 97 ? 1 : 2; } /* MIG slices */
  }
}

void gen_ref_fill(const gen_ref_params *p, uint64_t first, size_t count, uint32_t *ns_id,
                  int64_t *const *req, uint8_t *req_mask, int threads) {
  (void)threads;
  const uint32_t sbase = p->set * 16u;
  const uint64_t step = p->n_total ? UINT64_MAX / p->n_total : 0;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(nthreads(threads))
#endif
  for (long long k = 0; k < (long long)count; ++k) {
    const uint64_t i = first + (uint64_t)k;
    const uint64_t u = p->grouped ? i * step : gen_ref_hash(p->seed, sbase + 0, i);
    uint32_t rank;
    if (p->dist == GEN_DIST_ZIPF_CDF) rank = upper_bound_u64(p->cdf, p->ns_count, u);
    else rank = (uint32_t)(((u >> 32) * (uint64_t)p->ns_count) >> 32);
    ns_id[k] = (uint32_t)(((uint64_t)rank * p->scramble_a + p->scramble_b) % p->ns_count);
    for (uint32_t r = 0; r < p->R; ++r)
      req[r][k] = gen_value(p->kind[r], gen_ref_hash(p->seed, sbase + 1 + r, i));
    if (req_mask) {
      const uint64_t h = gen_ref_hash(p->seed, sbase + 8, i);
      unsigned m = 0xFF;
      for (uint32_t r = 0; r < p->R; ++r)
        if (((h >> (8 * r)) & 0xFF) < 8) m &= ~(1u << r); /* ~3 % of pods omit resource r */
      req_mask[k] = (uint8_t)m;
    }
  }
}

void gen_ref_policy_fill(uint64_t seed, uint64_t first, size_t count, uint8_t *op,
                         uint16_t *flags, int threads) {
  (void)threads;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(nthreads(threads))
#endif
  for (long long k = 0; k < (long long)count; ++k) {
    const uint64_t h = gen_ref_hash(seed, 64, first + (uint64_t)k);
    op[k] = (uint8_t)(h & 3);
    flags[k] = (uint16_t)((h >> 8) & 0x3FF);
  }
}

uint64_t gen_ref_checksum_u8(const uint8_t *v, uint64_t first, size_t count) {
  uint64_t s = 0;
  for (size_t k = 0; k < count; ++k) s += sm64((first + k) ^ ((uint64_t)v[k] << 56));
  return s;
}

/* ------------------------------------------------------------------ typed (narrow) columns
 * The same two functions over narrow host columns (the compact form the product's *_typed entry points
 * take): element types u8/u16/u32/i32/i64, value = widened element << shift (wrapping), ns ids u16 or
 * u32.  A block of pods is widened into a small thread-local buffer that stays in L1, then the loops
 * above run on it unchanged, so the typed forms decide exactly what the plain ones decide on the
 * widened columns. */
enum { QT_BLOCK = 2048 };

static inline int64_t load_typed(const void *col, uint32_t type, size_t i) {
  switch (type) {
    case QUOTA_REF_T_U32: return (int64_t)((const uint32_t *)col)[i];
    case QUOTA_REF_T_U16: return (int64_t)((const uint16_t *)col)[i];
    case QUOTA_REF_T_U8:  return (int64_t)((const uint8_t *)col)[i];
    case QUOTA_REF_T_I32: return (int64_t)((const int32_t *)col)[i];
    default:              return ((const int64_t *)col)[i];
  }
}

static void widen_block(const void *ns_id, uint32_t ns_type, const void *const *cols, const uint8_t *type,
                        const uint8_t *shift, uint32_t R, size_t i0, size_t m, uint32_t *ns_out,
                        int64_t (*q_out)[QT_BLOCK]) {
  if (ns_type == QUOTA_REF_T_U16) for (size_t k = 0; k < m; ++k) ns_out[k] = ((const uint16_t *)ns_id)[i0 + k];
  else for (size_t k = 0; k < m; ++k) ns_out[k] = ((const uint32_t *)ns_id)[i0 + k];
  for (uint32_t r = 0; r < R; ++r) {
    const unsigned sh = shift[r] & 63u;
    switch (type[r]) { /* one tight loop per element type so the compiler can vectorise the widening */
      case QUOTA_REF_T_U32: for (size_t k = 0; k < m; ++k) q_out[r][k] = (int64_t)((uint64_t)((const uint32_t *)cols[r])[i0 + k] << sh); break;
      case QUOTA_REF_T_U16: for (size_t k = 0; k < m; ++k) q_out[r][k] = (int64_t)((uint64_t)((const uint16_t *)cols[r])[i0 + k] << sh); break;
      case QUOTA_REF_T_U8:  for (size_t k = 0; k < m; ++k) q_out[r][k] = (int64_t)((uint64_t)((const uint8_t *)cols[r])[i0 + k] << sh); break;
      case QUOTA_REF_T_I32: for (size_t k = 0; k < m; ++k) q_out[r][k] = (int64_t)((uint64_t)(int64_t)((const int32_t *)cols[r])[i0 + k] << sh); break;
      default:              for (size_t k = 0; k < m; ++k) q_out[r][k] = (int64_t)((uint64_t)((const int64_t *)cols[r])[i0 + k] << sh); break;
    }
  }
  (void)load_typed;
}

/* workspace: threads * R * NS u64 (private tables), or NULL to allocate per call */
void quota_ref_usage_reduce_typed(const void *ns_id, uint32_t ns_type, const void *const *cols,
                                  const uint8_t *type, const uint8_t *shift, size_t n, uint32_t NS,
                                  uint32_t R, int64_t *used, int threads, uint64_t *workspace) {
  const size_t tab = (size_t)R * NS;
  const int T = nthreads(threads);
  uint64_t *priv = workspace ? workspace : (uint64_t *)malloc((size_t)T * tab * sizeof(uint64_t));
  const long long blocks = (long long)((n + QT_BLOCK - 1) / QT_BLOCK);
#ifdef _OPENMP
#pragma omp parallel num_threads(T)
#endif
  {
#ifdef _OPENMP
    uint64_t *mine = priv + (size_t)omp_get_thread_num() * tab;
#else
    uint64_t *mine = priv;
#endif
    memset(mine, 0, tab * sizeof(uint64_t));
    uint32_t ns[QT_BLOCK];
    int64_t q[QUOTA_REF_MAX_R][QT_BLOCK];
#ifdef _OPENMP
#pragma omp for schedule(static)
#endif
    for (long long b = 0; b < blocks; ++b) {
      const size_t i0 = (size_t)b * QT_BLOCK, m = n - i0 < QT_BLOCK ? n - i0 : QT_BLOCK;
      widen_block(ns_id, ns_type, cols, type, shift, R, i0, m, ns, q);
      for (size_t k = 0; k < m; ++k) {
        if (ns[k] >= NS) continue;
        for (uint32_t r = 0; r < R; ++r) mine[(size_t)r * NS + ns[k]] += (uint64_t)q[r][k];
      }
    }
#ifdef _OPENMP
#pragma omp for schedule(static)
#endif
    for (long long k = 0; k < (long long)tab; ++k) {
      uint64_t s = 0;
      for (int t = 0; t < T; ++t) s += priv[(size_t)t * tab + k];
      used[k] = (int64_t)s;
    }
  }
  if (!workspace) free(priv);
}

void quota_ref_quota_eval_typed(const void *ns_id, uint32_t ns_type, const void *const *cols,
                                const uint8_t *type, const uint8_t *shift, const uint8_t *req_mask,
                                size_t n, uint32_t NS, uint32_t R, const int64_t *used,
                                const int64_t *hard, const uint8_t *hard_mask, uint8_t compute_mask,
                                uint8_t *verdict, int threads) {
  const long long blocks = (long long)((n + QT_BLOCK - 1) / QT_BLOCK);
#ifdef _OPENMP
#pragma omp parallel num_threads(nthreads(threads))
#endif
  {
    uint32_t ns[QT_BLOCK];
    int64_t q[QUOTA_REF_MAX_R][QT_BLOCK];
    const int64_t *qp[QUOTA_REF_MAX_R];
    for (uint32_t r = 0; r < QUOTA_REF_MAX_R; ++r) qp[r] = q[r];
#ifdef _OPENMP
#pragma omp for schedule(static)
#endif
    for (long long b = 0; b < blocks; ++b) {
      const size_t i0 = (size_t)b * QT_BLOCK, m = n - i0 < QT_BLOCK ? n - i0 : QT_BLOCK;
      widen_block(ns_id, ns_type, cols, type, shift, R, i0, m, ns, q);
      for (size_t k = 0; k < m; ++k)
        verdict[i0 + k] = eval_one(ns[k], qp, req_mask ? req_mask[i0 + k] : 0xFF, k, NS, R, used, hard,
                                   hard_mask, compute_mask);
    }
  }
}

/* plain-column usage reduction with a caller-owned workspace (threads * R * NS u64): the all-cores
 * baseline without a calloc of fresh pages inside every timed call */
void quota_ref_usage_reduce_ws(const uint32_t *ns_id, const int64_t *const *req, size_t n, uint32_t NS,
                               uint32_t R, int64_t *used, int threads, uint64_t *workspace) {
  uint8_t type[QUOTA_REF_MAX_R], shift[QUOTA_REF_MAX_R];
  const void *cols[QUOTA_REF_MAX_R];
  for (uint32_t r = 0; r < R; ++r) { type[r] = QUOTA_REF_T_I64; shift[r] = 0; cols[r] = req[r]; }
  quota_ref_usage_reduce_typed(ns_id, QUOTA_REF_T_U32, cols, type, shift, n, NS, R, used, threads, workspace);
}
