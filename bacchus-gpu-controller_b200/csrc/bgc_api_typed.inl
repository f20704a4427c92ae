// bgc_api_typed.inl — host side of the narrow-column entry points (bgc_*_typed, include/bgc.h).
// Textually included at the end of bgc_api.cu: it uses that file's context struct and helpers.

namespace {

struct TypedArgs {
  const void* ns = nullptr;
  uint32_t ns_type = T_U32;
  TCol col[kMaxR];
};

int check_typed(bgc_ctx* ctx, const void* ns_id, int ns_type, const bgc_col* req, size_t n, TypedArgs* out) {
  NEED(ns_type == BGC_T_U32 || ns_type == BGC_T_U16, "ns_type must be BGC_T_U32 or BGC_T_U16");
  NEED(n == 0 || (ns_id && req), "null column pointer");
  NEED(ctx->R <= kMaxR8, "the typed entry points emit u8 verdicts: R <= 4");
  out->ns = ns_id;
  out->ns_type = (uint32_t)ns_type;
  for (int r = 0; r < ctx->R && n; ++r) {
    NEED(req[r].data, "null request column");
    NEED(req[r].type <= BGC_T_I32, "unknown column type");
    NEED(req[r].shift < 64, "shift must be 0..63");
    out->col[r].p = req[r].data;
    out->col[r].type = req[r].type;
    out->col[r].shift = req[r].shift;
  }
  return BGC_OK;
}

// every column aligned for its PODS-wide vector load?  all of them 16-byte aligned (bulk prefetch)?
void typed_alignment(const TypedArgs& a, int R, bool* vec, bool* pf_ok) {
  const int pods = R <= 4 ? 4 : 2;
  *vec = aligned(a.ns, type_width(a.ns_type) * pods);
  *pf_ok = aligned(a.ns, 16);
  for (int r = 0; r < R; ++r) {
    *vec = *vec && aligned(a.col[r].p, type_width(a.col[r].type) * pods);
    *pf_ok = *pf_ok && aligned(a.col[r].p, 16);
  }
}

template <int R>
int launch_k1_typed_r(bgc_ctx* ctx, cudaStream_t st, const TypedArgs& a, size_t n) {
  TypedCols<R> in;
  for (int r = 0; r < R; ++r) in.c[r] = a.col[r];
  in.ns = a.ns;
  in.ns_type = a.ns_type;
  bool vec, pf_ok;
  typed_alignment(a, R, &vec, &pf_ok);
  const size_t slot_bytes = k1_slot_bytes(ctx, R);
  const bool direct = (size_t)ctx->NS * slot_bytes <= kSmemBudget && ctx->k1_forced != BGC_K1_SMEM_HASH;
  const uint32_t cap = direct ? ctx->NS : (uint32_t)(kSmemMaxDynamic / (slot_bytes + 4));
  const size_t smem = slot_bytes * cap + (direct ? 0 : (size_t)cap * 4);
  const int threads = 1024;
  const size_t groups = (n + 1) / 2;
  int grid = ctx->sm_count;
  if ((size_t)grid * threads > groups) grid = (int)((groups + threads - 1) / threads);
  if (grid < 1) grid = 1;
  const uint32_t pf = pf_ok ? (uint32_t)(ctx->k1_prefetch >= 0 ? ctx->k1_prefetch : (direct ? 0 : 1)) : 0u;
  if (direct) {
    CU(cudaFuncSetAttribute(k1_scatter_typed<R, SINK_DIRECT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemMaxDynamic));
    k1_scatter_typed<R, SINK_DIRECT><<<grid, threads, smem, st>>>(in, n, ctx->d_used, ctx->NS, cap, ctx->wide_mask, pf, vec ? 1u : 0u);
    ctx->last_k1 = "typed_scatter_smem_direct";
  } else {
    CU(cudaFuncSetAttribute(k1_scatter_typed<R, SINK_HASH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemMaxDynamic));
    k1_scatter_typed<R, SINK_HASH><<<grid, threads, smem, st>>>(in, n, ctx->d_used, ctx->NS, cap, ctx->wide_mask, pf, vec ? 1u : 0u);
    ctx->last_k1 = "typed_scatter_smem_hash";
  }
  ctx->head_stale = true;
  ctx->tm.launches++;
  CU(cudaGetLastError());
  return BGC_OK;
}

int launch_k1_typed(bgc_ctx* ctx, cudaStream_t st, const TypedArgs& a, size_t n) {
  if (n == 0) return BGC_OK;
  switch (ctx->R) {
    case 1: return launch_k1_typed_r<1>(ctx, st, a, n);
    case 2: return launch_k1_typed_r<2>(ctx, st, a, n);
    case 3: return launch_k1_typed_r<3>(ctx, st, a, n);
    default: return launch_k1_typed_r<4>(ctx, st, a, n);
  }
}

template <int R, bool HAS_MASK>
int launch_k2_typed_inst(bgc_ctx* ctx, cudaStream_t st, const TypedArgs& a, const uint8_t* mask, size_t n, uint8_t* verdict) {
  TypedCols<R> in;
  for (int r = 0; r < R; ++r) in.c[r] = a.col[r];
  in.ns = a.ns;
  in.ns_type = a.ns_type;
  bool vec, pf_ok;
  typed_alignment(a, R, &vec, &pf_ok);
  vec = vec && aligned(verdict, 4) && (!HAS_MASK || aligned(mask, 4));
  if (ctx->head_stale) {
    const int g = grid_for((size_t)ctx->NS * head_words(R), 256, ctx->sm_count * 4);
    k_build_headroom<<<g, 256, 0, st>>>(ctx->d_used, ctx->d_hard, ctx->d_hard_mask, ctx->head, ctx->NS, R);
    ctx->tm.launches++;
    CU(cudaGetLastError());
    ctx->head_stale = false;
  }
  const int grid = grid_for((n + 3) / 4, kK2Threads, ctx->sm_count * 32);
  const uint32_t pf = pf_ok ? (uint32_t)(ctx->k2_prefetch >= 0 ? ctx->k2_prefetch : 1) : 0u;
  k2_quota_eval_typed<R, HAS_MASK, uint8_t><<<grid, kK2Threads, 0, st>>>(in, mask, n, ctx->head, ctx->d_hard_mask, ctx->NS,
                                                                        ctx->compute_mask, verdict, pf, vec ? 1u : 0u);
  ctx->tm.launches++;
  CU(cudaGetLastError());
  return BGC_OK;
}

int launch_k2_typed(bgc_ctx* ctx, cudaStream_t st, const TypedArgs& a, const uint8_t* mask, size_t n, uint8_t* verdict) {
  if (n == 0) return BGC_OK;
#define K2T(RR)                                                                              \
  case RR:                                                                                   \
    return mask ? launch_k2_typed_inst<RR, true>(ctx, st, a, mask, n, verdict)               \
                : launch_k2_typed_inst<RR, false>(ctx, st, a, mask, n, verdict);
  switch (ctx->R) {
    K2T(1) K2T(2) K2T(3)
    default: return mask ? launch_k2_typed_inst<4, true>(ctx, st, a, mask, n, verdict)
                         : launch_k2_typed_inst<4, false>(ctx, st, a, mask, n, verdict);
  }
#undef K2T
}

// Host-pointer form of both: the same chunked three-stream pipeline as the int64 entry points (H2D,
// compute, D2H of consecutive 4 Mi-pod chunks overlap), copying every column at its own width into the
// int64-sized staging slots.  verdict == nullptr: usage reduction; otherwise quota evaluation.
int typed_host_pipeline(bgc_ctx* ctx, const TypedArgs& a, const uint8_t* req_mask, size_t n, uint8_t* verdict) {
  const int R = ctx->R;
  int rc;
  if ((rc = ensure_slots(ctx, n, R))) return rc;
  ctx->ev_next = 0;
  Spans sp;
  const size_t chunk = ctx->slot_pods;
  const bool single = n <= chunk;
  cudaStream_t hs = single ? ctx->own_stream : ctx->h2d_stream;
  cudaStream_t ds = single ? ctx->own_stream : ctx->d2h_stream;
  const size_t nsw = type_width(a.ns_type);
  size_t k = 0;
  for (size_t off = 0; off < n; off += chunk, ++k) {
    Slot& s = ctx->slot[k % kSlots];
    const size_t m = (n - off < chunk) ? n - off : chunk;
    cudaEvent_t ea = next_event(ctx), eb = next_event(ctx), ec = next_event(ctx), ed = next_event(ctx),
                ee = next_event(ctx), ef = next_event(ctx);
    if (!ef) return fail(ctx, BGC_E_CUDA, "event pool exhausted");
    if (s.used && !single) CU(cudaStreamWaitEvent(hs, s.free_ev, 0));
    CU(cudaEventRecord(ea, hs));
    TypedArgs d = a;  // the same columns, in the staging slot
    CU(cudaMemcpyAsync(s.ns, static_cast<const char*>(a.ns) + off * nsw, m * nsw, cudaMemcpyHostToDevice, hs));
    d.ns = s.ns;
    for (int r = 0; r < R; ++r) {
      const size_t w = type_width(a.col[r].type);
      CU(cudaMemcpyAsync(s.req[r], static_cast<const char*>(a.col[r].p) + off * w, m * w, cudaMemcpyHostToDevice, hs));
      d.col[r].p = s.req[r];
    }
    if (verdict && req_mask) CU(cudaMemcpyAsync(s.mask, req_mask + off, m, cudaMemcpyHostToDevice, hs));
    CU(cudaEventRecord(eb, hs));
    if (!single) {
      CU(cudaEventRecord(s.h2d_done, hs));
      CU(cudaStreamWaitEvent(ctx->own_stream, s.h2d_done, 0));
    }
    CU(cudaEventRecord(ec, ctx->own_stream));
    if (verdict) rc = launch_k2_typed(ctx, ctx->own_stream, d, req_mask ? s.mask : nullptr, m, s.out0);
    else rc = launch_k1_typed(ctx, ctx->own_stream, d, m);
    if (rc) return rc;
    CU(cudaEventRecord(ed, ctx->own_stream));
    sp.h2d.push_back(ea); sp.h2d.push_back(eb);
    sp.k.push_back(ec); sp.k.push_back(ed);
    if (verdict) {
      if (!single) {
        CU(cudaEventRecord(s.k_done, ctx->own_stream));
        CU(cudaStreamWaitEvent(ds, s.k_done, 0));
      }
      CU(cudaEventRecord(ee, ds));
      CU(cudaMemcpyAsync(verdict + off, s.out0, m, cudaMemcpyDeviceToHost, ds));
      CU(cudaEventRecord(ef, ds));
      if (!single) CU(cudaEventRecord(s.free_ev, ds));
      sp.d2h.push_back(ee); sp.d2h.push_back(ef);
    } else if (!single) {
      CU(cudaEventRecord(s.free_ev, ctx->own_stream));
    }
    s.used = !single;
  }
  CU(cudaStreamSynchronize(ctx->own_stream));
  if (verdict) CU(cudaStreamSynchronize(ds));
  return finish_timings(ctx, sp);
}

}  // namespace

extern "C" {

int bgc_usage_reduce_typed_dev(bgc_ctx* ctx, const void* d_ns_id, int ns_type, const bgc_col* d_req, size_t n,
                               unsigned flags, int64_t* d_used_out) {
  if (!ctx) return BGC_E_BADARG;
  if (!ctx->d_used) return fail(ctx, BGC_E_STATE, "quota table not set");
  TypedArgs a;
  int rc;
  if ((rc = check_typed(ctx, d_ns_id, ns_type, d_req, n, &a))) return rc;
  CU(cudaSetDevice(ctx->device));
  if (!(flags & BGC_REDUCE_ACCUMULATE) && (rc = clear_used(ctx, ctx->stream))) return rc;
  if ((rc = launch_k1_typed(ctx, ctx->stream, a, n))) return rc;
  if (d_used_out) return export_used(ctx, ctx->stream, reinterpret_cast<long long*>(d_used_out));
  return BGC_OK;
}

int bgc_quota_eval_typed_dev(bgc_ctx* ctx, const void* d_ns_id, int ns_type, const bgc_col* d_req,
                             const uint8_t* d_req_mask, size_t n, uint8_t* d_verdict) {
  if (!ctx) return BGC_E_BADARG;
  if (!ctx->d_used) return fail(ctx, BGC_E_STATE, "quota table not set");
  NEED(n == 0 || d_verdict, "null verdict pointer");
  TypedArgs a;
  int rc;
  if ((rc = check_typed(ctx, d_ns_id, ns_type, d_req, n, &a))) return rc;
  CU(cudaSetDevice(ctx->device));
  return launch_k2_typed(ctx, ctx->stream, a, d_req_mask, n, d_verdict);
}

int bgc_usage_reduce_typed(bgc_ctx* ctx, const void* ns_id, int ns_type, const bgc_col* req, size_t n, unsigned flags,
                           int64_t* used_out) {
  if (!ctx) return BGC_E_BADARG;
  if (!ctx->d_used) return fail(ctx, BGC_E_STATE, "quota table not set");
  TypedArgs a;
  int rc;
  if ((rc = check_typed(ctx, ns_id, ns_type, req, n, &a))) return rc;
  CU(cudaSetDevice(ctx->device));
  ctx->tm.h2d_ns = ctx->tm.kernel_ns = ctx->tm.d2h_ns = 0;
  if (ctx->stream != ctx->own_stream) CU(cudaStreamSynchronize(ctx->stream));
  if (wants_bounce(ctx, ns_id, n)) {  // pageable caller: pinned bounce ring; pieces after the first accumulate
    BounceCol cols[kBounceMaxCols];
    cols[0] = {ns_id, type_width(a.ns_type)};
    for (int r = 0; r < ctx->R; ++r) cols[1 + r] = {req[r].data, type_width(req[r].type)};
    rc = bounced(ctx, cols, 1 + ctx->R, nullptr, 0, n, [&](const void* const* in, void* const*, size_t m, bool first) {
      bgc_col c2[kMaxR];
      for (int r = 0; r < ctx->R; ++r) { c2[r] = req[r]; c2[r].data = in[1 + r]; }
      return bgc_usage_reduce_typed(ctx, in[0], ns_type, c2, m, first ? flags : (flags | BGC_REDUCE_ACCUMULATE), nullptr);
    });
    if (rc) return rc;
    return used_out ? bgc_usage_get(ctx, used_out) : BGC_OK;
  }
  if (!(flags & BGC_REDUCE_ACCUMULATE) && (rc = clear_used(ctx, ctx->own_stream))) return rc;
  if (n && (rc = typed_host_pipeline(ctx, a, nullptr, n, nullptr))) return rc;
  if (used_out) {
    CU(cudaMemcpyAsync(used_out, ctx->d_used, (size_t)ctx->NS * ctx->R * 8, cudaMemcpyDeviceToHost, ctx->own_stream));
  }
  CU(cudaStreamSynchronize(ctx->own_stream));
  return BGC_OK;
}

int bgc_quota_eval_typed(bgc_ctx* ctx, const void* ns_id, int ns_type, const bgc_col* req, const uint8_t* req_mask,
                         size_t n, uint8_t* verdict) {
  if (!ctx) return BGC_E_BADARG;
  if (!ctx->d_used) return fail(ctx, BGC_E_STATE, "quota table not set");
  NEED(n == 0 || verdict, "null verdict pointer");
  TypedArgs a;
  int rc;
  if ((rc = check_typed(ctx, ns_id, ns_type, req, n, &a))) return rc;
  CU(cudaSetDevice(ctx->device));
  ctx->tm.h2d_ns = ctx->tm.kernel_ns = ctx->tm.d2h_ns = 0;
  if (ctx->stream != ctx->own_stream) CU(cudaStreamSynchronize(ctx->stream));
  if (n == 0) return BGC_OK;
  if (wants_bounce(ctx, ns_id, n)) {  // pageable caller: pinned bounce ring
    BounceCol cols[kBounceMaxCols];
    int nc = 0;
    cols[nc++] = {ns_id, type_width(a.ns_type)};
    for (int r = 0; r < ctx->R; ++r) cols[nc++] = {req[r].data, type_width(req[r].type)};
    if (req_mask) cols[nc++] = {req_mask, 1};
    const BounceOut outs[1] = {{verdict, 1}};
    return bounced(ctx, cols, nc, outs, 1, n, [&](const void* const* in, void* const* out, size_t m, bool) {
      bgc_col c2[kMaxR];
      for (int r = 0; r < ctx->R; ++r) { c2[r] = req[r]; c2[r].data = in[1 + r]; }
      return bgc_quota_eval_typed(ctx, in[0], ns_type, c2, req_mask ? static_cast<const uint8_t*>(in[1 + ctx->R]) : nullptr, m,
                                  static_cast<uint8_t*>(out[0]));
    });
  }
  return typed_host_pipeline(ctx, a, req_mask, n, verdict);
}

}  // extern "C"
