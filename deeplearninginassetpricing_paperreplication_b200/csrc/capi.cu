// extern "C" surface of libdlap_b200.so (declared in include/dlap.h): argument checks and
// kernel launches only.  No allocation, no synchronisation, no torch types.
#include "../../include/dlap.h"
#include "cell_f32.cuh"
#include "grad_finish.cuh"
#include "macro_role.cuh"

namespace dlap {
// launchers implemented next to their kernels
cudaError_t launch_cell_fwd_f32(const CellParams& P, int grid, cudaStream_t st);
cudaError_t launch_cell_bwd_f32(const CellParams& P, int grid, bool sdf, bool mom, cudaStream_t st);
cudaError_t launch_cell_fwd_tc(const CellParams& P, const MacroParams* MP, int grid, bool cooperative, cudaStream_t st);
cudaError_t launch_cell_bwd_tc(const CellParams& P, int grid, bool sdf, bool mom, cudaStream_t st);
size_t smem_bwd_bytes(int F);
cudaError_t launch_panel_index(const uint8_t* mask, int T, int N, int* index, cudaStream_t st);
cudaError_t launch_panel_image(const CellParams& P, uint8_t* image, int sms, cudaStream_t st);
size_t panel_image_tile_bytes();
cudaError_t launch_panel_pull(const float* x_host, const int* index, int T, int N, int F, float* x_dev, int sms, bool clear,
                              cudaStream_t st);
cudaError_t launch_panel_ingest(const void* raw, int raw_f64, long long cells, int F, float thr, float* x, float* R,
                                uint8_t* mask, int ctas, cudaStream_t st);
cudaError_t launch_macro_zscore(const float* raw, int T, int M, float* mean, float* stdv, int stats_given, int pairwise,
                                float* out, cudaStream_t st);
cudaError_t launch_step_reduce_c(const dlap_plan* plan, const float* u, const float* q, const float* R, const uint8_t* mask,
                                 const int* index, int weighted, int flags, const float* n_bar_in, int phase_lo, int phase_hi,
                                 float* row_stats, float* pr, float* scalars, float* w_out, double* col_part, float* e,
                                 float* t_i, float* G, double* loss_part, double* dS2, double* row_part,
                                 unsigned int* bar, float* xbuf, int T_series, int month0, int month_stride,
                                 double* row_sums, int n_total, const dlap_peers* peers, cudaStream_t st);
cudaError_t launch_loss_finish(const double*, int, int, int, float*, float*, double*, unsigned int*, float*,
                               cudaStream_t);
cudaError_t launch_grad_finish(const FinishParams& P, cudaStream_t st);
cudaError_t launch_weights(const float*, const uint8_t*, const int*, int, int, int, int, float*, cudaStream_t);
void fill_macro(MacroParams& P, const dlap_plan* plan, int T, const float* macro, const dlap_tensors* p, const float* h0,
                const float* c0);
cudaError_t launch_macro_fwd(const MacroParams& P, cudaStream_t st);
cudaError_t launch_macro_bwd(const MacroParams& P, const FinishParams& FIN, cudaStream_t st);
cudaError_t launch_const_bias(const dlap_tensors* p, int T, float* a_t, float* c_t, cudaStream_t st);
size_t macro_fwd_smem(int T, int M, int H);
void set_bwd_tc_timeline(long long* p);
void set_fwd_tc_timeline(long long* p);
void set_bptt_timeline(long long* p);
void set_reduce_timeline(long long* p);
size_t macro_bwd_smem(int T, int H);
}  // namespace dlap

using namespace dlap;

#define REQ(p) do { if (!(p)) return DLAP_E_NULL; } while (0)
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return (int)e_; } while (0)

static int check_dims(const dlap_dims& d) {
  if (d.T < 1 || d.N < 1 || d.F < 1 || d.F > FMAX) return DLAP_E_SHAPE;
  if (d.M < 0 || (d.M > 0 && (d.H < 1 || d.H > HMAX))) return DLAP_E_SHAPE;
  if ((long long)d.T * d.N >= (1ll << 32)) return DLAP_E_SHAPE;      // cell numbers (dropout key, packed offsets) are 32-bit
  if ((long long)d.N * d.F * 4 >= (1ll << 32)) return DLAP_E_SHAPE;   // byte offsets inside a month are 32-bit
  return DLAP_OK;
}

static int check_plan(const dlap_plan* p) {
  REQ(p);
  int rc = check_dims(p->dims);
  if (rc) return rc;
  if (p->tpm != (p->dims.N + TILE - 1) / TILE || p->grid_cell < 1 || p->sms < 1 || p->grid_cell > p->sms ||
      p->col_splits < 1 || p->col_splits > p->dims.T)
    return DLAP_E_PLAN;
  return DLAP_OK;
}

extern "C" {

int dlap_abi_version(void) { return DLAP_ABI_VERSION; }

int dlap_debug_timeline(long long* stamps) {
  set_bwd_tc_timeline(stamps);
  set_fwd_tc_timeline(stamps ? stamps + 256 * 16 : nullptr);
  set_bptt_timeline(stamps ? stamps + 3 * 256 * 16 : nullptr);
  set_reduce_timeline(stamps ? stamps + 3 * 256 * 16 + 2048 : nullptr);
  return DLAP_OK;
}

const char* dlap_error_string(int code) {
  switch (code) {
    case DLAP_OK: return "ok";
    case DLAP_E_NULL: return "required pointer is NULL";
    case DLAP_E_SHAPE: return "dimension outside the supported range (F<=64, H<=8, T,N>=1)";
    case DLAP_E_ALIGN: return "pointer alignment";
    case DLAP_E_PLAN: return "plan does not match the dims";
    case DLAP_E_PRECISION: return "unknown precision selector";
    case DLAP_E_NODEVICE: return "no usable CUDA device";
    case DLAP_E_SMEM: return "configuration needs more shared memory than an SM offers";
    case DLAP_E_LIMIT: return "size exceeds the fixed capacity of this no-allocation entry point "
                              "(see DLAP_NOALLOC_MAX_STOCKS in dlap.h)";
    default: return code > 0 ? cudaGetErrorString((cudaError_t)code) : "unknown dlap error";
  }
}

int dlap_make_plan(const dlap_dims* dims, int device, dlap_plan* out) {
  REQ(dims); REQ(out);
  int rc = check_dims(*dims);
  if (rc) return rc;
  int sms = 0, smem_max = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || sms < 1) {
    cudaGetLastError();
    return DLAP_E_NODEVICE;
  }
  CU(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
  if ((size_t)smem_max < smem_bwd_bytes(dims->F)) return DLAP_E_SMEM;
  if (dims->M > 0 && ((size_t)smem_max < macro_fwd_smem(dims->T, dims->M, dims->H) ||
                      (size_t)smem_max < macro_bwd_smem(dims->T, dims->H))) return DLAP_E_SMEM;
  dlap_plan p;
  p.dims = *dims;
  const int T = dims->T, N = dims->N, F = dims->F;
  p.tpm = (N + TILE - 1) / TILE;
  const long long ntiles = (long long)T * p.tpm;
  p.sms = sms;
  p.grid_cell = (int)(ntiles < sms ? ntiles : sms);
  p.cooperative = 1;
  // reductions: aim for ~4 CTAs per SM.  col_partial holds exactly four CTAs of 256 threads per SM (64 registers),
  // so its grid must not exceed 4 * sms: a few CTAs more would run as a second wave on an otherwise idle GPU and
  // double the kernel (round down, not up)
  const int nb = (N + 255) / 256, want = 4 * sms;
  int cs = want / nb; if (cs > T) cs = T; if (cs < 1) cs = 1;
  p.col_splits = cs;
  p.ldw1 = F + (dims->M > 0 ? dims->H : 0);
  p.ldwq = dims->M + F;
  p.col_part_bytes = (int64_t)p.col_splits * (KM + 2) * N * 8;
  p.loss_part_bytes = (int64_t)((N + 63) / 64) * 2 * 8;
  p.row_part_bytes = (int64_t)T * 8 * 4 * 8;
  p.bar_bytes = (int64_t)(1 + T) * 4;
  p.wpart_bytes = (int64_t)p.grid_cell * wp_stride(round4(F)) * 4;
  p.dadc_part_bytes = (int64_t)T * (p.tpm > 4 ? p.tpm : 4) * 72 * 4;
  p.index_bytes = (int64_t)idx_ints(T, N) * 4;
  p.image_bytes = F <= 46 ? (int64_t)T * p.tpm * (int64_t)panel_image_tile_bytes() : 0;
  p.lstm_save_bytes = (int64_t)T * 6 * (dims->H > 0 ? dims->H : 1) * 4;
  p.dpre_bytes = (int64_t)T * 4 * (dims->H > 0 ? dims->H : 1) * 4;
  *out = p;
  return DLAP_OK;
}

static int fill_cell(CellParams& P, const dlap_plan* plan, const float* x, const float* R, const uint8_t* mask,
                     const dlap_tensors* w, const dlap_macro_io* mio) {
  const dlap_dims& d = plan->dims;
  P = CellParams{};
  P.x = x; P.R = R; P.mask = mask;
  P.T = d.T; P.N = d.N; P.F = d.F;
  P.x_bytes = (long long)d.T * d.N * d.F * 4;
  P.W1 = w->w1; P.ldw1 = plan->ldw1; P.W2 = w->w2; P.b2 = w->b2; P.W3 = w->w3; P.b3 = w->b3;
  P.Wq = w->wq; P.ldwq = plan->ldwq; P.wq_off = d.M;
  // month t of the panel is month month0 + t * month_stride of the series the tables run over
  const int stride = mio->month_stride > 0 ? mio->month_stride : 1;
  const int series = mio->series_T > 0 ? mio->series_T : d.T;
  if (mio->month0 < 0 || (long long)mio->month0 + (long long)(d.T - 1) * stride >= series) return DLAP_E_SHAPE;
  P.a_t = mio->a_t + (size_t)mio->month0 * D1;
  P.c_t = mio->c_t ? mio->c_t + (size_t)mio->month0 * KM : nullptr;
  P.months_base = mio->month0;
  P.months_stride = stride;
  P.tpm = plan->tpm;
  return DLAP_OK;
}

static int set_dropout(CellParams& P, float p, uint64_t seed, const uint64_t* seed_dev) {
  P.seed_ptr = reinterpret_cast<const uint32_t*>(seed_dev);
  if (!(p >= 0.f) || p >= 1.f) return DLAP_E_SHAPE;
  uint32_t thr = (uint32_t)(p * 65536.f + 0.5f);
  if (thr > 65535u) thr = 65535u;
  P.drop_thr = thr;
  P.drop_scale = thr ? 65536.f / (float)(65536u - thr) : 1.f;
  P.seed_lo = (uint32_t)seed;
  P.seed_hi = (uint32_t)(seed >> 32);
  return DLAP_OK;
}

int dlap_macro_fwd(const dlap_plan* plan, const float* macro, const dlap_tensors* params, const float* h0,
                   const float* c0, float* lstm_save, float* a_t, float* c_t, float* hn, float* cn, void* stream) {
  int rc = check_plan(plan);
  if (rc) return rc;
  REQ(params); REQ(a_t); REQ(c_t); REQ(params->b1); REQ(params->bq);
  if (plan->dims.M == 0) {
    CU(launch_const_bias(params, plan->dims.T, a_t, c_t, (cudaStream_t)stream));
    return DLAP_OK;
  }
  REQ(macro); REQ(lstm_save); REQ(hn); REQ(cn);
  REQ(params->w_ih); REQ(params->w_hh); REQ(params->b_ih); REQ(params->b_hh); REQ(params->w1); REQ(params->wq);
  MacroParams P;
  fill_macro(P, plan, plan->dims.T, macro, params, h0, c0);
  P.save = lstm_save; P.a_t = a_t; P.c_t = c_t; P.hn = hn; P.cn = cn;
  CU(launch_macro_fwd(P, (cudaStream_t)stream));
  return DLAP_OK;
}

int dlap_panel_index(const dlap_plan* plan, const uint8_t* mask, int32_t* index, void* stream) {
  int rc = check_plan(plan);
  if (rc) return rc;
  REQ(mask); REQ(index);
  if (reinterpret_cast<uintptr_t>(index) & 15) return DLAP_E_ALIGN;
  CU(launch_panel_index(mask, plan->dims.T, plan->dims.N, index, (cudaStream_t)stream));
  return DLAP_OK;
}

int dlap_panel_pull(const dlap_plan* plan, const float* x_host, const int32_t* index, float* x_dev, int ctas,
                    int clear, void* stream) {
  int rc = check_plan(plan);
  if (rc) return rc;
  REQ(x_host); REQ(index); REQ(x_dev);
  if ((reinterpret_cast<uintptr_t>(x_host) | reinterpret_cast<uintptr_t>(x_dev)) & 3) return DLAP_E_ALIGN;
  if (ctas < 0) return DLAP_E_SHAPE;
  CU(launch_panel_pull(x_host, index, plan->dims.T, plan->dims.N, plan->dims.F, x_dev,
                       ctas > 0 ? ctas : plan->grid_cell, clear != 0, (cudaStream_t)stream));
  return DLAP_OK;
}

int dlap_panel_ingest(const dlap_plan* plan, const void* raw, int raw_f64, float valid_above, float* x, float* R,
                      uint8_t* mask, int ctas, void* stream) {
  int rc = check_plan(plan);
  if (rc) return rc;
  REQ(raw); REQ(x); REQ(R); REQ(mask);
  if ((reinterpret_cast<uintptr_t>(raw) | reinterpret_cast<uintptr_t>(x)) & 15) return DLAP_E_ALIGN;   // float4 chunks
  if (reinterpret_cast<uintptr_t>(R) & 3) return DLAP_E_ALIGN;
  if (ctas < 0 || (raw_f64 != 0 && raw_f64 != 1)) return DLAP_E_SHAPE;
  CU(launch_panel_ingest(raw, raw_f64, (long long)plan->dims.T * plan->dims.N, plan->dims.F, valid_above, x, R, mask,
                         ctas > 0 ? ctas : plan->sms * 8, (cudaStream_t)stream));
  return DLAP_OK;
}

int dlap_macro_zscore(const float* macro_raw, int T, int M, float* mean, float* std, int stats_given, int pairwise,
                      float* out, void* stream) {
  if (T < 0 || M < 0) return DLAP_E_SHAPE;
  if (T == 0 || M == 0) return DLAP_OK;
  REQ(macro_raw); REQ(mean); REQ(std); REQ(out);
  CU(launch_macro_zscore(macro_raw, T, M, mean, std, stats_given != 0, pairwise != 0, out, (cudaStream_t)stream));
  return DLAP_OK;
}

int dlap_panel_image(const dlap_plan* plan, const float* x, const int32_t* index, void* image, void* stream) {
  int rc = check_plan(plan);
  if (rc) return rc;
  REQ(x); REQ(index); REQ(image);
  if (plan->dims.F > 46) return DLAP_E_SHAPE;
  if ((reinterpret_cast<uintptr_t>(x) & 3) || (reinterpret_cast<uintptr_t>(image) & 15)) return DLAP_E_ALIGN;
  CellParams P{};
  P.x = x; P.T = plan->dims.T; P.N = plan->dims.N; P.F = plan->dims.F; P.index = index;
  CU(launch_panel_image(P, static_cast<uint8_t*>(image), plan->sms, (cudaStream_t)stream));
  return DLAP_OK;
}

int dlap_cell_fwd(const dlap_plan* plan, const float* x, const uint8_t* mask, const int32_t* index, const void* x_image,
                  const dlap_tensors* params, const dlap_macro_io* mio, float* u, float* q, int precision,
                  float dropout_p, uint64_t dropout_seed, const uint64_t* dropout_seed_dev, void* stream) {
  int rc = check_plan(plan);
  if (rc) return rc;
  REQ(x); REQ(mask); REQ(params); REQ(mio); REQ(mio->a_t); REQ(u);
  REQ(params->w1); REQ(params->w2); REQ(params->b2); REQ(params->w3); REQ(params->b3);
  if (q) { REQ(mio->c_t); REQ(params->wq); }
  if (reinterpret_cast<uintptr_t>(x) & 3) return DLAP_E_ALIGN;
  if (precision != DLAP_PREC_FP32 && precision != DLAP_PREC_BF16X3) return DLAP_E_PRECISION;
  const bool tc = precision == DLAP_PREC_BF16X3;
  if (tc) {
    if (plan->dims.F > 46) return DLAP_E_SHAPE;      // the ones column lives at k = 46
    REQ(index);
  }
  cudaStream_t st = (cudaStream_t)stream;
  const dlap_dims& d = plan->dims;
  CellParams P;
  rc = fill_cell(P, plan, x, nullptr, mask, params, mio);
  if (rc) return rc;
  P.u = u; P.q = q;
  rc = set_dropout(P, dropout_p, dropout_seed, dropout_seed_dev);
  if (rc) return rc;
  P.index = tc ? index : nullptr;
  if (x_image && (reinterpret_cast<uintptr_t>(x_image) & 15)) return DLAP_E_ALIGN;
  P.ximg = tc ? static_cast<const uint8_t*>(x_image) : nullptr;
  const int series = mio->series_T > 0 ? mio->series_T : d.T;
  // ---- the macro side
  MacroParams MP;
  bool fused = false;
  if (!mio->macro && mio->tables_ready) {
    // a_t / c_t are the caller's
  } else if (d.M == 0) {
    REQ(params->b1);
    if (q) REQ(params->bq);
    REQ(mio->c_t);
    CU(launch_const_bias(params, series, mio->a_t, mio->c_t, st));
  } else if (mio->macro) {
    REQ(mio->lstm_save); REQ(mio->c_t); REQ(mio->hn); REQ(mio->cn); REQ(params->b1); REQ(params->bq);
    REQ(params->w_ih); REQ(params->w_hh); REQ(params->b_ih); REQ(params->b_hh); REQ(params->wq);
    fill_macro(MP, plan, series, mio->macro, params, mio->h0, mio->c0);
    MP.save = mio->lstm_save; MP.a_t = mio->a_t; MP.c_t = mio->c_t; MP.hn = mio->hn; MP.cn = mio->cn;
    // the recurrence as a CTA of the per-cell grid: tensor-core kernel, H = 4 (the reference's default and every
    // shipped configuration), and a second SM to run it on
    fused = tc && d.H == 4 && plan->sms >= 2 && mio->sync != nullptr;
    if (fused) {
      if (reinterpret_cast<uintptr_t>(mio->sync) & 3) return DLAP_E_ALIGN;
      MP.months_ready = mio->sync;
    } else {
      CU(launch_macro_fwd(MP, st));
    }
  } else {
    return DLAP_E_NULL;       // macro input expected, or tables that are ready
  }
  if (tc) {
    P.cyc_sh = 2;      // blocks of four tiles: measured best of 1, 2, 4, 8 (smaller: more month changes per CTA;
                       // larger: the last CTAs wait longer for their first month, and the blocks balance worse)
    int grid = plan->grid_cell;
    if (fused) {
      P.months_ready = mio->sync;
      if (grid > plan->sms - 1) grid = plan->sms - 1;      // one SM is the recurrence's
    }
    CU(launch_cell_fwd_tc(P, fused ? &MP : nullptr, grid, plan->cooperative != 0, st));
    return DLAP_OK;
  }
  CU(launch_cell_fwd_f32(P, plan->grid_cell, st));
  return DLAP_OK;
}

int dlap_step_reduce(const dlap_plan* plan, const float* u, const float* q, const float* R, const uint8_t* mask,
                     const int32_t* index, int weighted_loss, int flags, const float* n_bar_in, int phase_lo, int phase_hi,
                     float* row_stats, float* pr, float* scalars, float* w_out, double* col_part, float* e,
                     float* t_i, float* G, double* loss_part, double* dS2, double* row_part, uint32_t* bar,
                     const dlap_macro_io* series, float* xbuf, double* row_sums, int n_stocks_total,
                     const dlap_peers* peers, void* stream) {
  int rc = check_plan(plan);
  if (rc) return rc;
  const bool stock_shard = (flags & DLAP_REDUCE_STOCK_SHARD) != 0;
  if (phase_lo < 0 || phase_hi > (stock_shard ? 5 : 4) || phase_lo > phase_hi) return DLAP_E_SHAPE;
  if (peers && !stock_shard) return DLAP_E_SHAPE;
  if (stock_shard) {
    REQ(dS2);
    if (xbuf || n_stocks_total < plan->dims.N) return DLAP_E_SHAPE;
    if (peers && peers->world > 1) {        // both exchanges inside the one launch
      if (peers->world > 8 || peers->rank < 0 || peers->rank >= peers->world || peers->capacity < plan->dims.T)
        return DLAP_E_SHAPE;
      REQ(peers->epoch);
      for (int r = 0; r < peers->world; ++r) REQ(peers->buf[r]);
      if (phase_lo != 0 || phase_hi != 5) return DLAP_E_SHAPE;
    } else {                                // the caller's all-reduces sit between phases 0 | 1..4 | 5
      REQ(row_sums);
      if (!((phase_lo == 0 && phase_hi == 0) || (phase_lo == 1 && phase_hi == 4) || (phase_lo == 5 && phase_hi == 5)))
        return DLAP_E_SHAPE;
    }
  }
  REQ(mask); REQ(R); REQ(row_stats); REQ(scalars); REQ(col_part); REQ(bar);
  if (reinterpret_cast<uintptr_t>(bar) & 3) return DLAP_E_ALIGN;
  if (phase_lo <= 1) REQ(u);
  if (phase_lo <= 1 || xbuf) REQ(pr);
  int T_series = plan->dims.T, month0 = 0, month_stride = 1;
  if (xbuf) {
    REQ(series);
    month_stride = series->month_stride > 0 ? series->month_stride : 1;
    T_series = series->series_T > 0 ? series->series_T : plan->dims.T;
    month0 = series->month0;
    if (month0 < 0 || (long long)month0 + (long long)(plan->dims.T - 1) * month_stride >= T_series) return DLAP_E_SHAPE;
    if (reinterpret_cast<uintptr_t>(xbuf) & 3) return DLAP_E_ALIGN;
  }
  if (phase_hi >= 3) { REQ(e); REQ(t_i); REQ(G); REQ(loss_part); }
  CU(launch_step_reduce_c(plan, u, q, R, mask, index, weighted_loss, flags, n_bar_in, phase_lo, phase_hi, row_stats, pr,
                          scalars, w_out, col_part, e, t_i, G, loss_part, dS2, row_part, bar, xbuf, T_series, month0,
                          month_stride, row_sums, n_stocks_total, (peers && peers->world > 1) ? peers : nullptr,
                          (cudaStream_t)stream));
  return DLAP_OK;
}

int dlap_loss_finish(const dlap_plan* plan, const double* col_part, int nsplit, float* e, float* t_i,
                     double* loss_part, uint32_t* counter, float* scalars, void* stream) {
  int rc = check_plan(plan);
  if (rc) return rc;
  REQ(col_part); REQ(e); REQ(t_i); REQ(loss_part); REQ(counter); REQ(scalars);
  if (nsplit < 1) return DLAP_E_SHAPE;
  CU(launch_loss_finish(col_part, nsplit, plan->dims.N, plan->dims.N, e, t_i, loss_part, counter, scalars,
                        (cudaStream_t)stream));
  return DLAP_OK;
}

int dlap_cell_bwd(const dlap_plan* plan, const float* x, const float* R, const uint8_t* mask, const int32_t* index,
                  const void* x_image, const dlap_tensors* params, const dlap_macro_io* mio, const float* row_stats, const double* dS2,
                  const float* gc, const float* gu, const float* dpr_cot, const float* G, const float* scalars,
                  int weighted_loss, int want_sdf, int want_mom, float* wpart, float* dadc_part, int precision,
                  float dropout_p, uint64_t dropout_seed, const uint64_t* dropout_seed_dev, const float* res_b,
                  const float* u_fwd, void* stream) {
  int rc = check_plan(plan);
  if (rc) return rc;
  REQ(x); REQ(R); REQ(mask); REQ(params); REQ(mio); REQ(mio->a_t); REQ(mio->c_t); REQ(row_stats); REQ(dS2); REQ(G);
  if (res_b) REQ(u_fwd);
  REQ(scalars); REQ(wpart); REQ(dadc_part);
  REQ(params->w1); REQ(params->w2); REQ(params->b2); REQ(params->w3); REQ(params->b3);
  if (want_mom) REQ(params->wq);
  if (!want_sdf && !want_mom) return DLAP_E_SHAPE;
  if (precision != DLAP_PREC_FP32 && precision != DLAP_PREC_BF16X3) return DLAP_E_PRECISION;
  CellParams P;
  rc = fill_cell(P, plan, x, R, mask, params, mio);
  if (rc) return rc;
  P.row_stats = row_stats; P.dS2 = dS2; P.dpr_cot = dpr_cot;
  P.cot[0] = gc ? gc : scalars + SC_ZERO; P.cot[1] = gu ? gu : scalars + SC_ZERO; P.ncot = 2;
  P.G = G; P.scalars = scalars; P.weighted = weighted_loss;
  P.wpart = wpart; P.dadc_part = dadc_part;
  P.res_b = want_sdf ? res_b : nullptr; P.u_fwd = u_fwd;
  rc = set_dropout(P, dropout_p, dropout_seed, dropout_seed_dev);
  if (rc) return rc;
  if (precision == DLAP_PREC_BF16X3) {
    if (plan->dims.F > 46) return DLAP_E_SHAPE;
    REQ(index);
    P.index = index;
    if (x_image && (reinterpret_cast<uintptr_t>(x_image) & 15)) return DLAP_E_ALIGN;
    P.ximg = static_cast<const uint8_t*>(x_image);
    CU(launch_cell_bwd_tc(P, plan->grid_cell, want_sdf != 0, want_mom != 0, (cudaStream_t)stream));
    return DLAP_OK;
  }
  CU(launch_cell_bwd_f32(P, plan->grid_cell, want_sdf != 0, want_mom != 0, (cudaStream_t)stream));
  return DLAP_OK;
}

// what the per-cell backward kernel left behind, and where its sums go
static int fill_finish(FinishParams& F, const dlap_plan* plan, const float* wpart, const float* dadc_part,
                       const int32_t* index, int precision, int want_sdf, int want_mom, const dlap_tensors* grads,
                       float* dadc, int dadc_row_stride) {
  if (precision != DLAP_PREC_FP32 && precision != DLAP_PREC_BF16X3) return DLAP_E_PRECISION;
  // fp32 kernel: one da_t|dc_t partial per dense tile; tensor-core kernel: one per (month, CTA) segment
  const bool tc = precision == DLAP_PREC_BF16X3;
  if (tc) REQ(index);
  REQ(wpart); REQ(dadc_part); REQ(grads); REQ(dadc);
  if (dadc_row_stride < 1) return DLAP_E_SHAPE;
  if (want_sdf) { REQ(grads->w1); REQ(grads->w2); REQ(grads->b2); REQ(grads->w3); REQ(grads->b3); }
  if (want_mom) REQ(grads->wq);
  F = FinishParams{};
  F.wpart = wpart; F.ncta = plan->grid_cell; F.F = plan->dims.F; F.have_sdf = want_sdf; F.have_mom = want_mom;
  F.dadc_part = dadc_part; F.T = plan->dims.T; F.tpm = plan->tpm; F.index = tc ? index : nullptr;
  F.gW1 = grads->w1; F.ldw1 = plan->ldw1; F.gW2 = grads->w2; F.gb2 = grads->b2; F.gW3 = grads->w3; F.gb3 = grads->b3;
  F.gWq = grads->wq; F.ldwq = plan->ldwq; F.wq_off = plan->dims.M;
  F.dadc = dadc; F.dadc_stride = dadc_row_stride;
  return DLAP_OK;
}

int dlap_grad_finish(const dlap_plan* plan, const float* wpart, const float* dadc_part, const int32_t* index,
                     int precision, int want_sdf, int want_mom, const dlap_tensors* grads, float* dadc,
                     int dadc_row_stride, void* stream) {
  int rc = check_plan(plan);
  if (rc) return rc;
  FinishParams F;
  rc = fill_finish(F, plan, wpart, dadc_part, index, precision, want_sdf, want_mom, grads, dadc, dadc_row_stride);
  if (rc) return rc;
  CU(launch_grad_finish(F, (cudaStream_t)stream));
  return DLAP_OK;
}

int dlap_macro_bwd(const dlap_plan* plan, const float* macro, const dlap_tensors* params, const float* h0,
                   const float* c0, const float* lstm_save, float* dadc, float* dpre, int want_sdf, int want_mom,
                   const dlap_tensors* grads, const float* wpart, const float* dadc_part, const int32_t* index,
                   int precision, uint32_t* counter, const dlap_peers* peers, void* stream) {
  int rc = check_plan(plan);
  if (rc) return rc;
  REQ(params); REQ(dadc); REQ(grads);
  if (want_sdf) REQ(grads->b1);
  if (want_mom) REQ(grads->bq);
  if (plan->dims.M > 0) {
    REQ(macro);
    if (want_sdf) {
      REQ(lstm_save); REQ(dpre); REQ(params->w_hh); REQ(params->w1);
      REQ(grads->w_ih); REQ(grads->w_hh); REQ(grads->b_ih); REQ(grads->b_hh); REQ(grads->w1);
    }
    if (want_mom) REQ(grads->wq);
  }
  FinishParams F{};
  if (wpart || dadc_part) {
    rc = fill_finish(F, plan, wpart, dadc_part, index, precision, want_sdf, want_mom, grads, dadc, 1);
    if (rc) return rc;
    REQ(counter);
    if (reinterpret_cast<uintptr_t>(counter) & 3) return DLAP_E_ALIGN;
    F.counter = counter; F.cooperative = plan->cooperative;
    if (peers && peers->world > 1) {
      if (!wpart || !dadc_part) return DLAP_E_NULL;
      if (peers->world > 8 || peers->rank < 0 || peers->rank >= peers->world || peers->capacity < 1) return DLAP_E_SHAPE;
      if (!(want_sdf && plan->dims.M > 0)) return DLAP_E_SHAPE;      // the exchange lives in the BPTT kernel
      REQ(peers->epoch);
      F.world = peers->world; F.rank = peers->rank; F.gcap = peers->capacity; F.gepoch = peers->epoch;
      for (int r = 0; r < peers->world; ++r) { REQ(peers->buf[r]); F.gpeer[r] = static_cast<float*>(peers->buf[r]); }
    }
  } else if (peers && peers->world > 1) {
    return DLAP_E_NULL;
  }
  MacroParams P;
  fill_macro(P, plan, plan->dims.T, macro, params, h0, c0);
  P.save = const_cast<float*>(lstm_save); P.dadc = dadc; P.dpre = dpre;
  P.have_sdf = want_sdf; P.have_mom = want_mom;
  P.gWih = grads->w_ih; P.gWhh = grads->w_hh; P.gbih = grads->b_ih; P.gbhh = grads->b_hh;
  P.gW1 = grads->w1; P.gb1 = grads->b1; P.gWq = grads->wq; P.gbq = grads->bq;
  CU(launch_macro_bwd(P, F, (cudaStream_t)stream));
  return DLAP_OK;
}

int dlap_weights(const dlap_plan* plan, const float* u, const uint8_t* mask, const int32_t* index, int normalized,
                 int demean, float* w_out, void* stream) {
  int rc = check_plan(plan);
  if (rc) return rc;
  REQ(u); REQ(mask); REQ(w_out);
  CU(launch_weights(u, mask, index, plan->dims.T, plan->dims.N, normalized, demean, w_out, (cudaStream_t)stream));
  return DLAP_OK;
}

}  // extern "C"
