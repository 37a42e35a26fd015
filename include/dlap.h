/* dlap.h -- C ABI of libdlap_b200.so: the adversarial-SDF training step of
 * omroot/DeepLearningInAssetPricing_PaperReplication as hand-written sm_100a kernels.
 *
 * The reference has no FFI of its own (pure Python): its boundary is the class
 * API of src/model.py.  Each entry point below replaces the torch op sequence of the cited
 * reference lines; the Python mirror of the class API
 * (deeplearninginassetpricing_paperreplication_b200/model.py) is a thin ctypes caller of these.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer on the current CUDA device unless it says "host";
 *     all floating-point data is fp32, row-major, contiguous; `mask` is torch.bool bytes (0/1);
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); calls only
 *     enqueue work, they never synchronise and never allocate;
 *   - return value: 0 = ok, negative = argument error (DLAP_E_*), positive = cudaError_t;
 *   - all buffers are caller-owned; sizes come from dlap_make_plan().  Inputs are not mutated;
 *   - no CPU fallback: on a machine without a CUDA device every compute call returns an error.
 *
 * Shapes: T months, N stocks, F characteristics (<= 64), M macro series, H LSTM units (<= 8),
 * hidden widths 64/64 and 8 moments (the reference defaults, src/model.py:304-344).
 */
#ifndef DLAP_H
#define DLAP_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DLAP_ABI_VERSION 3

enum {
  DLAP_OK = 0,
  DLAP_E_NULL = -1,       /* a required pointer is NULL */
  DLAP_E_SHAPE = -2,      /* dimension out of the supported range */
  DLAP_E_ALIGN = -3,      /* pointer not aligned as required (16 B for x, 4 B for floats) */
  DLAP_E_PLAN = -4,       /* plan does not belong to these dims */
  DLAP_E_PRECISION = -5,  /* unknown precision selector */
  DLAP_E_NODEVICE = -6,   /* no usable CUDA device */
  DLAP_E_SMEM = -7,       /* configuration needs more shared memory than one SM has */
  DLAP_E_LIMIT = -8       /* size beyond the fixed capacity of a no-allocation entry point */
};

/* arithmetic of the per-cell contractions */
enum {
  DLAP_PREC_FP32 = 0,     /* exact fp32 FFMA on CUDA cores */
  DLAP_PREC_BF16X3 = 1    /* tcgen05 BF16 MMAs on hi/lo-split operands, 3 products per fp32 product,
                             fp32 accumulation in TMEM; needs F <= 46 and the valid-cell index of
                             dlap_panel_index: only unmasked stock-months are visited */
};

typedef struct dlap_dims {
  int32_t T, N, F, M, H;  /* M = 0: no macro input (src/model.py:243-257, macro_features=None) */
} dlap_dims;

/* Launch geometry and scratch sizes for one (dims, device).  Filled by dlap_make_plan. */
typedef struct dlap_plan {
  dlap_dims dims;
  int32_t tpm;            /* tiles (128 stocks) per month */
  int32_t grid_cell;      /* persistent CTAs of the per-cell kernels (= SM count) */
  int32_t col_splits;     /* month chunks of the per-stock (column) reduction */
  int32_t row_splits;     /* stock chunks of the per-month (row) reduction of dS */
  int32_t ldw1;           /* row pitch of fc_layers.0.weight  = F + H (F when M = 0) */
  int32_t ldwq;           /* row pitch of moment output_proj.weight = M + F */
  int64_t row_part_bytes;   /* double [T, tpm, 4 slots, 4]  */
  int64_t col_part_bytes;   /* double [col_splits, 10, N]   */
  int64_t loss_part_bytes;  /* double [ceil(N/64), 2]       */
  int64_t ds_part_bytes;    /* double [T, row_splits]       */
  int64_t wpart_bytes;      /* float  [grid_cell, wstride]  */
  int64_t dadc_part_bytes;  /* float  [T, max(tpm,4), 72]   */
  int64_t lstm_save_bytes;  /* float  [T, 6H]               */
  int64_t dpre_bytes;       /* float  [T, 4H]               */
  int64_t index_bytes;      /* int32  valid-cell index (dlap_panel_index), 16-B aligned */
} dlap_plan;

/* The twelve parameter tensors in the reference's state_dict order.
 * LSTM pointers may be NULL when M = 0.  Used const for parameters, mutable for gradients. */
typedef struct dlap_tensors {
  float *w_ih, *w_hh, *b_ih, *b_hh;   /* sdf_net.macro_lstm.lstm.{weight_ih,weight_hh,bias_ih,bias_hh}_l0 */
  float *w1, *b1;                     /* sdf_net.fc_layers.0.{weight [64,F+H], bias [64]} */
  float *w2, *b2;                     /* sdf_net.fc_layers.3.{weight [64,64], bias [64]} */
  float *w3, *b3;                     /* sdf_net.output_proj.{weight [1,64], bias [1]} */
  float *wq, *bq;                     /* moment_net.output_proj.{weight [8,M+F], bias [8]} */
} dlap_tensors;

int dlap_abi_version(void);
const char* dlap_error_string(int code);

/* Geometry for a panel shape on CUDA device `device` (host-side, no GPU work). */
int dlap_make_plan(const dlap_dims* dims, int device, dlap_plan* out);

/* MacroLSTM.forward (src/model.py:50-73, nn.LSTM at :39-45) plus the two per-month biases that
 * replace the tile-over-N concats: a_t = W1[:,F:].h_t + b1 (src/model.py:253-255, :211) and
 * c_t = Wq[:,:M].macro_t + bq (src/model.py:514-516, :152).  h0/c0 [H] may be NULL (zero state).
 * Outputs: lstm_save [T,6H] (h,c,i,f,g,o), a_t [T,64], c_t [T,8], hn/cn [H].  With M = 0 only
 * a_t = b1 and c_t = bq are produced.  xproj_scratch: kept for ABI stability ([T,4H] floats until ABI 2;
 * is not used any more, may be NULL).  One launch of a single CTA: eleven warps project the months'
 * inputs, one warp runs the recurrence behind them, one turns h_t into a_t rows behind that.
 * months_ready (device, one int32, may be NULL): progress counter of the recurrence.  The caller zeroes it
 * on a stream ordered before this call AND before the dlap_cell_fwd that is to consume it; the kernel
 * then keeps *months_ready = number of leading months whose a_t and c_t rows are written.  It lets
 * dlap_cell_fwd run beside the recurrence on another stream instead of after it (the reference computes
 * the whole LSTM first, src/model.py:247-255; the result is the same). */
int dlap_macro_fwd(const dlap_plan* plan, const float* macro, const dlap_tensors* params,
                   const float* h0, const float* c0, float* lstm_save, float* a_t, float* c_t,
                   float* hn, float* cn, float* xproj_scratch, int32_t* months_ready, void* stream);

/* Valid-cell index of the panel: for every month the ascending list of stocks with mask != 0, cut into
 * tiles of <= 128 cells (layout: csrc/cell_f32.cuh).  The reference runs both nets over every cell and
 * multiplies by the mask afterwards (src/model.py:261-279, :363, :376-382, :428), so a masked cell
 * contributes exactly 0 to every weight, loss and gradient; the BF16X3 kernels use this list to skip
 * those cells altogether (u and g are then written for unmasked cells only, left-packed per month).  Integer work, bit-exact;
 * the mask contract is src/data_loader.py:50-65.  Rebuild whenever the mask bytes change.  Two launches. */
int dlap_panel_index(const dlap_plan* plan, const uint8_t* mask, int32_t* index, void* stream);

/* Panel ingest without a staging copy (the per-epoch `batch[k].to(device)` of src/train.py:69-75 for the
 * characteristics): x_host is the [T,N,F] panel in page-locked host memory that the device can address
 * (cudaHostAlloc / cudaHostRegister; torch: tensor.pin_memory()), x_dev the device panel.  The rows of unmasked
 * stock-months (index: dlap_panel_index of the mask, already on the device) are read over the bus by the
 * kernel itself; masked rows are not transferred and arrive as zeros (the loader zero-fills them anyway,
 * src/data_loader.py:61-65).  One memset + one launch of `ctas` CTAs of 256 threads (0: plan->grid_cell; a
 * handful saturates PCIe).  clear = 0 skips the memset: the caller has zeroed x_dev itself (a loader that
 * double-buffers its panels clears a buffer as soon as its step is done, while the GPU has SMs to spare,
 * instead of in front of the pull), or does not care what the masked rows hold.  To overlap the pull of the next panel with a step, lower plan->grid_cell of the
 * plan the per-cell kernels run with by the same number: they take one CTA per SM and all of its registers. */
int dlap_panel_pull(const dlap_plan* plan, const float* x_host, const int32_t* index, float* x_dev, int ctas, int clear,
                    void* stream);

/* SDFNetwork.forward body on every stock-month (src/model.py:251-271) and MomentNetwork.forward
 * (src/model.py:132-161 with the input of :514-518): u = MLP(x | a_t) [T,N];
 * g = tanh(Wqx.x + c_t) * R * m [8,T,N] (g NULL: moment net skipped, phase 'unconditional');
 * row_part = per-tile sums of m, m u, m u R, m R.  R may be NULL (weights-only inference).
 * dropout_p > 0 (model.train(), src/model.py:213, src/train.py:66): each hidden unit of each
 * stock-month is dropped with probability round(p*65536)/65536 by a counter-based generator keyed
 * on (dropout_seed, stock-month, layer, unit); pass the same seed to dlap_cell_bwd.  dropout_seed_dev (device,
 * one uint64, may be NULL) overrides dropout_seed when given: the seed of a step that is replayed from a CUDA
 * graph has to live in device memory.
 * index: dlap_panel_index output, required for BF16X3 (row_part is then per valid-cell tile; u and g are
 * written for unmasked cells only and left-packed: the j-th unmasked stock of month t at [t*N + j]); ignored
 * (may be NULL) for FP32 (dense u, g).
 * months_ready (may be NULL; BF16X3 only, DLAP_E_PRECISION otherwise): the progress counter handed to a
 * dlap_macro_fwd that was ENQUEUED BEFORE this call on another stream and may still be running: the kernel
 * waits for a month's a_t / c_t row before it touches the month's first tile, visits the tiles in an
 * order that follows the months, and leaves one SM to the recurrence.  a_t and c_t must not be written
 * by anyone else meanwhile.  months_base: month 0 of this panel is month months_base of the series the
 * recurrence runs over (a month shard passes its first month and its slab of a_t / c_t; 0 otherwise).
 * With NULL, a_t and c_t are complete when the kernel starts (stream order) and months_base is ignored.
 * months_stride (>= 1; BF16X3 only when != 1): month t of this panel uses row t * months_stride of a_t and
 * c_t (and is month months_base + t * months_stride of the recurrence) -- a rank that owns every n-th
 * month of a series passes the full tables, offset to its first month, and n. */
int dlap_cell_fwd(const dlap_plan* plan, const float* x, const float* R, const uint8_t* mask,
                  const int32_t* index, const dlap_tensors* params, const float* a_t, const float* c_t, float* u,
                  float* g, double* row_part, int precision, float dropout_p, uint64_t dropout_seed,
                  const uint64_t* dropout_seed_dev, const int32_t* months_ready, int months_base, int months_stride,
                  void* stream);

/* Per-month statistics: N_t, mu_t (src/model.py:276-278), P_t, portfolio return pr_t and
 * S_t = 1 + pr_t (src/model.py:359-367, :380), Rbar_t; N_bar (src/model.py:360) and the Sharpe
 * ratio of pr (src/model.py:551) into scalars[3], scalars[2].  n_bar_in (device, 1 float) overrides
 * N_bar when the panel is one month-shard of a larger one; NULL otherwise.  precision: the selector
 * and index dlap_cell_fwd ran with (they fix the layout of row_part). */
int dlap_row_finish(const dlap_plan* plan, const double* row_part, const int32_t* index, int precision,
                    int weighted_loss, const float* n_bar_in, float* row_stats, float* pr, float* scalars,
                    void* stream);

/* Per-stock sums over the months of each chunk: sum_t g_k S_t, sum_t R m S_t, sum_t m
 * (src/model.py:376, :382, :428) into col_part [col_splits,10,N]; fused: the demeaned weights
 * w = m (u - mu_t) (src/model.py:271-279) into w_out [T,N] when non-NULL.  g may be NULL.  u and g are
 * read at unmasked cells only.  index: the valid-cell index when u and g come from the BF16X3 dlap_cell_fwd
 * (they are then left-packed: the j-th unmasked stock of month t sits at [t*N + j]); NULL for FP32 (dense). */
int dlap_col_partial(const dlap_plan* plan, const float* g, const float* R, const uint8_t* mask,
                     const float* u, const float* row_stats, const int32_t* index, double* col_part,
                     float* w_out, void* stream);

/* e_ki = colsum / T_i, L_c = mean_k mean_i e_ki^2, L_u = mean_i e0_i^2 (src/model.py:382-385,
 * :428-431) -> e [9,N], t_i [N], scalars[0] = L_c, scalars[1] = L_u.  `nsplit` chunks are summed
 * in order (1 after an all-reduce across month shards).  counter: one zero-initialised uint32. */
int dlap_loss_finish(const dlap_plan* plan, const double* col_part, int nsplit, float* e, float* t_i,
                     double* loss_part, uint32_t* counter, float* scalars, void* stream);

/* Same result as dlap_loss_finish, for callers that cannot allocate device memory.  No counterpart
 * in the reference: it exists for hosts that run on a fixed, pre-partitioned memory budget, that sit
 * in a latency-bound loop where cudaMalloc/cudaFree (driver locks, device synchronisation) are
 * forbidden, or that only ever receive raw pointers from another runtime.  The two scratch buffers
 * (`loss_part`, `counter`) live in an array inside the library image instead: ~66 KB on every GPU that
 * loads the library, for the life of the process.  Limits and rules:
 *   - at most DLAP_NOALLOC_MAX_STOCKS stocks (plan->dims.N); beyond that the call returns DLAP_E_LIMIT;
 *   - the first call on each device looks up the array's device address (one-time set-up with that
 *     device current); make that first call before starting any stream capture.  "That device" is the
 *     CURRENT device of the calling thread (cudaGetDevice), not the device `stream` belongs to: make the
 *     device that owns `stream` and the buffers current before the call, as for every entry point here;
 *   - every call on a device shares the one array and it keeps the previous call's contents (the call
 *     itself clears the one word that must start at zero); two calls that overlap on the GPU - different
 *     streams, or different threads - race on it.  Serialise them.
 * It calls no cudaMalloc/cudaFree and no allocating library routine.  Inputs/outputs as dlap_loss_finish. */
#define DLAP_NOALLOC_MAX_STOCKS 262144
int dlap_loss_finish_noalloc(const dlap_plan* plan, const double* col_part, int nsplit, float* e,
                             float* t_i, float* scalars, void* stream);

/* Backward of the two losses (autograd at src/train.py:84): G_ki = gc 2 e_ki/(K N T_i),
 * G0_i = gu 2 e0_i/(N T_i); gc/gu are device scalars = dL/dL_c, dL/dL_u (NULL = 0). -> G [9,N] */
int dlap_loss_coeff(const dlap_plan* plan, const float* e, const float* t_i, const float* gc,
                    const float* gu, float* G, void* stream);

/* dS_t partials = sum_i (sum_k G_ki g_kti + G0_i R m) -> dS_part [T,row_splits].  index: as for
 * dlap_col_partial (layout of g). */
int dlap_ds_rows(const dlap_plan* plan, const float* g, const float* R, const uint8_t* mask,
                 const float* G, const int32_t* index, double* dS_part, void* stream);

/* Per-cell backward with activation recompute (autograd of src/model.py:261-279
 * and :152-153).  want_sdf / want_mom select the parameter groups (requires_grad semantics of
 * src/train.py:309-312).  dpr_cot [T] = extra cotangent of portfolio_returns or NULL.
 * Outputs: wpart (per-CTA weight-gradient partials), dadc_part (per-tile da_t | dc_t).
 * index: as for dlap_cell_fwd. */
int dlap_cell_bwd(const dlap_plan* plan, const float* x, const float* R, const uint8_t* mask,
                  const int32_t* index, const dlap_tensors* params, const float* a_t, const float* c_t,
                  const float* row_stats, const double* dS_part, const float* dpr_cot,
                  const float* G, const float* scalars, int weighted_loss, int want_sdf,
                  int want_mom, float* wpart, float* dadc_part, int precision, float dropout_p,
                  uint64_t dropout_seed, const uint64_t* dropout_seed_dev, void* stream);

/* Fixed-order sum of the partials into the reference's gradient layouts: grads->w1[:, :F], w2, b2,
 * w3, b3 (= 0, demeaning), wq[:, M:]; and dadc [T,72] = da_t | dc_t summed over stocks.
 * precision, index: what dlap_cell_bwd ran with (they fix the layout of dadc_part: one partial per
 * tile for FP32, one per (month, CTA) segment for BF16X3). */
int dlap_grad_finish(const dlap_plan* plan, const float* wpart, const float* dadc_part, const int32_t* index,
                     int precision, int want_sdf, int want_mom, const dlap_tensors* grads, float* dadc,
                     void* stream);

/* Month-level backward: dh_t = W1[:,F:]^T da_t, LSTM BPTT, and the month-summed gradients
 * grads->{w_ih,w_hh,b_ih,b_hh}, w1[:, F:], b1, wq[:, :M], bq.  dpre [T,4H] is scratch. */
int dlap_macro_bwd(const dlap_plan* plan, const float* macro, const dlap_tensors* params,
                   const float* h0, const float* c0, const float* lstm_save, const float* dadc,
                   float* dpre, int want_sdf, int want_mom, const dlap_tensors* grads, void* stream);

/* get_weights (src/model.py:565-594): w = m (u - mu_t); normalized != 0 divides each month by
 * clamp(sum_i |w| m, 1e-8) (the Python loop of :586-590).  index: as for dlap_col_partial (layout of u). */
int dlap_weights(const dlap_plan* plan, const float* u, const uint8_t* mask, const float* row_stats,
                 const int32_t* index, int normalized, float* w_out, void* stream);

/* Diagnostics (no counterpart in the reference): when `stamps` is a device buffer of 3*256*16 int64, every
 * following tensor-core dlap_cell_bwd (first third of the buffer: epilogue and MMA warps per tile; last third:
 * one gather warp per batch) and dlap_cell_fwd (second third) launch makes CTA 0 record clock64() at fixed
 * points of each of its first 256 tiles / batches (see the DLAP_STAMP sites in csrc/cell_bwd_tc.cu,
 * csrc/cell_fwd_tc.cu and gather_role in csrc/cell_tc.cuh).  NULL switches it off.  scripts/bwd_timeline.py
 * turns the stamps into per-phase cycle counts. */
int dlap_debug_timeline(long long* stamps);

/* Hardware self-test of the tensor-core plumbing (no counterpart in the reference): one CTA computes
 * D[128,N] = A.B^T (bf16 operands, fp32 accumulate in TMEM) from two operand images laid out by the
 * caller in the library's no-swizzle canonical layout, with caller-given descriptor strides (bytes) and
 * majors (0 = K-major, 1 = MN-major).  tests/ pin the layout conventions of csrc/tc_sm100.cuh with it. */
int dlap_umma_selftest(const void* a_img, int a_bytes, const void* b_img, int b_bytes, int M, int N, int K,
                       int a_mn_major, int b_mn_major, int a_lbo, int a_sbo, int a_kstep, int b_lbo,
                       int b_sbo, int b_kstep, float* D, void* stream);

/* As dlap_umma_selftest, but the A operand [128,K] (row-major bf16 in global memory) is placed in
 * tensor memory by its owning lanes and the MMA reads it from there (A-from-TMEM form). repeat > 1 and a
 * non-NULL cycles turn it into a timing run.  No counterpart in the reference. */
int dlap_umma_ts_selftest(const void* a_rowmajor, const void* b_img, int b_bytes, int N, int K,
                          int b_mn_major, int b_lbo, int b_sbo, int b_kstep, int repeat, float* D,
                          long long* cycles, void* stream);

/* Timing twin of dlap_umma_selftest (no counterpart in the reference): the K loop is issued `repeat`
 * times back to back on one CTA; *cycles (device int64) receives the SM cycles until completion. */
int dlap_umma_cycles(const void* a_img, int a_bytes, const void* b_img, int b_bytes, int M, int N, int K,
                     int a_mn_major, int b_mn_major, int a_lbo, int a_sbo, int a_kstep, int b_lbo, int b_sbo,
                     int b_kstep, int repeat, float* D, long long* cycles, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DLAP_H */
