/* dlap.h -- C ABI of libdlap_b200.so: the adversarial-SDF training step of
 * omroot/DeepLearningInAssetPricing_PaperReplication as hand-written sm_100a kernels.
 *
 * The reference has no FFI of its own (pure Python): its boundary is the class
 * API of src/model.py.  Each entry point below replaces the torch op sequence of the cited
 * reference lines; the Python mirror of the class API
 * (deeplearninginassetpricing_paperreplication_b200/model.py) is a thin ctypes caller of these.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer on the current CUDA device unless it says "host" (structs that are
 *     passed by pointer -- plan, tensors, macro_io -- are host structs holding device pointers);
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

#define DLAP_ABI_VERSION 5

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
  DLAP_PREC_FP32 = 0,     /* exact fp32 FFMA on CUDA cores; dense u / q */
  DLAP_PREC_BF16X3 = 1    /* tcgen05 BF16 MMAs on hi/lo-split operands, 3 products per fp32 product,
                             fp32 accumulation in TMEM; needs F <= 46 and the valid-cell index of
                             dlap_panel_index: only unmasked stock-months are visited; u / q left-packed */
};

typedef struct dlap_dims {
  int32_t T, N, F, M, H;  /* M = 0: no macro input (src/model.py:243-257, macro_features=None) */
} dlap_dims;

/* Launch geometry and scratch sizes for one (dims, device).  Filled by dlap_make_plan. */
typedef struct dlap_plan {
  dlap_dims dims;
  int32_t tpm;            /* tiles (128 stocks) per month */
  int32_t sms;            /* SMs the kernels of this plan spread over (the device's; a caller that overlaps another
                             kernel, e.g. dlap_panel_pull, lowers it together with grid_cell) */
  int32_t grid_cell;      /* persistent CTAs of the per-cell kernels (<= sms) */
  int32_t cooperative;    /* 1: kernels that synchronise across their grid (dlap_step_reduce, the fused dlap_cell_fwd)
                             are launched with cudaLaunchCooperativeKernel, which refuses a grid that cannot be
                             co-resident; 0: plain launches of the same grids (needed when another kernel is meant to
                             run beside them on the SMs left free: a cooperative grid does not share the device) */
  int32_t col_splits;     /* month chunks of the per-stock (column) sums */
  int32_t ldw1;           /* row pitch of fc_layers.0.weight  = F + H (F when M = 0) */
  int32_t ldwq;           /* row pitch of moment output_proj.weight = M + F */
  int64_t col_part_bytes;   /* double [col_splits, 10, N]   */
  int64_t loss_part_bytes;  /* double [ceil(N/64), 2]       */
  int64_t row_part_bytes;   /* double [T, 8, 4]: partial sums of dlap_step_reduce's per-month phases */
  int64_t bar_bytes;        /* uint32 [1 + T]: dlap_step_reduce's barrier word and per-month counters */
  int64_t wpart_bytes;      /* float  [grid_cell, wstride]  */
  int64_t dadc_part_bytes;  /* float  [T, max(tpm,4), 72]   */
  int64_t lstm_save_bytes;  /* float  [T, 6H]               */
  int64_t dpre_bytes;       /* float  [T, 4H]               */
  int64_t index_bytes;      /* int32  valid-cell index (dlap_panel_index), 16-B aligned */
  int64_t image_bytes;      /* bytes  X images of every tile the index can hold (dlap_panel_image), 16-B aligned;
                               0 when F > 46 (no tensor-core path) */
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

/* The macro side of a per-cell call.  The per-month biases that replace the reference's two tile-over-N concats,
 * a_t = W1[:,F:].h_t + b1 (src/model.py:253-255, :211) and c_t = Wq[:,:M].macro_t + bq (src/model.py:514-516, :152),
 * live in tables over a SERIES of series_T months; month t of the panel is month month0 + t * month_stride of that
 * series (a month shard that owns every n-th month passes its first month and n; otherwise 0 and 1, and series_T =
 * the plan's T).
 *   dlap_cell_fwd: macro != NULL: the call also runs MacroLSTM.forward (src/model.py:50-73) over the whole series
 *     and fills lstm_save [series_T,6H] (h,c,i,f,g,o), a_t [series_T,64], c_t [series_T,8], hn, cn [H] (h0 / c0 [H]
 *     may be NULL: zero state).  macro == NULL: with tables_ready != 0, a_t and c_t are INPUTS, complete when the
 *     call starts (a caller that builds the per-month biases itself, e.g. a model without the LSTM:
 *     a_t = W1[:,F:].macro_t + b1, src/model.py:200-205; the plan then has M = 0, and w1 / wq hold the
 *     characteristics' columns only); otherwise M must be 0 and the call fills them with b1 / bq.
 *   dlap_cell_bwd: reads a_t, c_t, month0, month_stride only.
 * sync (device, two int32, zero when the call starts and zero again when it has finished; may be NULL): BF16X3 with
 * macro and H = 4 only.  When given, the LSTM runs on one CTA of the per-cell kernel's own grid, beside the per-cell
 * work, and publishes the months as they complete (the reference computes the whole LSTM first, src/model.py:247-255;
 * the result is the same, bit for bit); sync[0] counts the published months.  NULL: the LSTM is a launch of its own
 * in front of the per-cell kernel. */
typedef struct dlap_macro_io {
  const float* macro;
  int32_t series_T, month0, month_stride, tables_ready;
  const float *h0, *c0;
  float *lstm_save, *a_t, *c_t, *hn, *cn;
  int32_t* sync;
} dlap_macro_io;

/* Stock shards on GPUs that can address each other's memory (NVLink / NVSwitch; torch: symmetric memory): the
 * exchanges of a step happen INSIDE the kernels instead of in collectives between them.  Every rank writes its
 * partial sums into its own exchange buffer, raises a flag at every peer and waits for everybody's (flags carry the
 * launch number and only ever grow), then adds all ranks' partial sums in rank order -- so every rank holds the same
 * bits.  buf[r]: rank r's exchange buffer as THIS device addresses it, all zero before the first call; epoch: zero-
 * initialised uint32 words in this rank's own memory.  All ranks must make the same calls in the same order; a rank
 * that waits more than 2 s for a peer traps.  One dlap_peers per entry point:
 *   dlap_step_reduce  buf: (6 capacity + 2) doubles then 16 uint32, capacity >= the plan's T months; epoch: 1 word;
 *   dlap_macro_bwd    buf: 2 capacity floats then 8 uint32, capacity >= the number of gradient elements the per-cell
 *                     kernel leaves (64*64 + 72*round4(F) + 128 + 72 T); epoch: 2 words.
 * (TMA multicast is not involved: plain loads and stores to peer addresses.) */
typedef struct dlap_peers {
  int32_t world, rank, capacity;   /* 2 .. 8 ranks */
  void* buf[8];
  uint32_t* epoch;
} dlap_peers;

int dlap_abi_version(void);
const char* dlap_error_string(int code);

/* Geometry for a panel shape on CUDA device `device` (host-side, no GPU work). */
int dlap_make_plan(const dlap_dims* dims, int device, dlap_plan* out);

/* MacroLSTM.forward on its own (src/model.py:50-73, nn.LSTM at :39-45) plus the two per-month bias tables; the
 * series is the plan's T months.  h0/c0 [H] may be NULL (zero state).  Outputs: lstm_save [T,6H] (h,c,i,f,g,o),
 * a_t [T,64], c_t [T,8], hn/cn [H].  With M = 0 only a_t = b1 and c_t = bq are produced.  One launch of a single
 * CTA: eleven warps project the months' inputs, one warp runs the recurrence behind them, one turns h_t into a_t
 * rows behind that. */
int dlap_macro_fwd(const dlap_plan* plan, const float* macro, const dlap_tensors* params,
                   const float* h0, const float* c0, float* lstm_save, float* a_t, float* c_t,
                   float* hn, float* cn, void* stream);

/* Valid-cell index of the panel: for every month the ascending list of stocks with mask != 0, cut into
 * tiles of <= 128 cells (layout: csrc/cell_f32.cuh).  The reference runs both nets over every cell and
 * multiplies by the mask afterwards (src/model.py:261-279, :363, :376-382, :428), so a masked cell
 * contributes exactly 0 to every weight, loss and gradient; the BF16X3 kernels use this list to skip
 * those cells altogether (u and q are then written for unmasked cells only, left-packed per month).  Integer work, bit-exact;
 * the mask contract is src/data_loader.py:50-65.  Rebuild whenever the mask bytes change.  Two launches. */
int dlap_panel_index(const dlap_plan* plan, const uint8_t* mask, int32_t* index, void* stream);

/* The panel's characteristics in the tensor-core kernels' operand format, once per panel (the reference re-reads
 * the same x every epoch, src/train.py:45-103; SURVEY.md 8 f4 "a bf16-split copy of x").  For every tile of the
 * valid-cell index (<= 128 listed stocks of one month) `image` receives, at [tile * 24768), the 24,768 bytes the
 * BF16X3 kernels otherwise build in shared memory per tile and step: the rows of x split into bf16 hi + lo
 * (x = hi + lo up to 2^-18 |x|) in the no-swizzle canonical K-major layout, a ones column for the first-layer
 * biases, rows past the end of a tile repeating its last row.  image: 16-B aligned, at least 24768 * index[0] bytes
 * (index[0]: the tiles the index lists; plan->image_bytes is the bound that holds for any mask).  Requires F <= 46.  Rebuild whenever x or the index changes.  One launch; reads x once (listed rows only), HBM-bound. */
int dlap_panel_image(const dlap_plan* plan, const float* x, const int32_t* index, void* image, void* stream);

/* Panel ingest without a staging copy (the per-epoch `batch[k].to(device)` of src/train.py:69-75 for the
 * characteristics): x_host is the [T,N,F] panel in page-locked host memory that the device can address
 * (cudaHostAlloc / cudaHostRegister; torch: tensor.pin_memory()), x_dev the device panel.  The rows of unmasked
 * stock-months (index: dlap_panel_index of the mask, already on the device) are read over the bus by the
 * kernel itself; masked rows are not transferred and arrive as zeros (the loader zero-fills them anyway,
 * src/data_loader.py:61-65).  One memset + one launch of `ctas` CTAs of 256 threads (0: plan->grid_cell; a
 * handful saturates PCIe).  clear = 0 skips the memset: the caller has zeroed x_dev itself (a loader that
 * double-buffers its panels clears a buffer as soon as its step is done, while the GPU has SMs to spare,
 * instead of in front of the pull), or does not care what the masked rows hold.  To overlap the pull of the next
 * panel with a step, lower plan->sms and plan->grid_cell of the plan the step runs with by the same number and
 * clear plan->cooperative: the per-cell kernels take one CTA per SM and all of its registers. */
int dlap_panel_pull(const dlap_plan* plan, const float* x_host, const int32_t* index, float* x_dev, int ctas, int clear,
                    void* stream);

/* The loader's array work on the device (AssetPricingDataset.__init__, src/data_loader.py:42-66), one launch:
 * raw is the `data` array of the characteristics file, [T,N,F+1] with column 0 the return (src/data_loader.py:45-46),
 * fp32 (raw_f64 = 0) or fp64 (raw_f64 = 1: rounded to fp32 first, the `.astype(np.float32)` of :45-46), 16-B aligned,
 * in device memory OR in page-locked host memory the device can address (the kernel then reads it over the bus: the
 * upload, the mask and the zero-fill are one pass, nothing is staged).  Outputs, all on the device:
 *   mask[t,n] = R > valid_above and every characteristic > valid_above   (src/data_loader.py:50-57; a NaN fails '>';
 *               the reference's threshold is the fp32 value of -99.99 + 1)
 *   R[t,n], x[t,n,:] = the raw values where mask, +0.0 elsewhere          (src/data_loader.py:60-65)
 * Bit-exact (comparisons and copies only).  4(F+1) B read and 4(F+1)+1 B written per stock-month, coalesced on both
 * sides through shared memory.  ctas: CTAs of 256 threads (0: 8 per SM of the plan). */
int dlap_panel_ingest(const dlap_plan* plan, const void* raw, int raw_f64, float valid_above, float* x, float* R,
                      uint8_t* mask, int ctas, void* stream);

/* Macro z-score of the loader (src/data_loader.py:76-86): macro_raw [T,M] fp32 on the device, row-major (after the
 * macro_idx selection of :72-73).  stats_given = 0 (training file): mean[m] = sum_t x / T, std[m] = sqrt(sum_t
 * (x-mean)^2 / T) + 1e-8 are computed and stored (:78-79); stats_given = 1 (validation / test file): mean, std are
 * read (:81-82).  out = (x - mean) / std (:84).  fp32, each operation rounded separately, the sums in numpy's order
 * for the array the loader holds: pairwise = 0, the file's own C-contiguous array: row after row; pairwise = 1, the
 * column-major result of `macro_features[:, macro_idx]`: numpy's pairwise scheme down each column (8 interleaved
 * accumulators up to 128 terms, recursive halves beyond).  Bit-identical to the reference's arrays either way.
 * One launch, one thread per column. */
int dlap_macro_zscore(const float* macro_raw, int T, int M, float* mean, float* std, int stats_given, int pairwise,
                      float* out, void* stream);

/* SDFNetwork.forward body on every stock-month (src/model.py:251-271) and MomentNetwork.forward
 * (src/model.py:132-161 with the input of :514-518), and -- see dlap_macro_io -- the macro LSTM in front of them:
 * u = MLP(x | a_t) [T,N];  q = tanh(Wqx.x + c_t) [8,T,N] (q NULL: moment net skipped, phase 'unconditional').
 * Everything that sums over stocks or months is dlap_step_reduce's.
 * dropout_p > 0 (model.train(), src/model.py:213, src/train.py:66): each hidden unit of each
 * stock-month is dropped with probability round(p*65536)/65536 by a counter-based generator keyed
 * on (dropout_seed, stock-month, layer, unit); pass the same seed to dlap_cell_bwd.  dropout_seed_dev (device,
 * one uint64, may be NULL) overrides dropout_seed when given: the seed of a step that is replayed from a CUDA
 * graph has to live in device memory.
 * index: dlap_panel_index output, required for BF16X3 (u and q are then written for unmasked cells only and
 * left-packed: the j-th unmasked stock of month t at [t*N + j]); ignored (may be NULL) for FP32 (dense u, q).
 * x_image: dlap_panel_image output for this x and index, or NULL (BF16X3 only; ignored for FP32).  With it the
 * kernel fetches each tile's operand image with one bulk copy instead of gathering and converting 128 rows of x
 * (x is then not read at all); the results are bit-identical either way.
 * Launches: BF16X3 with macro, H = 4 and mio->sync: ONE (the recurrence is a CTA of the per-cell grid); otherwise the
 * macro kernel (or the constant-bias fill when M = 0), then the per-cell kernel. */
int dlap_cell_fwd(const dlap_plan* plan, const float* x, const uint8_t* mask, const int32_t* index,
                  const void* x_image, const dlap_tensors* params, const dlap_macro_io* mio, float* u, float* q, int precision,
                  float dropout_p, uint64_t dropout_seed, const uint64_t* dropout_seed_dev, void* stream);

/* Every sum of the step over stocks (per month) and over months (per stock), one launch whose CTAs step through
 * the phases phase_lo .. phase_hi with grid-wide barriers in between (csrc/step_reduce.cu):
 *   0  N_t, mu_t (src/model.py:276-278), P_t, Rbar_t from u and R                          -> row_stats [T,8]
 *   1  N_bar (src/model.py:360; n_bar_in, device, overrides it when the panel is a month shard), portfolio return
 *      pr_t and S_t = 1 + pr_t (src/model.py:359-367, :380), Sharpe ratio of pr (src/model.py:551)
 *                                -> row_stats, pr [T], scalars[2] = Sharpe, scalars[3] = N_bar, scalars[7] = 0;
 *      per stock and month chunk: sum_t q_k R S_t, sum_t R m S_t, sum_t m (src/model.py:376, :382, :428)
 *                                                      -> col_part [col_splits,10,N]
 *      and the demeaned weights w = m (u - mu_t) (src/model.py:271-279)      -> w_out [T,N] when non-NULL
 *   2  the month chunks summed into chunk 0 of col_part (month shards: into xbuf, see below)
 *   3  e_ki = colsum / T_i (src/model.py:382, :428) -> e [9,N], t_i [N]; the loss coefficients for UNIT cotangents,
 *      G_ki = 2 e_ki / (K N T_i), G_8i = 2 e0_i / (N T_i) (autograd at src/train.py:84)
 *                              -> G [N,12], one row per stock: G_0i .. G_7i, G_8i, three floats of padding
 *   4  L_c = mean_k mean_i e_ki^2, L_u = mean_i e0_i^2 (src/model.py:385, :429-431) -> scalars[0], scalars[1];
 *      dS2 [T,2] (double): dS_t for unit cotangents of L_c and of L_u; dS2 NULL: no backward pass will follow
 * flags: DLAP_REDUCE_NO_DEMEAN: normalize_w = False (src/model.py:274): w = m u, mu_t and Rbar_t are reported as 0;
 * DLAP_REDUCE_Q_DENSE: q is a dense [8,T,N] tensor even though u is left-packed (moment outputs that did not come
 * from dlap_cell_fwd: a moment network with hidden layers, evaluated by the caller).
 * Month shards (xbuf != NULL; series: the shard's place in the series -- series_T, month0, month_stride, nothing
 * else is read): phase 2 writes what has to cross the wire into xbuf, fp32 [10 N + series_T]: the per-stock sums
 * of this rank's months, then the series' portfolio returns with this rank's months filled in and zeros elsewhere.
 * The caller sum-all-reduces xbuf between a call for phases 0..2 and a call for phases 3..4; phase 3 then reads the
 * per-stock sums from it and writes the whole series' pr [series_T] and its Sharpe ratio.
 * Stock shards (flags & DLAP_REDUCE_STOCK_SHARD: this rank owns a slice of the stocks and all months; every sum
 * over stocks is a partial sum, every sum over months is complete): three calls with two small all-reduces in
 * between.  Phases 0..0 leave the months' raw sums in row_sums (double [T,4]: count, sum u, sum u R, sum R); the
 * caller sum-all-reduces it; phases 1..4 turn it into row_stats and go on as usual with n_stocks_total stocks in
 * the loss means, and leave the sums of e^2 behind the dS sums: dS2 is double [T,2] + [2] (required); the caller
 * sum-all-reduces it; phase 5 writes L_c and L_u.  With `peers` (see dlap_peers) ONE call for phases 0..5 does it all,
 * the two exchanges over peer memory; row_sums is then not used and dS2 is an output only.
 * Single device: phases 0..4 in one launch.  q may be NULL (phase 'unconditional').  index: the valid-cell index
 * when u and q come from the BF16X3 dlap_cell_fwd, NULL for FP32.  row_part: scratch (row_part_bytes; NULL: the
 * per-month phases run one CTA per month).  bar (device, bar_bytes, all zero when the call starts and all zero
 * again when it has finished): the barrier word and one counter per month. */
#define DLAP_REDUCE_NO_DEMEAN 1
#define DLAP_REDUCE_Q_DENSE 2
#define DLAP_REDUCE_STOCK_SHARD 4
int dlap_step_reduce(const dlap_plan* plan, const float* u, const float* q, const float* R, const uint8_t* mask,
                     const int32_t* index, int weighted_loss, int flags, const float* n_bar_in, int phase_lo, int phase_hi,
                     float* row_stats, float* pr, float* scalars, float* w_out, double* col_part, float* e,
                     float* t_i, float* G, double* loss_part, double* dS2, double* row_part, uint32_t* bar,
                     const dlap_macro_io* series, float* xbuf, double* row_sums, int n_stocks_total,
                     const dlap_peers* peers, void* stream);

/* The loss stage on its own, for a caller that holds per-stock sums from elsewhere: e_ki = colsum / T_i,
 * L_c = mean_k mean_i e_ki^2, L_u = mean_i e0_i^2 (src/model.py:382-385, :428-431) -> e [9,N], t_i [N],
 * scalars[0] = L_c, scalars[1] = L_u.  `nsplit` chunks [nsplit,10,N] are summed in order.  counter: one
 * zero-initialised uint32. */
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

/* Per-cell backward with activation recompute (autograd of src/model.py:261-279
 * and :152-153).  gc / gu (device scalars, NULL = 0) = dL/dL_c, dL/dL_u: dS_t = gc dS2[t][0] + gu dS2[t][1]
 * (+ dpr_cot[t], the extra cotangent of portfolio_returns, or NULL), and the moment side uses gc G_k.
 * want_sdf / want_mom select the parameter groups (requires_grad semantics of src/train.py:309-312).
 * mio: a_t, c_t, month0, month_stride as for dlap_cell_fwd (nothing else is read).
 * Outputs: wpart (per-CTA weight-gradient partials), dadc_part (per-tile da_t | dc_t).
 * index, x_image: as for dlap_cell_fwd.
 * res_b, u_fwd (both NULL: off): the gradient of the residual-loss regulariser (compute_residual_loss,
 * src/model.py:435-483, added to the loss at :543-548).  With w = m (u - mu_t) and beta_t = R.w / w.w it reaches u as
 * a_t (R - Rbar_t) + res_b[t] (u - mu_t): the caller adds a_t to dpr_cot (divided by N_bar / N_t when weighted_loss)
 * and passes res_b [T] and the forward's u (left-packed for BF16X3, dense for FP32). */
int dlap_cell_bwd(const dlap_plan* plan, const float* x, const float* R, const uint8_t* mask,
                  const int32_t* index, const void* x_image, const dlap_tensors* params, const dlap_macro_io* mio,
                  const float* row_stats, const double* dS2, const float* gc, const float* gu,
                  const float* dpr_cot, const float* G, const float* scalars, int weighted_loss, int want_sdf,
                  int want_mom, float* wpart, float* dadc_part, int precision, float dropout_p,
                  uint64_t dropout_seed, const uint64_t* dropout_seed_dev, const float* res_b, const float* u_fwd,
                  void* stream);

/* Fixed-order sum of the partials into the reference's gradient layouts: grads->w1[:, :F], w2, b2,
 * w3, b3 (= 0, demeaning), wq[:, M:]; and da_t | dc_t summed over stocks: row t of the panel goes to row
 * t * dadc_row_stride of dadc [.,72].  precision, index: what dlap_cell_bwd ran with (they fix the layout of
 * dadc_part: one partial per tile for FP32, one per (month, CTA) segment for BF16X3).  A launch of its own for
 * callers that exchange the sums before the month-level backward (month shards); otherwise hand the partials to
 * dlap_macro_bwd. */
int dlap_grad_finish(const dlap_plan* plan, const float* wpart, const float* dadc_part, const int32_t* index,
                     int precision, int want_sdf, int want_mom, const dlap_tensors* grads, float* dadc,
                     int dadc_row_stride, void* stream);

/* Month-level backward: dh_t = W1[:,F:]^T da_t, LSTM BPTT, and the month-summed gradients
 * grads->{w_ih,w_hh,b_ih,b_hh}, w1[:, F:], b1, wq[:, :M], bq.  dpre [T,4H] is scratch.
 * wpart / dadc_part non-NULL (with the index and precision dlap_cell_bwd ran with; the plan is then also the
 * per-cell plan): the call also does dlap_grad_finish's work -- dadc [T,72] becomes an output -- on extra CTAs of
 * the BPTT kernel: the recurrence's CTA stages its tables meanwhile and waits for the da_t | dc_t rows, the
 * weight-gradient sums run beside the serial chain; with plan->cooperative those CTAs then go on to the month-summed
 * gradients as soon as the chain is through (one launch instead of two).  counter (device, four uint32, zero when
 * the call starts and zero again when it has finished; required with partials).  peers (stock shards over peer memory, see dlap_peers;
 * NULL otherwise; needs the partials): the sums of the partials are this rank's share only, and the same kernel adds
 * all ranks' shares -- the backward pass' one exchange -- before the recurrence starts.  One launch (partials and
 * plan->cooperative), else two. */
int dlap_macro_bwd(const dlap_plan* plan, const float* macro, const dlap_tensors* params, const float* h0,
                   const float* c0, const float* lstm_save, float* dadc, float* dpre, int want_sdf,
                   int want_mom, const dlap_tensors* grads, const float* wpart, const float* dadc_part,
                   const int32_t* index, int precision, uint32_t* counter, const dlap_peers* peers, void* stream);

/* get_weights (src/model.py:565-594): w = m (u - mu_t), mu_t the month's mean of u over unmasked stocks (demean = 0:
 * w = m u, normalize_w = False, src/model.py:274); normalized != 0 divides each month by clamp(sum_i |w| m, 1e-8)
 * (the Python loop of :586-590).  index: the layout of u (as for dlap_step_reduce). */
int dlap_weights(const dlap_plan* plan, const float* u, const uint8_t* mask, const int32_t* index, int normalized,
                 int demean, float* w_out, void* stream);

/* clip_grad_norm_ + optimizer.step() of the training loop (src/train.py:87-94: torch.nn.utils.clip_grad_norm_ over
 * the scope's parameters, then torch.optim.Adam with its defaults: no weight decay, no amsgrad) over up to
 * DLAP_ADAM_MAX_TENSORS tensors in ONE launch: the 2-norm of all gradients -> *grad_norm (device, may be NULL); with
 * max_norm > 0 every gradient is scaled IN PLACE by min(max_norm / (norm + 1e-6), 1); then exp_avg, exp_avg_sq and
 * the parameters are updated with Adam's arithmetic, op for op.  step[k] (device, one int64 per tensor: the number of
 * updates that tensor has had -- torch counts per parameter, and one without a gradient in some step lags behind)
 * is incremented, so the call can be replayed from a CUDA graph.  The training step's twelve tensors are 12,233
 * numbers: one CTA. */
#define DLAP_ADAM_MAX_TENSORS 16
typedef struct dlap_adam_tensors {
  int32_t count;
  float* param[DLAP_ADAM_MAX_TENSORS];
  float* grad[DLAP_ADAM_MAX_TENSORS];
  float* exp_avg[DLAP_ADAM_MAX_TENSORS];
  float* exp_avg_sq[DLAP_ADAM_MAX_TENSORS];
  int64_t* step[DLAP_ADAM_MAX_TENSORS];
  int64_t numel[DLAP_ADAM_MAX_TENSORS];
} dlap_adam_tensors;
int dlap_clip_adam(const dlap_adam_tensors* tensors, float max_norm, float lr, float beta1, float beta2, float eps,
                   float* grad_norm, void* stream);

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
