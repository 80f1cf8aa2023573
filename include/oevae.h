/* oevae.h -- C ABI of liboevae.so: the B200-native OE-VAE hot path.
 *
 * The reference (dominiquegodin/ATLAS-VAE) is pure Python on TensorFlow-Keras and has NO FFI
 * layer of its own; the drop-in boundary is the Python contract between OE-VAE/vae.py and
 * OE-VAE/models.py / utils.py / plots.py.  atlas_vae_b200/{models,utils,cuts}.py keeps that
 * Python surface and binds the entry points below through ctypes (atlas_vae_b200/_ffi.py).
 * Every entry point cites the reference code it replaces (paths relative to /root/reference).
 *
 * Conventions
 *   - return 0 on success, a negative OEVAE_E* code on failure; text via oevae_last_error().
 *     Nothing throws or aborts; CUDA errors are mapped to OEVAE_ECUDA.
 *   - every data pointer is a DEVICE pointer owned by the caller unless the name ends in _host.
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*); no hidden host sync
 *     except in the *_host getters/setters.
 *   - a handle is not thread-safe.  One handle per (process, GPU).
 *   - matrices are row-major fp32; kernels are Keras [in,out]; the flat parameter vector is
 *     Keras trainable_weights order (models.py:22-28,43-48): encoder hidden (kernel,bias)...,
 *     dense_mean, dense_log_var, decoder hidden..., dense_output.
 */
#ifndef OEVAE_H
#define OEVAE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define OEVAE_VERSION 200
#if defined(__GNUC__)
#define OEVAE_API __attribute__((visibility("default")))
#else
#define OEVAE_API
#endif

enum { OEVAE_OK = 0, OEVAE_EINVAL = -1, OEVAE_ECUDA = -2, OEVAE_ENOMEM = -3, OEVAE_EUNSUPPORTED = -4 };

/* OE_type of models.py:72-105 */
enum { OEVAE_OE_KLD = 0, OEVAE_OE_MSE = 1, OEVAE_OE_MAE = 2, OEVAE_OE_MSE_MARGIN = 3, OEVAE_OE_MAE_MARGIN = 4 };

/* arithmetic of the Dense contractions (value 2 is unassigned: a kind::tf32 mode was planned and dropped -- bf16x3 reaches
 * fp32-grade accuracy at the same MMA cost per product) */
enum { OEVAE_PREC_FP32 = 0,      /* CUDA-core fp32 FMA: the reference-dtype mode and parity anchor   */
       OEVAE_PREC_BF16X3 = 1,    /* tcgen05 kind::f16, operands split hi+lo bf16, 3 MMAs, fp32 accum */
       OEVAE_PREC_BF16 = 3 };    /* tcgen05 kind::f16 on bf16-rounded operands, fp32 accumulate      */

/* where the reparameterisation noise comes from (models.py:9-15) */
enum { OEVAE_NOISE_FIXED = 0,    /* reference seed=0 behaviour: same eps tensor for every call of a shape */
       OEVAE_NOISE_PER_STEP = 1, /* fresh Philox4x32-10 draw per call (reference seed=None) */
       OEVAE_NOISE_EXTERNAL = 2  /* caller passes eps explicitly (parity tests) */ };

typedef struct oevae_handle* oevae_t;

typedef struct {
    int32_t  input_dim;          /* D = n_dims*n_const (+HLVs)          vae.py:72-76            */
    int32_t  n_hidden;           /* len(FC_layers)-1                    models.py:22            */
    int32_t  hidden[8];          /* FC_layers[:-1]                                              */
    int32_t  latent_dim;         /* FC_layers[-1]                       models.py:27-28         */
    int32_t  oe_type;            /* OEVAE_OE_*                                                  */
    int32_t  precision;          /* OEVAE_PREC_*                                                */
    int32_t  noise_mode;         /* OEVAE_NOISE_*                                               */
    int32_t  device;             /* CUDA ordinal                                                */
    uint64_t seed;               /* Philox key                                                  */
    int64_t  max_batch;          /* workspace is sized for this many jets per call (grown on demand) */
} oevae_config;

/* ---- lifecycle: VariationalAutoEncoder.__init__ (models.py:55-61) ------------------------- */
OEVAE_API int         oevae_create(const oevae_config* cfg, oevae_t* out);
OEVAE_API int         oevae_destroy(oevae_t h);
OEVAE_API int         oevae_version(void);
OEVAE_API const char* oevae_last_error(oevae_t h);           /* h may be NULL: last error of a failed create */
OEVAE_API int64_t     oevae_n_params(oevae_t h);
/* (name-less) layout: writes 2*(2*n_hidden+3) offsets [kernel_off, bias_off] per Dense, returns count */
OEVAE_API int         oevae_param_layout(oevae_t h, int64_t* offsets, int32_t* dims_in, int32_t* dims_out, int32_t cap);

/* OE_type used by the next train/eval calls (get_losses takes it per call, models.py:108) */
OEVAE_API int oevae_set_oe_type(oevae_t h, int32_t oe_type);

/* ---- weights / optimiser state: save_weights / load_weights (models.py:227, vae.py:101,141) */
OEVAE_API int oevae_set_params_host(oevae_t h, const float* flat_host, int64_t n);
OEVAE_API int oevae_get_params_host(oevae_t h, float* flat_host, int64_t n);
OEVAE_API int oevae_set_opt_state_host(oevae_t h, const float* m_host, const float* v_host, int64_t n, int64_t step);
OEVAE_API int oevae_get_opt_state_host(oevae_t h, float* m_host, float* v_host, int64_t n, int64_t* step);
/* fresh optimiser (m = v = 0, step = 0): the reference builds a new tf.keras.optimizers.Adam in every train_model call (models.py:138) */
OEVAE_API int oevae_reset_opt_state(oevae_t h);
OEVAE_API float* oevae_params_device(oevae_t h);              /* device address of the flat parameters (read-only for callers) */
/* Philox row counter of row 0 of the next calls: a data-parallel rank passes its offset inside the global
 * batch so that the noise of a jet does not depend on how the batch is sharded. */
OEVAE_API int oevae_set_noise_offset(oevae_t h, uint64_t row_offset);

/* ---- training step: train_model batch body (models.py:180-190) ----------------------------
 * get_losses (models.py:108-132) + tape.gradient of the SUM over the batch (models.py:183) +
 * clip_values on the gradient (models.py:185) + Keras Adam (models.py:138,186).
 *   x_bkg [B,D], w_bkg [B];  x_ood [B,D] / w_ood [B] nullable together (then the OE term is 0)
 *   eps    nullable unless noise_mode==EXTERNAL; [n_eps,B,L], n_eps in {1, number of passes}
 *   per_jet_out nullable [4,B]: rec*w, beta*kld*w, lamb*oe*w_ood, total  (models.py:132)
 *   sums_acc    nullable double[4]: the four batch SUMS are ADDED to it (keras.metrics.Mean numerators,
 *               models.py:187-190); the caller zeroes it per epoch and divides by the jet count.
 */
OEVAE_API int oevae_train_step(oevae_t h, const float* x_bkg, const float* w_bkg,
                     const float* x_ood, const float* w_ood,
                     const float* eps, int32_t n_eps, int64_t B,
                     float lr, float beta, float lamb, float margin,
                     float* per_jet_out, double* sums_acc, void* stream);

/* forward + backward only: leaves d(sum_i total_i)/d(theta) (un-clipped) in grad_out [P] so that a
 * data-parallel caller can all-reduce(SUM) it before oevae_apply_update. */
OEVAE_API int oevae_loss_and_grad(oevae_t h, const float* x_bkg, const float* w_bkg,
                        const float* x_ood, const float* w_ood,
                        const float* eps, int32_t n_eps, int64_t B,
                        float beta, float lamb, float margin,
                        float* per_jet_out, double* sums_acc, float* grad_out, void* stream);

/* ---- data parallel over NVLink peer memory (replaces jet-ID's MirroredStrategy semantics, jet-ID/classifier.py:136-138,
 * for --n_gpus N; one process per GPU on one node).  oevae_dp_export allocates this rank's receive buffer and writes
 * its 64-byte cudaIpcMemHandle_t; the caller gathers the handles of all ranks (any transport), passes the
 * world x 64 bytes to oevae_dp_connect and synchronises the ranks once (a barrier) before the first step.
 * oevae_train_step_dp is oevae_train_step for one rank's slice of the global batch: its update kernel sums this rank's
 * partial gradients, pushes the sums into every peer's receive buffer (P2P stores, value and sequence tag in one 16-byte
 * store) and applies clip + Adam to the SUM over ranks (models.py:183), added in rank order -- identically on every rank.
 * Every rank must take part in every such step, in the same order.  B == 0 is allowed there:
 * a rank whose shard of a short last batch is empty still joins the exchange, with a zero gradient.
 * OEVAE_EUNSUPPORTED from oevae_dp_connect (network too large for a co-resident launch): use oevae_loss_and_grad +
 * all-reduce + oevae_apply_update. */
OEVAE_API int oevae_dp_export(oevae_t h, void* handle_out64);
OEVAE_API int oevae_dp_connect(oevae_t h, int32_t rank, int32_t world, const void* handles);
OEVAE_API int oevae_train_step_dp(oevae_t h, const float* x_bkg, const float* w_bkg,
                        const float* x_ood, const float* w_ood,
                        const float* eps, int32_t n_eps, int64_t B,
                        float lr, float beta, float lamb, float margin,
                        float* per_jet_out, double* sums_acc, void* stream);

/* the SUM over ranks of the last oevae_train_step_dp's gradient, before clip_values (host copy; synchronises): lets a
 * data-parallel caller check the exchanged gradient against oevae_loss_and_grad of the global batch */
OEVAE_API int oevae_dp_last_grad_host(oevae_t h, float* grad_host, int64_t n);

/* clip_values(grad) + Adam with the given (possibly all-reduced) gradient; increments the step. */
OEVAE_API int oevae_apply_update(oevae_t h, const float* grad, float lr, void* stream);

/* get_losses without the tape: validation loop (models.py:201-211) and the public get_losses(). */
OEVAE_API int oevae_eval_losses(oevae_t h, const float* x_bkg, const float* w_bkg,
                      const float* x_ood, const float* w_ood,
                      const float* eps, int32_t n_eps, int64_t B,
                      float beta, float lamb, float margin,
                      float* per_jet_out, double* sums_acc, void* stream);

/* ---- model call surface: vae.encoder(X) / vae(X) (models.py:30-36,62-69) ------------------ */
OEVAE_API int oevae_encode(oevae_t h, const float* x, int64_t N, const float* eps,
                 float* mu, float* log_var, float* z, void* stream);

/* ---- anomaly scoring: model.predict x n_iter (vae.py:163-166) + loss_function MSE/MAE/KLD/JSD
 * (utils.py:346-375) + latent_loss (utils.py:378-386) + filtering flag (utils.py:532-538), one pass.
 * Outputs are nullable [N] (xhat: nullable [N,D]).  eps: nullable [n_iter,N,L] (EXTERNAL mode). */
OEVAE_API int oevae_score(oevae_t h, const float* x, int64_t N, int32_t n_iter, const float* eps,
                float* mse, float* mae, float* kld, float* jsd, float* latent,
                uint8_t* finite, float* xhat, void* stream);

/* ---- handle-free kernels ---------------------------------------------------------------- */
/* jets_sorting (utils.py:564-571) + load_data slice/pad/n_dims (utils.py:89-97,115-119) + get_4v
 * (utils.py:300-308) + affine scaler apply (utils.py:406-417).  raw_dtype: 0 fp32, 1 fp16.
 * sort: 1 = pT-sort first (jets_sorting); 0 = keep the stored order (get_4v on sorted jets).
 * raw [N,4*n_raw] (E,px,py,pz); x_out [N,n_dims*n_const]; pt_out, m_out nullable [N];
 * center/scale nullable double[n_dims*n_const]: fp32(fp32(x-center)/scale), evaluated in fp64 like sklearn's
 * in-place transform of a float32 array (RobustScaler keeps float64 center_/scale_). */
OEVAE_API int oevae_preprocess(const void* raw, int32_t raw_dtype, int64_t N, int32_t n_raw, int32_t n_const,
                     int32_t n_dims, const double* center, const double* scale, int32_t sort,
                     float* x_out, float* pt_out, float* m_out, int32_t device, void* stream);

/* fp16 ingest: the reference's files store the constituents as float16 (tools/merging.py:31,51-53); load_data widens them to
 * float32 (utils.py:89-97) and apply_scaling transforms them (utils.py:406-417).  raw_f16 [N,D] (already sorted / sliced) ->
 * x [N,D] = fp32(fp32(fp32(raw) - center) / scale), center / scale nullable as in oevae_preprocess.  Bit-identical to
 * oevae_affine on the widened array, at half the upload bytes. */
OEVAE_API int oevae_ingest_f16(const void* raw_f16, float* x, int64_t N, int32_t D, const double* center, const double* scale,
                     int32_t device, void* stream);

/* apply_scaling for affine scalers (utils.py:406-417) on an already sliced [N,D] array; y may alias x */
OEVAE_API int oevae_affine(const float* x, float* y, int64_t N, int32_t D, const double* center, const double* scale,
                 int32_t device, void* stream);

/* apply_scaling for the NON-affine scalers of fit_scaler (utils.py:393-396,411) on an [N,D] float32 array; y may alias x.
 * oevae_quantile_normal: QuantileTransformer(output_distribution='normal').transform -- quantiles_t: DEVICE double[D][n_q] (column f's
 *   quantiles_[:, f] contiguous), references: DEVICE double[n_q], clip_min / clip_max = norm.ppf(1e-7 - eps) / norm.ppf(1 - (1e-7 - eps));
 *   bit-identical to scikit-learn's float32 in-place transform up to the last bit of the float64 logarithm in the tails.
 * oevae_power_yj: PowerTransformer(method='yeo-johnson', standardize=True).transform for a transformer fitted on float32 data
 *   (float32 lambdas_; mean / scale = its StandardScaler's mean_ / scale_ as float32, nullable) -- float32 log1p / expm1, so equal
 *   to scikit-learn's result to a few float32 ulps, not bit for bit. */
OEVAE_API int oevae_quantile_normal(const float* x, float* y, int64_t N, int32_t D, const double* quantiles_t, const double* references, int32_t n_q,
                          double clip_min, double clip_max, int32_t device, void* stream);
OEVAE_API int oevae_power_yj(const float* x, float* y, int64_t N, int32_t D, const float* lambdas, const float* mean, const float* scale,
                   int32_t device, void* stream);

/* loss_function (utils.py:346-375) between two given arrays P, Q [N,D]; every output nullable [N]:
 * MSE, MAE, MARE, KLD, JSD, X-S */
OEVAE_API int oevae_pair_metrics(const float* P, const float* Q, int64_t N, int32_t D, float* mse, float* mae, float* mare,
                       float* kld, float* jsd, float* xs, int32_t device, void* stream);

/* loss_mapping global predicates (plots.py:15-20): minmax_out double[3] = {min, max, #NaN} over
 * rows with keep[i] (keep nullable); the caller initialises it to {+inf, -inf, 0} */
OEVAE_API int oevae_minmax(const float* v, const uint8_t* keep, int64_t N, double* minmax_out, int32_t device, void* stream);
/* loss_mapping elementwise branch (0..4) applied in place */
OEVAE_API int oevae_loss_map(float* v, int64_t N, int32_t branch, int32_t device, void* stream);

/* make_cut mask (plots.py:259) + per-class histogram of plot_distributions (plots.py:697-715):
 * class 0: JZW == -1, class 1: JZW >= 0.  A jet enters iff keep[i] (nullable) and score[i] > cut
 * (score nullable => no cut).  edges: double[n_edges] (np.histogram semantics: half-open bins, last bin
 * closed).  counts int64 [2,n_edges-1] and wsums double [2,n_edges-1] are ADDED to.
 * wscale / density: the weight scalings of plots.py:706-712 applied per jet in the reference's rounding sequence:
 * w1 = fp32(w * wscale) (normalize: wscale = fp32(100) / fp32(sum of the uncut sample's weights), else 1),
 * w2 = fp32(fp64(w1) / bin width) when density != 0; wsums accumulates the scaled weights. */
OEVAE_API int oevae_hist(const float* var, const float* w, const float* jzw, const float* score, const uint8_t* keep,
               int64_t N, float cut, const double* edges, int32_t n_edges, float wscale, int32_t density,
               int64_t* counts, double* wsums, int32_t device, void* stream);
/* fp64 sum of a float32 vector over the rows with keep[i] (keep nullable), ADDED to *sum_acc (device): np.sum(sample['weights']) of
 * plots.py:707 */
OEVAE_API int oevae_wsum(const float* w, const uint8_t* keep, int64_t N, double* sum_acc, int32_t device, void* stream);

/* ---- weighted ROC curve and cut thresholds on the device (the step right after scoring) ------------------------------------
 * oevae_roc: get_rates (plots.py:228-232) = sklearn.metrics.roc_curve(y_true, X_loss, pos_label, sample_weight) with its default
 * drop_intermediate=True, bit-identical to scikit-learn's float64 arithmetic (stable descending sort, zero-weight samples
 * dropped, np.cumsum order; roc_kernels.cuh).  score [N] float32 (finite), y_true [N] uint8, w nullable [N] float32 (=> 1).
 * fpr, tpr, thr: DEVICE double arrays of capacity N + 1; *n_points_host receives the number of points written (the first one
 * is (0, 0, inf)).  Allocates its workspace and synchronises the stream.
 * oevae_roc_cuts: over the points with fpr != 0, rates in percent (plots.py:230): entry e < n_eff = argmin |fpr - eff[e]|
 * (make_cut, plots.py:256-258), entry n_eff = argmax tpr / fpr over tpr > min_tpr (best_threshold 'gain', plots.py:237-246).
 * out_host: double[4 (n_eff + 1)] = {threshold, value, index in the curve, fpr there} per entry.
 * oevae_roc_searchsorted: thresholds[min(searchsorted(fpr, eff, 'right'), len - 1)] over the same points (bump_scan's cut grid,
 * plots.py:272-282). */
OEVAE_API int oevae_roc(const float* score, const uint8_t* y_true, int32_t pos_label, const float* w, int64_t N,
              double* fpr, double* tpr, double* thr, int64_t* n_points_host, int32_t device, void* stream);
OEVAE_API int oevae_roc_cuts(const double* fpr, const double* tpr, const double* thr, int64_t n_points, const double* eff_host, int32_t n_eff,
                   double min_tpr, double* out_host, int32_t device, void* stream);
OEVAE_API int oevae_roc_searchsorted(const double* fpr, const double* thr, int64_t n_points, const double* eff_host, int32_t n_q,
                           double* thr_out_host, double* eff_out_host, int32_t device, void* stream);

/* ---- mass decorrelation (plots.py:54-85, --decorrelation ON) -----------------------------------------------------------------
 * oevae_digitize_count: counts[clip(np.digitize(var, edges), 1, n_edges - 1) - 1] += 1 over the rows with keep[i] (nullable): the
 *   occupancy get_bins (plots.py:54-63) merges bins on; edges DEVICE double[n_edges], counts DEVICE uint32[n_edges - 1] (ADDED to).
 * oevae_mass_deco: loss[i] <- empirical CDF (cum_distribution, plots.py:64-66, scipy interp1d linear rule) of the BACKGROUND scores of
 *   jet i's (mass, pT) cell, in place; the pT edges of mass bin k are pt_edges[pt_off[k] .. pt_off[k + 1]) (DEVICE; int32 pt_off
 *   [n_m_edges]).  Allocates its workspace and synchronises the stream. */
OEVAE_API int oevae_digitize_count(const float* var, const uint8_t* keep, int64_t N, const double* edges, int32_t n_edges, uint32_t* counts,
                         int32_t device, void* stream);
OEVAE_API int oevae_mass_deco(float* loss, const float* m, const float* pt, const uint8_t* is_bkg, int64_t N, const double* m_edges,
                    int32_t n_m_edges, const double* pt_edges, const int32_t* pt_off, int32_t device, void* stream);

/* ---- OoD <-> background pairing (the step right before training, OE-VAE/utils.py:193-227) -------------------------------------
 * oevae_argsort_f32: np.argsort(OoD_sample['m']) (utils.py:195; stable) -> order_out int64 [N] (device).
 * oevae_ood_pairing: for every background jet, the index (into the m-SORTED OoD arrays) of one OoD jet drawn uniformly from the
 *   (m, pT) window around it, the window widened x2 alternately in pT and in m, starting at 10 x 10 GeV, until it is not empty
 *   (utils.py:198-210).  Draw: Philox4x32-10 keyed by (seed, jet) -- the reference's np.random.choice stream under fork() is not
 *   reproducible; pick_first != 0 returns the first candidate instead (tests).  widths_out: nullable float [n_bkg][2] = final
 *   (m_width, pt_width).  idx_out int64 [n_bkg] (device). */
OEVAE_API int oevae_argsort_f32(const float* v, int64_t N, int64_t* order_out, int32_t device, void* stream);
/* oevae_argsort_f32_static: NO COUNTERPART IN THE REFERENCE -- added for callers that may not
 *   allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same result as
 *   oevae_argsort_f32, but the radix-sort workspace is a file-scope static __device__ array of the library (about 4.1 MiB, held by
 *   every GPU that loads the library until exit), so the call allocates and frees nothing and does not synchronise: order_out is
 *   ready in stream order.  N > OEVAE_ARGSORT_STATIC_MAX -> OEVAE_EINVAL with a message naming the limit.  The first call on a device
 *   resolves the array's device address (cudaGetSymbolAddress); make that call outside a capture.  All calls on one GPU share the
 *   array: calls that overlap on the GPU (different streams, or a replayed graph next to a live call) race -- keep them in one
 *   stream order. */
#define OEVAE_ARGSORT_STATIC_MAX 262144
OEVAE_API int oevae_argsort_f32_static(const float* v, int64_t N, int64_t* order_out, int32_t device, void* stream);
OEVAE_API int oevae_ood_pairing(const float* m_bkg, const float* pt_bkg, int64_t n_bkg, const float* m_ood_sorted, const float* pt_ood_sorted,
                      int64_t n_ood, uint64_t seed, int32_t pick_first, int64_t* idx_out, float* widths_out, int32_t device, void* stream);

/* ---- train_model's batch loop over one HOST-resident load (models.py:172-190), one call per group of batches ----------------
 * Replaces the implicit per-step host->device copies of models.py:180-182: batches [first_batch, first_batch + n_batches) of
 * `batch_size` rows are copied host->device on an internal copy stream, up to five batches ahead of the compute stream, and stepped with
 * oevae_train_step (update 0) or oevae_train_step_dp (update 1; this rank's contiguous shard [n rank / world, n (rank+1) / world)
 * of every global batch, an empty shard joins the exchange).  Host arrays should be page-locked (oevae_host_register).
 *   x_dtype 1: x_host / xo_host hold float16 rows as the reference's files do (tools/merging.py:31); they are widened and scaled on
 *              the device (oevae_ingest_f16).  x_dtype 0 with a scaler: oevae_affine on the device.  center / scale: DEVICE double[D].
 *   x_resident (+ w_resident, xo_resident, wo_resident): nullable DEVICE buffers that receive the float32 batches, batch b at row
 *              b * resident_stride (stride >= the largest shard): the load then stays on the device and later epochs step on it
 *              without another copy.  Null: a six-slot ring inside the handle.
 *   sums_acc   DEVICE double[4] (as oevae_train_step); sums_trace_host: nullable page-locked double[n_batches][4] that receives
 *              sums_acc after every step (device->host copy in stream order): the per-step metric read of models.py:187-190. */
typedef struct {
    const void*  x_host;  const float* w_host;
    const void*  xo_host; const float* wo_host;       /* OoD arrays: both null => no OE term */
    int32_t      x_dtype;                             /* 0 float32, 1 float16 */
    int32_t      update;                              /* 0 single rank, 1 peer-memory data parallel */
    int64_t      n_rows, batch_size, first_batch, n_batches;
    int32_t      rank, world;                         /* row shard of every batch: [n rank / world, n (rank+1) / world); world 1 = all rows */
    int64_t      noise_row0;                          /* added to the Philox row counter (a rank that feeds its own pre-sharded load
                                                         passes rank * batch_size so that ranks do not share noise) */
    const double* center; const double* scale;
    float*       x_resident; float* w_resident; float* xo_resident; float* wo_resident;
    int64_t      resident_stride;
    float        lr, beta, lamb, margin;
    double*      sums_acc; double* sums_trace_host;
    int64_t      h2d_bytes;                           /* out: bytes copied host->device by this call */
} oevae_host_load;
OEVAE_API int oevae_train_batches_host(oevae_t h, oevae_host_load* load, void* stream);

/* _BatchStager (host loads -> device batches, replacing the implicit per-step H2D of models.py:180): page-lock a
 * host array in place.  Returns 0 when newly registered (release it with oevae_host_unregister), 1 when the range
 * was already page-locked (nothing to release), OEVAE_ECUDA otherwise; never leaves a CUDA error pending. */
OEVAE_API int oevae_host_register(void* host_ptr, int64_t bytes);
OEVAE_API int oevae_host_unregister(void* host_ptr);

/* number of kernels this library has launched since load (bench.py "gpu_launches") */
OEVAE_API int64_t oevae_launch_count(void);

/* per-kernel CUDA-event timing for bench.py's roofline leg: enable, run steps, then read
 * "name,launches,total_ms" lines (sorted by total time).  Adds two event records per launch while on. */
OEVAE_API int oevae_profile_enable(int32_t on);
OEVAE_API int oevae_profile_report(char* buf, int32_t cap);

#ifdef __cplusplus
}
#endif
#endif
