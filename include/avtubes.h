/*
 * avtubes.h -- C ABI of libavtubes.so (sm_100a CUDA kernels for the audio-visual
 * similarity + hard-way localisation head and the cIoU/AUC evaluation).
 *
 * The reference (tonymisic/audio-visual-tubes) has no FFI: the path sits behind
 * Python call signatures.  Each entry point below names the
 * reference lines it replaces.  All pointers are DEVICE pointers unless the name
 * ends in _host; all tensors are caller-owned, contiguous and 16-byte aligned.
 * Every call only enqueues work on `stream` (a cudaStream_t passed as void*);
 * no call synchronises the device, allocates persistent memory, throws or exits.
 *
 * Return value: 0 = ok, negative = error (avt_last_error() gives the text, per thread).
 *
 * Layouts
 *   v        [N_v, C, P]  fp32   raw visual maps, NCHW with H*W flattened (P = 196); fp16 / bf16
 *                                accepted by the tensor-core path (v_dtype = AVT_DTYPE_*)
 *   a / a_hat[N_a, C]     fp32   pooled audio embeddings, raw / L2-normalised
 *   pos_idx  [N_v]        int32  column of a_hat that is row n's positive, or NULL
 *                                for pos[n] = col_offset + n (model.py:115 has
 *                                pos = arange(B); col_offset serves row-sharded ranks)
 *   logits   [N_v, L]     fp32   L = N_a + 1 + use_neg  (model.py:143-146)
 */
#ifndef AVTUBES_H_
#define AVTUBES_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AVT_OK                 0
#define AVT_ERR_INVALID       -1   /* bad shape / null pointer / unsupported size   */
#define AVT_ERR_CUDA          -2   /* a CUDA runtime or driver call failed          */
#define AVT_ERR_WORKSPACE     -3   /* workspace too small (see *_workspace_bytes)   */
#define AVT_ERR_UNSUPPORTED   -4   /* device is not sm_100 / feature not built      */

/* precision of the all-pairs contraction (model.py:125) */
#define AVT_PRECISION_FP32     0   /* CUDA-core fp32 FMA, the reference's own arithmetic */
#define AVT_PRECISION_F16      1   /* tcgen05 kind::f16: IEEE fp16 operands (11-bit significand; the
                                      operands are L2-normalised, the cell gradients carry a
                                      power-of-two scale), fp32 accumulate in tensor memory */

/* element type of the visual maps handed to the head: fp32 as the reference's fp32 backbones emit
 * them, or fp16 / bf16 when the backbone ran under autocast (AVT_PRECISION_F16 only) */
#define AVT_DTYPE_F32          0
#define AVT_DTYPE_F16          1
#define AVT_DTYPE_BF16         2

/* ground-truth rasterisation modes (utils.py:241-309) */
#define AVT_GT_VGGSS           0   /* union of boxes            -> weights {0,1},   den 1 */
#define AVT_GT_FLICKR          1   /* sum of annotators, clip 2 -> weights {0,1,2}, den 2 */
#define AVT_GT_FLICKR_FRAME    2   /* raw sum of boxes          -> counts,          den 1 */

int         avt_version(void);
const char* avt_last_error(void);
/* 1 if the current device can run the library (compute capability 10.x), else 0 */
int         avt_device_ok(void);

/* ---- measurement aids (used by bench.py; process-global, single caller) ------------------
 * avt_launch_count: kernels this library has launched since it was loaded.
 * avt_profile_enable(1) makes every entry point bracket its stages with CUDA events on the
 * stream it launches on; avt_profile_read(slot) synchronises those events and returns their
 * summed duration and the number of intervals since the last avt_profile_enable call.
 * Slots: 0 visual prep, 1 similarity (forward contraction + pooling), 2 row finish,
 * 3 row gradients, 4 backward cell-gradient contraction, 5 backward dv contraction,
 * 6 backward d_ahat contraction, 7 normalise Jacobian, 8 evaluation. */
#define AVT_PROF_PREP      0
#define AVT_PROF_SIM       1
#define AVT_PROF_FINISH    2
#define AVT_PROF_ROWGRAD   3
#define AVT_PROF_BWD_G     4
#define AVT_PROF_BWD_DV    5
#define AVT_PROF_BWD_DA    6
#define AVT_PROF_DVFINISH  7
#define AVT_PROF_EVAL      8
long long avt_launch_count(void);
int       avt_profile_enable(int on);
int       avt_profile_read(int slot, double* total_ms, int* intervals);

/* ---- audio embedding normalisation ---------------------------------------------
 * F.normalize(aud, dim=1)                        model.py:122 (tube :31)
 * a_hat[k,:] = a[k,:] / max(||a[k,:]||, 1e-12);  rnorm_a[k] = 1 / max(||a[k,:]||, 1e-12) */
int avt_audio_normalize(const float* a, int N_a, int C,
                        float* a_hat, float* rnorm_a, void* stream);
/* adjoint: da[k,:] = rnorm_a[k] * (d_ahat[k,:] - a_hat[k,:] * <a_hat[k,:], d_ahat[k,:]>) */
int avt_audio_normalize_backward(const float* a_hat, const float* rnorm_a,
                                 const float* d_ahat, int N_a, int C,
                                 float* da, void* stream);

/* ---- similarity + hard-way loss head --------------------------------------------
 * Replaces model.py:117 (F.normalize(img)), :124-125 (both einsums), :128-135
 * (tri-map sigmoids), :138-146 (masked pooling, diagonal mask, logits), :148-152
 * (weighted_A) and nn.CrossEntropyLoss(logits, 0) (train_hardway_1frame.py:113,131);
 * tube model: model.py:35,47-59.  The N_v x N_a x P similarity tensor is never
 * written to memory.
 *
 * Outputs
 *   A, Pos, Neg  [N_v, P]        model.py:124,128,131/133  (returned maps)
 *   logits       [N_v, L]
 *   loss_rows    [N_v]           logsumexp(logits[n]) - logits[n,0]
 *   loss_sum     [1] or NULL     sum_n loss_rows[n] (deterministic order)
 *   weighted_A   [N_v, P] or NULL model.py:148-152
 * Saved for backward (all required)
 *   rnorm_v [N_v, P], sim [N_v, N_a] (pooled, before mask), den [N_v, N_a]
 *
 * Tube mode, col_mult = T > 1 (train_3D.py:129-130, train_hardway.py:128-129
 * repeat every spectrogram T times, so model.py:47-51 scores (bT)^2 cells of which only
 * b*(bT) are distinct): pass the b DISTINCT audio rows (N_a = b), pos_idx[n] = the audio row of
 * frame n (required), and col_mult = T.  Audio row k then stands for T identical logit columns;
 * exactly one of them, in the row's positive audio, carries the 1 - mask_scale factor of
 * model.py:47,57.  loss_rows / loss_sum are those of the reference's [N_v, N_a*T + 2] logits;
 * `logits` holds the COMPACT row [sim1, N_a masked-as-usual columns, sim2] (the host mirror
 * expands it), and d_ahat of the backward is the gradient wrt the distinct rows (the sum over
 * the copies, what autograd gives through `repeat`).  dlogits, if given in tube mode, is
 * compact too: per audio row the sum over its T copies with the mask factor already applied.
 * 1/T of the similarity flops, 1/T of the audio-backbone work upstream. */
size_t avt_head_workspace_bytes(int N_v, int N_a, int C, int P, int precision, int backward);

/* stages of avt_head_forward (0 = both).  AVT_STAGE_PREP is the one pass over the visual maps
 * (1/||v||, the positive map A, mean_c, the fp16 operand copy); it reads a_hat ONLY at the rows'
 * positives, so a row-sharded caller may run it with just its own audio rows (N_a = rows held,
 * col_offset = 0) while the all-gather of the other ranks' rows is still in flight, then call
 * AVT_STAGE_SIM (audio operand, contraction, pooling, row finish) with the full a_hat.
 * fuse_grad_scale != 0 (with AVT_STAGE_SIM; needs a backward-sized workspace): the fused training
 * step -- the row-finish kernel also emits the row gradients of  fuse_grad_scale * sum_n loss_rows[n],
 * and avt_head_backward is then called with AVT_BWD_ROWGRAD_DONE (one launch less on the
 * critical path; dloss_scale / dlogits / dA / dWeighted / dloss_rows must be 0 / NULL). */
#define AVT_STAGE_PREP         1
#define AVT_STAGE_SIM          2
#define AVT_BWD_ROWGRAD_DONE   1

int avt_head_forward(const void* v, int v_dtype, const float* a_hat,
                     int N_v, int N_a, int C, int P,
                     const int32_t* pos_idx, int col_offset,
                     float eps, float eps2, float tau, float temp, float mask_scale,
                     int trimap, int use_neg, int precision, int col_mult,
                     float* A, float* Pos, float* Neg, float* logits,
                     float* loss_rows, float* loss_sum, float* weighted_A,
                     float* rnorm_v, float* sim, float* den,
                     int stages, float fuse_grad_scale,
                     void* workspace, size_t workspace_bytes, void* stream);

/* Backward of the head (autograd of model.py:117-146 + CrossEntropyLoss).
 * Gradient of   dloss_scale * sum_n loss_rows[n]  +  <dloss_rows, loss_rows>  +  <dlogits, logits>
 *               +  <dA, A>  +  <dWeighted, weighted_A>
 * (dloss_rows [N_v], dlogits, dA, dWeighted [N_v, P] may be NULL; dWeighted is what train_hardway.py:138-142
 * sends back through weighted_A = mean_c(v_hat * normalize(Pos)), model.py:148-152: it
 * reaches v through every channel alike and through Pos -> A; needs the workspace the forward
 * ran with, which still holds mean_c v_hat).  Writes dv [N_v, C, P] (gradient wrt the RAW maps) and
 * d_ahat [N_a, C] (gradient wrt the NORMALISED audio rows, summed over this call's
 * N_v rows; finish with avt_audio_normalize_backward, after a reduce-scatter when
 * rows are sharded over ranks).
 * dv_stream (cudaStream_t or NULL): when given, the dv contraction is enqueued on THAT stream
 * (ordered after the stages it depends on) and d_ahat completes on `stream` without waiting for
 * it -- a sharded caller starts its reduce-scatter of d_ahat under the dv kernel and orders
 * dv_stream before whatever consumes dv.  NULL: everything is complete in `stream` order. */
int avt_head_backward(const void* v, int v_dtype, const float* a_hat,
                      int N_v, int N_a, int C, int P,
                      const int32_t* pos_idx, int col_offset,
                      float eps, float eps2, float tau, float temp, float mask_scale,
                      int trimap, int use_neg, int precision, int col_mult,
                      const float* A, const float* logits, const float* loss_rows,
                      const float* rnorm_v, const float* sim, const float* den,
                      float dloss_scale, const float* dloss_rows,
                      const float* dlogits, const float* dA,
                      const float* dWeighted,
                      float* dv, float* d_ahat,
                      int flags, void* dv_stream,
                      void* workspace, size_t workspace_bytes, void* stream);

/* ---- evaluation -------------------------------------------------------------------
 * One launch for the reference's per-clip glue, train_hardway_1frame.py:198-207
 * (== test.py:97-107): cv2.resize 14->224 INTER_LINEAR, utils.normalize_img(-h)
 * (utils.py:234-239), pred = 1 - h, threshold = sorted(pred)[25088], binarise with a
 * strict '>', Evaluator.cal_CIOU(pred, gt, 0.5) (utils.py:209-214).
 *
 *   maps14     [N, 14, 14] fp32
 *   gt_dense   [N, 224, 224] uint8 weights (gt = weight / gt_den), or NULL to rasterise
 *   boxes      [N, max_boxes, 4] int16 pixel rectangles (xmin, ymin, xmax, ymax) as
 *              utils.py:254,267 derive them, box_count [N] uint8; used when gt_dense NULL
 *   gt_mode    AVT_GT_*           (fixes gt_den: 1, 2, 1)
 *   up224      [N, 224, 224] fp32 or NULL     the upsampled map (train_hardway_1frame.py:198)
 *   mask       [N, 224, 224] uint8 or NULL    the binary prediction (:203-204)
 *   inter_w, union_w [N] int32    integer numerator / denominator in weight units:
 *              ciou = inter_w / union_w exactly as utils.py:212-214 computes in float64
 *   ciou       [N] float64        (NaN when union_w == 0, as numpy)
 *   workspace  AVT_EVAL_WORKSPACE_BYTES device bytes owned by this call until it completes (the
 *              work-ticket word; zeroed by the library on `stream`), so concurrent launches,
 *              captured graphs and failed launches share no state                        */
#define AVT_EVAL_WORKSPACE_BYTES 64
int avt_eval_maps(const float* maps14, int N,
                  const uint8_t* gt_dense,
                  const int16_t* boxes, const uint8_t* box_count, int max_boxes, int gt_mode,
                  float* up224, uint8_t* mask,
                  int32_t* inter_w, int32_t* union_w, double* ciou,
                  void* workspace, void* stream);

/* Selection statistics of avt_eval_maps since the last reset (host array of 2):
 *   out2[0] = maps whose median-rank threshold (train_hardway_1frame.py:202, np.sort(...)[25088])
 *             needed the generic counted-bisection path instead of the sampled bracket,
 *   out2[1] = maps evaluated.  Synchronises the device; diagnostics only. */
int avt_eval_stats(unsigned long long* out2, int reset);

/* Evaluator.cal_CIOU(infer, gtmap, thres) (utils.py:209-214) on N ready-made 224x224 fp32
 * maps: infer_map = (infer >= thres); gt = gt_w / gt_den with gt_w uint8 weights.
 * inter_w / union_w / ciou as in avt_eval_maps. */
int avt_ciou_counts(const float* infer, float thres, const uint8_t* gt_w, int gt_den, int N,
                    int32_t* inter_w, int32_t* union_w, double* ciou, void* stream);

/* utils.py:241-274 / 276-309: boxes -> dense weight map [N,224,224] uint8 */
int avt_rasterise_gt(const int16_t* boxes, const uint8_t* box_count, int N, int max_boxes,
                     int gt_mode, uint8_t* gt_dense, void* stream);

/* train_hardway_1frame.py:208-215 / utils.py:216-229:
 *   counts[i] = #{ciou >= 0.05*i}, i = 0..20 (float64 compare, NaN never counts);
 *   out[0] = cIoU@0.5 = counts-style mean(ciou >= 0.5); out[1] = trapezoid AUC.
 * counts [21] int32, out [2] float64 (device). */
int avt_auc(const double* ciou, int N, int32_t* counts, double* out, void* stream);

/* utils.py:311-318 mTC: mean over consecutive frames of cal_CIOU(mask[i], mask[i+1], 0.5).
 * masks [T,224,224] uint8 {0,1}; pair_inter, pair_union [T-1] int32; out [1] float32. */
int avt_mtc(const uint8_t* masks, int T, int32_t* pair_inter, int32_t* pair_union,
            float* out, void* stream);

/* avt_mtc for callers that can not allocate (code under CUDA graph capture, serving loops, harnesses
 * that forbid cudaMalloc).  This entry point has NO counterpart in the reference.
 * Same masks / T / out and the same result as avt_mtc, but the pair
 * counts live in a fixed device array inside the library (2 x AVT_MTC_STATIC_MAX_FRAMES int32 = 32 KB,
 * held by every GPU that loads the library until the process exits); the call allocates nothing.
 * T > AVT_MTC_STATIC_MAX_FRAMES returns AVT_ERR_INVALID (use avt_mtc with caller scratch).  Every call
 * on a device shares that array: calls whose kernels may overlap on the GPU (different streams or
 * host threads) race with each other -- serialise them.  Nothing in the array has to be cleared between
 * calls (each call writes every entry it later reads). */
#define AVT_MTC_STATIC_MAX_FRAMES 4096
int avt_mtc_static(const uint8_t* masks, int T, float* out, void* stream);

/* The per-video metrics of the PerFrameLabels test loop, train_hardway.py:163-203: video v owns
 * frames [v*F, (v+1)*F) of ciou [V*F] float64 and masks [V*F,224,224] uint8 (both as
 * avt_eval_maps wrote them).  out [V,3] float64 (device):
 *   out[v][0] = #(ciou >= 0.5) / F                                   train_hardway.py:200
 *   out[v][1] = sklearn auc of the video's 21-point success curve    :193-199
 *   out[v][2] = mTC(preds of the video)  utils.py:311-318 (float32 accumulation; NaN if F == 1)
 * pair_inter, pair_union [V*F-1] int32 scratch (consecutive-mask counts). */
int avt_video_metrics(const double* ciou, const uint8_t* masks, int V, int F,
                      int32_t* pair_inter, int32_t* pair_union, double* out, void* stream);

/* ---- audio front-end ------------------------------------------------
 * datasets/dataloader.py:86-96 for B clips in one launch: tile a short clip to n_keep samples
 * (:88-91), clip to [-1, 1] (:92-93), scipy.signal.spectrogram(x, fs, nperseg=512, noverlap=1)
 * (:94; hop 511, per-segment mean removal, window, one-sided 'density' PSD, 257 bins),
 * log(S + log_eps) (:95) and Normalize(0, std) (:72,96).  fp64 arithmetic like scipy, fp32 out
 * (the training loops call .float(), train_hardway_1frame.py:127).
 *   samples  [B, stride] fp32 mono PCM in [-1, 1] (exact for <= 24-bit WAV)
 *   lengths  [B] int32 valid samples per clip (tiled when < n_keep), or NULL = all n_keep
 *   n_keep   samplerate * 10;   window [512] float64 (device), scipy's get_window(('tukey', .25), 512)
 *   scale    1 / (samplerate * sum(window^2));   log_eps 1e-7;   std 12
 *   out      [B, 257, T] fp32,  T = avt_spectrogram_frames(n_keep) = (n_keep - 1) / 511 */
int avt_spectrogram_frames(int n_keep);
int avt_log_spectrogram(const float* samples, const int32_t* lengths, int B, long long stride,
                        int n_keep, const double* window, double scale, double log_eps, double std,
                        float* out, int T, void* stream);
/* the same from 16-bit PCM as it sits in a WAV file: sample / 32768, what sf.read returns for
 * PCM_16 (datasets/dataloader.py:86); half the bytes to upload and to read */
int avt_log_spectrogram_pcm16(const int16_t* samples, const int32_t* lengths, int B, long long stride,
                              int n_keep, const double* window, double scale, double log_eps, double std,
                              float* out, int T, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* AVTUBES_H_ */
