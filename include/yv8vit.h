/* yv8vit.h -- C-ABI of libyv8vit.so: the B200-native (sm_100a) second-stage hot
 * path of Voyager0587/yolov8-vit.
 *
 * The reference has no FFI/plugin layer for this path: its boundary is plain
 * Python function calls.  Each entry point below therefore
 * cites the reference Python/markdown location whose behaviour it replaces.
 * The ctypes binding that sits on top is yolov8_vit_b200/_lib.py; the
 * reference-side stub a maintainer would add is shown in INTEGRATION.md.
 *
 * Conventions
 *   - every pointer is caller-owned DEVICE memory unless the name ends in _host;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - calls only enqueue work (no device synchronisation) unless stated;
 *   - no hidden allocation: scratch comes from the caller via `ws`, sized by the
 *     matching *_workspace_bytes() query (yv_vit_create is the one allocator:
 *     it owns the packed bf16 weights);
 *   - return 0 on success, a negative YV_E* code on failure; yv_last_error()
 *     returns the thread-local message of the last failure.
 */
#ifndef YV8VIT_H_
#define YV8VIT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define YV_OK 0
#define YV_EINVAL (-22)   /* bad argument                                  */
#define YV_ENOMEM (-12)   /* workspace too small / allocation failed       */
#define YV_ECUDA (-5)     /* CUDA runtime / driver error (see last_error)  */
#define YV_ENODEV (-19)   /* no sm_100 device                              */

typedef void *yv_stream_t;

/* --- library ------------------------------------------------------------- */
int yv_version(void);                 /* 100*major + minor                   */
const char *yv_last_error(void);      /* thread-local, never NULL            */
/* 0 iff device `dev` has compute capability 10.x (the only target).        */
int yv_check_device(int dev);

/* --- a1: decode + confidence filter --------------------------------------
 * Replaces the detector wrapper's per-image post-processing loop:
 * docs/YOLO_TensorRT_Technical.md:243-267 (per-anchor max over class scores,
 * cxcywh -> xyxy), YOLOTensorRT_yolodet_py_解读.md:89-94 (un-letterbox,
 * `if score < 0.35: continue`).
 *   head   [B, 4+nc, A] f32, rows cx,cy,w,h then class probabilities
 *   keep iff !(score < conf); label = first arg-max; xyxy in fp32
 *   (x -/+ 0.5*w), then if unletterbox: (v - dwdh) / ratio.
 * Outputs, per image b, compacted in ASCENDING ANCHOR ORDER:
 *   boxes [B,cap,4] f32, scores [B,cap] f32, labels [B,cap] i32,
 *   anchors [B,cap] i32, counts [B] i32 = entries WRITTEN = min(found, cap).
 *   `cap` is a capacity bound, not a top-k: with cap < found the list is cut
 *   in anchor order (the lowest anchors stay), so size it for the worst case
 *   (cap = A never truncates).  A NaN class score wins the arg-max like
 *   torch.max (first NaN's index) and passes the filter (!(NaN < conf)).
 * One kernel launch (+ one memset of the per-image tile counters).          */
size_t yv_decode_filter_workspace_bytes(int B, int nc, int A);
int yv_decode_filter(const float *head, int B, int nc, int A, float conf,
                     int unletterbox, float ratio, float dw, float dh, int cap,
                     float *boxes, float *scores, int32_t *labels,
                     int32_t *anchors, int32_t *counts, void *ws,
                     size_t ws_bytes, yv_stream_t stream);

/* --- f1: the same, starting from the Detect head's RAW output ------------------
 * Fuses the head decode upstream of a1 (docs/YOLO_TensorRT_Technical.md:13-30
 * make_anchors, :71-77 DFL: softmax over 16 bins x arange, anchors -/+ distances,
 * x strides; YOLOTensorRT_yolodet_py_解读.md:64-92) into the filter kernel.
 *   raw [B, 64+nc, A] f32: 4 x 16 DFL bin logits, then nc class LOGITS, anchors of
 *   the 3 pyramid levels back to back (A = sum h_l*w_l), as the head's
 *   torch.cat([x_l.view(B, no, -1)], 2) lays them out.
 *   level_hw_host [3][2] (h, w) and level_stride_host [3] are HOST arrays.
 *   score = sigmoid(max logit), label = first arg-max, keep iff !(score < conf);
 *   box xyxy = ((gx+.5 - l) s, (gy+.5 - t) s, (gx+.5 + r) s, (gy+.5 + b) s).
 * The DFL rows are read for kept anchors only.  Outputs / workspace as a1
 * (yv_decode_filter_workspace_bytes(B, nc, A)).  Scores and boxes match the
 * torch fp32 reference to 1e-5 relative (expf), integer outputs exactly unless a
 * score lies within that distance of `conf`.                                    */
int yv_decode_filter_raw(const float *raw, int B, int nc,
                         const int32_t *level_hw_host,
                         const float *level_stride_host, float conf,
                         int unletterbox, float ratio, float dw, float dh,
                         int cap, float *boxes, float *scores, int32_t *labels,
                         int32_t *anchors, int32_t *counts, void *ws,
                         size_t ws_bytes, yv_stream_t stream);

/* --- f2: letterbox + blob of the detector input ---------------------------------
 * The wrapper's per-image host preprocessing (YOLOTensorRT_yolodet_py_解读.md:
 * 67-78): letterbox(bgr, (W, H)) -> cvtColor(BGR2RGB) -> blob (/255, CHW), for a
 * batch of equally sized images.  images [B,H,W,3] u8 (swap_rb = 1 for BGR input),
 * blob [B,3,out_h,out_w] f32.  The resize is OpenCV's 8-bit INTER_LINEAR fixed-point
 * arithmetic, bit-exact; padding value 114.  yv_letterbox_geometry (host, no
 * device work) returns the ratio / (dw, dh) the wrapper feeds to the un-letterbox
 * step, the resized size and its top-left offset.                                 */
int yv_letterbox_geometry(int H, int W, int out_h, int out_w, double *ratio,
                          double *dw, double *dh, int32_t *unpad_wh,
                          int32_t *left_top);
int yv_letterbox_blob(const uint8_t *images, int B, int H, int W, int swap_rb,
                      int out_h, int out_w, float *blob, yv_stream_t stream);

/* --- a2: custom_nms ---------------------------------------------------------
 * Replaces custom_nms(boxes, scores, iou_threshold) (README.md:62-83 =
 * tech.md:72-93) with torchvision.ops.box_iou's fp32 arithmetic, batched:
 * image b owns boxes[b, 0:n_b], n_b = counts ? min(counts[b], stride) : n_all.
 * Order: descending score, ties -> lower index first (stable).
 * Survivor rule: iou < thr (strict, fp32; NaN suppressed).  max_keep > 0 stops
 * after that many keeps (engine top-k, test.ipynb:22); <= 0 = unlimited.
 *   keep [B,stride] i32: ORIGINAL indices in descending-score order
 *   keep_counts [B] i32
 * Bit-exact against the reference for any input without NaN scores.
 * n_b <= 65536.  One kernel launch, one CTA per image.                      */
size_t yv_nms_workspace_bytes(int B, int stride);
int yv_nms_batched(const float *boxes, const float *scores,
                   const int32_t *counts, int B, int stride, int n_all,
                   float iou_thr, int max_keep, int32_t *keep,
                   int32_t *keep_counts, void *ws, size_t ws_bytes,
                   yv_stream_t stream);

/* yv_nms_batched (custom_nms rule) WITHOUT a caller workspace and without any
 * allocation: the kernel's scratch lives in a 4 MiB __device__ array inside the
 * library (held by every GPU that loads the library until exit), for callers
 * that cannot allocate -- code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc.  This entry point has NO counterpart in the
 * reference.
 * Limit: yv_nms_workspace_bytes(B, stride) <= YV_NMS_NOALLOC_BYTES, i.e. about
 * 65 000 boxes per call (B x stride x 56 B); larger calls return YV_EINVAL with a
 * message naming both sizes.  It calls no cudaMalloc / cudaFree / allocating
 * library routine (and, unlike the other entry points, is not seen by the launch
 * profiler, whose event pool may grow).  All calls on a device share the one
 * array: calls that may overlap on the GPU (different streams without an
 * ordering between them) RACE -- serialise them.  The array keeps the last
 * call's contents; the kernel writes every element before it reads it, so
 * nothing needs clearing.  Outputs and semantics are those of yv_nms_batched.   */
#define YV_NMS_NOALLOC_BYTES (4u << 20)
int yv_nms_batched_noalloc(const float *boxes, const float *scores,
                           const int32_t *counts, int B, int stride, int n_all,
                           float iou_thr, int max_keep, int32_t *keep,
                           int32_t *keep_counts, yv_stream_t stream);

/* Same kernel with the survivor rule selectable (row f4): YV_NMS_RULE_CUSTOM is
 * custom_nms (suppress unless iou < thr); YV_NMS_RULE_STANDARD is
 * torchvision.ops.nms / the engine's EfficientNMS plugin (suppress iff iou > thr;
 * docs/YOLO_TensorRT_Technical.md:108-150: iou .65, conf .25, top-100).  The
 * class-aware form offsets the boxes per label first (torchvision batched_nms).  */
#define YV_NMS_RULE_CUSTOM 0
#define YV_NMS_RULE_STANDARD 1
int yv_nms_batched_rule(const float *boxes, const float *scores,
                        const int32_t *counts, int B, int stride, int n_all,
                        float iou_thr, int max_keep, int rule, int32_t *keep,
                        int32_t *keep_counts, void *ws, size_t ws_bytes,
                        yv_stream_t stream);

/* --- a3: box rounding + dilation + clamp -----------------------------------
 * README.md:86-124 in the integer form of utils/trainClass.py:76-91.
 * d = ((max-min)*num)/den (floor), mode 1 ("infer_half") halves it again;
 * x0=max(0,x_min-dx) x1=min(W,x_max+dx), same in y; valid = x1>x0 && y1>y0.
 * boxes_i32 [n,4] -> out_boxes [n,4], out_valid [n].                        */
#define YV_DILATE_FULL 0
#define YV_DILATE_INFER_HALF 1
int yv_dilate_boxes(const int32_t *boxes_i32, int n, int W, int H, int num,
                    int den, int mode, int32_t *out_boxes, int32_t *out_valid,
                    yv_stream_t stream);

/* Gathers the kept detections of every image into dense per-image records and
 * a compacted crop list (the wrapper's per-detection loop,
 * YOLOTensorRT_yolodet_py_解读.md:93-100 + the authors' dilation step):
 * for image b, rank k < min(keep_counts[b], max_det):
 *   box = round_half_even(boxes[b, keep[b,k]]) -> int32 (D6, 解读.md:97),
 *   crop_box = dilate+clamp(box), valid as in yv_dilate_boxes.
 * det_* are [B,max_det,...]; det_crop [B,max_det] is the crop index or -1.
 * crop_box [crop_cap,4], crop_img [crop_cap] (image index), crop_det
 * [crop_cap] (b*max_det+k), n_crops [1] (total valid; entries beyond crop_cap
 * dropped).  Crops are ordered by (image, rank).                            */
int yv_build_crops(const float *boxes, const float *scores,
                   const int32_t *labels, const int32_t *anchors,
                   const int32_t *keep, const int32_t *keep_counts, int B,
                   int stride, int max_det, int W, int H, int num, int den,
                   int mode, int32_t *det_box, int32_t *det_crop_box,
                   float *det_score, int32_t *det_label, int32_t *det_anchor,
                   int32_t *det_count, int32_t *det_crop, int32_t *crop_box,
                   int32_t *crop_img, int32_t *crop_det, int crop_cap,
                   int32_t *n_crops, yv_stream_t stream);

/* --- a4: crop + resize + normalise -----------------------------------------
 * utils/trainClass.py:92 (PIL crop), app.py:39-42 (A.Resize + A.Normalize
 * mean=.5 std=.5), utils/trainClass.py:266 (HWC->CHW).
 *   images [B,H,W,3] u8 RGB; crop_box [n,4] i32 (x0,y0,x1,y1, inside the image,
 *   non-empty); crop_img [n] i32.
 *   interp 0 = nearest (cv2 INTER_NEAREST rule, fp64 index), 1 = bilinear
 *   (half-pixel centres); value = (p - 127.5) * (1/127.5) in fp32.
 *   layout 0: patch-major bf16 [n*(out/patch)^2, 3*patch*patch], column =
 *             c*patch^2 + iy*patch + ix (the ViT patch-embed GEMM operand);
 *   layout 1: the reference's fp32 CHW [n,3,out,out].
 * n_dev (optional, may be NULL) is a device int: only min(n, *n_dev) crops
 * are produced.  out is 224 in the reference; out % patch == 0, patch % 8 == 0. */
#define YV_INTERP_NEAREST 0
#define YV_INTERP_BILINEAR 1
#define YV_LAYOUT_PATCH_BF16 0
#define YV_LAYOUT_CHW_F32 1
int yv_crop_resize(const uint8_t *images, int B, int H, int W,
                   const int32_t *crop_box, const int32_t *crop_img, int n,
                   const int32_t *n_dev, int interp, int out_size, int patch,
                   int layout, void *out, yv_stream_t stream);

/* --- a5: ViT-B/{16,8} + Network_Wrapper head ------------------------------
 * utils/utils.py:59-87 (Network_Wrapper: timm ViT -> ReLU -> Linear(1000,128)
 * -> ReLU -> Linear(128,num_classes)); timm vit_base_patch{16,8}_224
 * architecture (pre-norm, LN eps 1e-6, GELU, class token, learned
 * pos-embed, final norm, class-token pooling).  bf16 operands, fp32
 * accumulation (tcgen05 / TMEM), fp32 LayerNorm and softmax statistics.
 * GELU: timm's nn.GELU is the erf form x * Phi(x); the fc1 epilogue
 * evaluates the TANH form 0.5 x (1 + tanh(sqrt(2/pi) (x + 0.044715 x^3)))
 * with tanh.approx.f32 -- a deliberate arithmetic deviation (|diff| <= 4.8e-4
 * absolute per activation, below the bf16 rounding of the stored value) that
 * is part of the 1e-2 logit tolerance, not a bit-exact restatement.          */
typedef struct yv_vit yv_vit_t;
typedef struct {
  int32_t img, patch, dim, depth, heads, mlp, head_out, hidden, num_classes;
  float eps;
} yv_vit_config;

int yv_vit_create(const yv_vit_config *cfg, yv_vit_t **out);
void yv_vit_destroy(yv_vit_t *m);
/* Copies one parameter (fp32, DEVICE pointer, `numel` elements) into the
 * model, converting to the packed bf16 / fp32 layout.  Names are the
 * Network_Wrapper.state_dict() keys ("model.blocks.3.attn.qkv.weight",
 * "fc.1.bias", ...).  Synchronous on `stream`.                               */
int yv_vit_set_param(yv_vit_t *m, const char *name, const float *data,
                     size_t numel, yv_stream_t stream);
/* number of parameters still unset (0 = ready) */
int yv_vit_missing_params(const yv_vit_t *m);
/* Classification only reads the class-token row of each crop (timm's token
 * pooling, reached through utils/utils.py:71-72), and after the last attention
 * no operation mixes rows: with enable != 0 the proj / residual / norm2 / MLP
 * of the LAST block run on the class rows alone.  Logits and top-1 are
 * bit-identical to the full computation (every GEMM row is independent);
 * 2.1 of the 35.1 GFLOP per ViT-B/16 crop are not executed.  Off by default:
 * the benchmarked forward computes every row of every block.  A per-model
 * flag read at the start of yv_vit_forward: set it before forwards are issued,
 * not concurrently with one (default stream / LayerNorm modes only; the fp32-
 * stream and LayerNorm-kernel A/B modes ignore it and compute every row).    */
int yv_vit_last_block_cls_rows(yv_vit_t *m, int enable);
size_t yv_vit_workspace_bytes(const yv_vit_t *m, int max_crops);
/* patches: layout-0 output of yv_crop_resize for n crops.
 * logits [n,num_classes] f32, top1 [n] i32 (first arg-max), hidden (optional,
 * may be NULL) [n,head_out] f32 backbone output.                            */
int yv_vit_forward(yv_vit_t *m, const void *patches_bf16, int n,
                   float *logits, int32_t *top1, float *hidden, void *ws,
                   size_t ws_bytes, yv_stream_t stream);

/* --- building blocks of a5, exported for the parity tests ------------------ */
/* D[M,N] = A[M,K] * W[N,K]^T (+bias[N]) (+GELU) (+residual[M,N]); A, W bf16
 * row-major; D bf16 (out_type 0), f32 (1) or f16 (2).  tcgen05.mma, TMEM
 * accumulators, TMA-fed.  K % 64 == 0, N % 8 == 0.  residual (or NULL) is bf16
 * [M,N] for bf16 / f32 outputs; for an f16 output it is f16 and MAY ALIAS D --
 * the ViT's residual-stream update x += A W^T + b, whose 32 x 32 boxes travel
 * by TMA in both directions (no GELU with an f16 output).                     */
#define YV_EPI_NONE 0
#define YV_EPI_GELU 1
#define YV_OUT_BF16 0
#define YV_OUT_F32 1
#define YV_OUT_F16 2
int yv_gemm_bf16(const void *A, const void *W, const float *bias,
                 const void *residual, void *D, int M, int N, int K, int epi,
                 int out_type, yv_stream_t stream);
/* y = LayerNorm(x) * g + b over the last dim; x,y bf16 [rows,dim]; fp32 stats */
int yv_layernorm_bf16(const void *x, const float *gamma, const float *beta,
                      void *y, int rows, int dim, float eps,
                      yv_stream_t stream);
/* D = [GELU](Linear(LayerNorm(x))) for an f16 x [M,K] (the ViT's residual
 * stream), W bf16 [N,K], D bf16 [M,N] -- the form the ViT blocks run
 * (timm vision_transformer Block: norm1 -> attn.qkv, norm2 -> mlp.fc1) with the
 * LayerNorm FOLDED into the GEMM: the tensor cores multiply the raw f16 rows by
 * fp16(g (.) W) and the epilogue applies rstd_m (acc - mu_m colsum_n) + c_n,
 * c = bias + beta . W^T.  Three launches: weight fold (the ViT does it once per
 * model), row statistics (reads x once, writes 8 B per row), GEMM.  ws holds
 * the folded weights and statistics: yv_linear_ln_f16_ws_bytes(M, N, K).
 * K % 256 == 0, 256 <= K <= 1024, N % 16 == 0.                                 */
size_t yv_linear_ln_f16_ws_bytes(int M, int N, int K);
int yv_linear_ln_f16(const void *x, const void *W, const float *bias,
                     const float *gamma, const float *beta, void *D, int M,
                     int N, int K, float eps, int epi, void *ws, size_t ws_bytes,
                     yv_stream_t stream);
/* x += A W^T + bias on the f16 residual stream x [M,N] IN PLACE (A bf16 [M,K], W
 * bf16 [N,K]) -- attn.proj / mlp.fc2 of a timm Block followed by its residual
 * add -- and, when stats != NULL, the (mean, 1/sqrt(var + eps)) pair of every
 * UPDATED row into stats [M,2] f32 as a by-product of the GEMM epilogue (partial
 * sums per 64 columns in ws, then a combine kernel), i.e. the statistics the next
 * yv_linear_ln_f16-style GEMM needs without another pass over x.
 * N % 64 == 0, N <= 1024; ws: yv_residual_linear_f16_ws_bytes(M, N).           */
size_t yv_residual_linear_f16_ws_bytes(int M, int N);
int yv_residual_linear_f16(const void *A, const void *W, const float *bias,
                           void *x, int M, int N, int K, float eps, float *stats,
                           void *ws, size_t ws_bytes, yv_stream_t stream);
/* Scheduling knob: the number of SMs a GEMM launch / a ViT-B/16 attention launch
 * may occupy (persistent grids; 0 = all SMs, the default).  With budgets that add
 * up to the SM count a GEMM and an attention launch issued on two streams run side
 * by side on disjoint SMs.  Process-wide, not thread-safe against running
 * launches: set it between forwards.  GEMM budget must be even (CTA pairs).     */
int yv_tune_sm_budget(int gemm_sms, int attention_sms);
/* softmax(Q K^T / sqrt(dh)) V for qkv [n, T, 3, heads, dh] bf16 -> out
 * [n, T, heads*dh] bf16 (timm's qkv packing).                               */
int yv_attention_bf16(const void *qkv, void *out, int n, int T, int heads,
                      int dh, yv_stream_t stream);

/* --- K9: result records ------------------------------------------------------
 * One fixed-size record per (image, rank): 8 x f32
 *   x0,y0,x1,y1 (rounded box), cls (ViT top-1, or -1), score, anchor, valid.
 * This is the buffer the final NCCL gather ships and the
 * source of main()'s (name, cls, conf, xmin, ymin, xmax, ymax) tuples
 * (YOLOTensorRT_yolodet_py_解读.md:103-116).                                  */
int yv_pack_results(const int32_t *det_box, const float *det_score,
                    const int32_t *det_anchor, const int32_t *det_count,
                    const int32_t *det_crop, const int32_t *top1, int B,
                    int max_det, float *records, yv_stream_t stream);

/* --- measurement hooks (bench.py) -------------------------------------------
 * The library counts every kernel launch per kernel class.  With profiling on,
 * each launch is additionally bracketed by CUDA events on its own stream, so
 * bench.py can report per-kernel device time, launches and algorithmic work
 * (bytes for the HBM-bound kernels, flops for the GEMM) for the roofline.     */
int yv_prof_enable(int on);                 /* resets the counters              */
int yv_prof_kinds(void);
const char *yv_prof_kind_name(int kind);
int yv_prof_read(double *ms, long long *launches, double *work); /* syncs      */
/* Debug: per-tile clock64() stamps of CTA 0 of the pipelined attention kernel
 * (8 slots per tile, first 64 tiles).  enable=1 arms the next launches;
 * host_out (HOST pointer, n <= 512 entries) receives the stamps; enable=0 frees. */
int yv_debug_attn_trace(int enable, unsigned long long *host_out, int n);

#ifdef __cplusplus
}
#endif
#endif /* YV8VIT_H_ */
