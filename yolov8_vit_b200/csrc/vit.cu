// vit.cu -- a5: ViT-B/{16,8} + Network_Wrapper head forward, composed from the tcgen05 GEMM
// (gemm_sm100.cu) and the kernels in vit_ops.cu / attention.cu.
//
// Reference: utils/utils.py:59-87 (Network_Wrapper, build_model); the backbone arithmetic is
// timm's vit_base_patch{16,8}_224 (third-party, unpinned, restated from its published
// architecture: pre-norm blocks, LayerNorm eps 1e-6, GELU (evaluated in its tanh form, see gemm_sm100.cu), class token, learned position
// embedding, final norm, class-token pooling, linear head).
//
// Data layout in HBM (per pass of Cc crops, M = Cc * T token rows):
//   x    [M, dim]   f16   residual stream.  fp16, not bf16: 11 significant bits, so the 24 residual adds of the
//                         stack stay below the bf16 rounding of the GEMM operands (saturating conversion, no inf);
//                         proj / fc2 update it IN PLACE in their epilogue (x += y W^T + b, TMA in and out)
//   y    [M, dim]   bf16  LayerNorm output / attention output (GEMM A operands)
//   qkv  [M, 3 dim] bf16  timm packing (q | k | v, head-major)
//   h    [M, mlp]   bf16  GELU(fc1)
// = 18 B per stream element per pass (2.8 KB... per token: 13.8 KB), LayerNorm moves 4 B per element.
// YV_VIT_STREAM=f32 selects the round-1 layout instead (A/B and high-precision mode): x32 [M, dim] f32 stream,
// proj / fc2 write bf16 deltas d / d2 which the NEXT LayerNorm folds into x32 (22 B per element per block).
// Weights are packed once to bf16 [out, in] (K-major = the UMMA B operand as stored by torch).
#include "common.cuh"

#include <cstdlib>
#include "gemm_sm100.cuh"
#include "vit_ops.cuh"

#include <string>
#include <unordered_map>
#include <vector>

namespace {

struct Param {
  size_t off;       // element offset into the bf16 or f32 arena
  size_t numel;
  bool bf16;
  bool set;
  bool derived = false;   // computed by finalize() from other parameters (folded LayerNorm weights)
};

__global__ void f32_to_bf16_kernel(const float *__restrict__ in, __nv_bfloat16 *__restrict__ out, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t step = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += step) out[i] = __float2bfloat16_rn(in[i]);
}

// pos_table[0] = cls + pos[0]; pos_table[t] = pos[t] + patch_bias  (t >= 1)
__global__ void pos_table_kernel(const float *__restrict__ pos, const float *__restrict__ cls,
                                 const float *__restrict__ pbias, float *__restrict__ table, int T, int dim) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= T * dim) return;
  const int t = i / dim, e = i % dim;
  table[i] = pos[i] + (t == 0 ? cls[e] : pbias[e]);
}

__global__ void gather_rows_f32_kernel(const float *__restrict__ in, size_t in_stride, float *__restrict__ out,
                                       int rows, int dim) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * dim) return;
  out[i] = in[(size_t)(i / dim) * in_stride + (i % dim)];
}

}  // namespace

struct yv_vit {
  yv_vit_config cfg;
  int T;                      // tokens per crop
  int pdim;                   // 3 * patch * patch
  std::unordered_map<std::string, Param> params;
  __nv_bfloat16 *wb = nullptr;   // bf16 arena
  float *wf = nullptr;           // f32 arena
  size_t nb = 0, nf = 0;
  size_t pos_table_off = 0;      // f32 arena: [T, dim]
  bool finalized = false;
  bool last_block_cls = false;   // yv_vit_last_block_cls_rows: proj / MLP of the last block on the class-token rows only

  const __nv_bfloat16 *B(const std::string &k) const { return wb + params.at(k).off; }
  const float *F(const std::string &k) const { return wf + params.at(k).off; }
};

namespace {

void add_param(yv_vit *m, const std::string &name, size_t numel, bool bf16, bool derived = false) {
  Param p;
  p.numel = numel;
  p.bf16 = bf16;
  p.set = false;
  p.derived = derived;
  if (bf16) { p.off = m->nb; m->nb += (numel + 63) / 64 * 64; }      // 128-byte aligned rows for TMA
  else { p.off = m->nf; m->nf += (numel + 63) / 64 * 64; }
  m->params[name] = p;
}

struct VitWs {
  void *x;                      // residual stream: f16 [M, dim] (default) or f32
  __nv_bfloat16 *y, *d, *d2, *qkv, *h, *cls_ln;      // d / d2: fp32-stream mode only
  float *hidden;
  float *stats;                 // folded-LayerNorm mode: (mean, rstd) per token row
  float *part;                  // ... and the partial statistics of the proj / fc2 epilogues, [dim / 64][M] float2
  size_t total;
};

// residual stream precision (read once per process)
bool vit_stream_f32() {
  static const bool f32 = []() { const char *e = getenv("YV_VIT_STREAM"); return e && (e[0] == 'f' || e[0] == 'F') && e[1] == '3'; }();
  return f32;
}

// LayerNorm folded into the qkv / fc1 GEMMs (default for the fp16 stream; YV_VIT_LNFOLD=0 runs the LayerNorm kernels)
bool vit_lnfold() {
  static const bool on = []() { const char *e = getenv("YV_VIT_LNFOLD"); return !(e && e[0] == '0'); }();
  return on && !vit_stream_f32();
}

// Row statistics of the folded LayerNorms: by default they come out of the proj / fc2 GEMM epilogues (partial sums per
// 64-column group + a tiny combine kernel); YV_VIT_STATS=kernel runs the stream-reading statistics kernel instead (A/B).
bool vit_stats_fused() {
  static const bool on = []() { const char *e = getenv("YV_VIT_STATS"); return !(e && (e[0] == 'k' || e[0] == 'K')); }();
  return on;
}

VitWs vit_ws(const yv_vit *m, void *base, int chunk) {
  const yv_vit_config &c = m->cfg;
  const size_t M = (size_t)chunk * m->T;
  const bool f32 = vit_stream_f32();
  size_t o = 0;
  char *b = (char *)base;
  VitWs w;
  w.x = (void *)(b + o);               o = yv_align_up(o + M * c.dim * (f32 ? 4 : 2), 1024);
  w.y = (__nv_bfloat16 *)(b + o);      o = yv_align_up(o + M * c.dim * 2, 1024);
  w.d = w.d2 = nullptr;
  if (f32) {
    w.d = (__nv_bfloat16 *)(b + o);    o = yv_align_up(o + M * c.dim * 2, 1024);
    w.d2 = (__nv_bfloat16 *)(b + o);   o = yv_align_up(o + M * c.dim * 2, 1024);
  }
  w.qkv = (__nv_bfloat16 *)(b + o);    o = yv_align_up(o + M * 3 * c.dim * 2, 1024);
  w.h = (__nv_bfloat16 *)(b + o);      o = yv_align_up(o + M * c.mlp * 2, 1024);
  w.cls_ln = (__nv_bfloat16 *)(b + o); o = yv_align_up(o + (size_t)chunk * c.dim * 2, 1024);
  w.hidden = (float *)(b + o);         o = yv_align_up(o + (size_t)chunk * c.head_out * 4, 1024);
  w.stats = (float *)(b + o);          o = yv_align_up(o + M * 2 * 4, 1024);
  w.part = (float *)(b + o);           o = yv_align_up(o + M * (size_t)(c.dim / 64) * 2 * 4, 1024);
  w.total = o;
  return w;
}

// Largest number of crops per pass of the block stack (2.7 MB of workspace per crop for ViT-B/16).  One pass over
// 1072 crops instead of 4 x 256 + 48 measured +1.5 % on the whole step (450 -> 90 launches, no 48-crop remainder
// pass); passes are balanced when n exceeds the limit.  YV_VIT_CHUNK overrides (timing).
int vit_default_chunk() {
  static const int c = []() { const char *e = getenv("YV_VIT_CHUNK"); int v = e ? atoi(e) : 0; return v > 0 ? v : 2048; }();
  return c;
}
#define VIT_DEFAULT_CHUNK vit_default_chunk()

int finalize(yv_vit *m, cudaStream_t s) {
  if (m->finalized) return YV_OK;
  const int miss = yv_vit_missing_params(m);
  if (miss) {
    yv_set_error("yv_vit_forward: %d parameters were never set", miss);
    return YV_EINVAL;
  }
  const int n = m->T * m->cfg.dim;
  YvProf prof(YV_K_MISC, s, 0.0);
  pos_table_kernel<<<(n + 255) / 256, 256, 0, s>>>(m->F("model.pos_embed"), m->F("model.cls_token"),
                                                   m->F("model.patch_embed.proj.bias"), m->wf + m->pos_table_off,
                                                   m->T, m->cfg.dim);
  YV_LAUNCH_CHECK();
  if (vit_lnfold()) {
    const int D = m->cfg.dim;
    for (int i = 0; i < m->cfg.depth; ++i) {
      const std::string p = "model.blocks." + std::to_string(i) + ".";
      int rc;
      if ((rc = yv_fold_ln_launch(m->B(p + "attn.qkv.weight"), m->F(p + "attn.qkv.bias"), m->F(p + "norm1.weight"), m->F(p + "norm1.bias"),
                                  3 * D, D, m->wb + m->params.at(p + "norm1.fold.w").off, m->wf + m->params.at(p + "norm1.fold.s").off,
                                  m->wf + m->params.at(p + "norm1.fold.c").off, s))) return rc;
      if ((rc = yv_fold_ln_launch(m->B(p + "mlp.fc1.weight"), m->F(p + "mlp.fc1.bias"), m->F(p + "norm2.weight"), m->F(p + "norm2.bias"),
                                  m->cfg.mlp, D, m->wb + m->params.at(p + "norm2.fold.w").off, m->wf + m->params.at(p + "norm2.fold.s").off,
                                  m->wf + m->params.at(p + "norm2.fold.c").off, s))) return rc;
    }
  }
  m->finalized = true;
  return YV_OK;
}

}  // namespace

YV_EXPORT int yv_vit_create(const yv_vit_config *cfg, yv_vit_t **out) {
  YV_REQUIRE(cfg && out, "yv_vit_create: null pointer");
  const yv_vit_config &c = *cfg;
  YV_REQUIRE(c.img > 0 && c.patch >= 8 && c.patch % 8 == 0 && c.img % c.patch == 0, "yv_vit_create: bad img/patch %d/%d", c.img, c.patch);
  YV_REQUIRE(c.dim % 256 == 0 && c.dim <= 1024 && c.heads > 0 && c.dim == c.heads * 64, "yv_vit_create: dim=%d heads=%d (need dim %% 256 == 0, head dim 64)", c.dim, c.heads);
  YV_REQUIRE(c.mlp % 64 == 0 && c.head_out % 8 == 0 && c.hidden >= 1 && c.num_classes >= 1 && c.depth >= 1, "yv_vit_create: bad mlp/head sizes");
  YV_REQUIRE((3 * c.patch * c.patch) % 64 == 0, "yv_vit_create: 3*patch^2 must be a multiple of 64");
  yv_vit *m = new yv_vit();
  m->cfg = c;
  const int g = c.img / c.patch;
  m->T = g * g + 1;
  m->pdim = 3 * c.patch * c.patch;
  const size_t D = c.dim;
  add_param(m, "model.cls_token", D, false);
  add_param(m, "model.pos_embed", (size_t)m->T * D, false);
  add_param(m, "model.patch_embed.proj.weight", D * m->pdim, true);
  add_param(m, "model.patch_embed.proj.bias", D, false);
  for (int i = 0; i < c.depth; ++i) {
    const std::string p = "model.blocks." + std::to_string(i) + ".";
    add_param(m, p + "norm1.weight", D, false);
    add_param(m, p + "norm1.bias", D, false);
    add_param(m, p + "attn.qkv.weight", 3 * D * D, true);
    add_param(m, p + "attn.qkv.bias", 3 * D, false);
    add_param(m, p + "attn.proj.weight", D * D, true);
    add_param(m, p + "attn.proj.bias", D, false);
    add_param(m, p + "norm2.weight", D, false);
    add_param(m, p + "norm2.bias", D, false);
    add_param(m, p + "mlp.fc1.weight", (size_t)c.mlp * D, true);
    add_param(m, p + "mlp.fc1.bias", c.mlp, false);
    add_param(m, p + "mlp.fc2.weight", D * c.mlp, true);
    add_param(m, p + "mlp.fc2.bias", D, false);
    if (vit_lnfold()) {            // fp16 g (.) W (kept in the 16-bit arena), column sums, folded bias: finalize()
      add_param(m, p + "norm1.fold.w", 3 * D * D, true, true);
      add_param(m, p + "norm1.fold.s", 3 * D, false, true);
      add_param(m, p + "norm1.fold.c", 3 * D, false, true);
      add_param(m, p + "norm2.fold.w", (size_t)c.mlp * D, true, true);
      add_param(m, p + "norm2.fold.s", c.mlp, false, true);
      add_param(m, p + "norm2.fold.c", c.mlp, false, true);
    }
  }
  add_param(m, "model.norm.weight", D, false);
  add_param(m, "model.norm.bias", D, false);
  add_param(m, "model.head.weight", (size_t)c.head_out * D, true);
  add_param(m, "model.head.bias", c.head_out, false);
  add_param(m, "fc.1.weight", (size_t)c.hidden * c.head_out, false);
  add_param(m, "fc.1.bias", c.hidden, false);
  add_param(m, "fc.3.weight", (size_t)c.num_classes * c.hidden, false);
  add_param(m, "fc.3.bias", c.num_classes, false);
  m->pos_table_off = m->nf;
  m->nf += ((size_t)m->T * D + 63) / 64 * 64;
  cudaError_t e = cudaMalloc(&m->wb, m->nb * sizeof(__nv_bfloat16));
  if (e == cudaSuccess) e = cudaMalloc(&m->wf, m->nf * sizeof(float));
  if (e != cudaSuccess) {
    if (m->wb) cudaFree(m->wb);
    delete m;
    yv_cuda_fail(e, "cudaMalloc(ViT weights)", __FILE__, __LINE__);
    return YV_ENOMEM;
  }
  *out = m;
  return YV_OK;
}

YV_EXPORT void yv_vit_destroy(yv_vit_t *m) {
  if (!m) return;
  cudaFree(m->wb);
  cudaFree(m->wf);
  delete m;
}

YV_EXPORT int yv_vit_set_param(yv_vit_t *m, const char *name, const float *data, size_t numel, yv_stream_t stream) {
  YV_REQUIRE(m && name && data, "yv_vit_set_param: null pointer");
  auto it = m->params.find(name);
  YV_REQUIRE(it != m->params.end(), "yv_vit_set_param: unknown parameter '%s'", name);
  Param &p = it->second;
  YV_REQUIRE(!p.derived, "yv_vit_set_param: '%s' is derived, not settable", name);
  YV_REQUIRE(p.numel == numel, "yv_vit_set_param: '%s' has %zu elements, expected %zu", name, numel, p.numel);
  cudaStream_t s = (cudaStream_t)stream;
  if (p.bf16) {
    const int grid = (int)((numel + 255) / 256 < 4096 ? (numel + 255) / 256 : 4096);
    f32_to_bf16_kernel<<<grid, 256, 0, s>>>(data, m->wb + p.off, numel);
    YV_LAUNCH_CHECK();
  } else {
    YV_CUDA(cudaMemcpyAsync(m->wf + p.off, data, numel * sizeof(float), cudaMemcpyDeviceToDevice, s));
  }
  YV_CUDA(cudaStreamSynchronize(s));
  p.set = true;
  m->finalized = false;
  return YV_OK;
}

YV_EXPORT int yv_vit_last_block_cls_rows(yv_vit_t *m, int enable) {
  YV_REQUIRE(m, "yv_vit_last_block_cls_rows: null model");
  m->last_block_cls = enable != 0;
  return YV_OK;
}

YV_EXPORT int yv_vit_missing_params(const yv_vit_t *m) {
  if (!m) return -1;
  int miss = 0;
  for (const auto &kv : m->params) miss += (kv.second.set || kv.second.derived) ? 0 : 1;
  return miss;
}

YV_EXPORT size_t yv_vit_workspace_bytes(const yv_vit_t *m, int max_crops) {
  if (!m || max_crops <= 0) return 0;
  const int chunk = max_crops < VIT_DEFAULT_CHUNK ? max_crops : VIT_DEFAULT_CHUNK;
  return vit_ws(m, nullptr, chunk).total;
}

YV_EXPORT int yv_vit_forward(yv_vit_t *m, const void *patches_bf16, int n, float *logits, int32_t *top1,
                             float *hidden, void *ws, size_t ws_bytes, yv_stream_t stream) {
  YV_REQUIRE(m && n >= 0, "yv_vit_forward: bad argument");
  if (n == 0) return YV_OK;
  YV_REQUIRE(patches_bf16 && logits && top1 && ws, "yv_vit_forward: null pointer");
  cudaStream_t s = (cudaStream_t)stream;
  int rc = finalize(m, s);
  if (rc) return rc;
  const yv_vit_config &c = m->cfg;
  // largest chunk the workspace allows
  int chunk = n < VIT_DEFAULT_CHUNK ? n : VIT_DEFAULT_CHUNK;
  while (chunk > 1 && vit_ws(m, nullptr, chunk).total > ws_bytes) chunk = (chunk + 1) / 2;
  if (vit_ws(m, nullptr, chunk).total > ws_bytes) {
    yv_set_error("yv_vit_forward: workspace %zu too small (one crop needs %zu)", ws_bytes, vit_ws(m, nullptr, 1).total);
    return YV_ENOMEM;
  }
  chunk = (n + (n + chunk - 1) / chunk - 1) / ((n + chunk - 1) / chunk);      // balanced passes
  const VitWs w = vit_ws(m, ws, chunk);
  const int T = m->T, D = c.dim, NP = T - 1;
  const float *pos_table = m->wf + m->pos_table_off;

  const bool f32 = vit_stream_f32();
  const int XT = f32 ? 1 : 2;                            // yv_ln_launch element type of the stream
  for (int c0 = 0; c0 < n; c0 += chunk) {
    const int cc = (n - c0) < chunk ? (n - c0) : chunk;
    const int M = cc * T;
    const __nv_bfloat16 *patches = (const __nv_bfloat16 *)patches_bf16 + (size_t)c0 * NP * m->pdim;
    GemmEpilogue ep;
    // patch embedding: [cc*NP, pdim] x [D, pdim]^T -> token rows (+ position, + bias) of the residual stream
    ep = GemmEpilogue{w.x, nullptr, nullptr, pos_table, NP, false, f32 ? 1 : 2, 0};
    if ((rc = yv_gemm_launch(patches, m->B("model.patch_embed.proj.weight"), cc * NP, D, m->pdim, ep, s))) return rc;
    if ((rc = yv_cls_rows_launch(w.x, !f32, pos_table, cc, T, D, s))) return rc;
    if (!f32) {
      // fp16 residual stream, updated in place by the proj / fc2 epilogues (residual boxes in and sums out by TMA).
      // Default: the two LayerNorms of a block are FOLDED into the qkv / fc1 GEMMs -- the GEMM reads the fp16 stream
      // itself against fp16 g (.) W and its epilogue applies rstd (acc - mu colsum) + c; a statistics kernel reads the
      // stream once (2 B per element, no write).  YV_VIT_LNFOLD=0: LayerNorm kernels (2 B in + 2 B out) into w.y.
      const bool fold = vit_lnfold();
      const bool fused = fold && vit_stats_fused();          // statistics out of the proj / fc2 epilogues
      const bool cls_tail = m->last_block_cls && fused && c.depth >= 1;
      const __nv_bfloat16 *x16 = (const __nv_bfloat16 *)w.x;      // 16-bit rows; the instruction descriptor says f16
      for (int i = 0; i < c.depth; ++i) {
        const std::string p = "model.blocks." + std::to_string(i) + ".";
        if (fold) {
          // block 0 reads the rows the patch embedding wrote; later blocks got their statistics from fc2's epilogue
          if (!fused || i == 0) { if ((rc = yv_row_stats_launch(w.x, D, w.stats, M, D, c.eps, s))) return rc; }
          ep = GemmEpilogue{w.qkv, m->F(p + "norm1.fold.c"), nullptr, nullptr, 0, false, 0, 0, w.stats, m->F(p + "norm1.fold.s")};
          if ((rc = yv_gemm_launch(x16, m->B(p + "norm1.fold.w"), M, 3 * D, D, ep, s))) return rc;
        } else {
          if ((rc = yv_ln_launch(w.x, XT, D, nullptr, 0, false, m->F(p + "norm1.weight"), m->F(p + "norm1.bias"), w.y, M, D, c.eps, s))) return rc;
          ep = GemmEpilogue{w.qkv, m->F(p + "attn.qkv.bias"), nullptr, nullptr, 0, false, 0, 0};
          if ((rc = yv_gemm_launch(w.y, m->B(p + "attn.qkv.weight"), M, 3 * D, D, ep, s))) return rc;
        }
        if ((rc = yv_attention_launch(w.qkv, w.y, cc, T, c.heads, D / c.heads, s))) return rc;
        if (cls_tail && i + 1 == c.depth) {
          // Only the class-token row of a crop reaches the final norm and the heads, and after the last attention no
          // operation mixes rows any more: proj, the residual adds, norm2 and the MLP of the LAST block run on the cc class
          // rows alone (gathered into the dead qkv buffer).  Every row of a GEMM is computed independently of the others
          // (same K order, same epilogue), so the logits are bit-identical to the full computation (tested).
          char *tb = (char *)w.qkv;
          void *x_cls = tb;                                  tb += yv_align_up((size_t)cc * D * 2, 1024);
          __nv_bfloat16 *y_cls = (__nv_bfloat16 *)tb;        tb += yv_align_up((size_t)cc * D * 2, 1024);
          __nv_bfloat16 *h_cls = (__nv_bfloat16 *)tb;
          YV_CUDA(cudaMemcpy2DAsync(x_cls, (size_t)D * 2, w.x, (size_t)T * D * 2, (size_t)D * 2, cc, cudaMemcpyDeviceToDevice, s));
          YV_CUDA(cudaMemcpy2DAsync(y_cls, (size_t)D * 2, w.y, (size_t)T * D * 2, (size_t)D * 2, cc, cudaMemcpyDeviceToDevice, s));
          ep = GemmEpilogue{x_cls, m->F(p + "attn.proj.bias"), x_cls, nullptr, 0, false, 2, 2};
          ep.stats_part = w.part;
          if ((rc = yv_gemm_launch(y_cls, m->B(p + "attn.proj.weight"), cc, D, D, ep, s))) return rc;
          if ((rc = yv_stats_finalize_launch(w.part, cc, D, c.eps, w.stats, s))) return rc;
          ep = GemmEpilogue{h_cls, m->F(p + "norm2.fold.c"), nullptr, nullptr, 0, true, 0, 0, w.stats, m->F(p + "norm2.fold.s")};
          if ((rc = yv_gemm_launch((const __nv_bfloat16 *)x_cls, m->B(p + "norm2.fold.w"), cc, c.mlp, D, ep, s))) return rc;
          ep = GemmEpilogue{x_cls, m->F(p + "mlp.fc2.bias"), x_cls, nullptr, 0, false, 2, 2};
          if ((rc = yv_gemm_launch(h_cls, m->B(p + "mlp.fc2.weight"), cc, D, c.mlp, ep, s))) return rc;
          if ((rc = yv_ln_launch(x_cls, XT, (size_t)D, nullptr, 0, false, m->F("model.norm.weight"), m->F("model.norm.bias"), w.cls_ln, cc, D, c.eps, s))) return rc;
          break;
        }
        ep = GemmEpilogue{w.x, m->F(p + "attn.proj.bias"), w.x, nullptr, 0, false, 2, 2};
        if (fused) ep.stats_part = w.part;
        if ((rc = yv_gemm_launch(w.y, m->B(p + "attn.proj.weight"), M, D, D, ep, s))) return rc;
        if (fold) {
          if (fused) { if ((rc = yv_stats_finalize_launch(w.part, M, D, c.eps, w.stats, s))) return rc; }
          else if ((rc = yv_row_stats_launch(w.x, D, w.stats, M, D, c.eps, s))) return rc;
          ep = GemmEpilogue{w.h, m->F(p + "norm2.fold.c"), nullptr, nullptr, 0, true, 0, 0, w.stats, m->F(p + "norm2.fold.s")};
          if ((rc = yv_gemm_launch(x16, m->B(p + "norm2.fold.w"), M, c.mlp, D, ep, s))) return rc;
        } else {
          if ((rc = yv_ln_launch(w.x, XT, D, nullptr, 0, false, m->F(p + "norm2.weight"), m->F(p + "norm2.bias"), w.y, M, D, c.eps, s))) return rc;
          ep = GemmEpilogue{w.h, m->F(p + "mlp.fc1.bias"), nullptr, nullptr, 0, true, 0, 0};
          if ((rc = yv_gemm_launch(w.y, m->B(p + "mlp.fc1.weight"), M, c.mlp, D, ep, s))) return rc;
        }
        ep = GemmEpilogue{w.x, m->F(p + "mlp.fc2.bias"), w.x, nullptr, 0, false, 2, 2};
        const bool next = fused && i + 1 < c.depth;          // the last block feeds the class-row LayerNorm instead
        if (next) ep.stats_part = w.part;
        if ((rc = yv_gemm_launch(w.h, m->B(p + "mlp.fc2.weight"), M, D, c.mlp, ep, s))) return rc;
        if (next && (rc = yv_stats_finalize_launch(w.part, M, D, c.eps, w.stats, s))) return rc;
      }
      // final norm on the class-token rows only, then the heads
      if (!cls_tail && (rc = yv_ln_launch(w.x, XT, (size_t)T * D, nullptr, 0, false, m->F("model.norm.weight"), m->F("model.norm.bias"), w.cls_ln, cc, D, c.eps, s))) return rc;
    } else {
      // fp32 residual stream (round-1 layout): the residual adds are fused into the LayerNorms, so proj / fc2 store
      // plain bf16 deltas.  The fp32 stream is rewritten ONCE per block: norm2 normalises x32 + d_attn on the fly
      // without writing it back, and the next norm1 folds (x32 + d_attn) + d_mlp in and writes x32 -- the same fp32
      // sums in the same order as two separate updates.
      bool have = false;                                 // deltas of the previous block pending?
      for (int i = 0; i < c.depth; ++i) {
        const std::string p = "model.blocks." + std::to_string(i) + ".";
        if ((rc = yv_ln2_launch(w.x, XT, D, have ? w.d : nullptr, have ? w.d2 : nullptr, D, true, m->F(p + "norm1.weight"), m->F(p + "norm1.bias"), w.y, M, D, c.eps, s))) return rc;
        ep = GemmEpilogue{w.qkv, m->F(p + "attn.qkv.bias"), nullptr, nullptr, 0, false, 0, 0};
        if ((rc = yv_gemm_launch(w.y, m->B(p + "attn.qkv.weight"), M, 3 * D, D, ep, s))) return rc;
        if ((rc = yv_attention_launch(w.qkv, w.y, cc, T, c.heads, D / c.heads, s))) return rc;
        ep = GemmEpilogue{w.d, m->F(p + "attn.proj.bias"), nullptr, nullptr, 0, false, 0, 0};
        if ((rc = yv_gemm_launch(w.y, m->B(p + "attn.proj.weight"), M, D, D, ep, s))) return rc;
        if ((rc = yv_ln_launch(w.x, XT, D, w.d, D, false, m->F(p + "norm2.weight"), m->F(p + "norm2.bias"), w.y, M, D, c.eps, s))) return rc;
        ep = GemmEpilogue{w.h, m->F(p + "mlp.fc1.bias"), nullptr, nullptr, 0, true, 0, 0};
        if ((rc = yv_gemm_launch(w.y, m->B(p + "mlp.fc1.weight"), M, c.mlp, D, ep, s))) return rc;
        ep = GemmEpilogue{w.d2, m->F(p + "mlp.fc2.bias"), nullptr, nullptr, 0, false, 0, 0};
        if ((rc = yv_gemm_launch(w.h, m->B(p + "mlp.fc2.weight"), M, D, c.mlp, ep, s))) return rc;
        have = true;
      }
      // final norm on the class-token rows only (+ both deltas of the last block), 1000-way head, wrapper MLP + argmax
      if ((rc = yv_ln2_launch(w.x, XT, (size_t)T * D, w.d, w.d2, (size_t)T * D, false, m->F("model.norm.weight"), m->F("model.norm.bias"), w.cls_ln, cc, D, c.eps, s))) return rc;
    }
    ep = GemmEpilogue{w.hidden, m->F("model.head.bias"), nullptr, nullptr, 0, false, 1, 0};
    if ((rc = yv_gemm_launch(w.cls_ln, m->B("model.head.weight"), cc, c.head_out, D, ep, s))) return rc;
    if ((rc = yv_wrapper_head_launch(w.hidden, c.head_out, m->F("fc.1.weight"), m->F("fc.1.bias"), m->F("fc.3.weight"),
                                     m->F("fc.3.bias"), cc, c.head_out, c.hidden, c.num_classes,
                                     logits + (size_t)c0 * c.num_classes, top1 + c0, s)))
      return rc;
    if (hidden) {
      const int tot = cc * c.head_out;
      YvProf prof(YV_K_MISC, s, 0.0);
      gather_rows_f32_kernel<<<(tot + 255) / 256, 256, 0, s>>>(w.hidden, c.head_out, hidden + (size_t)c0 * c.head_out, cc, c.head_out);
      YV_LAUNCH_CHECK();
    }
  }
  return YV_OK;
}

// LayerNorm -> Linear as one operator (the unit the ViT blocks are built from); see include/yv8vit.h.
YV_EXPORT size_t yv_linear_ln_f16_ws_bytes(int M, int N, int K) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  return yv_align_up((size_t)N * K * 2, 1024) + 2 * yv_align_up((size_t)N * 4, 1024) + yv_align_up((size_t)M * 8, 1024);
}

YV_EXPORT int yv_linear_ln_f16(const void *x, const void *W, const float *bias, const float *gamma, const float *beta,
                               void *D, int M, int N, int K, float eps, int epi, void *ws, size_t ws_bytes,
                               yv_stream_t stream) {
  YV_REQUIRE(x && W && bias && gamma && beta && D && ws, "yv_linear_ln_f16: null pointer");
  YV_REQUIRE(epi == YV_EPI_NONE || epi == YV_EPI_GELU, "yv_linear_ln_f16: unknown epilogue %d", epi);
  YV_REQUIRE(M >= 1 && N >= 1 && ws_bytes >= yv_linear_ln_f16_ws_bytes(M, N, K), "yv_linear_ln_f16: workspace too small");
  cudaStream_t s = (cudaStream_t)stream;
  char *b = (char *)ws;
  void *wf16 = b;                      b += yv_align_up((size_t)N * K * 2, 1024);
  float *colsum = (float *)b;          b += yv_align_up((size_t)N * 4, 1024);
  float *cf = (float *)b;              b += yv_align_up((size_t)N * 4, 1024);
  float *stats = (float *)b;
  int rc;
  if ((rc = yv_fold_ln_launch((const __nv_bfloat16 *)W, bias, gamma, beta, N, K, wf16, colsum, cf, s))) return rc;
  if ((rc = yv_row_stats_launch(x, (size_t)K, stats, M, K, eps, s))) return rc;
  GemmEpilogue ep{D, cf, nullptr, nullptr, 0, epi == YV_EPI_GELU, 0, 0, stats, colsum};
  return yv_gemm_launch((const __nv_bfloat16 *)x, (const __nv_bfloat16 *)wf16, M, N, K, ep, s);
}

// x += A W^T + b on an fp16 stream x [M,N] (in place) with the row statistics of the UPDATED rows as a by-product of
// the GEMM epilogue -- the other half of a ViT block next to yv_linear_ln_f16; see include/yv8vit.h.
YV_EXPORT size_t yv_residual_linear_f16_ws_bytes(int M, int N) {
  if (M <= 0 || N <= 0) return 0;
  return yv_align_up((size_t)M * (size_t)((N + 63) / 64) * 8, 1024);
}

YV_EXPORT int yv_residual_linear_f16(const void *A, const void *W, const float *bias, void *x, int M, int N, int K,
                                     float eps, float *stats, void *ws, size_t ws_bytes, yv_stream_t stream) {
  YV_REQUIRE(A && W && x, "yv_residual_linear_f16: null pointer");
  YV_REQUIRE(M >= 1 && N >= 64 && N % 64 == 0 && N <= 1024, "yv_residual_linear_f16: N=%d (need a multiple of 64 up to 1024)", N);
  YV_REQUIRE(!stats || (ws && ws_bytes >= yv_residual_linear_f16_ws_bytes(M, N)), "yv_residual_linear_f16: workspace too small");
  cudaStream_t s = (cudaStream_t)stream;
  GemmEpilogue ep{x, bias, x, nullptr, 0, false, 2, 2};
  if (stats) ep.stats_part = (float *)ws;
  int rc = yv_gemm_launch(A, W, M, N, K, ep, s);
  if (rc || !stats) return rc;
  return yv_stats_finalize_launch((const float *)ws, M, N, eps, stats, s);
}
