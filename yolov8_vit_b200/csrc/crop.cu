// crop.cu -- a3/a4/K9: box rounding + dilation + clamp, crop list construction,
// fused crop + resize + normalise gather (patch-major bf16 for the ViT, or the reference's
// fp32 CHW), and the per-image result records.
//
// Reference behaviour: README.md:86-124 + utils/trainClass.py:70-93 (dilation / crop),
// app.py:39-42 (A.Resize(224,224,INTER_NEAREST) + A.Normalize(.5,.5)),
// utils/trainClass.py:266 (HWC->CHW), YOLOTensorRT_yolodet_py_解读.md:93-116 (per-detection loop,
// result tuples).
#include "common.cuh"

namespace {

// floor division for a possibly negative numerator (python //), den > 0
__device__ __forceinline__ int floordiv(int a, int d) {
  int q = a / d, r = a % d;
  return (r != 0 && ((r < 0) != (d < 0))) ? q - 1 : q;
}

__device__ __forceinline__ bool dilate_clamp(int x_min, int y_min, int x_max, int y_max, int W, int H,
                                             int num, int den, int mode, int4 &o) {
  int dx = floordiv((x_max - x_min) * num, den);
  int dy = floordiv((y_max - y_min) * num, den);
  if (mode == YV_DILATE_INFER_HALF) { dx = floordiv(dx, 2); dy = floordiv(dy, 2); }
  o.z = min(W, x_max + dx);
  o.x = max(0, x_min - dx);
  o.w = min(H, y_max + dy);
  o.y = max(0, y_min - dy);
  return o.z > o.x && o.w > o.y;
}

__global__ void dilate_kernel(const int4 *__restrict__ in, int n, int W, int H, int num, int den, int mode,
                              int4 *__restrict__ out, int32_t *__restrict__ valid) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int4 b = in[i], o;
  bool ok = dilate_clamp(b.x, b.y, b.z, b.w, W, H, num, den, mode, o);
  out[i] = o;
  valid[i] = ok ? 1 : 0;
}

// torch.round(x).to(int32): round half to even, then C-style truncating cast (values are integral)
__device__ __forceinline__ int round_i32(float v) { return (int)rintf(v); }

constexpr int BC_T = 512;
constexpr int BC_K = 4;      // detections per lane in flight (4 x 32 = 128 >= the usual max_det of 100)

// One CTA (the crop list is one global prefix over images).  A warp owns an image at a time and walks its
// max_det slots 32 x BC_K at a time with all loads of a lane issued back to back, so an image costs one
// memory round trip instead of max_det dependent ones; stores are coalesced over k.  Pass 1 writes the
// per-detection arrays and counts the valid crops per image, a block scan turns the counts into offsets,
// pass 2 ranks the valid crops inside the image with ballots and writes the compacted (image, rank)-ordered list.
__global__ void __launch_bounds__(BC_T)
build_crops_kernel(const float4 *__restrict__ boxes, const float *__restrict__ scores,
                   const int32_t *__restrict__ labels, const int32_t *__restrict__ anchors,
                   const int32_t *__restrict__ keep, const int32_t *__restrict__ keep_counts, int B,
                   int stride, int max_det, int W, int H, int num, int den, int mode,
                   int4 *__restrict__ det_box, int4 *__restrict__ det_crop_box,
                   float *__restrict__ det_score, int32_t *__restrict__ det_label,
                   int32_t *__restrict__ det_anchor, int32_t *__restrict__ det_count,
                   int32_t *__restrict__ det_crop, int4 *__restrict__ crop_box,
                   int32_t *__restrict__ crop_img, int32_t *__restrict__ crop_det, int crop_cap,
                   int32_t *__restrict__ n_crops) {
  __shared__ int s_warp[BC_T / 32];
  __shared__ int s_cnt[BC_T], s_off[BC_T];
  __shared__ int s_carry;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_carry = 0;
  __syncthreads();
  for (int b0 = 0; b0 < B; b0 += BC_T) {
    const int nb = min(BC_T, B - b0);
    // ---- pass 1: per-detection arrays + valid-crop count per image
    for (int bi = warp; bi < nb; bi += BC_T / 32) {
      const int b = b0 + bi;
      const int nk = min(keep_counts[b], max_det);
      const size_t ib = (size_t)b * stride;
      int nvalid = 0;
      for (int k0 = 0; k0 < max_det; k0 += 32 * BC_K) {
        int id[BC_K];
        float4 f[BC_K];
        float sc[BC_K];
        int lb[BC_K], an[BC_K];
#pragma unroll
        for (int u = 0; u < BC_K; ++u) {
          const int k = k0 + u * 32 + lane;
          id[u] = k < nk ? keep[ib + k] : -1;
        }
#pragma unroll
        for (int u = 0; u < BC_K; ++u) {
          if (id[u] >= 0) {
            f[u] = boxes[ib + id[u]];
            sc[u] = scores[ib + id[u]];
            lb[u] = labels ? labels[ib + id[u]] : 0;
            an[u] = anchors ? anchors[ib + id[u]] : id[u];
          }
        }
#pragma unroll
        for (int u = 0; u < BC_K; ++u) {
          const int k = k0 + u * 32 + lane;
          bool ok = false;
          if (k < max_det) {
            const size_t o = (size_t)b * max_det + k;
            if (id[u] >= 0) {
              const int4 q = make_int4(round_i32(f[u].x), round_i32(f[u].y), round_i32(f[u].z), round_i32(f[u].w));
              int4 c;
              ok = dilate_clamp(q.x, q.y, q.z, q.w, W, H, num, den, mode, c);
              det_box[o] = q;
              det_crop_box[o] = c;
              det_score[o] = sc[u];
              det_label[o] = lb[u];
              det_anchor[o] = an[u];
              det_crop[o] = ok ? 0 : -1;   // provisional: rank filled in by pass 2
            } else {
              det_box[o] = make_int4(0, 0, 0, 0);
              det_crop_box[o] = make_int4(0, 0, 0, 0);
              det_score[o] = 0.f;
              det_label[o] = -1;
              det_anchor[o] = -1;
              det_crop[o] = -1;
            }
          }
          nvalid += __popc(__ballot_sync(0xffffffffu, ok));
        }
      }
      if (lane == 0) {
        det_count[b] = nk;
        s_cnt[bi] = nvalid;
      }
    }
    __syncthreads();
    // ---- block exclusive scan of the per-image counts (thread tid <-> image b0 + tid)
    const int nvalid = tid < nb ? s_cnt[tid] : 0;
    int incl = nvalid;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      // s_warp has BC_T / 32 entries: lanes past that neither read nor write it
      int v = lane < BC_T / 32 ? s_warp[lane] : 0, inc2 = v;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, inc2, d);
        if (lane >= d) inc2 += t;
      }
      if (lane < BC_T / 32) s_warp[lane] = inc2 - v;
    }
    __syncthreads();
    const int excl = s_carry + s_warp[warp] + incl - nvalid;
    s_off[tid] = excl;
    __syncthreads();
    if (tid == BC_T - 1) s_carry = excl + nvalid;     // block total carried into the next BC_T images
    // ---- pass 2: rank the valid crops of each image, write the compacted list
    for (int bi = warp; bi < nb; bi += BC_T / 32) {
      const int b = b0 + bi;
      const int nk = min(keep_counts[b], max_det);
      int off = s_off[bi];
      for (int k0 = 0; k0 < nk; k0 += 32 * BC_K) {
        int4 c[BC_K];
        bool ok[BC_K];
#pragma unroll
        for (int u = 0; u < BC_K; ++u) {
          const int k = k0 + u * 32 + lane;
          const size_t o = (size_t)b * max_det + k;
          ok[u] = k < nk && det_crop[o] == 0;         // this lane's own pass-1 store
          if (ok[u]) c[u] = det_crop_box[o];
        }
#pragma unroll
        for (int u = 0; u < BC_K; ++u) {
          const int k = k0 + u * 32 + lane;
          const size_t o = (size_t)b * max_det + k;
          const unsigned m = __ballot_sync(0xffffffffu, ok[u]);
          if (ok[u]) {
            const int dst = off + __popc(m & lanemask_lt());
            if (dst < crop_cap) {
              crop_box[dst] = c[u];
              crop_img[dst] = b;
              crop_det[dst] = (int32_t)o;
              det_crop[o] = dst;
            } else {
              det_crop[o] = -1;
            }
          }
          off += __popc(m);
        }
      }
    }
    __syncthreads();
  }
  if (tid == 0) n_crops[0] = s_carry;
}

// ----------------------------------------------------------------------------- crop gather
struct AxisTap { int i0, i1; float f; };

// bilinear, half-pixel centres, fp32, mirrors oracle.hotpath.crop_resize_norm.axis()
__device__ __forceinline__ AxisTap bilinear_tap(int d, int n, int out) {
  const float scale = __fdiv_rn((float)n, (float)out);
  float s = __fsub_rn(__fmul_rn(__fadd_rn((float)d, 0.5f), scale), 0.5f);
  s = fmaxf(s, 0.f);
  const float fl = floorf(s);
  AxisTap t;
  t.f = __fsub_rn(s, fl);
  int i0 = (int)fl;
  i0 = min(i0, n - 1);
  t.i0 = i0;
  t.i1 = min(i0 + 1, n - 1);
  return t;
}
// cv2 INTER_NEAREST: sx = min(floor(x * (1.0 / (dst / src))), src - 1), evaluated in fp64
__device__ __forceinline__ int nearest_tap(int d, int n, int out) {
  const double fx = (double)out / (double)n;
  const double ifx = 1.0 / fx;
  int s = (int)floor((double)d * ifx);
  return min(s, n - 1);
}

__device__ __forceinline__ float normalise(float v) {
  return __fmul_rn(__fsub_rn(v, 127.5f), (float)(1.0 / 127.5));
}

// exact u8 -> fp32 without the conversion pipe: the byte becomes the low mantissa bits of 2^23
__device__ __forceinline__ float byte_f32(unsigned word, int k) {
  return __fsub_rn(__uint_as_float(__byte_perm(word, 0x4B000000u, 0x7440 | k)), 8388608.f);
}
// Six consecutive bytes (two RGB pixels) starting at byte offset o of a 4-byte-aligned buffer:
// lo = bytes 0..3, hi = bytes 4..5.  Three aligned word loads + two funnel shifts instead of six
// byte loads; the last 12 bytes of the buffer take the byte path so nothing is read past `total`.
__device__ __noinline__ void fetch6_tail(const uint8_t *__restrict__ base, size_t o, size_t total, unsigned &lo,
                                         unsigned &hi) {
  unsigned b[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) b[k] = (o + k < total) ? (unsigned)base[o + k] : 0u;
  lo = b[0] | (b[1] << 8) | (b[2] << 16) | (b[3] << 24);
  hi = b[4] | (b[5] << 8);
}
__device__ __forceinline__ void fetch6(const uint8_t *__restrict__ base, size_t o, size_t total, unsigned &lo,
                                       unsigned &hi) {
  if (o + 12 <= total) {
    const unsigned *w = reinterpret_cast<const unsigned *>(base) + (o >> 2);
    const unsigned w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2);
    const unsigned sh = ((unsigned)o & 3u) * 8u;
    lo = __funnelshift_r(w0, w1, sh);
    hi = __funnelshift_r(w1, w2, sh);
  } else {
    fetch6_tail(base, o, total, lo, hi);
  }
}

constexpr int CR_T = 256;
constexpr int CR_PX = 8;     // output pixels per thread (one 16-byte bf16 store per channel)

// One thread = 8 consecutive output x pixels of one output row, all 3 channels.
// Thread order inside a crop is (output row, 8-px segment): a warp's 256 pixels lie in one or two output rows, so every
// load instruction of the warp touches two to four source rows.  (Round 1 ordered the threads (patch, row-in-patch,
// segment) so that a warp's 16-byte stores were contiguous in the patch-major output -- 512 B per channel -- but then
// each load instruction touched the 16 source rows of a patch and the kernel was bound by L1 wavefronts: 0.72 ms per
// 1072 crops against 0.47 ms in row order, where a warp's stores are 32-byte pieces 1536 B apart.)
template <int INTERP, int LAYOUT, bool WORDS>
__global__ void __launch_bounds__(CR_T)
crop_resize_kernel(const uint8_t *__restrict__ images, int H, int W, const int4 *__restrict__ crop_box,
                   const int32_t *__restrict__ crop_img, int n, const int32_t *__restrict__ n_dev,
                   int out, int patch, size_t total, void *__restrict__ dst, int rowmap) {
  const int crop = blockIdx.y;
  if (crop >= n || (n_dev && crop >= *n_dev)) return;
  const int segs_per_patch = patch * patch / CR_PX;
  const int gp = out / patch;
  const int work = gp * gp * segs_per_patch;
  const int t = blockIdx.x * CR_T + threadIdx.x;
  if (t >= work) return;
  int p, iy, ix0, oy, ox0;
  if (rowmap) {
    // row order: a warp covers 256 consecutive pixels of one or two OUTPUT rows, so its taps fall into two to four
    // source rows (a handful of cache lines per load instruction) instead of the 16 rows of a patch
    const int segs_per_orow = out / CR_PX;
    oy = t / segs_per_orow;
    ox0 = (t - oy * segs_per_orow) * CR_PX;
    p = (oy / patch) * gp + ox0 / patch;
    iy = oy % patch;
    ix0 = ox0 % patch;
  } else {
    p = t / segs_per_patch;
    const int sp = t % segs_per_patch, segs_per_row = patch / CR_PX;
    iy = sp / segs_per_row, ix0 = (sp % segs_per_row) * CR_PX;
    oy = (p / gp) * patch + iy, ox0 = (p % gp) * patch + ix0;
  }

  const int4 bx = crop_box[crop];
  const int cw = bx.z - bx.x, ch = bx.w - bx.y;
  const uint8_t *img = images + (size_t)crop_img[crop] * H * W * 3;
  float v[3][CR_PX];

  if (INTERP == YV_INTERP_NEAREST) {
    const int sy = bx.y + nearest_tap(oy, ch, out);
    const uint8_t *row = img + ((size_t)sy * W) * 3;
#pragma unroll
    for (int j = 0; j < CR_PX; ++j) {
      const int sx = bx.x + nearest_tap(ox0 + j, cw, out);
      const uint8_t *px = row + (size_t)sx * 3;
#pragma unroll
      for (int c = 0; c < 3; ++c) v[c][j] = normalise(byte_f32(px[c], 0));
    }
  } else {
    const AxisTap ty = bilinear_tap(oy, ch, out);
    const size_t img_row = (size_t)crop_img[crop] * H;
    const size_t r0 = (img_row + bx.y + ty.i0) * W * 3, r1 = (img_row + bx.y + ty.i1) * W * 3;
#pragma unroll
    for (int j = 0; j < CR_PX; ++j) {
      const AxisTap tx = bilinear_tap(ox0 + j, cw, out);
      const size_t xa = (size_t)(bx.x + tx.i0) * 3;
      float a0[3], a1[3], b0[3], b1[3];      // a = left tap, b = right tap; 0/1 = upper/lower row
      if (WORDS) {
        unsigned lo0, hi0, lo1, hi1;
        fetch6(images, r0 + xa, total, lo0, hi0);
        fetch6(images, r1 + xa, total, lo1, hi1);
        const bool same = tx.i1 == tx.i0;     // right edge of the crop: both taps are the same pixel
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          a0[c] = byte_f32(lo0, c);
          a1[c] = byte_f32(lo1, c);
        }
        b0[0] = byte_f32(lo0, 3); b0[1] = byte_f32(hi0, 0); b0[2] = byte_f32(hi0, 1);
        b1[0] = byte_f32(lo1, 3); b1[1] = byte_f32(hi1, 0); b1[2] = byte_f32(hi1, 1);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          b0[c] = same ? a0[c] : b0[c];
          b1[c] = same ? a1[c] : b1[c];
        }
      } else {
        const size_t xb = (size_t)(bx.x + tx.i1) * 3;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          a0[c] = byte_f32(images[r0 + xa + c], 0); b0[c] = byte_f32(images[r0 + xb + c], 0);
          a1[c] = byte_f32(images[r1 + xa + c], 0); b1[c] = byte_f32(images[r1 + xb + c], 0);
        }
      }
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float top = __fadd_rn(a0[c], __fmul_rn(tx.f, __fsub_rn(b0[c], a0[c])));
        const float bot = __fadd_rn(a1[c], __fmul_rn(tx.f, __fsub_rn(b1[c], a1[c])));
        v[c][j] = normalise(__fadd_rn(top, __fmul_rn(ty.f, __fsub_rn(bot, top))));
      }
    }
  }

  if (LAYOUT == YV_LAYOUT_PATCH_BF16) {
    const int pp = patch * patch;
    __nv_bfloat16 *o = (__nv_bfloat16 *)dst + ((size_t)crop * gp * gp + p) * (size_t)(3 * pp) + iy * patch + ix0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      uint4 pk;
      __nv_bfloat162 h0 = __floats2bfloat162_rn(v[c][0], v[c][1]);
      __nv_bfloat162 h1 = __floats2bfloat162_rn(v[c][2], v[c][3]);
      __nv_bfloat162 h2 = __floats2bfloat162_rn(v[c][4], v[c][5]);
      __nv_bfloat162 h3 = __floats2bfloat162_rn(v[c][6], v[c][7]);
      pk.x = *reinterpret_cast<unsigned *>(&h0);
      pk.y = *reinterpret_cast<unsigned *>(&h1);
      pk.z = *reinterpret_cast<unsigned *>(&h2);
      pk.w = *reinterpret_cast<unsigned *>(&h3);
      *reinterpret_cast<uint4 *>(o + (size_t)c * pp) = pk;
    }
  } else {
    float *o = (float *)dst + (((size_t)crop * 3) * out + oy) * out + ox0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float4 *q = reinterpret_cast<float4 *>(o + (size_t)c * out * out);
      q[0] = make_float4(v[c][0], v[c][1], v[c][2], v[c][3]);
      q[1] = make_float4(v[c][4], v[c][5], v[c][6], v[c][7]);
    }
  }
}

__global__ void pack_results_kernel(const int4 *__restrict__ det_box, const float *__restrict__ det_score,
                                    const int32_t *__restrict__ det_anchor, const int32_t *__restrict__ det_count,
                                    const int32_t *__restrict__ det_crop, const int32_t *__restrict__ top1, int B,
                                    int max_det, float4 *__restrict__ rec) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * max_det) return;
  const int b = i / max_det, k = i % max_det;
  const bool have = k < det_count[b];
  float4 r0 = make_float4(0.f, 0.f, 0.f, 0.f), r1 = make_float4(-1.f, 0.f, -1.f, 0.f);
  if (have) {
    const int4 q = det_box[i];
    const int c = det_crop[i];
    r0 = make_float4((float)q.x, (float)q.y, (float)q.z, (float)q.w);
    r1 = make_float4((c >= 0 && top1) ? (float)top1[c] : -1.f, det_score[i], (float)det_anchor[i], 1.f);
  }
  rec[2 * (size_t)i] = r0;
  rec[2 * (size_t)i + 1] = r1;
}

}  // namespace

YV_EXPORT int yv_dilate_boxes(const int32_t *boxes_i32, int n, int W, int H, int num, int den, int mode,
                              int32_t *out_boxes, int32_t *out_valid, yv_stream_t stream) {
  YV_REQUIRE(n >= 0 && den > 0 && num >= 0 && W > 0 && H > 0, "yv_dilate_boxes: bad argument");
  YV_REQUIRE(mode == YV_DILATE_FULL || mode == YV_DILATE_INFER_HALF, "yv_dilate_boxes: unknown mode %d", mode);
  if (n == 0) return YV_OK;
  YV_REQUIRE(boxes_i32 && out_boxes && out_valid, "yv_dilate_boxes: null pointer");
  YvProf prof(YV_K_MISC, (cudaStream_t)stream, 0.0);
  dilate_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>((const int4 *)boxes_i32, n, W, H, num, den, mode,
                                                                   (int4 *)out_boxes, out_valid);
  YV_LAUNCH_CHECK();
  return YV_OK;
}

YV_EXPORT int yv_build_crops(const float *boxes, const float *scores, const int32_t *labels,
                             const int32_t *anchors, const int32_t *keep, const int32_t *keep_counts, int B,
                             int stride, int max_det, int W, int H, int num, int den, int mode,
                             int32_t *det_box, int32_t *det_crop_box, float *det_score, int32_t *det_label,
                             int32_t *det_anchor, int32_t *det_count, int32_t *det_crop, int32_t *crop_box,
                             int32_t *crop_img, int32_t *crop_det, int crop_cap, int32_t *n_crops,
                             yv_stream_t stream) {
  YV_REQUIRE(B >= 1 && stride >= 1 && max_det >= 1 && crop_cap >= 1 && den > 0 && num >= 0 && W > 0 && H > 0,
             "yv_build_crops: bad argument");
  YV_REQUIRE(mode == YV_DILATE_FULL || mode == YV_DILATE_INFER_HALF, "yv_build_crops: unknown mode %d", mode);
  YV_REQUIRE(boxes && scores && keep && keep_counts && det_box && det_crop_box && det_score && det_label &&
                 det_anchor && det_count && det_crop && crop_box && crop_img && crop_det && n_crops,
             "yv_build_crops: null pointer");
  YvProf prof(YV_K_BUILD, (cudaStream_t)stream, 0.0);
  build_crops_kernel<<<1, BC_T, 0, (cudaStream_t)stream>>>(
      (const float4 *)boxes, scores, labels, anchors, keep, keep_counts, B, stride, max_det, W, H, num, den, mode,
      (int4 *)det_box, (int4 *)det_crop_box, det_score, det_label, det_anchor, det_count, det_crop, (int4 *)crop_box,
      crop_img, crop_det, crop_cap, n_crops);
  YV_LAUNCH_CHECK();
  return YV_OK;
}

YV_EXPORT int yv_crop_resize(const uint8_t *images, int B, int H, int W, const int32_t *crop_box,
                             const int32_t *crop_img, int n, const int32_t *n_dev, int interp, int out_size,
                             int patch, int layout, void *out, yv_stream_t stream) {
  YV_REQUIRE(B >= 1 && H >= 1 && W >= 1 && n >= 0, "yv_crop_resize: bad shape");
  YV_REQUIRE(patch >= 8 && patch % 8 == 0 && out_size % patch == 0, "yv_crop_resize: need patch %% 8 == 0 and out %% patch == 0 (out=%d patch=%d)", out_size, patch);
  YV_REQUIRE(interp == YV_INTERP_NEAREST || interp == YV_INTERP_BILINEAR, "yv_crop_resize: unknown interp %d", interp);
  YV_REQUIRE(layout == YV_LAYOUT_PATCH_BF16 || layout == YV_LAYOUT_CHW_F32, "yv_crop_resize: unknown layout %d", layout);
  if (n == 0) return YV_OK;
  YV_REQUIRE(n <= 65535, "yv_crop_resize: n=%d exceeds 65535 crops per call", n);
  YV_REQUIRE(images && crop_box && crop_img && out, "yv_crop_resize: null pointer");
  YV_REQUIRE(((uintptr_t)out % 16) == 0, "yv_crop_resize: out must be 16-byte aligned");
  const int work = out_size * out_size / CR_PX;
  dim3 grid((work + CR_T - 1) / CR_T, n);
  cudaStream_t s = (cudaStream_t)stream;
  const int4 *cb = (const int4 *)crop_box;
  YvProf prof(YV_K_CROP, s, (double)n * out_size * out_size * 3 * (layout == YV_LAYOUT_PATCH_BF16 ? 2 : 4));
  // aligned word loads need a 4-byte-aligned image base (any CUDA allocation); otherwise byte loads
  const size_t total = (size_t)B * H * W * 3;
  const bool words = ((uintptr_t)images % 4) == 0;
  // thread order inside a crop: output rows (default; measured 0.47 against 0.72 ms per 1072 crops) or YV_CROP_MAP=patch
  static const int rowmap = []() { const char *e = getenv("YV_CROP_MAP"); return e ? (e[0] != 'p') : 1; }();
#define YV_CR(I, L)                                                                                              \
  do {                                                                                                           \
    if (words) crop_resize_kernel<I, L, true><<<grid, CR_T, 0, s>>>(images, H, W, cb, crop_img, n, n_dev, out_size, patch, total, out, rowmap); \
    else crop_resize_kernel<I, L, false><<<grid, CR_T, 0, s>>>(images, H, W, cb, crop_img, n, n_dev, out_size, patch, total, out, rowmap); \
  } while (0)
  if (interp == YV_INTERP_NEAREST) {
    if (layout == YV_LAYOUT_PATCH_BF16) YV_CR(YV_INTERP_NEAREST, YV_LAYOUT_PATCH_BF16);
    else YV_CR(YV_INTERP_NEAREST, YV_LAYOUT_CHW_F32);
  } else {
    if (layout == YV_LAYOUT_PATCH_BF16) YV_CR(YV_INTERP_BILINEAR, YV_LAYOUT_PATCH_BF16);
    else YV_CR(YV_INTERP_BILINEAR, YV_LAYOUT_CHW_F32);
  }
#undef YV_CR
  YV_LAUNCH_CHECK();
  return YV_OK;
}

YV_EXPORT int yv_pack_results(const int32_t *det_box, const float *det_score, const int32_t *det_anchor,
                              const int32_t *det_count, const int32_t *det_crop, const int32_t *top1, int B,
                              int max_det, float *records, yv_stream_t stream) {
  YV_REQUIRE(B >= 1 && max_det >= 1, "yv_pack_results: bad shape");
  YV_REQUIRE(det_box && det_score && det_anchor && det_count && det_crop && records, "yv_pack_results: null pointer");
  const int n = B * max_det;
  YvProf prof(YV_K_MISC, (cudaStream_t)stream, 0.0);
  pack_results_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
      (const int4 *)det_box, det_score, det_anchor, det_count, det_crop, top1, B, max_det, (float4 *)records);
  YV_LAUNCH_CHECK();
  return YV_OK;
}
