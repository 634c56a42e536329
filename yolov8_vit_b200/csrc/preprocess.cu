// preprocess.cu -- row f2: letterbox + BGR->RGB + /255 + HWC->CHW "blob" of the detector input on the GPU.
//
// Replaces the wrapper's per-image host preprocessing (YOLOTensorRT_yolodet_py_解读.md:67-78:
// letterbox(bgr, (W, H)) = cv2.resize(INTER_LINEAR) + copyMakeBorder(114), cvtColor(BGR2RGB), blob = /255,
// CHW).  The resize reproduces OpenCV's 8-bit INTER_LINEAR arithmetic bit for bit:
//   fx = float((dx + 0.5) * scale - 0.5) (double inside), sx = floor(fx), fx -= sx; the x taps clamp the
//   fraction to 0 at the borders, the y taps keep the fraction and clamp the ROW indices;
//   coefficients = cvRound(float * 2048) as shorts; horizontal pass in int32, vertical pass
//   ((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2.
// One thread per output pixel (x fastest): 3 coalesced plane stores, source taps through L1/L2.
// HBM-bound: B*H*W*3 bytes read + B*3*Ho*Wo*4 bytes written.
#include "common.cuh"

#include <cmath>

namespace {

struct LbGeom {
  int uw, uh, left, top;       // resized size and its offset inside the padded output
  double sx, sy;               // source / resized scale factors (inv_scale of cv2.resize)
  int resize;                  // 0 = the source already has the unpadded size (the wrapper skips cv2.resize)
};

__device__ __forceinline__ void lb_tap_x(int d, double scale, int n, int &i0, int &i1, int &a0, int &a1) {
  float f = (float)__dadd_rn(__dmul_rn((double)d + 0.5, scale), -0.5);   // no FMA contraction
  int s = (int)floorf(f);
  f -= (float)s;
  if (s < 0) { s = 0; f = 0.f; }
  if (s >= n - 1) { s = n - 1; f = 0.f; }
  i0 = s;
  i1 = min(s + 1, n - 1);
  a0 = __float2int_rn((1.f - f) * 2048.f);
  a1 = __float2int_rn(f * 2048.f);
}
__device__ __forceinline__ void lb_tap_y(int d, double scale, int n, int &i0, int &i1, int &b0, int &b1) {
  float f = (float)__dadd_rn(__dmul_rn((double)d + 0.5, scale), -0.5);   // no FMA contraction
  const int s = (int)floorf(f);
  f -= (float)s;
  i0 = min(max(s, 0), n - 1);
  i1 = min(max(s + 1, 0), n - 1);
  b0 = __float2int_rn((1.f - f) * 2048.f);
  b1 = __float2int_rn(f * 2048.f);
}

__global__ void __launch_bounds__(256)
letterbox_blob_kernel(const uint8_t *__restrict__ images, int B, int H, int W, int swap_rb, int Ho, int Wo,
                      LbGeom g, float *__restrict__ blob) {
  const size_t total = (size_t)B * Ho * Wo;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int x = (int)(idx % Wo);
    const int y = (int)((idx / Wo) % Ho);
    const int b = (int)(idx / ((size_t)Wo * Ho));
    int v[3] = {114, 114, 114};
    const int rx = x - g.left, ry = y - g.top;
    if (rx >= 0 && rx < g.uw && ry >= 0 && ry < g.uh) {
      const uint8_t *img = images + (size_t)b * H * W * 3;
      if (!g.resize) {
        const uint8_t *p = img + ((size_t)ry * W + rx) * 3;
        v[0] = p[0]; v[1] = p[1]; v[2] = p[2];
      } else {
        int x0, x1, a0, a1, y0, y1, b0, b1;
        lb_tap_x(rx, g.sx, W, x0, x1, a0, a1);
        lb_tap_y(ry, g.sy, H, y0, y1, b0, b1);
        const uint8_t *r0 = img + (size_t)y0 * W * 3, *r1 = img + (size_t)y1 * W * 3;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const int h0 = (int)r0[x0 * 3 + c] * a0 + (int)r0[x1 * 3 + c] * a1;
          const int h1 = (int)r1[x0 * 3 + c] * a0 + (int)r1[x1 * 3 + c] * a1;
          v[c] = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
        }
      }
    }
    float *o = blob + (size_t)b * 3 * Ho * Wo + (size_t)y * Wo + x;
    const size_t plane = (size_t)Ho * Wo;
    o[0] = __fdiv_rn((float)v[swap_rb ? 2 : 0], 255.f);
    o[plane] = __fdiv_rn((float)v[1], 255.f);
    o[2 * plane] = __fdiv_rn((float)v[swap_rb ? 0 : 2], 255.f);
  }
}

}  // namespace

// the wrapper's letterbox geometry (Python round() = half to even = nearbyint in the default mode)
YV_EXPORT int yv_letterbox_geometry(int H, int W, int out_h, int out_w, double *ratio, double *dw, double *dh,
                                    int32_t *unpad_wh, int32_t *left_top) {
  YV_REQUIRE(H >= 1 && W >= 1 && out_h >= 1 && out_w >= 1, "yv_letterbox_geometry: bad size");
  const double r = std::fmin((double)out_h / H, (double)out_w / W);
  const int uw = (int)std::nearbyint(W * r), uh = (int)std::nearbyint(H * r);
  const double pw = (out_w - uw) / 2.0, ph = (out_h - uh) / 2.0;
  if (ratio) *ratio = r;
  if (dw) *dw = pw;
  if (dh) *dh = ph;
  if (unpad_wh) { unpad_wh[0] = uw; unpad_wh[1] = uh; }
  if (left_top) { left_top[0] = (int)std::nearbyint(pw - 0.1); left_top[1] = (int)std::nearbyint(ph - 0.1); }
  return YV_OK;
}

YV_EXPORT int yv_letterbox_blob(const uint8_t *images, int B, int H, int W, int swap_rb, int out_h, int out_w,
                                float *blob, yv_stream_t stream) {
  YV_REQUIRE(B >= 0 && H >= 1 && W >= 1 && out_h >= 1 && out_w >= 1, "yv_letterbox_blob: bad shape");
  if (B == 0) return YV_OK;
  YV_REQUIRE(images && blob, "yv_letterbox_blob: null pointer");
  double r, pw, ph;
  int32_t un[2], lt[2];
  yv_letterbox_geometry(H, W, out_h, out_w, &r, &pw, &ph, un, lt);
  YV_REQUIRE(un[0] >= 1 && un[1] >= 1 && lt[0] >= 0 && lt[1] >= 0 && lt[0] + un[0] <= out_w && lt[1] + un[1] <= out_h,
             "yv_letterbox_blob: degenerate letterbox %dx%d -> %dx%d", W, H, out_w, out_h);
  LbGeom g;
  g.uw = un[0]; g.uh = un[1]; g.left = lt[0]; g.top = lt[1];
  g.sx = (double)W / un[0];
  g.sy = (double)H / un[1];
  g.resize = !(un[0] == W && un[1] == H);
  cudaStream_t s = (cudaStream_t)stream;
  const size_t total = (size_t)B * out_h * out_w;
  const int blocks = (int)std::min<size_t>((total + 255) / 256, (size_t)yv_num_sms() * 16);
  YvProf prof(YV_K_MISC, s, (double)B * H * W * 3.0 + (double)total * 12.0);
  letterbox_blob_kernel<<<blocks, 256, 0, s>>>(images, B, H, W, swap_rb ? 1 : 0, out_h, out_w, g, blob);
  YV_LAUNCH_CHECK();
  return YV_OK;
}
