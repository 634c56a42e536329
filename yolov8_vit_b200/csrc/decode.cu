// decode.cu -- K1: YOLOv8 head decode + confidence filter + order-preserving compaction.
//
// Replaces the per-anchor loop of the detector wrapper's post-processing
// (docs/YOLO_TensorRT_Technical.md:243-267; YOLOTensorRT_yolodet_py_解读.md:89-94).
//
// Layout: head [B, 4+nc, A] fp32, channel-major.  A warp reads 32 x VEC consecutive
// anchors of one channel row per load (LDG.128, streaming / no L1 allocate), so every
// request is a fully coalesced 512 B run; each thread keeps UNROLL independent row loads
// in flight.  A tile is NT*VEC anchors of one image.  Kept anchors are ranked inside the
// tile with popc + warp shuffle scan (+ a 4-entry cross-warp scan) and staged tile-locally;
// the LAST tile of an image to finish (threadfence + atomic ticket, no spinning) scans the
// tile counts and moves the staged candidates to their final, ascending-anchor positions.
// HBM-bound: algorithmic bytes = B*(4+nc)*A*4 read (+ 28 B per kept candidate).
#include "common.cuh"

namespace {

constexpr int DEC_NT = 128;      // threads per tile
constexpr int DEC_UNROLL = 8;    // class rows in flight per thread
constexpr int DEC_GS = 256;      // tiles per image handled by the parallel gather

template <int VEC> struct RowVec;
template <> struct RowVec<4> {
  float v[4];
  __device__ __forceinline__ void load(const float *p) {
    float4 t = ld_stream_f4(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
};
template <> struct RowVec<1> {
  float v[1];
  __device__ __forceinline__ void load(const float *p) { v[0] = ld_stream_f1(p); }
};

struct DecodeStage {          // tile-local staging, SoA, capacity = tiles*TILE per image
  float4 *box;
  float *score;
  int32_t *label;
  int32_t *anchor;
  int32_t *tile_count;        // [B*tiles]
  unsigned *ticket;           // [B], zero before launch, reset by the last tile
};

// Raw-head mode (row f1): the tensor is the Detect head's concatenated raw output [B, 4*16+nc, A] --
// 4 x 16 DFL bin logits then nc class logits per anchor, anchors of the 3 levels back to back
// (docs/YOLO_TensorRT_Technical.md:13-30 make_anchors, :71-77 DFL decode; 解读.md:64-92).
struct RawLevels {
  int start[3];      // first anchor of each level
  int width[3];      // feature-map width of each level
  float stride[3];
};
constexpr int DFL_BINS = 16;

template <int VEC, bool RAW>
__global__ void __launch_bounds__(DEC_NT)
decode_filter_kernel(const float *__restrict__ head, int nc, int A, float conf, int unletterbox,
                     float ratio, float dw, float dh, int cap, DecodeStage st,
                     float4 *__restrict__ out_box, float *__restrict__ out_score,
                     int32_t *__restrict__ out_label, int32_t *__restrict__ out_anchor,
                     int32_t *__restrict__ out_count, RawLevels lv) {
  constexpr int TILE = DEC_NT * VEC;
  constexpr int BOXROWS = RAW ? 4 * DFL_BINS : 4;
  const int tiles = gridDim.x;
  const int tile = blockIdx.x, b = blockIdx.y;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int a0 = tile * TILE + tid * VEC;
  const float *hp = head + (size_t)b * (size_t)(BOXROWS + nc) * (size_t)A;
  // VEC == 4 requires A % 4 == 0, so a thread's 4 anchors are all in range or all out.
  const bool in_range = a0 < A;

  float best[VEC];
  int lab[VEC];
  RowVec<VEC> bx[4];
#pragma unroll
  for (int j = 0; j < VEC; ++j) { best[j] = 0.f; lab[j] = 0; }

  if (in_range) {
    const float *cp = hp + (size_t)BOXROWS * A + a0;
    {  // class 0 initialises the running max.  Update rule = torch.max(dim): first max wins on ties, and a NaN
       // score wins over every number and is never replaced (the first NaN's index is returned):
       // !(v <= best) is "v > best or unordered", best == best locks a NaN that is already held
      RowVec<VEC> r;
      r.load(cp);
#pragma unroll
      for (int j = 0; j < VEC; ++j) best[j] = r.v[j];
    }
    int c = 1;
    for (; c + DEC_UNROLL <= nc; c += DEC_UNROLL) {
      RowVec<VEC> r[DEC_UNROLL];
#pragma unroll
      for (int u = 0; u < DEC_UNROLL; ++u) r[u].load(cp + (size_t)(c + u) * A);
#pragma unroll
      for (int u = 0; u < DEC_UNROLL; ++u)
#pragma unroll
        for (int j = 0; j < VEC; ++j)
          if (!(r[u].v[j] <= best[j]) && best[j] == best[j]) { best[j] = r[u].v[j]; lab[j] = c + u; }
    }
    for (; c < nc; ++c) {
      RowVec<VEC> r;
      r.load(cp + (size_t)c * A);
#pragma unroll
      for (int j = 0; j < VEC; ++j)
        if (!(r.v[j] <= best[j]) && best[j] == best[j]) { best[j] = r.v[j]; lab[j] = c; }
    }
    if (!RAW) {
#pragma unroll
      for (int k = 0; k < 4; ++k) bx[k].load(hp + (size_t)k * A + a0);
    } else {
      // class probabilities are sigmoid(logit): monotonic, so the arg-max is taken on the logits and only
      // the winner is squashed (torch.sigmoid in fp32)
#pragma unroll
      for (int j = 0; j < VEC; ++j) best[j] = __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-best[j])));
    }
  }

  // D4: keep iff not (score < conf)
  unsigned flags = 0;
  if (in_range) {
#pragma unroll
    for (int j = 0; j < VEC; ++j)
      if (!(best[j] < conf)) flags |= 1u << j;
  }
  const int cnt = __popc(flags);

  // tile-local exclusive rank: warp shuffle scan + cross-warp scan
  int incl = cnt;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += t;
  }
  __shared__ int s_wsum[DEC_NT / 32];
  __shared__ int s_last;
  if (lane == 31) s_wsum[warp] = incl;
  __syncthreads();
  int woff = 0, total = 0;
#pragma unroll
  for (int w = 0; w < DEC_NT / 32; ++w) {
    int t = s_wsum[w];
    if (w < warp) woff += t;
    total += t;
  }
  int rank = woff + incl - cnt;

  const size_t sbase = ((size_t)b * tiles + tile) * TILE;
  if (flags) {
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      if (flags & (1u << j)) {
        float x0, y0, x1, y1;
        if (!RAW) {
          float cx = bx[0].v[j], cy = bx[1].v[j], w = bx[2].v[j], h = bx[3].v[j];
          float hw = __fmul_rn(0.5f, w), hh = __fmul_rn(0.5f, h);
          x0 = __fsub_rn(cx, hw); y0 = __fsub_rn(cy, hh);
          x1 = __fadd_rn(cx, hw); y1 = __fadd_rn(cy, hh);
        } else {
          // DFL only for KEPT anchors (sparse): distance = sum_i i * softmax(bins)_i per side, then
          // (anchor -/+ distance) * stride, anchor = cell index + 0.5
          const int a = a0 + j;
          const int l = a >= lv.start[2] ? 2 : (a >= lv.start[1] ? 1 : 0);
          const int cell = a - lv.start[l];
          const int gy = cell / lv.width[l], gx = cell - gy * lv.width[l];
          const float ax = (float)gx + 0.5f, ay = (float)gy + 0.5f, sd = lv.stride[l];
          float d[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            float e[DFL_BINS], m = -INFINITY;
#pragma unroll
            for (int i = 0; i < DFL_BINS; ++i) {
              e[i] = __ldg(hp + (size_t)(k * DFL_BINS + i) * A + a);
              m = fmaxf(m, e[i]);
            }
            float den = 0.f, num = 0.f;
#pragma unroll
            for (int i = 0; i < DFL_BINS; ++i) {
              e[i] = expf(e[i] - m);
              den += e[i];
            }
#pragma unroll
            for (int i = 0; i < DFL_BINS; ++i) num = fmaf(__fdiv_rn(e[i], den), (float)i, num);
            d[k] = num;
          }
          x0 = __fmul_rn(__fsub_rn(ax, d[0]), sd); y0 = __fmul_rn(__fsub_rn(ay, d[1]), sd);
          x1 = __fmul_rn(__fadd_rn(ax, d[2]), sd); y1 = __fmul_rn(__fadd_rn(ay, d[3]), sd);
        }
        if (unletterbox) {
          x0 = __fdiv_rn(__fsub_rn(x0, dw), ratio);
          x1 = __fdiv_rn(__fsub_rn(x1, dw), ratio);
          y0 = __fdiv_rn(__fsub_rn(y0, dh), ratio);
          y1 = __fdiv_rn(__fsub_rn(y1, dh), ratio);
        }
        st.box[sbase + rank] = make_float4(x0, y0, x1, y1);
        st.score[sbase + rank] = best[j];
        st.label[sbase + rank] = lab[j];
        st.anchor[sbase + rank] = a0 + j;
        ++rank;
      }
    }
  }

  // publish the tile, take a ticket; the last tile of the image gathers.
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    st.tile_count[b * tiles + tile] = total;
    __threadfence();
    unsigned prev = atomicAdd(&st.ticket[b], 1u);
    s_last = (prev == (unsigned)(tiles - 1));
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();

  // Gather, parallel form: tile counts -> exclusive offsets (warp 0), then every staged candidate is one
  // independent copy (binary search of its tile), so the last tile pays two memory latencies instead of one
  // per tile (the serial per-tile walk was a 17 us tail per image: 4.4 -> 5.5 TB/s at the bench shape).
  // Staged reads go through L2 (.cg).  Serial walk kept for > DEC_GS tiles.
  __shared__ int s_cnt[DEC_GS], s_off[DEC_GS], s_total;
  int off = 0;
  if (tiles <= DEC_GS) {
    for (int t = tid; t < tiles; t += DEC_NT) s_cnt[t] = __ldcg(&st.tile_count[b * tiles + t]);
    __syncthreads();
    if (tid < 32) {
      int carry = 0;
      for (int base = 0; base < tiles; base += 32) {
        const int v = base + lane < tiles ? s_cnt[base + lane] : 0;
        int inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          int u = __shfl_up_sync(0xffffffffu, inc, d);
          if (lane >= d) inc += u;
        }
        if (base + lane < tiles) s_off[base + lane] = carry + inc - v;
        carry += __shfl_sync(0xffffffffu, inc, 31);
      }
      if (lane == 0) s_total = carry;
    }
    __syncthreads();
    off = s_total;
    const int lim = off < cap ? off : cap;
    for (int idx = tid; idx < lim; idx += DEC_NT) {
      int lo = 0, hi = tiles - 1;
      while (lo < hi) {                                 // last tile whose offset is <= idx (it is non-empty)
        const int mid = (lo + hi + 1) >> 1;
        if (s_off[mid] <= idx) lo = mid; else hi = mid - 1;
      }
      const size_t src = ((size_t)b * tiles + lo) * TILE + (idx - s_off[lo]);
      const size_t o = (size_t)b * cap + idx;
      out_box[o] = __ldcg(&st.box[src]);
      out_score[o] = __ldcg(&st.score[src]);
      out_label[o] = __ldcg(&st.label[src]);
      out_anchor[o] = __ldcg(&st.anchor[src]);
    }
  } else {
    for (int t = 0; t < tiles; ++t) {
      const int n = __ldcg(&st.tile_count[b * tiles + t]);
      const size_t src = ((size_t)b * tiles + t) * TILE;
      for (int i = tid; i < n; i += DEC_NT) {
        const int dst = off + i;
        if (dst < cap) {
          const size_t o = (size_t)b * cap + dst;
          out_box[o] = __ldcg(&st.box[src + i]);
          out_score[o] = __ldcg(&st.score[src + i]);
          out_label[o] = __ldcg(&st.label[src + i]);
          out_anchor[o] = __ldcg(&st.anchor[src + i]);
        }
      }
      off += n;
    }
  }
  if (tid == 0) {
    out_count[b] = off < cap ? off : cap;   // what was written: with cap < candidates the list is cut in anchor order
    st.ticket[b] = 0u;   // ready for the next launch on this workspace
  }
}

struct DecodeLayout {
  int vec, tile, tiles;
  size_t per_img;       // staged entries per image
  size_t o_box, o_score, o_label, o_anchor, o_tcount, o_ticket, total;
};

DecodeLayout decode_layout(int B, int A) {
  DecodeLayout L;
  L.vec = (A % 4 == 0) ? 4 : 1;
  L.tile = DEC_NT * L.vec;
  L.tiles = (A + L.tile - 1) / L.tile;
  L.per_img = (size_t)L.tiles * L.tile;
  size_t n = (size_t)B * L.per_img, o = 0;
  L.o_box = o;    o = yv_align_up(o + n * sizeof(float4), 256);
  L.o_score = o;  o = yv_align_up(o + n * sizeof(float), 256);
  L.o_label = o;  o = yv_align_up(o + n * sizeof(int32_t), 256);
  L.o_anchor = o; o = yv_align_up(o + n * sizeof(int32_t), 256);
  L.o_tcount = o; o = yv_align_up(o + (size_t)B * L.tiles * sizeof(int32_t), 256);
  L.o_ticket = o; o = yv_align_up(o + (size_t)B * sizeof(unsigned), 256);
  L.total = o;
  return L;
}

}  // namespace

YV_EXPORT size_t yv_decode_filter_workspace_bytes(int B, int nc, int A) {
  (void)nc;
  if (B <= 0 || A <= 0) return 0;
  return decode_layout(B, A).total;
}

YV_EXPORT int yv_decode_filter(const float *head, int B, int nc, int A, float conf, int unletterbox,
                               float ratio, float dw, float dh, int cap, float *boxes, float *scores,
                               int32_t *labels, int32_t *anchors, int32_t *counts, void *ws,
                               size_t ws_bytes, yv_stream_t stream) {
  YV_REQUIRE(B >= 0 && nc >= 1 && A >= 1 && cap >= 1, "yv_decode_filter: bad shape B=%d nc=%d A=%d cap=%d", B, nc, A, cap);
  if (B == 0) return YV_OK;
  YV_REQUIRE(B <= 65535, "yv_decode_filter: B=%d exceeds 65535 (gridDim.y)", B);
  YV_REQUIRE(head && boxes && scores && labels && anchors && counts && ws, "yv_decode_filter: null pointer");
  YV_REQUIRE(((uintptr_t)head % 16) == 0 && ((uintptr_t)boxes % 16) == 0, "yv_decode_filter: head/boxes must be 16-byte aligned");
  const DecodeLayout L = decode_layout(B, A);
  if (ws_bytes < L.total) {
    yv_set_error("yv_decode_filter: workspace %zu < required %zu", ws_bytes, L.total);
    return YV_ENOMEM;
  }
  cudaStream_t s = (cudaStream_t)stream;
  char *w = (char *)ws;
  DecodeStage st;
  st.box = (float4 *)(w + L.o_box);
  st.score = (float *)(w + L.o_score);
  st.label = (int32_t *)(w + L.o_label);
  st.anchor = (int32_t *)(w + L.o_anchor);
  st.tile_count = (int32_t *)(w + L.o_tcount);
  st.ticket = (unsigned *)(w + L.o_ticket);
  YV_CUDA(cudaMemsetAsync(st.ticket, 0, (size_t)B * sizeof(unsigned), s));
  dim3 grid(L.tiles, B);
  YvProf prof(YV_K_DECODE, s, (double)B * (4 + nc) * A * 4.0);
  const RawLevels none{};
  if (L.vec == 4)
    decode_filter_kernel<4, false><<<grid, DEC_NT, 0, s>>>(head, nc, A, conf, unletterbox, ratio, dw, dh, cap, st,
                                                           (float4 *)boxes, scores, labels, anchors, counts, none);
  else
    decode_filter_kernel<1, false><<<grid, DEC_NT, 0, s>>>(head, nc, A, conf, unletterbox, ratio, dw, dh, cap, st,
                                                           (float4 *)boxes, scores, labels, anchors, counts, none);
  YV_LAUNCH_CHECK();
  return YV_OK;
}

YV_EXPORT int yv_decode_filter_raw(const float *raw, int B, int nc, const int32_t *level_hw_host,
                                   const float *level_stride_host, float conf, int unletterbox, float ratio,
                                   float dw, float dh, int cap, float *boxes, float *scores, int32_t *labels,
                                   int32_t *anchors, int32_t *counts, void *ws, size_t ws_bytes,
                                   yv_stream_t stream) {
  YV_REQUIRE(level_hw_host && level_stride_host, "yv_decode_filter_raw: null level description");
  RawLevels lv;
  int A = 0;
  for (int l = 0; l < 3; ++l) {
    const int h = level_hw_host[2 * l], w = level_hw_host[2 * l + 1];
    YV_REQUIRE(h >= 1 && w >= 1 && level_stride_host[l] > 0.f, "yv_decode_filter_raw: bad level %d (%d x %d)", l, h, w);
    lv.start[l] = A;
    lv.width[l] = w;
    lv.stride[l] = level_stride_host[l];
    A += h * w;
  }
  YV_REQUIRE(B >= 0 && nc >= 1 && cap >= 1, "yv_decode_filter_raw: bad shape B=%d nc=%d cap=%d", B, nc, cap);
  if (B == 0) return YV_OK;
  YV_REQUIRE(B <= 65535, "yv_decode_filter_raw: B=%d exceeds 65535 (gridDim.y)", B);
  YV_REQUIRE(raw && boxes && scores && labels && anchors && counts && ws, "yv_decode_filter_raw: null pointer");
  YV_REQUIRE(((uintptr_t)raw % 16) == 0 && ((uintptr_t)boxes % 16) == 0, "yv_decode_filter_raw: raw/boxes must be 16-byte aligned");
  const DecodeLayout L = decode_layout(B, A);
  if (ws_bytes < L.total) {
    yv_set_error("yv_decode_filter_raw: workspace %zu < required %zu", ws_bytes, L.total);
    return YV_ENOMEM;
  }
  cudaStream_t s = (cudaStream_t)stream;
  char *w = (char *)ws;
  DecodeStage st;
  st.box = (float4 *)(w + L.o_box);
  st.score = (float *)(w + L.o_score);
  st.label = (int32_t *)(w + L.o_label);
  st.anchor = (int32_t *)(w + L.o_anchor);
  st.tile_count = (int32_t *)(w + L.o_tcount);
  st.ticket = (unsigned *)(w + L.o_ticket);
  YV_CUDA(cudaMemsetAsync(st.ticket, 0, (size_t)B * sizeof(unsigned), s));
  dim3 grid(L.tiles, B);
  // algorithmic bytes: the class rows of every anchor (the 64 DFL rows are read for kept anchors only)
  YvProf prof(YV_K_DECODE, s, (double)B * nc * A * 4.0);
  if (L.vec == 4)
    decode_filter_kernel<4, true><<<grid, DEC_NT, 0, s>>>(raw, nc, A, conf, unletterbox, ratio, dw, dh, cap, st,
                                                          (float4 *)boxes, scores, labels, anchors, counts, lv);
  else
    decode_filter_kernel<1, true><<<grid, DEC_NT, 0, s>>>(raw, nc, A, conf, unletterbox, ratio, dw, dh, cap, st,
                                                          (float4 *)boxes, scores, labels, anchors, counts, lv);
  YV_LAUNCH_CHECK();
  return YV_OK;
}
