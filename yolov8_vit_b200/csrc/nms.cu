// nms.cu -- K2+K3: batched custom_nms (README.md:62-83 = tech.md:72-93), one CTA per image.
//
// Phase 1  stable descending sort of the scores: LSD radix sort, 8-bit digits, on an
//          order-preserving key transform of the fp32 score; every warp owns a contiguous
//          segment, ranks its elements with __match_any_sync (stable inside the warp) and a
//          (digit, warp) prefix makes it stable across warps.  Ties therefore come out in
//          ascending original index -- torch.argsort(descending=True, stable=True), D5.
// Phase 2  greedy sweep over the SURVIVOR LIST (the reference's `sorted_indices`), 64 heads
//          at a time: the 64x64 suppression bit-matrix of the heads is built in shared
//          memory (one 64-bit row per head, 8 lanes per row, OR-reduced with shuffles), one
//          warp resolves the greedy dependency with ffs over the rows, and the kept heads are
//          then applied to the rest of the list, which is stream-compacted (warp ballot +
//          per-warp prefix) into the other list buffer -- so the work is output sensitive,
//          sum over passes of |survivors| x |kept in pass|, like the reference's loop, not
//          the N^2/2 of a full bitmask.
// IoU arithmetic = torchvision.ops.box_iou in fp32, no FMA contraction, IEEE division;
// survivors need iou < thr (strict), NaN is suppressed.  Integer outputs, bit-exact.
//
// Large candidate lists (the 30 000-box stress case, BASELINE.json configs[4]): a CLUSTER of NMS_CLUSTER CTAs per
// image.  The O(survivors x kept) part -- applying the kept heads of a pass to the rest of the list -- is split
// over the CTAs of the cluster (contiguous slices of the list, so the order is preserved); every CTA resolves the
// 64 heads redundantly (4096 IoUs, the same deterministic arithmetic, so nothing has to be broadcast), the CTAs
// exchange their survivor counts through distributed shared memory and write their compacted slices at the
// resulting offsets of the ONE global list.  Two cluster barriers per pass.
#include "common.cuh"
#include <cstdlib>
#include <mutex>

namespace {

constexpr int NMS_T = 512;
constexpr int NMS_W = NMS_T / 32;
constexpr int NMS_MAX_N = 65536;
constexpr int NMS_HEAD = 64;
constexpr int NMS_MAX_ITERS = (NMS_MAX_N / NMS_W + 31) / 32;   // ballot words per warp per pass
constexpr int NMS_CLUSTER = 4;          // CTAs per image for long lists (32 images x 4 = 128 of 148 SMs in cfg5)
constexpr int NMS_CLUSTER_MIN_N = 4096; // shorter lists stay on one CTA (two cluster barriers per pass would dominate)

__device__ __forceinline__ unsigned cluster_rank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// release / acquire at cluster scope: global and shared writes before it are visible to the cluster after it
__device__ __forceinline__ void cluster_barrier() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// store into the same shared-memory variable of CTA `cta` of the cluster
__device__ __forceinline__ void st_cluster_s32(int *p, unsigned cta, int v) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "st.shared::cluster.s32 [ra], %2;\n\t}"
      ::"r"((unsigned)__cvta_generic_to_shared(p)), "r"(cta), "r"(v)
      : "memory");
}

__device__ __forceinline__ float max_t(float a, float b) {   // torch.max: NaN propagates
  return (a != a) ? a : ((b != b) ? b : (a > b ? a : b));
}
__device__ __forceinline__ float min_t(float a, float b) {
  return (a != a) ? a : ((b != b) ? b : (a < b ? a : b));
}
__device__ __forceinline__ float area_of(const float4 &p) {
  return __fmul_rn(__fsub_rn(p.z, p.x), __fsub_rn(p.w, p.y));
}
// true iff box q is SUPPRESSED by kept box p.  rule 0 (custom_nms, README.md:77-79): not (iou < thr);
// rule 1 (torchvision.ops.nms / EfficientNMS, row f4): iou > thr (equality and NaN are kept).
__device__ __forceinline__ bool suppresses(const float4 &p, float area_p, const float4 &q, float thr, int rule = 0) {
  float area_q = area_of(q);
  float w = __fsub_rn(min_t(p.z, q.z), max_t(p.x, q.x));
  float h = __fsub_rn(min_t(p.w, q.w), max_t(p.y, q.y));
  w = (w < 0.f) ? 0.f : w;          // clamp(min=0); NaN stays NaN
  h = (h < 0.f) ? 0.f : h;
  float inter = __fmul_rn(w, h);
  float uni = __fsub_rn(__fadd_rn(area_p, area_q), inter);
  float iou = __fdiv_rn(inter, uni);
  return rule ? (iou > thr) : !(iou < thr);
}

__device__ __forceinline__ unsigned score_key_desc(float s) {
  unsigned u = __float_as_uint(s);
  if ((u << 1) == 0u) u = 0u;                                   // -0.0 == +0.0 for the sort
  unsigned asc = (u & 0x80000000u) ? ~u : (u | 0x80000000u);    // ascending-order key
  return ~asc;                                                  // ascending radix = descending score
}

struct NmsWs {
  unsigned *key[2];
  int32_t *idx[2];
  int32_t *lidx[2];
  float4 *lbox[2];
};

// CL = CTAs per image (1, or NMS_CLUSTER launched as a thread-block cluster)
template <int CL>
__global__ void __launch_bounds__(NMS_T)
nms_kernel(const float4 *boxes, const float *scores,
           const int32_t *__restrict__ counts, int stride, int n_all, float thr, int max_keep,
           int32_t *keep, int32_t *__restrict__ keep_counts, NmsWs ws, int rule) {
  const int b = blockIdx.x / CL;
  const int rank = CL > 1 ? (int)cluster_rank() : 0;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int n = counts ? counts[b] : n_all;
  n = n < 0 ? 0 : (n > stride ? stride : n);
  const size_t base = (size_t)b * stride;
  boxes += base;
  scores += base;
  keep += base;
  unsigned *key[2] = {ws.key[0] + base, ws.key[1] + base};
  int32_t *idx[2] = {ws.idx[0] + base, ws.idx[1] + base};
  int32_t *lidx[2] = {ws.lidx[0] + base, ws.lidx[1] + base};
  float4 *lbox[2] = {ws.lbox[0] + base, ws.lbox[1] + base};

  __shared__ unsigned s_whist[NMS_W][256];
  __shared__ unsigned s_dbase[256];
  __shared__ int s_skip;
  __shared__ unsigned long long s_row[NMS_HEAD];
  __shared__ float4 s_hbox[NMS_HEAD];
  __shared__ int32_t s_hidx[NMS_HEAD];
  __shared__ float4 s_kbox[NMS_HEAD];
  __shared__ float s_karea[NMS_HEAD];
  __shared__ int s_k;
  __shared__ unsigned s_ballot[NMS_W][NMS_MAX_ITERS];
  __shared__ int s_wcnt[NMS_W];
  __shared__ int s_ccnt[CL];            // survivors of every CTA of the cluster in this pass (written by its owner)

  if (n == 0) {                         // (uniform over the cluster: no barrier is skipped by a part of it)
    if (tid == 0 && rank == 0) keep_counts[b] = 0;
    return;
  }

  // ------------------------------------------------------------------ phase 1: sort (CTA 0 of the cluster)
  if (rank == 0) {
  for (int i = tid; i < n; i += NMS_T) {
    key[0][i] = score_key_desc(scores[i]);
    idx[0][i] = i;
  }
  __syncthreads();
  int cur = 0;
  {
    const int seg = ((n + NMS_W - 1) / NMS_W + 31) & ~31;   // per-warp segment, multiple of 32
    const int w_beg = min(n, warp * seg), w_end = min(n, (warp + 1) * seg);
    for (int shift = 0; shift < 32; shift += 8) {
      for (int i = tid; i < NMS_W * 256; i += NMS_T) (&s_whist[0][0])[i] = 0u;
      if (tid == 0) s_skip = 0;
      __syncthreads();
      const unsigned *kin = key[cur];
      const int32_t *iin = idx[cur];
      for (int i0 = w_beg; i0 < w_end; i0 += 32) {
        const int i = i0 + lane;
        const unsigned d = (i < w_end) ? ((kin[i] >> shift) & 255u) : (256u + lane);
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        if (i < w_end && (peers & lanemask_lt()) == 0u) s_whist[warp][d] += __popc(peers);
        __syncwarp();
      }
      __syncthreads();
      // (digit, warp) exclusive prefix: thread d < 256 walks the warps of its digit
      if (tid < 256) {
        unsigned run = 0;
#pragma unroll
        for (int w = 0; w < NMS_W; ++w) {
          unsigned t = s_whist[w][tid];
          s_whist[w][tid] = run;
          run += t;
        }
        s_dbase[tid] = run;
        if (run == (unsigned)n) s_skip = 1;      // every key has this digit: pass is a no-op
      }
      __syncthreads();
      if (s_skip) { __syncthreads(); continue; }
      if (warp == 0) {                           // exclusive scan of the 256 digit totals
        unsigned v[8], sum = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) { v[j] = s_dbase[lane * 8 + j]; sum += v[j]; }
        unsigned incl = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          unsigned t = __shfl_up_sync(0xffffffffu, incl, d);
          if (lane >= d) incl += t;
        }
        unsigned run = incl - sum;
#pragma unroll
        for (int j = 0; j < 8; ++j) { s_dbase[lane * 8 + j] = run; run += v[j]; }
      }
      __syncthreads();
      unsigned *kout = key[cur ^ 1];
      int32_t *iout = idx[cur ^ 1];
      for (int i0 = w_beg; i0 < w_end; i0 += 32) {
        const int i = i0 + lane;
        const bool valid = i < w_end;
        unsigned k = 0;
        int32_t id = 0;
        if (valid) { k = kin[i]; id = iin[i]; }
        const unsigned d = valid ? ((k >> shift) & 255u) : (256u + lane);
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        const unsigned lower = peers & lanemask_lt();
        unsigned pos = 0;
        if (valid) {
          pos = s_dbase[d] + s_whist[warp][d] + __popc(lower);
          kout[pos] = k;
          iout[pos] = id;
        }
        __syncwarp();
        if (valid && lower == 0u) s_whist[warp][d] += __popc(peers);
        __syncwarp();
      }
      cur ^= 1;
      __syncthreads();
    }
  }

  // sorted candidate list: (original index, box) in descending-score order
  for (int i = tid; i < n; i += NMS_T) {
    const int32_t id = idx[cur][i];
    lidx[0][i] = id;
    lbox[0][i] = boxes[id];
  }
  }  // rank == 0
  __syncthreads();
  if (CL > 1) cluster_barrier();        // the list is in global memory for every CTA of the cluster

  // ------------------------------------------------------------------ phase 2: greedy sweep
  const int kmax = (max_keep > 0 && max_keep < n) ? max_keep : n;
  int lc = 0, m = n, nkeep = 0;
  while (m > 0 && nkeep < kmax) {
    const int hb = m < NMS_HEAD ? m : NMS_HEAD;
    const int32_t *li = lidx[lc];
    const float4 *lb = lbox[lc];
    if (tid < hb) {
      s_hbox[tid] = lb[tid];
      s_hidx[tid] = li[tid];
    }
    __syncthreads();
    {  // 64x64 suppression rows: 8 lanes per row, 8 columns per lane, upper triangle only
      const int r = tid >> 3, part = tid & 7;
      unsigned long long bits = 0ull;
      if (r < hb) {
        const float4 p = s_hbox[r];
        const float ap = area_of(p);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int c = part * 8 + j;
          if (c > r && c < hb && suppresses(p, ap, s_hbox[c], thr, rule)) bits |= 1ull << c;
        }
      }
      bits |= __shfl_xor_sync(0xffffffffu, bits, 1);
      bits |= __shfl_xor_sync(0xffffffffu, bits, 2);
      bits |= __shfl_xor_sync(0xffffffffu, bits, 4);
      if (part == 0) s_row[r] = bits;
    }
    __syncthreads();
    if (warp == 0) {  // greedy resolve (warp-uniform), then emit the kept heads
      unsigned long long alive = (hb == 64) ? ~0ull : ((1ull << hb) - 1ull);
      unsigned long long kept = 0ull;
      int k = 0;
      const int room = kmax - nkeep;
      while (alive && k < room) {
        const int i = __ffsll((long long)alive) - 1;
        kept |= 1ull << i;
        ++k;
        alive &= ~(s_row[i] | (1ull << i));
      }
      for (int h = lane; h < hb; h += 32) {
        if (kept & (1ull << h)) {
          const int slot = __popcll(kept & ((1ull << h) - 1ull));
          if (rank == 0) keep[nkeep + slot] = s_hidx[h];
          const float4 p = s_hbox[h];
          s_kbox[slot] = p;
          s_karea[slot] = area_of(p);
        }
      }
      if (lane == 0) s_k = k;
    }
    __syncthreads();
    const int k = s_k;
    nkeep += k;
    const int rem = m - hb;
    if (rem == 0 || nkeep >= kmax) break;

    // apply the k kept heads to the rest of the list and compact the survivors: this CTA's slice of the rest,
    // cut into one contiguous segment per warp
    const int cseg = CL > 1 ? (((rem + CL - 1) / CL + 31) & ~31) : rem;
    const int c_beg = min(rem, rank * cseg), c_len = min(rem, (rank + 1) * cseg) - c_beg;
    const int seg = ((c_len + NMS_W - 1) / NMS_W + 31) & ~31;
    const int w_beg = hb + c_beg + min(c_len, warp * seg), w_end = hb + c_beg + min(c_len, (warp + 1) * seg);
    int wcnt = 0, it = 0;
    for (int i0 = w_beg; i0 < w_end; i0 += 32, ++it) {
      const int i = i0 + lane;
      bool live = i < w_end;
      if (live) {
        const float4 q = lb[i];
        for (int t = 0; t < k; ++t)
          if (suppresses(s_kbox[t], s_karea[t], q, thr, rule)) { live = false; break; }
      }
      const unsigned bal = __ballot_sync(0xffffffffu, live);
      if (lane == 0) s_ballot[warp][it] = bal;
      wcnt += __popc(bal);
    }
    if (lane == 0) s_wcnt[warp] = wcnt;
    __syncthreads();
    int off = 0, m_new = 0;
#pragma unroll
    for (int w = 0; w < NMS_W; ++w) {
      const int t = s_wcnt[w];
      if (w < warp) off += t;
      m_new += t;
    }
    if (CL > 1) {                       // cluster-wide offsets: every CTA publishes its count to all of them
      if (tid < CL) st_cluster_s32(&s_ccnt[rank], (unsigned)tid, m_new);
      cluster_barrier();
      int coff = 0, tot = 0;
#pragma unroll
      for (int c = 0; c < CL; ++c) {
        const int t = s_ccnt[c];
        if (c < rank) coff += t;
        tot += t;
      }
      off += coff;
      m_new = tot;
    }
    int32_t *lo = lidx[lc ^ 1];
    float4 *bo = lbox[lc ^ 1];
    it = 0;
    for (int i0 = w_beg; i0 < w_end; i0 += 32, ++it) {
      const unsigned bal = s_ballot[warp][it];
      if (bal & (1u << lane)) {
        const int pos = off + __popc(bal & lanemask_lt());
        lo[pos] = li[i0 + lane];
        bo[pos] = lb[i0 + lane];
      }
      off += __popc(bal);
    }
    __syncthreads();
    if (CL > 1) cluster_barrier();      // the new list is complete (and s_ccnt may be rewritten)
    lc ^= 1;
    m = m_new;
  }
  if (tid == 0 && rank == 0) keep_counts[b] = nkeep;
}

struct NmsLayout {
  size_t o_key[2], o_idx[2], o_lidx[2], o_lbox[2], total;
};
NmsLayout nms_layout(int B, int stride) {
  NmsLayout L;
  size_t n = (size_t)B * stride, o = 0;
  for (int i = 0; i < 2; ++i) { L.o_lbox[i] = o; o = yv_align_up(o + n * sizeof(float4), 256); }
  for (int i = 0; i < 2; ++i) { L.o_key[i] = o; o = yv_align_up(o + n * sizeof(unsigned), 256); }
  for (int i = 0; i < 2; ++i) { L.o_idx[i] = o; o = yv_align_up(o + n * sizeof(int32_t), 256); }
  for (int i = 0; i < 2; ++i) { L.o_lidx[i] = o; o = yv_align_up(o + n * sizeof(int32_t), 256); }
  L.total = o;
  return L;
}

}  // namespace

YV_EXPORT size_t yv_nms_workspace_bytes(int B, int stride) {
  if (B <= 0 || stride <= 0) return 0;
  return nms_layout(B, stride).total;
}

YV_EXPORT int yv_nms_batched_rule(const float *boxes, const float *scores, const int32_t *counts, int B,
                                  int stride, int n_all, float iou_thr, int max_keep, int rule, int32_t *keep,
                                  int32_t *keep_counts, void *ws, size_t ws_bytes, yv_stream_t stream) {
  YV_REQUIRE(rule == 0 || rule == 1, "yv_nms_batched: unknown rule %d", rule);
  YV_REQUIRE(B >= 0 && stride >= 1, "yv_nms_batched: bad shape B=%d stride=%d", B, stride);
  if (B == 0) return YV_OK;
  YV_REQUIRE(boxes && scores && keep && keep_counts && ws, "yv_nms_batched: null pointer");
  YV_REQUIRE(((uintptr_t)boxes % 16) == 0, "yv_nms_batched: boxes must be 16-byte aligned");
  YV_REQUIRE(counts || (n_all >= 0 && n_all <= stride), "yv_nms_batched: n_all=%d outside [0, stride=%d]", n_all, stride);
  YV_REQUIRE(stride <= NMS_MAX_N, "yv_nms_batched: stride=%d exceeds %d candidates per image", stride, NMS_MAX_N);
  const NmsLayout L = nms_layout(B, stride);
  if (ws_bytes < L.total) {
    yv_set_error("yv_nms_batched: workspace %zu < required %zu", ws_bytes, L.total);
    return YV_ENOMEM;
  }
  char *w = (char *)ws;
  NmsWs W;
  for (int i = 0; i < 2; ++i) {
    W.key[i] = (unsigned *)(w + L.o_key[i]);
    W.idx[i] = (int32_t *)(w + L.o_idx[i]);
    W.lidx[i] = (int32_t *)(w + L.o_lidx[i]);
    W.lbox[i] = (float4 *)(w + L.o_lbox[i]);
  }
  YvProf prof(YV_K_NMS, (cudaStream_t)stream, counts ? 0.0 : (double)B * n_all * 20.0);
  // long lists: a cluster of NMS_CLUSTER CTAs per image (YV_NMS_CLUSTER=0 / 1 forces one CTA / the cluster, A/B timing)
  static const int force = []() { const char *e = getenv("YV_NMS_CLUSTER"); return e ? atoi(e) : -1; }();
  // (with per-image `counts` the list lengths live on the device -- the detector path, hundreds of candidates per
  // image -- and one CTA per image is right; the cluster is for callers that state a long n_all)
  const bool cluster = force >= 0 ? force != 0 : (!counts && n_all >= NMS_CLUSTER_MIN_N);
  if (cluster) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)B * NMS_CLUSTER);
    cfg.blockDim = dim3(NMS_T);
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = NMS_CLUSTER; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    YV_CUDA(cudaLaunchKernelEx(&cfg, nms_kernel<NMS_CLUSTER>, (const float4 *)boxes, scores, counts, stride, n_all,
                               iou_thr, max_keep, keep, keep_counts, W, rule));
  } else {
    nms_kernel<1><<<B, NMS_T, 0, (cudaStream_t)stream>>>((const float4 *)boxes, scores, counts, stride, n_all,
                                                         iou_thr, max_keep, keep, keep_counts, W, rule);
  }
  YV_LAUNCH_CHECK();
  return YV_OK;
}

YV_EXPORT int yv_nms_batched(const float *boxes, const float *scores, const int32_t *counts, int B,
                             int stride, int n_all, float iou_thr, int max_keep, int32_t *keep,
                             int32_t *keep_counts, void *ws, size_t ws_bytes, yv_stream_t stream) {
  return yv_nms_batched_rule(boxes, scores, counts, B, stride, n_all, iou_thr, max_keep, YV_NMS_RULE_CUSTOM, keep,
                             keep_counts, ws, ws_bytes, stream);
}

// ------------------------------------------------------------------------------------ no-allocation variant
// yv_nms_batched_noalloc: the same kernels with their scratch (radix keys / indices and the two survivor lists,
// 56 B per box) in a file-scope __device__ array instead of a caller workspace, for callers that cannot allocate
// (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Every GPU that loads the library holds the 4 MiB until exit.
// In HOST code `g_nms_static_ws` is only a stand-in for the symbol, not a device address: the address is fetched
// once per device with cudaGetSymbolAddress (with that device current) and passed to the kernel.
static_assert(YV_NMS_NOALLOC_BYTES == (4u << 20), "include/yv8vit.h states 4 MiB");
static __device__ __align__(256) unsigned char g_nms_static_ws[YV_NMS_NOALLOC_BYTES];

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
void *g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};

// device address of the scratch array on the CURRENT device (looked up once per device)
int nms_static_ws(void **out) {
  int dev = -1;
  YV_CUDA(cudaGetDevice(&dev));
  YV_REQUIRE(dev >= 0 && dev < NOALLOC_MAX_DEVICES, "yv_nms_batched_noalloc: device %d out of range", dev);
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    void *p = nullptr;
    YV_CUDA(cudaGetSymbolAddress(&p, g_nms_static_ws));
    YV_REQUIRE(p != nullptr && ((uintptr_t)p % 256) == 0, "yv_nms_batched_noalloc: bad scratch address %p", p);
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return YV_OK;
}
}  // namespace

YV_EXPORT int yv_nms_batched_noalloc(const float *boxes, const float *scores, const int32_t *counts, int B,
                                     int stride, int n_all, float iou_thr, int max_keep, int32_t *keep,
                                     int32_t *keep_counts, yv_stream_t stream) {
  YV_REQUIRE(B >= 0 && stride >= 1, "yv_nms_batched_noalloc: bad shape B=%d stride=%d", B, stride);
  if (B == 0) return YV_OK;
  YV_REQUIRE(boxes && scores && keep && keep_counts, "yv_nms_batched_noalloc: null pointer");
  YV_REQUIRE(((uintptr_t)boxes % 16) == 0, "yv_nms_batched_noalloc: boxes must be 16-byte aligned");
  YV_REQUIRE(counts || (n_all >= 0 && n_all <= stride), "yv_nms_batched_noalloc: n_all=%d outside [0, stride=%d]", n_all, stride);
  YV_REQUIRE(stride <= NMS_MAX_N, "yv_nms_batched_noalloc: stride=%d exceeds %d candidates per image", stride, NMS_MAX_N);
  const NmsLayout L = nms_layout(B, stride);
  if (L.total > (size_t)YV_NMS_NOALLOC_BYTES) {
    yv_set_error("yv_nms_batched_noalloc: B=%d x stride=%d needs %zu B of scratch, the built-in array holds %u B "
                 "(about 65 000 boxes per call); use yv_nms_batched with a caller workspace",
                 B, stride, L.total, (unsigned)YV_NMS_NOALLOC_BYTES);
    return YV_EINVAL;
  }
  void *ws = nullptr;
  int rc = nms_static_ws(&ws);
  if (rc) return rc;
  char *w = (char *)ws;
  NmsWs W;
  for (int i = 0; i < 2; ++i) {
    W.key[i] = (unsigned *)(w + L.o_key[i]);
    W.idx[i] = (int32_t *)(w + L.o_idx[i]);
    W.lidx[i] = (int32_t *)(w + L.o_lidx[i]);
    W.lbox[i] = (float4 *)(w + L.o_lbox[i]);
  }
  // The kernel writes every scratch element before it reads it (sort keys, then the lists), so nothing has to be
  // cleared between calls.  No launch profiler here: its event pool may create events (an allocation).
  const bool cluster = !counts && n_all >= NMS_CLUSTER_MIN_N;      // the rule of yv_nms_batched
  if (cluster) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)B * NMS_CLUSTER);
    cfg.blockDim = dim3(NMS_T);
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = NMS_CLUSTER; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    YV_CUDA(cudaLaunchKernelEx(&cfg, nms_kernel<NMS_CLUSTER>, (const float4 *)boxes, scores, counts, stride, n_all,
                               iou_thr, max_keep, keep, keep_counts, W, (int)YV_NMS_RULE_CUSTOM));
  } else {
    nms_kernel<1><<<B, NMS_T, 0, (cudaStream_t)stream>>>((const float4 *)boxes, scores, counts, stride, n_all,
                                                         iou_thr, max_keep, keep, keep_counts, W, (int)YV_NMS_RULE_CUSTOM);
  }
  YV_LAUNCH_CHECK();
  return YV_OK;
}
