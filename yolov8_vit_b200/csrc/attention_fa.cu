// attention_fa.cu -- K7: softmax(Q K^T / sqrt(dh)) V for ANY sequence length on tcgen05 / TMEM, fed by TMA.
//
// timm Attention.forward (vit_base_patch{16,8}_224; reached by the reference through utils/utils.py:71-72):
// qkv [n, T, 3, heads, 64] -> per head softmax(q k^T * 64^-0.5) v.  One kernel serves ViT-B/16 (T = 197: two key
// blocks of 112 + 96) and the reference-literal ViT-B/8 (T = 785: six key blocks of 128 + one of 32).
//
// Flash-style: the keys are cut into blocks of <= 128 with an online softmax across the blocks, which makes the
// kernel independent of the sequence length.  What bounds a softmax on this machine is the MUFU pipe: 16 ex2 per
// clock per SM (measured, tools/ubench_tmem.cu, profiles/r02_ubench_tmem.md), i.e. 1664 clocks for the 26 624
// exponentials of a 128 x 208 tile, against ~530 clocks to read its fp32 scores TWICE from tensor memory
// (tcgen05.ld.32x32b.x16: 400 B/clock/SM; the .x32 shape only reaches 115) and ~830 clocks of tensor pipe.  The
// round-1 kernel ran at 3.7 k clocks per tile because all 16 softmax warps worked on ONE tile in lock step (max
// pass, barrier, exp pass: the MUFU pipe idles during the first half and nothing hides the dependent chain).  Here
// TWO query tiles are in flight per SM, each owned by its own group of 8 warps, so one tile's MUFU-bound exp pass
// overlaps the other's load / max / exchange / store.  The scores are read twice (max pass, exp pass): holding 64
// of them per thread between the passes was tried first and cost more (96-register cap at 640 threads: spills,
// S2R re-reads, setmaxnreg juggling) than the second read.
//
// One persistent CTA per SM, 20 warps.  A work item is a (crop, head, PAIR of 128-row query tiles): the pair shares
// every K / V block (loaded once).  Slot s in {0, 1} = query tile 2 pair + s, owned by softmax group s.
//   warp 0      TMA producer: Q tiles (2-stage ring over items), K / V blocks (4-stage ring)
//   warp 1      score-MMA issuer: S = Q K^T of entry e = (item, block, slot) into score buffer e % 3.  Tensor memory
//               holds three 128-column score buffers that rotate over the entries in issue order, plus one 64-column
//               O accumulator per slot = 512 columns, so S runs up to three entries ahead of the softmax.
//   warps 2, 3  P V issuers, one per slot: O_s (+)= P V once the group has written P (A operand from tensor memory).
//   warps 4-11  softmax group 0, warps 12-19 group 1: 4 TMEM lane quarters x 2 key halves.  Per block a thread
//               takes the maximum of its <= 64 scores (tcgen05.ld x16, FMNMX3), exchanges it with the thread that
//               holds the other half of the row (shared memory + a 64-thread named barrier), reads the scores
//               again, exponentiates (FFMA2 + MUFU.EX2), accumulates the row sum and writes P as packed bf16 back
//               INTO the head of its own score columns (P never touches shared memory).
//               Online softmax: the running maximum is only raised when a block exceeds it by more than 2^8
//               (P then stays <= 256, exact in bf16 / fp32); raising it rescales O in tensor memory
//               (tcgen05.ld / st, decided per warp, rare after the first block).  A finished tile is drained
//               (O / row sum -> bf16 -> global) inside the first block of the slot's NEXT tile, when its last
//               P V has long retired.
// Warps whose 32 query rows are all past T (the tail tile) only pass the barriers on.
//
// Which lengths run here (yv_attention_launch, attention_sm100.cu): every shape the round-1 pipelined kernel does
// not take -- ViT-B/8 and any T > 224.  For ViT-B/16 itself this kernel measured 0.43 - 0.47 ms per 268 crops
// against 0.38 ms for attention_pipe_kernel (tools/attn_bench.py), so that one stays the default there and
// YV_ATTN_B16=fa selects this kernel for A/B; the parity tests run both.
#include <cstdlib>
#include "common.cuh"
#include "gemm_sm100.cuh"
#include "sm100_ptx.cuh"
#include "vit_ops.cuh"

using namespace sm100;

namespace {

constexpr int FA_DH = 64;
constexpr int FA_THREADS = 20 * 32;                 // TMA, S issuer, 2 P V issuers + 2 groups x 8 softmax warps
constexpr int FA_TILE_BYTES = 128 * FA_DH * 2;      // one 128-row operand tile: 16 KB
constexpr int FA_Q_STAGES = 2;
// Two layouts of tensor / shared memory (template parameter WIDE):
//   general  key blocks of <= 128: three 128-column score buffers rotating over the entries + one O accumulator per
//            slot (3 x 128 + 2 x 64 = 512 columns); K / V ring of 4 stages x (16 + 16) KB.
//   WIDE     16 T' <= 208 keys and two query tiles (ViT-B/16: T = 197): ONE key block per tile, so a tile is one
//            entry (half the barrier / MMA-issue round trips, no online rescale) in a 208-column score buffer per
//            slot, and ONE O accumulator that the slots use in turn (2 x 208 + 64 = 480 columns); K / V ring of
//            2 stages x (26 + 26) KB.
constexpr int FA_MAX_KV_STAGES = 4;
constexpr int FA_MAX_SBUF = 3;
constexpr int FA_WIDE_KEYS = 208;
constexpr float FA_RESCALE_THR = 8.0f;              // log2 units

struct FaShared {
  uint64_t q_full[FA_Q_STAGES], q_empty[FA_Q_STAGES];
  uint64_t kv_full[FA_MAX_KV_STAGES], kv_empty[FA_MAX_KV_STAGES];
  uint64_t s_full[FA_MAX_SBUF], pv_done[FA_MAX_SBUF];
  uint64_t p_full[2][FA_MAX_SBUF];       // per slot: its waiter (the slot's P V issuer) sees EVERY phase of its barriers
  uint64_t o_done[2], o_free[2];
  uint32_t tmem_slot, pad;
  float xmax[2][2][2][128];          // [group][block parity][key half][row]
  float xsum[2][2][128];             // [group][key half][row]
};
constexpr int fa_kv_stages(bool wide) { return wide ? 2 : 4; }
constexpr int fa_kv_tile(bool wide) { return wide ? FA_WIDE_KEYS * FA_DH * 2 : FA_TILE_BYTES; }        // 26 KB / 16 KB
constexpr int fa_smem(bool wide) {
  return FA_Q_STAGES * 2 * FA_TILE_BYTES + fa_kv_stages(wide) * 2 * fa_kv_tile(wide) + (int)sizeof(FaShared) + 1024;
}

// MN-major B operand (V: keys x 64, 128-byte swizzled rows of dh): 8-key groups are 1024 B apart
__device__ __forceinline__ uint64_t fa_desc_mn_sw128(uint32_t smem_addr) {
  const uint32_t lo = ((smem_addr >> 4) & 0x3FFFu) | ((1024u >> 4) << 16);
  const uint32_t hi = (1024u >> 4) | (1u << 14) | (2u << 29);
  return ((uint64_t)hi << 32) | lo;
}
// kind::f16, D=f32, A=bf16 K-major (or from TMEM), B=bf16 MN-major
__host__ __device__ constexpr uint32_t fa_idesc_bmn(int M, int N) { return umma_idesc_bf16(M, N) | (1u << 16); }

__device__ __forceinline__ float fa_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fa_max3(float a, float b, float c) {
  float y;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(y) : "f"(a), "f"(b), "f"(c));
  return y;
}
__device__ __forceinline__ uint32_t fa_pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t *>(&h);
}
// mbar_wait with a name: a protocol bug prints which wait of which warp starved before the launch is failed
// (bounded spin, like sm100::mbar_wait -- a hang would cost a GPU box)
__device__ int fa_timeout_reports = 0;
__device__ __noinline__ void fa_wait_timeout(int tag, uint32_t parity) {
  if (atomicAdd(&fa_timeout_reports, 1) < 48)
    printf("attention_fa: wait %d (parity %u) timed out: cta %d warp %d lane %d\n", tag, parity, (int)blockIdx.x,
         (int)(threadIdx.x >> 5), (int)(threadIdx.x & 31));
}
// The wait carries a suspend-time hint: the warp sleeps inside try_wait until the phase completes (or the hint
// expires) instead of spinning -- every SM sub-partition hosts warps of BOTH groups, and a group that polls steals
// issue slots from the one that is computing (ncu: half of the executed instructions were wait loops).
__device__ __forceinline__ void fa_wait(uint64_t *bar, uint32_t parity, int tag) {
  uint32_t spins = 0, ok = 0;
  for (;;) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity), "r"(4000u)
        : "memory");
    if (ok) break;
    ++spins;
    if (spins == (1u << 18)) fa_wait_timeout(tag, parity);      // report, give the other starved waits time to report too
    if (spins > (1u << 22)) __trap();
  }
}
enum { W_Q_EMPTY = 1, W_KV_EMPTY, W_PV_DONE, W_Q_FULL, W_KV_FULL, W_P_FULL, W_O_FREE, W_S_FULL, W_O_DONE_BLK, W_O_DONE_DRAIN, W_O_FREE_DEAD };

struct FaShape {
  int T, heads;
  int nkb;            // key blocks
  int blk0, blk_last; // keys of the first nkb - 1 blocks / of the last block (multiples of 16, <= 128)
  int nq, npair;      // 128-row query tiles per (crop, head); pairs of them
};

// Drain of one finished tile by one softmax warp: O (this thread: 32 of the 64 head-dim columns of its row) / row sum
// -> bf16 -> global, and the warp's arrival on o_free.  Every warp of the group arrives once per tile, in the tile's
// own phase: a warp that has no rows in the tile runs ahead of the others, so it first waits until the previous
// tile's phase has completed (the MMA warp is waiting for the same phase before the slot's next first P V).
// `freed` is the o_free barrier the 8 warps of the group arrive on: the slot's own (general layout: its P V issuer
// waits for it before the slot's next tile) or the OTHER slot's (WIDE: the shared accumulator is handed over).
__device__ __forceinline__ void fa_drain(FaShared *bar, uint64_t *freed, int g, int lane, bool live, uint32_t ntile,
                                         uint32_t phase, uint32_t o_addr, float inv, __nv_bfloat16 *dst_row) {
  if (!live) {
    if (ntile > 1u) fa_wait(freed, (ntile - 2u) & 1u, W_O_FREE_DEAD);
    __syncwarp();
    if (lane == 0) mbar_arrive(freed);
    return;
  }
  fa_wait(&bar->o_done[g], phase, W_O_DONE_DRAIN);               // the tile's last P V has retired
  tc_fence_after();
  uint32_t o[32];                                  // two x16 loads: the x32 shape runs at a quarter of the x16 rate
  tmem_ld_32x16(o_addr, *reinterpret_cast<uint32_t(*)[16]>(&o[0]));
  tmem_ld_32x16(o_addr + 16, *reinterpret_cast<uint32_t(*)[16]>(&o[16]));
  tmem_ld_wait();
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive(freed);               // the accumulator may be overwritten by the next tile that uses it
  if (dst_row) {
    uint4 *dst = reinterpret_cast<uint4 *>(dst_row);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      uint4 u;
      u.x = fa_pack_bf16x2(__uint_as_float(o[8 * k + 0]) * inv, __uint_as_float(o[8 * k + 1]) * inv);
      u.y = fa_pack_bf16x2(__uint_as_float(o[8 * k + 2]) * inv, __uint_as_float(o[8 * k + 3]) * inv);
      u.z = fa_pack_bf16x2(__uint_as_float(o[8 * k + 4]) * inv, __uint_as_float(o[8 * k + 5]) * inv);
      u.w = fa_pack_bf16x2(__uint_as_float(o[8 * k + 6]) * inv, __uint_as_float(o[8 * k + 7]) * inv);
      dst[k] = u;
    }
  }
}

template <bool WIDE>
__global__ void __launch_bounds__(FA_THREADS, 1)
attention_fa_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV0,
                    const __grid_constant__ CUtensorMap tmKV1, __nv_bfloat16 *__restrict__ out, int n_items,
                    const FaShape sh, float scale_log2e) {
  extern __shared__ __align__(1024) uint8_t fa_raw[];
  uint8_t *smem = fa_raw;
  if ((smem_u32(smem) & 1023u) != 0u) __trap();                  // 128B-swizzled tiles need 1024-byte alignment
  uint8_t *sQ = smem;                                            // [stage][slot] 16 KB
  constexpr int FA_SBUF = WIDE ? 2 : 3;                          // score buffers (entry e uses buffer e % FA_SBUF)
  constexpr int FA_SCOLS = WIDE ? FA_WIDE_KEYS : 128;            // TMEM columns per score buffer
  constexpr int FA_OCOL = FA_SBUF * FA_SCOLS;                    // O accumulator(s) behind the score buffers
  constexpr int FA_KV_STAGES = fa_kv_stages(WIDE);
  constexpr int FA_KV_TILE = fa_kv_tile(WIDE);
  uint8_t *sKV = smem + FA_Q_STAGES * 2 * FA_TILE_BYTES;         // [stage] K | V
  FaShared *bar = reinterpret_cast<FaShared *>(sKV + FA_KV_STAGES * 2 * FA_KV_TILE);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int T = sh.T, NKB = sh.nkb, D = sh.heads * FA_DH;
  const int my_items = (n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmQ);
    prefetch_tmap(&tmKV0);
    prefetch_tmap(&tmKV1);
  }
  if (warp == 0 && lane == 0) {
    for (int i = 0; i < FA_Q_STAGES; ++i) { mbar_init(&bar->q_full[i], 1); mbar_init(&bar->q_empty[i], 1); }
    for (int i = 0; i < FA_MAX_KV_STAGES; ++i) { mbar_init(&bar->kv_full[i], 1); mbar_init(&bar->kv_empty[i], 2); }   // both P V issuers
    for (int i = 0; i < FA_MAX_SBUF; ++i) { mbar_init(&bar->s_full[i], 1); mbar_init(&bar->p_full[0][i], 8); mbar_init(&bar->p_full[1][i], 8); mbar_init(&bar->pv_done[i], 1); }   // p_full: 8 warps of a group
    for (int i = 0; i < 2; ++i) { mbar_init(&bar->o_done[i], 1); mbar_init(&bar->o_free[i], 8); }
    fence_barrier_init();
  }
  if (warp == 1) {                                               // (the whole warp; one lane of it issues the MMAs later)
    tmem_alloc(&bar->tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bar->tmem_slot;

  // item -> (crop, head, first query tile); slot 1 of the last pair of an odd tile count does not exist
  auto item_of = [&](int i, int &crop, int &head, int &tile0) {
    const int item = (int)blockIdx.x + i * (int)gridDim.x;
    const int ch = item / sh.npair, pair = item - ch * sh.npair;
    crop = ch / sh.heads;
    head = ch - crop * sh.heads;
    tile0 = 2 * pair;
  };
  auto keys_of = [&](int j) { return j == NKB - 1 ? sh.blk_last : sh.blk0; };

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      for (int i = 0; i < my_items; ++i) {
        int crop, head, tile0;
        item_of(i, crop, head, tile0);
        const int row0 = crop * T;
        const int qs = i & 1;
        const bool two = tile0 + 1 < sh.nq;
        fa_wait(&bar->q_empty[qs], (((uint32_t)i >> 1) & 1u) ^ 1u, W_Q_EMPTY);          // S MMAs of item i - 2 retired
        mbar_arrive_expect_tx(&bar->q_full[qs], (two ? 2u : 1u) * FA_TILE_BYTES);
        tma_load_2d(sQ + (qs * 2 + 0) * FA_TILE_BYTES, &tmQ, &bar->q_full[qs], head * FA_DH, row0 + tile0 * 128);
        if (two) tma_load_2d(sQ + (qs * 2 + 1) * FA_TILE_BYTES, &tmQ, &bar->q_full[qs], head * FA_DH, row0 + (tile0 + 1) * 128);
        for (int j = 0; j < NKB; ++j) {
          const uint32_t g = (uint32_t)(i * NKB + j);
          const int st = (int)(g % FA_KV_STAGES);
          fa_wait(&bar->kv_empty[st], ((g / FA_KV_STAGES) & 1u) ^ 1u, W_KV_EMPTY);       // P V MMAs of the stage's last user retired
          const CUtensorMap *tm = (j == NKB - 1) ? &tmKV1 : &tmKV0;
          mbar_arrive_expect_tx(&bar->kv_full[st], 2u * (uint32_t)keys_of(j) * FA_DH * 2u);
          uint8_t *sK = sKV + st * 2 * FA_KV_TILE;
          tma_load_2d(sK, tm, &bar->kv_full[st], D + head * FA_DH, row0 + j * sh.blk0);
          tma_load_2d(sK + FA_KV_TILE, tm, &bar->kv_full[st], 2 * D + head * FA_DH, row0 + j * sh.blk0);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== score-MMA issuer =====================
    // Entries e = (item i, block j, slot s) in the order ((i * NKB + j) * 2 + s); entry e uses score buffer e % 3, so
    // S runs up to three entries ahead of the softmax.  Entries of a slot that does not exist are passed through
    // (empty commits) so that every barrier sees one completion per entry and the phase of entry e is (e / 3) & 1
    // on every side.  The MMA issue is split over THREE threads (this one for S = Q K^T, one per slot for P V):
    // with one thread for everything the ~270 dependent instructions it needs per entry (mbarrier polls, uniform-
    // register descriptor arithmetic, 11 UTCHMMA) took ~2.3 k clocks per entry, 9 k per (crop, head) -- three
    // times the MUFU time of the softmax it feeds (ncu: the issuing warp never waited, the softmax warps did).
    // tcgen05.mma from different threads is not ordered, so S(e) waits for the COMMIT of P V(e - 3) (pv_done),
    // which read P from the buffer S(e) overwrites.
    if (elect_one()) {
      const uint32_t q_base = smem_u32(sQ), kv_base = smem_u32(sKV);
      int buf = 0, st = 0;
      uint32_t bph = 0u, kph = 0u, e = 0u;
      for (int i = 0; i < my_items; ++i) {
        bool two = sh.nq >= 2;
        if (sh.npair > 1) two = 2 * (((int)blockIdx.x + i * (int)gridDim.x) % sh.npair) + 1 < sh.nq;
        const uint32_t q_addr = q_base + (uint32_t)((i & 1) * 2) * FA_TILE_BYTES;
        for (int j = 0; j < NKB; ++j) {
          const uint32_t idesc = umma_idesc_bf16(128, keys_of(j));
          const uint64_t dk = umma_desc_k_sw128(kv_base + (uint32_t)st * 2 * FA_KV_TILE);
#pragma unroll
          for (int sl = 0; sl < 2; ++sl) {
            if (e >= (uint32_t)FA_SBUF) fa_wait(&bar->pv_done[buf], bph ^ 1u, W_PV_DONE);   // P V(e - 3) has read its P
            if (sl == 0 || two) {
              if (j == 0 && sl == 0) fa_wait(&bar->q_full[i & 1], ((uint32_t)i >> 1) & 1u, W_Q_FULL);
              if (sl == 0) fa_wait(&bar->kv_full[st], kph, W_KV_FULL);
              tc_fence_after();
              const uint64_t dq = umma_desc_k_sw128(q_addr + (uint32_t)sl * FA_TILE_BYTES);
              const uint32_t d = tmem_base + (uint32_t)buf * FA_SCOLS;
#pragma unroll
              for (int k = 0; k < FA_DH / 16; ++k)
                umma_bf16(d, dq + (uint64_t)(k * 2), dk + (uint64_t)(k * 2), idesc, (uint32_t)(k != 0));
            }
            umma_commit(&bar->s_full[buf]);
            if (j == NKB - 1 && sl == 1) umma_commit(&bar->q_empty[i & 1]);        // every S MMA of the item is issued
            if (++buf == FA_SBUF) { buf = 0; bph ^= 1u; }
            ++e;
          }
          if (++st == FA_KV_STAGES) { st = 0; kph ^= 1u; }
        }
      }
    }
  } else if (warp < 4) {
    // ===================== P V issuer of slot warp - 2 =====================
    if (elect_one()) {
      const int sl = warp - 2;
      const uint32_t kv_base = smem_u32(sKV);
      const uint32_t idesc_o = fa_idesc_bmn(128, FA_DH);
      const uint32_t o = tmem_base + FA_OCOL + (WIDE ? 0u : (uint32_t)sl * FA_DH);      // WIDE: one accumulator, used in turn
      // entry e = 2 * block + sl uses score buffer e % 3; the (slot, buffer) pair recurs every 3 blocks, so the phase
      // of p_full[sl][buf] is (block / 3) & 1.  (One p_full per buffer shared by both slots was wrong: a parity wait
      // is only exact if the waiter has seen the previous phase complete, and this thread would skip the other
      // slot's phases -- it ran through a barrier whose previous phase was still open.)
      int buf = sl, st = 0, k3 = 0;
      uint32_t bph = 0u, tiles = 0u;
      for (int i = 0; i < my_items; ++i) {
        bool exists = sl == 0 || sh.nq >= 2;
        if (sl == 1 && sh.npair > 1) exists = 2 * (((int)blockIdx.x + i * (int)gridDim.x) % sh.npair) + 1 < sh.nq;
        for (int j = 0; j < NKB; ++j) {
          fa_wait(&bar->p_full[sl][buf], bph, W_P_FULL);
          if (exists) {
            if (j == 0) {                            // O of the slot's previous tile has been drained
              // general: the group has drained the slot's previous tile.  WIDE: the OTHER group has drained the tile
              // that used the shared accumulator before this one (slot 1 follows slot 0 of the same item, slot 0
              // follows slot 1 of the previous item -- so slot 1 really waits for its first phase, slot 0 does not)
              fa_wait(&bar->o_free[sl], (WIDE && sl == 1) ? (tiles & 1u) : ((tiles & 1u) ^ 1u), W_O_FREE);
              ++tiles;
            }
            tc_fence_after();
            uint64_t dv = fa_desc_mn_sw128(kv_base + (uint32_t)st * 2 * FA_KV_TILE + FA_KV_TILE);
            uint32_t pc = tmem_base + (uint32_t)buf * FA_SCOLS;
            const int nch = keys_of(j) >> 4, hc = (nch + 1) >> 1;
            // P chunk k (16 keys = 8 packed columns) sits at the head of the score columns of the half that wrote it:
            // column 8 k for the first hc chunks, 16 hc + 8 (k - hc) for the rest
            for (int k = 0; k < nch; ++k, dv += 128) {                         // +16 keys = +2048 B = +128 in (addr >> 4)
              umma_bf16_ts(o, pc, dv, idesc_o, (uint32_t)((j | k) != 0));
              pc += (k + 1 == hc) ? (uint32_t)(8 * hc + 8) : 8u;
            }
            umma_commit(&bar->o_done[sl]);
            umma_commit(&bar->pv_done[buf]);         // the score buffer may be overwritten by S(e + 3)
            umma_commit(&bar->kv_empty[st]);         // (2 arrivals: both slots have consumed the block)
          } else {
            // a query tile that does not exist: plain arrivals.  (A tcgen05.commit from a thread that has not issued
            // any MMA yet -- the last pair of an odd tile count on a CTA with one item -- never arrived.)  No
            // reordering against this thread's earlier commits is possible: consecutive arrivals on either
            // barrier are separated by a full S / softmax / TMA dependency chain.
            mbar_arrive(&bar->pv_done[buf]);
            mbar_arrive(&bar->kv_empty[st]);
          }
          buf += 2;                                  // e += 2
          if (buf >= FA_SBUF) buf -= FA_SBUF;
          if (++k3 == (WIDE ? 1 : 3)) { k3 = 0; bph ^= 1u; }    // the (slot, buffer) pair recurs every 3 blocks (WIDE: every block)
          if (++st == FA_KV_STAGES) st = 0;
        }
      }
    }
  } else {
    // ===================== softmax groups =====================
    const int g = (warp - 4) >> 3;                   // group = slot
    const int h = ((warp - 4) >> 2) & 1;             // key half of the block
    const int q = warp & 3;                          // TMEM lane quarter a warp may access = warp id % 4; the four warps
                                                     // of a (group, half) are consecutive, so they cover all quarters
    const int trow = q * 32 + lane;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
    const uint32_t o_addr = lane_addr + FA_OCOL + (WIDE ? 0u : (uint32_t)g * FA_DH) + (uint32_t)h * 32;
    const uint64_t sc2 = pk2(scale_log2e, scale_log2e);
    const int bar_id = 1 + g * 4 + q;                // 64-thread named barrier of the two warps that share these rows
    uint32_t nblk = 0;                               // blocks of existing tiles this slot has seen (o_done phases, xmax parity)
    uint32_t ntile = 0;                              // existing tiles this slot has seen (o_free phases)
    uint64_t *freed = &bar->o_free[WIDE ? (g ^ 1) : g];
    // The drain of a finished tile (wait for its last P V, O -> registers -> bf16 -> global) is DEFERRED into the
    // first block of the slot's next tile: by then the MMA has long retired, so the group never sits out the
    // MMA-issue + tensor-pipe round trip at the end of a tile.
    bool pend = false, pend_live = false;            // a tile of this slot waits to be drained / this warp has rows in it
    float pend_inv = 0.f;
    uint32_t pend_phase = 0;                         // o_done phase of its last block
    __nv_bfloat16 *pend_dst = nullptr;               // this thread's output row (null: the row is past T)
    for (int i = 0; i < my_items; ++i) {
      int crop, head, tile0;
      item_of(i, crop, head, tile0);
      const int tile = tile0 + g;
      const bool exists = tile < sh.nq;
      const bool live = exists && tile * 128 + q * 32 < T;                     // warp-uniform: some row of this warp is a token
      float m_used = 0.f, l = 0.f;
      for (int j = 0; j < NKB; ++j) {
        const int e = (i * NKB + j) * 2 + g;
        const uint32_t sbuf = lane_addr + (uint32_t)(e % FA_SBUF) * FA_SCOLS;
        fa_wait(&bar->s_full[e % FA_SBUF], ((uint32_t)e / FA_SBUF) & 1u, W_S_FULL);
        if (live) {
        tc_fence_after();
        const int nch = keys_of(j) >> 4;                                       // 16-key chunks of the block
        const int c0 = h ? (nch + 1) >> 1 : 0, c1 = h ? nch : (nch + 1) >> 1;  // this thread's chunks (<= 4)
        const int key0 = j * sh.blk0;
        const int pbase = 16 * c0;                   // P of this half goes to the head of ITS OWN score columns: the other
                                                     // half still reads its scores (second pass) while this one stores
        // ---- pass 1: row maximum over this thread's chunks (tensor memory is read twice: 400 B/clock/SM make the
        // second read cheaper than the 64 registers -- spills, setmaxnreg -- a single read costs)
        float mx = -INFINITY;
        {
          uint32_t r[2][16];
          auto reduce = [&](const uint32_t(&v)[16], int c) {
            if (key0 + 16 * c + 16 > T) {                                      // keys >= T are padding
#pragma unroll
              for (int k = 0; k < 16; ++k)
                if (key0 + 16 * c + k < T) mx = fmaxf(mx, __uint_as_float(v[k]));
            } else {
#pragma unroll
              for (int k = 0; k < 16; k += 2) mx = fa_max3(mx, __uint_as_float(v[k]), __uint_as_float(v[k + 1]));
            }
          };
#pragma unroll 1
          for (int c = c0; c < c1; c += 2) {         // two chunks per trip, both loads in flight before one wait
            const bool two = c + 1 < c1;
            tmem_ld_32x16(sbuf + 16 * c, r[0]);
            if (two) tmem_ld_32x16(sbuf + 16 * c + 16, r[1]);
            tmem_ld_wait();
            reduce(r[0], c);
            if (two) reduce(r[1], c + 1);
          }
        }
        // ---- row maximum of the block: exchange with the thread holding the other key half
        bar->xmax[g][nblk & 1u][h][trow] = mx;
        asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory");
        mx = fmaxf(mx, bar->xmax[g][nblk & 1u][h ^ 1][trow]);
        // ---- online softmax: raise the running maximum only when the block exceeds it by more than 2^THR
        float o_scale = 1.f;                         // factor the O accumulator (and the row sum) take when it is raised
        bool rescale = false;                        // warp-uniform: some row of this warp raises its maximum
        if (j == 0) {
          m_used = mx;
        } else {
          const bool need = (mx - m_used) * scale_log2e > FA_RESCALE_THR;
          rescale = __any_sync(0xffffffffu, need);
          if (need) {
            o_scale = fa_ex2((m_used - mx) * scale_log2e);
            m_used = mx;
            l *= o_scale;
          }
        }
        // ---- pass 2: p = exp2((s - m) * scale); row sum; P -> tensor memory as packed bf16, chunk c0 + i at column
        // 16 c0 + 8 i: inside this thread's own score columns and never ahead of what it has already loaded.
        const float mneg = -m_used * scale_log2e;
        const uint64_t mn2 = pk2(mneg, mneg);
        uint64_t sum2 = pk2(0.f, 0.f);
        {
          uint32_t r[2][16];
          auto expo = [&](const uint32_t(&v)[16], int c) {
            float p[16];
#pragma unroll
            for (int k = 0; k < 16; k += 2) {
              float x0, x1;
              unpk2(fma2(pk2(__uint_as_float(v[k]), __uint_as_float(v[k + 1])), sc2, mn2), x0, x1);
              p[k] = fa_ex2(x0);
              p[k + 1] = fa_ex2(x1);
            }
            if (key0 + 16 * c + 16 > T) {
#pragma unroll
              for (int k = 0; k < 16; ++k)
                if (key0 + 16 * c + k >= T) p[k] = 0.f;
            }
            uint32_t w[8];
#pragma unroll
            for (int k = 0; k < 16; k += 2) {
              sum2 = add2(sum2, pk2(p[k], p[k + 1]));
              w[k >> 1] = fa_pack_bf16x2(p[k], p[k + 1]);
            }
            tmem_st_32x8(sbuf + pbase + 8 * (c - c0), w);
          };
#pragma unroll 1
          for (int c = c0; c < c1; c += 2) {
            const bool two = c + 1 < c1;
            tmem_ld_32x16(sbuf + 16 * c, r[0]);
            if (two) tmem_ld_32x16(sbuf + 16 * c + 16, r[1]);
            tmem_ld_wait();
            expo(r[0], c);
            if (two) expo(r[1], c + 1);
          }
        }
        {
          float s0, s1;
          unpk2(sum2, s0, s1);
          l += s0 + s1;
        }
          if (j > 0) {
            // The P V of the slot's previous block has retired (it was issued a whole block ago).  EVERY commit on
            // o_done is observed exactly once and in order -- here, or by the drain for a tile's last block: a
            // parity wait is only exact while the waiter is at most one phase behind the barrier (a version that
            // skipped this wait unless a rescale was due let a fast warp reach a tile's drain two phases early).
            fa_wait(&bar->o_done[g], (nblk - 1u) & 1u, W_O_DONE_BLK);
            if (rescale) {                           // rare after the first block: O *= 2^(old max - new max), in place
              tc_fence_after();
#pragma unroll
              for (int part = 0; part < 2; ++part) {
                uint32_t o[16];
                tmem_ld_32x16(o_addr + 16 * part, o);
                tmem_ld_wait();
#pragma unroll
                for (int k = 0; k < 16; ++k) o[k] = __float_as_uint(__uint_as_float(o[k]) * o_scale);
                tmem_st_32x8(o_addr + 16 * part, *reinterpret_cast<uint32_t(*)[8]>(&o[0]));
                tmem_st_32x8(o_addr + 16 * part + 8, *reinterpret_cast<uint32_t(*)[8]>(&o[8]));
              }
            }
          }
        }  // live
        // the slot's previous tile is drained here, in the first block of the next one (its last P V retired long
        // ago), and before this block's arrival: the MMA warp waits for o_free right after p_full
        if (j == 0 && pend && exists) {
          fa_drain(bar, freed, g, lane, pend_live, ntile, pend_phase, o_addr, pend_inv, pend_dst);
          pend = false;
        }
        if (live) {
          tmem_st_wait();
          tc_fence_before();                         // P (and a rescaled O) are in tensor memory before the MMA reads them
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar->p_full[g][e % FA_SBUF]);
        if (exists) ++nblk;
      }
      if (!exists) continue;
      ++ntile;
      pend = true;
      pend_live = live;
      if (live) {
        // row sum of the tile: exchange the two halves now (cheap), drain later
        bar->xsum[g][h][trow] = l;
        asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory");
        pend_inv = 1.0f / (l + bar->xsum[g][h ^ 1][trow]);
        pend_phase = (nblk - 1u) & 1u;
        const int tok = tile * 128 + trow;
        pend_dst = tok < T ? out + ((size_t)crop * T + tok) * D + (size_t)head * FA_DH + h * 32 : nullptr;
      }
      if (WIDE) {                                    // the other slot is waiting for the shared accumulator: drain now
        fa_drain(bar, freed, g, lane, pend_live, ntile, pend_phase, o_addr, pend_inv, pend_dst);
        pend = false;
      }
    }
    if (pend) fa_drain(bar, freed, g, lane, pend_live, ntile, pend_phase, o_addr, pend_inv, pend_dst);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

}  // namespace

int yv_attention_fa_launch(const __nv_bfloat16 *qkv, __nv_bfloat16 *out, int n, int T, int heads, int dh,
                           cudaStream_t s) {
  YV_REQUIRE(dh == FA_DH, "attention: head dim %d unsupported (need %d)", dh, FA_DH);
  YV_REQUIRE(n >= 1 && heads >= 1 && T >= 1, "attention: bad shape");
  YV_REQUIRE(((uintptr_t)qkv % 16) == 0 && ((uintptr_t)out % 16) == 0, "attention: operands must be 16-byte aligned");
  FaShape sh;
  sh.T = T;
  sh.heads = heads;
  const int t16 = (T + 15) & ~15, nch = t16 >> 4;
  const int nq = (T + 127) >> 7;
  // WIDE layout: all keys in one block per tile (<= 208) and exactly two query tiles, i.e. 129 <= T <= 208 (ViT-B/16)
  // (the default ViT-B/16 kernel is attention_sm100.cu's; this layout runs there with YV_ATTN_B16=fa)
  static const bool no_wide = []() { const char *e = getenv("YV_ATTN_WIDE"); return e && e[0] == '0'; }();
  const bool wide = !no_wide && t16 <= FA_WIDE_KEYS && nq == 2;
  if (wide || t16 <= 128) { sh.nkb = 1; sh.blk0 = sh.blk_last = t16; }
  else if (t16 <= 256) { sh.nkb = 2; sh.blk0 = 16 * ((nch + 1) >> 1); sh.blk_last = t16 - sh.blk0; }
  else { sh.nkb = (t16 + 127) / 128; sh.blk0 = 128; sh.blk_last = t16 - (sh.nkb - 1) * 128; }
  sh.nq = (T + 127) >> 7;
  sh.npair = (sh.nq + 1) >> 1;
  const long long items = (long long)n * heads * sh.npair;
  YV_REQUIRE(items * (long long)sh.nkb * 2 < (1ll << 30), "attention: too many work items");
  const int D = heads * FA_DH;
  const float scale_log2e = (1.0f / sqrtf((float)dh)) * 1.4426950408889634f;
  CUtensorMap tmQ, tmKV0, tmKV1;
  int rc = yv_tmap_bf16_2d(&tmQ, qkv, (uint64_t)n * T, (uint64_t)3 * D, 128u);
  if (rc) return rc;
  if ((rc = yv_tmap_bf16_2d(&tmKV0, qkv, (uint64_t)n * T, (uint64_t)3 * D, (uint32_t)sh.blk0))) return rc;
  if ((rc = yv_tmap_bf16_2d(&tmKV1, qkv, (uint64_t)n * T, (uint64_t)3 * D, (uint32_t)sh.blk_last))) return rc;
  rc = yv_device_once(YV_INIT_ATTN_FA, []() -> int {
    YV_CUDA(cudaFuncSetAttribute(attention_fa_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, fa_smem(false)));
    YV_CUDA(cudaFuncSetAttribute(attention_fa_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, fa_smem(true)));
    return YV_OK;
  });
  if (rc) return rc;
  YvProf prof(YV_K_ATTN, s, 4.0 * n * heads * (double)T * T * dh);
  const int grid = items < yv_num_sms() ? (int)items : yv_num_sms();
  if (wide) attention_fa_kernel<true><<<grid, FA_THREADS, fa_smem(true), s>>>(tmQ, tmKV0, tmKV1, out, (int)items, sh, scale_log2e);
  else attention_fa_kernel<false><<<grid, FA_THREADS, fa_smem(false), s>>>(tmQ, tmKV0, tmKV1, out, (int)items, sh, scale_log2e);
  YV_LAUNCH_CHECK();
  return YV_OK;
}
