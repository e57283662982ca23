// Fused ResidualStack on tcgen05 / TMEM (TTS/vocoder/layers/melgan.py:13-38): all n blocks
//     x <- S_b x + W2_b lrelu( W1_b (*) lrelu(x)[reflect, dilation 3^b] + b1_b ) + (b2_b + bs_b)
// of one upsampling stage run in ONE persistent kernel.  A work item is a chain of KT adjacent 128-row time
// tiles of one utterance; the stage input is read from HBM once, only the stack output goes back.
//
// Where every tensor lives while an item is in flight:
//   XA   (shared, canonical [CP/8][RS][8] bf16, double-buffered across items)  lrelu(x): A operand of the
//        dilated conv — each tap is the same tile at a different start row, so only the descriptor moves;
//   XR[t] (TMEM, bf16, CP/2 columns)  raw x of tile t: A operand of the 1x1 shortcut (A-from-TMEM MMA);
//   D1[t] (TMEM, fp32, CP columns)    dilated-conv accumulator.  Its FIRST MMA is the bias: A = a constant "ones"
//        operand in TMEM (8 columns, K0 = K1 = 1), B = the bias as a [CP x 16] operand in shared memory (K-row 0 =
//        bf16 high part, K-row 1 = bf16 remainder, so the sum carries ~16 mantissa bits), accumulate off — no
//        epilogue adds or pre-loads a bias (tcgen05.st runs at 128 B/cycle/SM, tests/cuda/epi_probe.cu: writing two
//        bias rows per tile and block cost as many cycles as the MMAs).  Its epilogue writes h = lrelu(D1) as bf16
//        over the low CP/2 columns of the same region -> A operand of the 1x1 conv, never touching smem;
//   D2[t] (TMEM, fp32, CP columns)    W2 h + S x accumulator = x of the next block.
//   W1 / W2S (shared)  weights of the current block, or of all blocks when they fit (then loaded once per CTA).
// Only lrelu(x_new) makes the round trip registers -> shared memory; h and x never leave the tensor-core side.
//
// Pipeline: the MMA thread issues, per block, the dilated conv of tiles 0..KT-1 and then the 1x1 pair of tiles
// 0..KT-1; tcgen05 executes in issue order, so epilogue 1 of tile t overlaps the dilated conv of tile t+1 and
// epilogue 2 of tile t overlaps the 1x1 MMAs of tile t+1.  Tile t of the next block starts as soon as tiles
// t-1, t, t+1 have been rewritten (per-tile mbarriers), not when the whole block is done.
//
// Sequence edges: ReflectionPad1d precedes every dilated conv.  Threads whose row lies within `guard` rows of
// an edge also write it to its mirror position; rows outside [0, L) are never written by their owners.  Tile
// (not sequence) boundaries get true neighbours: a chain recomputes sum_d halo rows per side.
//
// Warps: KT epilogue groups of 4*CS warps, then the bulk-copy producer (+ TMEM owner) and the MMA issuer as the
// two highest warp ids (the SMSP arbiter favours high ids, so the issuer is never starved).  Group t
// owns tile t (prologue, both epilogues of every block), one warp per TMEM lane quarter and channel split.
// Tiles therefore advance as independent chains: the latency of one tile's epilogue (barrier -> tcgen05.ld ->
// math -> tcgen05.st -> barrier) hides behind the MMAs and epilogues of the other tiles.
#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "common.cuh"
#include "tc_common.cuh"
#include "tc_conv.cuh"
#include "trace.cuh"

namespace ttsb {

using namespace tc;

namespace {

constexpr int TAPS = 3;
constexpr int MAX_KT = TC_STACK_MAX_KT;
// Ring mode: channels per streamed weight chunk.  A bulk copy costs its issuing warp ~540 cycles however small it
// is and only two are in flight per SM (tests/cuda/ingest_probe.cu: 12 KB chunks from one warp = 22 B/cycle/SM,
// 24 KB chunks from two warps = 90), so chunks are as large as the ring allows and two warps issue them.
constexpr int RING_KB = 64;
constexpr int RING_MAX_SLOTS = 12;
// Service warps after the epilogue groups: bulk-copy producer (+ TMEM owner), dilated-conv MMA issuer, and a third
// warp that is the pointwise (1x1 pair + fused tail) MMA issuer, or in ring mode the second weight producer.
// A fourth one counts finished items for the next kernel (done_out): the fence that publishes an item's stores
// stalls whoever executes it until they have drained, so it must not be an epilogue warp (measured: +7..10 % per kernel).
constexpr int SERVICE_THREADS = 128;


#ifdef TTSB_STACK_TRACE
#define DBG(bit) (p.debug & (bit))
#else
#define DBG(bit) false
#endif

struct Bars {
  uint64_t x_full[2], x_free[2];
  uint64_t w1_full[TC_STACK_MAX_BLOCKS], w2_full[TC_STACK_MAX_BLOCKS], w1_free, w2_free, wh_full;
  uint64_t xa_ready[MAX_KT], xr_ready[MAX_KT], d1_full[MAX_KT], h_ready[MAX_KT], d2_full[MAX_KT];
  uint64_t ring_full[RING_MAX_SLOTS], ring_empty[RING_MAX_SLOTS];
  uint64_t edge[4 * MAX_KT];   // fused head: warp w's first / last three accumulator rows are in the edge buffer
  uint32_t items_done[2];      // epilogue warps that have issued their stores, summed over the even / odd items (->
                               // the signal warp; tc_common.cuh smem_count_add: counters, not parity barriers, so the
                               // signal warp may lag its CTA by any number of items)
};

// TAIL: what follows the last block — 0 the stack output is stored, 1 fused output head, 2 fused transposed conv
// of the next stage.  A template parameter so each variant carries (and allocates registers for) its own code only.
template <int CP, int KT, int CS, int TAIL>
__global__ void __launch_bounds__(SERVICE_THREADS + KT * CS * 128, (KT * 5 * CP / 2 + 8 <= 256) ? 2 : 1)
    tc_stack_kernel(const __grid_constant__ TcStackParams p) {
  constexpr int NT = SERVICE_THREADS + KT * CS * 128;
  constexpr int GROUP_THREADS = CS * 128;
  constexpr int PL = CP / 8;          // 8-channel planes
  constexpr int KS = CP / 16;         // K = 16 steps per tap
  constexpr int TR = KT * 128;        // staged rows owned by the chain
  constexpr int CW = CP / CS;         // channels per epilogue thread
  // C = 256: D1 | D2 alone are the 512 TMEM columns.  Raw x then lives INSIDE D1, in its upper half (h takes the
  // lower one): the shortcut MMAs, which read it, are issued before the block's dilated conv, which overwrites it —
  // one issuer, so they execute in that order.  The D1 bias cannot be pre-loaded then (it would wipe x): the first
  // conv MMA overwrites and epilogue 1 adds the bias.  One XA buffer (two would leave no room for the weight ring).
  constexpr bool OVERLAY = CP > 192;
  constexpr int TCOLS = OVERLAY ? 2 * CP : 5 * CP / 2;   // TMEM columns per tile: D1 | D2 | XR
  constexpr int XR_OFF = OVERLAY ? CP / 2 : 2 * CP;      // column of raw x (bf16, CP / 2 columns) within the tile
  constexpr int NXA = OVERLAY ? 1 : 2;                   // staged x tiles
  constexpr int ONES_COL = KT * TCOLS;   // 8 columns behind the tiles: the constant A operand of the bias MMAs
  constexpr int ONES_W = CP > 96 ? 0 : 8;                // ring mode has no bias MMAs
  static_assert(KT * TCOLS + ONES_W <= 512, "TMEM budget");
  constexpr uint32_t TMEM_ALLOC = KT * TCOLS + ONES_W <= 128 ? 128 : KT * TCOLS + ONES_W <= 256 ? 256 : 512;   // <= 256: two CTAs share an SM
  static_assert(CW % 8 == 0, "CP must be a multiple of 16");
  // Ring mode (wide stages, one tile per CTA): a block's weights do not fit in shared memory, so they stream
  // through a ring of [RING_KB channels x CP columns] chunks in exactly the order the MMAs consume them.
  constexpr bool RING = CP > 96;
  // Wide stages: epilogue 2 publishes lrelu(x) (what the next dilated conv needs) first and raw x + the D2 bias
  // row afterwards behind a second barrier.  Narrow stages are instruction-issue bound; the extra barrier costs more
  // than the shorter critical path gains (measured), so there everything is published at once.
  constexpr bool SPLIT_E2 = CP >= 96;
  // ... and run epilogue 2 in 16-column chunks (see there); narrow stages do it all at once.
  constexpr bool CHUNKED_E2 = CP >= 96;
  // ... and stage the next item (raw x -> XR, lrelu(x) -> XA) while the fused tail's MMAs run.  On the narrow stages
  // the extra live state costs more than the overlap gains (measured: +4..6 %).
  constexpr bool EARLY_STAGE = TAIL != 0 && CP >= 96;
  static_assert(!RING || KT == 1, "ring mode handles one tile per CTA");
  constexpr int NK = CP / RING_KB;                       // chunks per tap
  constexpr uint32_t SLOT_BYTES = RING_KB * CP * 2;
  constexpr int NCH = (TAPS + 2) * NK;                   // chunks per block

  // Biases: every accumulator's first MMA is a bias MMA (see the header) — except in ring mode, whose shared memory
  // is full (two x tiles + four weight slots) and whose single tile per CTA is MMA-bound anyway: there the epilogues
  // pre-load the accumulators with the bias rows by tcgen05.st as before.
  constexpr bool BIAS_MMA = !RING;
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) Bars bars;
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float s_bias[RING ? TC_STACK_MAX_BLOCKS : 1][2][RING ? CP : 4];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int G = p.guard, RS = TR + 2 * G;
  constexpr bool head = TAIL == 1;
  constexpr bool up = TAIL == 2;                        // fused transposed conv of the next stage
  constexpr bool tail = head || up;                     // one more MMA round on lrelu(x_out) after the last block
  static_assert(!head || CS == 1, "the fused head has no channel split");
  static_assert(!up || !RING, "ring mode has no fused transposed conv");
  const int hh = head ? (p.head_taps - 1) / 2 : up ? 1 : 0;   // halo of the fused tail conv
  const int UW = p.up_n / CS;                           // tail-conv columns per epilogue thread
  const uint32_t up_bytes = up ? (uint32_t)2 * CP * p.up_n * 2 : 0;
  const int V = TR - 2 * (p.sum_d + hh);
  const int SD = p.sum_d + hh;                          // rows recomputed per chain side
  const int nb = p.nblocks;
  const uint32_t plane_bytes = (uint32_t)RS * 16;
  const uint32_t xa_bytes = (uint32_t)PL * plane_bytes;
  constexpr uint32_t W1_BYTES = TAPS * CP * CP * 2, W2_BYTES = 2 * CP * CP * 2;
  const int wsets = p.w_sets;  // nb: every block's weights stay resident; 1: reloaded per block; ring mode: slots
  uint8_t* sXA = smem;
  uint8_t* sW = smem + NXA * xa_bytes;   // [set][W1 | W2S]
  uint8_t* sWH = sW + (RING ? (size_t)wsets * SLOT_BYTES : (size_t)wsets * (W1_BYTES + W2_BYTES));   // head weights
  const uint32_t wh_bytes = head ? (uint32_t)CP * p.head_n * 2 : 0;
  // Bias operands (B of the bias MMAs): per block {b1, b2 + bs} as [CP rows x 16 B] planes (K-rows 0/1 = hi/lo), the
  // fused tail's bias, and one all-zero plane every descriptor points its second K-plane at (LBO = zero - start).
  const int tail_n = up ? p.up_n : 0;   // the head adds its bias in the epilogue
  uint8_t* sB = sWH + (head ? wh_bytes : (up && p.up_mode == 1) ? up_bytes : 0);
  uint8_t* sBtail = sB + (size_t)nb * 2 * CP * 16;
  uint8_t* sBzero = sBtail + (size_t)tail_n * 16;
  const uint32_t bzero = smem_u32(sBzero);
  // fused head: per epilogue warp, its first and last three accumulator rows (6 x 32 fp32) for the neighbouring warps
  float* sEdge = reinterpret_cast<float*>(sBzero + (size_t)(CP > tail_n ? CP : tail_n) * 16);
  // k24 mode (24 real channels in CP = 32): a constant plane of 128 rows {1, 1, 0 x 6}, the A half of the K-step that
  // carries the dilated conv's bias.  Above both XA buffers, so the offset from any XA row to it is positive.
  uint8_t* sOnes = reinterpret_cast<uint8_t*>(sEdge) + (head ? (size_t)4 * KT * 6 * 32 * 4 : 0);
  auto bias_desc = [&](const uint8_t* bp) {
    const uint32_t a = smem_u32(bp);
    return make_desc(a, bzero - a, 128);
  };

  const int n_items = p.n_chains * p.batch;
  auto xbuf = [](int it) { return NXA == 2 ? (it & 1) : 0; };               // which staged x tile item `it` uses
  auto xpar = [](int it) { return NXA == 2 ? ((it >> 1) & 1) : (it & 1); };  // phase parity of its x_full / x_free
  // Ring mode: global address of chunk `ch` of block `blk` in the order the MMAs consume them — dilated conv, W2, S;
  // with raw x inside D1 (OVERLAY) the shortcut S goes first.  p.w2s[blk] = [W2 | S].
  auto ring_src = [&](int blk, int ch) -> const __nv_bfloat16* {
    constexpr int NKc = CP / RING_KB;
    constexpr size_t CH = (size_t)RING_KB * CP;   // elements per chunk
    if (!OVERLAY)
      return ch < TAPS * NKc ? p.w1[blk] + (size_t)ch * CH : p.w2s[blk] + (size_t)(ch - TAPS * NKc) * CH;
    if (ch < NKc) return p.w2s[blk] + (size_t)(NKc + ch) * CH;
    if (ch < NKc + TAPS * NKc) return p.w1[blk] + (size_t)(ch - NKc) * CH;
    return p.w2s[blk] + (size_t)(ch - NKc - TAPS * NKc) * CH;
  };

  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars.x_full[i], 1);
      mbar_init(&bars.x_free[i], 1);
    }
    for (int i = 0; i < TC_STACK_MAX_BLOCKS; ++i) {
      mbar_init(&bars.w1_full[i], 1);
      mbar_init(&bars.w2_full[i], 1);
    }
    mbar_init(&bars.w1_free, 1);
    mbar_init(&bars.w2_free, 1);
    mbar_init(&bars.wh_full, 1);
    for (int i = 0; i < RING_MAX_SLOTS; ++i) {
      mbar_init(&bars.ring_full[i], 1);
      mbar_init(&bars.ring_empty[i], 1);
    }
    for (int i = 0; i < 4 * MAX_KT; ++i) mbar_init(&bars.edge[i], 32);
    bars.items_done[0] = bars.items_done[1] = 0;
    for (int t = 0; t < KT; ++t) {
      mbar_init(&bars.xa_ready[t], GROUP_THREADS);
      mbar_init(&bars.xr_ready[t], GROUP_THREADS);
      mbar_init(&bars.d1_full[t], 1);
      mbar_init(&bars.h_ready[t], GROUP_THREADS);
      mbar_init(&bars.d2_full[t], 1);
    }
    fence_mbar_init();
  }
  constexpr int PRODUCER_WARP = KT * CS * 4, MMA_WARP = PRODUCER_WARP + 1;
  constexpr int THIRD_WARP = MMA_WARP + 1;       // ring mode: issues the odd weight chunks; else: pointwise MMA issuer
  constexpr int SIGNAL_WARP = THIRD_WARP + 1;    // publishes finished items to the next kernel
  if (warp == PRODUCER_WARP) tmem_alloc(&tmem_base_s, TMEM_ALLOC);
  if (!BIAS_MMA) {
    for (int e = tid; e < nb * 2 * CP; e += NT) {
      const int blk = e / (2 * CP), r = e - blk * 2 * CP;
      s_bias[RING ? blk : 0][r / CP][RING ? r % CP : 0] = r < CP ? p.b1[blk][r] : p.b2s[blk][r - CP];
    }
  } else {
    auto put_bias = [](uint8_t* dst, float b) {
      const __nv_bfloat16 hi = __float2bfloat16(b);
      const __nv_bfloat16 lo = __float2bfloat16(b - __bfloat162float(hi));
      *reinterpret_cast<uint4*>(dst) =
          make_uint4((uint32_t)__bfloat16_as_ushort(hi) | ((uint32_t)__bfloat16_as_ushort(lo) << 16), 0u, 0u, 0u);
    };
    for (int e = tid; e < nb * 2 * CP; e += NT) {
      const int blk = e / (2 * CP), r = e - blk * 2 * CP;
      put_bias(sB + (size_t)e * 16, r < CP ? p.b1[blk][r] : p.b2s[blk][r - CP]);
    }
    for (int e = tid; e < tail_n; e += NT) put_bias(sBtail + (size_t)e * 16, p.up_b[e]);
    const int zn = CP > tail_n ? CP : tail_n;
    for (int e = tid; e < zn; e += NT) *reinterpret_cast<uint4*>(sBzero + (size_t)e * 16) = make_uint4(0u, 0u, 0u, 0u);
    if (p.k24)
      for (int e = tid; e < 128; e += NT) *reinterpret_cast<uint4*>(sOnes + (size_t)e * 16) = make_uint4(0x3F803F80u, 0u, 0u, 0u);
    fence_proxy_async();   // the MMAs read these through the async proxy
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_base_s;
  // With follow.done the previous kernel is followed segment by segment (the loader polls its counters) and nobody
  // waits for its whole grid; what this kernel writes is then never a buffer the previous kernel still reads.
  const bool follow = p.follow.done != nullptr;
  if (follow && !p.follow.early) {   // late form: the check precedes this CTA's trigger (follow.cuh follow_armed)
    if (tid == 0) follow_armed(p.follow);
    __syncthreads();
  }
  pdl_launch_dependents();   // every thread: the CTA counts as triggered once all of its threads have
  if (warp != PRODUCER_WARP && !follow) pdl_wait();   // the producer first prefetches weights (not written by any kernel)

  if (warp == PRODUCER_WARP) {
    // =============================== producer ===============================
    if (elect_one()) {
      FollowState fst;   // which segments of the previous kernel's output are known complete (follow mode)
      auto load_x = [&](int it) {
        const int item = blockIdx.x + it * gridDim.x;
        if (item >= n_items) return;
        const int buf = xbuf(it);
        if (it >= NXA) mbar_wait(&bars.x_free[buf], xpar(it - NXA));
        const int b = item / p.n_chains, chain = item - b * p.n_chains;
        if (follow) follow_wait(p.follow, b, chain * V - SD, chain * V - SD + TR, fst);
        const int g0 = chain * V - SD;
        mbar_expect_tx(&bars.x_full[buf], (uint32_t)PL * TR * 16);
        const __nv_bfloat16* gx = p.x + (int64_t)b * p.x_batch_stride + (int64_t)(p.x_halo + g0) * 8;
        uint8_t* dst = sXA + (size_t)buf * xa_bytes + (size_t)G * 16;
        for (int pl = 0; pl < PL; ++pl)
          bulk_g2s(dst + (size_t)pl * plane_bytes, gx + (int64_t)pl * p.x_plane_rows * 8, (uint32_t)TR * 16,
                   &bars.x_full[buf]);
      };
      if (head) {
        mbar_expect_tx(&bars.wh_full, wh_bytes);
        bulk_g2s(sWH, p.head_w, wh_bytes, &bars.wh_full);
      } else if (up && p.up_mode == 1) {   // tail weights resident for the life of the CTA
        mbar_expect_tx(&bars.wh_full, up_bytes);
        bulk_g2s(sWH, p.up_w, up_bytes, &bars.wh_full);
      }
      if (!follow) pdl_wait();   // the previous kernel's output is complete from here on
      load_x(0);
      const bool up_stream = up && p.up_mode == 2;   // tail weights take turns with W2S in its slot
      auto load_up = [&](uint32_t& n_w2) {
        mbar_wait(&bars.w2_free, (n_w2 - 1) & 1);
        mbar_expect_tx(&bars.w2_full[0], up_bytes);
        bulk_g2s(sW + W1_BYTES, p.up_w, up_bytes, &bars.w2_full[0]);
        ++n_w2;
      };
      int gb = 0;
      uint32_t n_w2 = 0;   // loads issued into the streamed W2S slot (blocks, plus the fused transposed conv)
      for (int it = 0; blockIdx.x + it * gridDim.x < n_items; ++it) {
        for (int blk = 0; blk < nb; ++blk, ++gb) {
          if (RING) {
            for (int ch = 0; ch < NCH; ++ch) {
              const uint32_t gc = (uint32_t)gb * NCH + ch;          // global chunk counter
              if (NXA == 2 && blk == 0 && ch == 0) load_x(it + 1);
              if (gc & 1) continue;                                 // odd chunks: the second producer warp
              const int slot = gc % wsets;
              mbar_wait(&bars.ring_empty[slot], ((gc / wsets) & 1) ^ 1);
              mbar_expect_tx(&bars.ring_full[slot], SLOT_BYTES);
              bulk_g2s(sW + (size_t)slot * SLOT_BYTES, ring_src(blk, ch), SLOT_BYTES, &bars.ring_full[slot]);
            }
            // one x buffer: the next tile can only come in once this item's last dilated conv has read the old one
            if (NXA == 1 && blk == nb - 1) load_x(it + 1);
            continue;
          }
          if (wsets == 1 || it == 0) {
            const int set = wsets == 1 ? 0 : blk;
            uint8_t* w1 = sW + (size_t)set * (W1_BYTES + W2_BYTES);
            if (wsets == 1 && gb > 0) mbar_wait(&bars.w1_free, (gb - 1) & 1);
            mbar_expect_tx(&bars.w1_full[set], W1_BYTES);
            bulk_g2s(w1, p.w1[blk], W1_BYTES, &bars.w1_full[set]);
            // the previous item's tail weights go in behind this item's first W1, so that load is never held up
            if (up_stream && blk == 0 && it > 0) load_up(n_w2);
            if (wsets == 1 && n_w2 > 0) mbar_wait(&bars.w2_free, (n_w2 - 1) & 1);
            mbar_expect_tx(&bars.w2_full[set], W2_BYTES);
            bulk_g2s(w1 + W1_BYTES, p.w2s[blk], W2_BYTES, &bars.w2_full[set]);
            ++n_w2;
          }
          if (blk == 0) load_x(it + 1);   // prefetch the next item's tile behind this item's first weights
        }
      }
      if (up_stream && n_w2 > 0) load_up(n_w2);   // tail weights of this CTA's last item
    }
  } else if (RING && warp == THIRD_WARP) {
    // ================ second weight producer (ring mode): the odd chunks of the same sequence ================
    if (elect_one()) {
      for (int it = 0, gb = 0; blockIdx.x + it * gridDim.x < n_items; ++it)
        for (int blk = 0; blk < nb; ++blk, ++gb)
          for (int ch = 0; ch < NCH; ++ch) {
            const uint32_t gc = (uint32_t)gb * NCH + ch;
            if (!(gc & 1)) continue;
            const int slot = gc % wsets;
            mbar_wait(&bars.ring_empty[slot], ((gc / wsets) & 1) ^ 1);
            mbar_expect_tx(&bars.ring_full[slot], SLOT_BYTES);
            bulk_g2s(sW + (size_t)slot * SLOT_BYTES, ring_src(blk, ch), SLOT_BYTES, &bars.ring_full[slot]);
          }
    }
  } else if (RING && warp == MMA_WARP) {
    // =============================== MMA issuer, ring mode ===============================
    // One issuer: the weight chunks arrive through the ring in exactly the order the MMAs consume them.
    // The whole warp walks the loops (uniform control flow keeps addresses in uniform registers, so an MMA
    // costs a couple of uniform adds); only the tcgen05 instructions themselves are issued by the elected lane.
    const bool leader = elect_one() != 0;
    constexpr uint32_t idesc = make_idesc_bf16(128, CP);
    constexpr uint32_t LBO_B = CP * 16;
    const uint32_t lbo_a = plane_bytes;
    const uint64_t a_tmpl = make_desc(0, lbo_a, 128), b_tmpl = make_desc(0, LBO_B, 128);
    int gb = 0;
    for (int it = 0; blockIdx.x + it * gridDim.x < n_items; ++it) {
      const int buf = xbuf(it);
      const uint32_t xa = smem_u32(sXA + (size_t)buf * xa_bytes);
      for (int blk = 0; blk < nb; ++blk, ++gb) {
        const uint32_t ph = gb & 1;
        const int d = p.dil[blk];
        uint32_t gc = (uint32_t)gb * NCH;
        // NK chunks of A-from-TMEM MMAs into D2: the 1x1 conv on h (A at column 0) or the shortcut on raw x (XR_OFF)
        auto ts_round = [&](uint32_t a_col0) {
          for (int kc = 0; kc < NK; ++kc, ++gc) {
            const int slot = gc % wsets;
            mbar_wait(&bars.ring_full[slot], (gc / wsets) & 1);
            tc_fence_after();
            if (leader) {
              const uint32_t wb = smem_u32(sW + (size_t)slot * SLOT_BYTES);
              const uint32_t acol = a_col0 + 8 * (kc * (RING_KB / 16));
#pragma unroll
              for (int kk = 0; kk < RING_KB / 16; ++kk)
                umma_bf16_ts(tbase + CP, acol + 8 * kk, b_tmpl | (uint64_t)((wb + 2 * kk * LBO_B) >> 4), idesc, 1);
              umma_commit(&bars.ring_empty[slot]);
            }
            __syncwarp();
          }
        };
        if (OVERLAY) {   // the shortcut first: raw x sits where the dilated conv is about to accumulate
          mbar_wait(&bars.xr_ready[0], ph);
          tc_fence_after();
          ts_round(tbase + XR_OFF);
        }
        // ---- dilated conv: D1 = sum_tap XA[rows + (tap-1) d] W1[tap]^T
        mbar_wait(&bars.xa_ready[0], ph);
        tc_fence_after();
        for (int j = 0; j < TAPS; ++j) {
          const uint32_t a_row = xa + (uint32_t)(G + (j - 1) * d) * 16;
          for (int kc = 0; kc < NK; ++kc, ++gc) {
            const int slot = gc % wsets;
            mbar_wait(&bars.ring_full[slot], (gc / wsets) & 1);
            tc_fence_after();
            if (leader) {
              const uint32_t wb = smem_u32(sW + (size_t)slot * SLOT_BYTES);
#pragma unroll
              for (int kk = 0; kk < RING_KB / 16; ++kk)
                umma_bf16(tbase, a_tmpl | (uint64_t)((a_row + (uint32_t)(kc * (RING_KB / 8) + 2 * kk) * lbo_a) >> 4),
                          b_tmpl | (uint64_t)((wb + 2 * kk * LBO_B) >> 4), idesc,
                          OVERLAY ? (uint32_t)(j + kc + kk > 0) : 1u);   // OVERLAY: D1 holds no bias row, the first MMA overwrites
              umma_commit(&bars.ring_empty[slot]);
            }
            __syncwarp();
          }
        }
        if (leader) {
          umma_commit(&bars.d1_full[0]);
          if (blk == nb - 1) umma_commit(&bars.x_free[buf]);
        }
        __syncwarp();
        // ---- 1x1 conv on h (+ 1x1 shortcut on raw x), A operands in TMEM
        mbar_wait(&bars.h_ready[0], ph);
        if (SPLIT_E2 && !OVERLAY) mbar_wait(&bars.xr_ready[0], ph);
        tc_fence_after();
        ts_round(tbase);
        if (!OVERLAY) ts_round(tbase + XR_OFF);
        if (leader) umma_commit(&bars.d2_full[0]);
        __syncwarp();
      }
    }
  } else if (!RING && warp == MMA_WARP) {
    // =============================== conv issuer: the dilated convs ===============================
    // Two issuing warps, so that a tile's 1x1 pair (and the fused tail) never queues behind the barrier waits of
    // the other tiles' dilated convs and vice versa: this warp follows xa_ready, the pointwise issuer follows
    // h_ready / xr_ready.  All cross-warp ordering is carried by mbarriers (tcgen05.commit tracks the MMAs of
    // the committing thread only, and each barrier is committed by the warp that issued what it guards).
    // The whole warp walks the loops (uniform control flow keeps addresses in uniform registers, so an MMA
    // costs a couple of uniform adds); only the tcgen05 instructions themselves are issued by the elected lane.
    // (Measured alternatives that lost: polling all tiles' barriers with one ballot of test_waits and issuing
    // whatever is ready — the polls cost more than the in-order waits; one CTA per SM with 5-6 tiles — this
    // warp's issue rate, ~120 cycles per MMA at N = 32, then bounds the stage.)
    const bool leader = elect_one() != 0;
    constexpr uint32_t idesc = make_idesc_bf16(128, CP);
    constexpr uint32_t LBO_B = CP * 16;
    const uint32_t lbo_a = plane_bytes;
    // descriptor templates: only the 14-bit start-address field changes between MMAs
    const uint64_t a_tmpl = make_desc(0, lbo_a, 128), b_tmpl = make_desc(0, LBO_B, 128);
    Tracer tr = make_tracer(p.trace, 0, leader);
    int gb = 0;
    uint32_t n_xa = 0;   // completed xa_ready rounds (one per block, plus one per item with a fused tail conv)
    for (int it = 0; blockIdx.x + it * gridDim.x < n_items; ++it) {
      const int buf = xbuf(it);
      const uint32_t xa = smem_u32(sXA + (size_t)buf * xa_bytes);
      for (int blk = 0; blk < nb; ++blk, ++gb) {
        const int set = wsets == 1 ? 0 : blk;
        const uint32_t w1 = smem_u32(sW + (size_t)set * (W1_BYTES + W2_BYTES));
        const int d = p.dil[blk];
        const uint64_t b1_desc = bias_desc(sB + (size_t)(2 * blk) * CP * 16);
        if (wsets == 1) mbar_wait(&bars.w1_full[0], gb & 1);
        else if (it == 0) mbar_wait(&bars.w1_full[set], 0);
        // Issue order: tap j of tile t reads rows shifted by (j - 1) d, so taps 0 and 1 need tiles <= t only and
        // start as soon as tile t has been rewritten; tap 2 follows once tile t + 1 is there.
        auto issue_taps = [&](int t, int j0, int j1) {
          const uint32_t dcol = tbase + t * TCOLS;
          if (leader) {
#pragma unroll
            for (int j = j0; j < j1; ++j) {
              const uint32_t a_row = xa + (uint32_t)(G + t * 128 + (j - 1) * d) * 16;
              const uint32_t w_tap = w1 + (uint32_t)j * (CP * CP * 2);
#pragma unroll
              for (int kk = 0; kk < KS; ++kk)
                umma_bf16(dcol, a_tmpl | (uint64_t)((a_row + 2 * kk * lbo_a) >> 4),
                          b_tmpl | (uint64_t)((w_tap + 2 * kk * LBO_B) >> 4), idesc, 1);
            }
          }
        };
        // k24 mode (CP = 32 carrying 24 real channels, the last MB-MelGAN stage): the 3 x 24 input channels of the three
        // taps are packed into FIVE K = 16 steps instead of six — (tap0 ch 0-15) (tap1 ch 0-7 | tap0 ch 16-23)
        // (tap1 ch 8-23) (tap2 ch 0-15) (tap2 ch 16-23 | ones plane x bias row) — the two halves of a K-step being any
        // two planes whose distance is the descriptor's leading-dimension offset.  The last step also delivers the
        // bias, so the bias MMA goes too: 5 shared-memory-bound MMAs per tile and block instead of 6 + 1 (these
        // MMAs' operand reads are what bounds the stage).  Weights: schedule_bf16.inl pack_w1_k24.
        auto issue_k24 = [&](int t, int half) {   // half 0: steps 0-2 (taps 0, 1; the first overwrites), 1: steps 3-4
          if (leader) {
            const uint32_t dcol = tbase + t * TCOLS;
            const uint32_t r0 = xa + (uint32_t)(G + t * 128 - d) * 16, r1 = r0 + (uint32_t)d * 16, r2 = r1 + (uint32_t)d * 16;
            auto mma = [&](uint32_t start, uint32_t lbo, int ks, uint32_t acc) {
              umma_bf16(dcol, make_desc(start, lbo, 128), b_tmpl | (uint64_t)((w1 + ks * 2 * LBO_B) >> 4), idesc, acc);
            };
            if (half == 0) {
              mma(r0, lbo_a, 0, 0);
              mma(r1, 2 * lbo_a - (uint32_t)d * 16, 1, 1);
              mma(r1 + lbo_a, lbo_a, 2, 1);
            } else {
              mma(r2, lbo_a, 3, 1);
              mma(r2 + 2 * lbo_a, smem_u32(sOnes) - (r2 + 2 * lbo_a), 4, 1);
            }
          }
        };
#pragma unroll 1
        for (int t = 0; t < KT; ++t) {
          mbar_wait(&bars.xa_ready[t], n_xa & 1);
          tc_fence_after();
          tr(100 + t);
          if (p.k24) {
            if (t > 0) {
              issue_k24(t - 1, 1);
              if (leader) umma_commit(&bars.d1_full[t - 1]);
            }
            issue_k24(t, 0);
          } else {
            if (t > 0) {
              issue_taps(t - 1, 2, 3);
              if (leader) umma_commit(&bars.d1_full[t - 1]);
            }
            if (leader) umma_bf16_ts(tbase + t * TCOLS, tbase + ONES_COL, b1_desc, idesc, 0);   // D1[t] <- b1 (overwrites)
            issue_taps(t, 0, 2);
          }
          __syncwarp();
          tr(110 + t);
        }
        if (p.k24) issue_k24(KT - 1, 1);
        else issue_taps(KT - 1, 2, 3);
        if (leader) {
          umma_commit(&bars.d1_full[KT - 1]);
          if (wsets == 1) umma_commit(&bars.w1_free);
          // XA is not read again for this item (the fused head reads lrelu(x_out) from TMEM)
          if (blk == nb - 1 && (!tail || head)) umma_commit(&bars.x_free[buf]);
        }
        __syncwarp();
        ++n_xa;
        tr(112);
      }
      if (tail) {
        // The tail round of xa_ready belongs to the pointwise issuer; pass it here too so that this warp's
        // parity never runs a full phase ahead of the barrier (a parity wait cannot tell phase n from n + 2).
#pragma unroll 1
        for (int t = 0; t < KT; ++t) mbar_wait(&bars.xa_ready[t], n_xa & 1);
        ++n_xa;
      }
    }
  } else if (!RING && warp == THIRD_WARP) {
    // =============================== pointwise issuer: 1x1 pair, fused tail ===============================
    const bool leader = elect_one() != 0;
    constexpr uint32_t idesc = make_idesc_bf16(128, CP);
    constexpr uint32_t LBO_B = CP * 16;
    const uint32_t lbo_a = plane_bytes;
    const uint64_t a_tmpl = make_desc(0, lbo_a, 128), b_tmpl = make_desc(0, LBO_B, 128);
    Tracer tr = make_tracer(p.trace, TRACE_ROLES - 1, leader);
    int gb = 0;
    uint32_t n_xa = 0;   // same count as the conv issuer's
    uint32_t n_w2 = 0;   // consumed fills of the streamed W2S slot
    for (int it = 0; blockIdx.x + it * gridDim.x < n_items; ++it) {
      const int buf = xbuf(it);
      const uint32_t xa = smem_u32(sXA + (size_t)buf * xa_bytes);
      for (int blk = 0; blk < nb; ++blk, ++gb) {
        const uint32_t ph = gb & 1;
        const int set = wsets == 1 ? 0 : blk;
        const uint32_t w2 = smem_u32(sW + (size_t)set * (W1_BYTES + W2_BYTES)) + W1_BYTES;
        const uint64_t b2_desc = bias_desc(sB + (size_t)(2 * blk + 1) * CP * 16);
        if (wsets == 1) mbar_wait(&bars.w2_full[0], n_w2++ & 1);
        else if (it == 0) mbar_wait(&bars.w2_full[set], 0);
        tr(113);
        // ---- D2[t] = S x (raw x in TMEM since the previous round's epilogue 2: these run while epilogue 1 works)
#pragma unroll 1
        for (int t = 0; t < KT; ++t) {
          if (SPLIT_E2) mbar_wait(&bars.xr_ready[t], ph);      // raw x + D2 bias row: the deferred half of epilogue 2
          else mbar_wait(&bars.xa_ready[t], n_xa & 1);         // published together with lrelu(x)
          tc_fence_after();
          const uint32_t tcol = tbase + t * TCOLS;
          if (leader) {
            umma_bf16_ts(tcol + CP, tbase + ONES_COL, b2_desc, idesc, 0);   // D2[t] <- b2 + bs (overwrites)
#pragma unroll
            for (int kk = 0; kk < KS; ++kk)
              umma_bf16_ts(tcol + CP, tcol + XR_OFF + 8 * kk,
                           b_tmpl | (uint64_t)((w2 + CP * CP * 2 + 2 * kk * LBO_B) >> 4), idesc, 1);
          }
          __syncwarp();
        }
        ++n_xa;
        // ---- ... + W2 h (h in TMEM, written by epilogue 1)
#pragma unroll 1
        for (int t = 0; t < KT; ++t) {
          mbar_wait(&bars.h_ready[t], ph);
          tc_fence_after();
          tr(120 + t);
          const uint32_t tcol = tbase + t * TCOLS;
          if (leader) {
#pragma unroll
            for (int kk = 0; kk < KS; ++kk)
              umma_bf16_ts(tcol + CP, tcol + 8 * kk, b_tmpl | (uint64_t)((w2 + 2 * kk * LBO_B) >> 4), idesc, 1);
            umma_commit(&bars.d2_full[t]);
          }
          __syncwarp();
          tr(130 + t);
        }
        if (leader && wsets == 1) umma_commit(&bars.w2_free);
        __syncwarp();
      }
      if (head) {
        // ---- fused output head, taps packed on N: P[t][j * ch + c] = sum_ci lrelu(x_out)[t][ci] Wh[c][ci][j], ONE
        // K = CP round per tile with A = lrelu(x_out) in TMEM (XR[t], written by the last epilogue 2): KS MMAs instead
        // of taps * KS shared-memory-read-bound N = 16 MMAs (34 % of this kernel's tensor time before).  The tap
        // shifts move to the output side: y[t][c] = sum_j P[t + j - hh][j * ch + c], summed by the epilogue.
        const uint32_t idesc_h = make_idesc_bf16(128, p.head_n);
        const uint32_t lbo_h = (uint32_t)p.head_n * 16;
        const uint64_t bh_tmpl = make_desc(0, lbo_h, 128);
        const uint32_t wh = smem_u32(sWH);
        if (it == 0) mbar_wait(&bars.wh_full, 0);
#pragma unroll 1
        for (int t = 0; t < KT; ++t) {
          mbar_wait(&bars.xa_ready[t], n_xa & 1);
          tc_fence_after();
          tr(150 + t);
          const uint32_t tcol = tbase + t * TCOLS;
          if (leader) {
#pragma unroll
            for (int kk = 0; kk < KS; ++kk)
              umma_bf16_ts(tcol, tcol + XR_OFF + 8 * kk, bh_tmpl | (uint64_t)((wh + 2 * kk * lbo_h) >> 4), idesc_h,
                           kk > 0);
            umma_commit(&bars.d1_full[t]);
          }
          __syncwarp();
        }
        ++n_xa;
      }
      if (up) {
        // ---- fused transposed conv of the next stage: D[t][0:N] = a[q] Wu[0]^T + a[q-1] Wu[1]^T, a = lrelu(x_out)
        const int N = p.up_n;
        const uint32_t idesc_u = make_idesc_bf16(128, N);
        const uint32_t lbo_u = (uint32_t)N * 16;
        const uint64_t bu_tmpl = make_desc(0, lbo_u, 128);
        const bool resident = p.up_mode == 1;
        const uint32_t wu = smem_u32(resident ? sWH : sW + W1_BYTES);
        if (!resident) mbar_wait(&bars.w2_full[0], n_w2++ & 1);
        else if (it == 0) mbar_wait(&bars.wh_full, 0);
#pragma unroll 1
        for (int t = 0; t < KT; ++t) {
          mbar_wait(&bars.xa_ready[t], n_xa & 1);   // taps reach back one row only: tile t needs tiles <= t
          tc_fence_after();
          tr(140 + t);
          if (leader) {
            umma_bf16_ts(tbase + t * TCOLS, tbase + ONES_COL, bias_desc(sBtail), idesc_u, 0);   // accumulator <- bias
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              const uint32_t a_row = xa + (uint32_t)(G + t * 128 - j) * 16;
              const uint32_t w_tap = wu + (uint32_t)j * (uint32_t)(CP * N * 2);
#pragma unroll
              for (int kk = 0; kk < KS; ++kk)
                umma_bf16(tbase + t * TCOLS, a_tmpl | (uint64_t)((a_row + 2 * kk * lbo_a) >> 4),
                          bu_tmpl | (uint64_t)((w_tap + 2 * kk * lbo_u) >> 4), idesc_u, 1);
            }
            umma_commit(&bars.d1_full[t]);
          }
          __syncwarp();
        }
        ++n_xa;
        if (leader) {
          if (!resident) umma_commit(&bars.w2_free);
          umma_commit(&bars.x_free[buf]);
        }
        __syncwarp();
      }
    }
  } else if (warp == SIGNAL_WARP) {
    // =============================== signal warp ===============================
    // Item by item: once every epilogue warp has issued its stores (items_done: their release add hands them to this
    // thread's acquire), make them visible device-wide and count the item for its utterance.  The next kernel's loader polls
    // that counter (follow.cuh follow_wait) instead of waiting for this whole grid.
    if (p.done_out && elect_one()) {
      for (int it = 0; blockIdx.x + it * gridDim.x < n_items; ++it) {
        const int item = blockIdx.x + it * gridDim.x;
        const int b = item / p.n_chains, chain = item - b * p.n_chains;
        smem_count_wait(&bars.items_done[it & 1], (uint32_t)((it >> 1) + 1) * (KT * CS * 4));
        __threadfence();
        atomicAdd(p.done_out + b * FOLLOW_SEGS + chain / p.count_units_per_seg, 1);
      }
    }
  } else {
    // =============================== epilogue warps ===============================
    const int q4 = warp & 3;                    // TMEM lane quarter this warp may touch
    const int t = warp / (4 * CS);              // the tile this warp's group owns
    const int cs = (warp / 4) % CS;             // which slice of the channels
    const int m = q4 * 32 + lane;               // row within the 128-row tile
    const uint32_t tlane = tbase + ((uint32_t)(q4 * 32) << 16) + t * TCOLS;
    const int c0 = cs * CW;                     // first channel of this thread
    const float slope = p.slope;
    const Slope2 slope2 = make_slope2(slope);
    const int L = p.L;
    const int pair_bar = 1 + t * 4 + q4;        // named barrier shared by the CS warps of a quarter
    // debug bit 512: issuer timelines only (the epilogue trace points perturb the kernel by ~20 %)
    Tracer tr = make_tracer(p.trace, 1 + t, m == 0 && cs == 0 && !DBG(512));
    Tracer trw = make_tracer(p.trace, 1 + KT + warp, lane == 0 && !DBG(512));   // per-warp arrival times
    uint32_t n_d1 = 0;                          // completed d1_full rounds (blocks, plus the fused head)
    if (BIAS_MMA && t == 0 && cs == 0) {
      // constant A operand of every bias MMA: K0 = K1 = 1.0 (bf16 pair in column 0), K2..15 = 0; published with the
      // first item's prologue (tcgen05.wait::st + fence + arrive), which every issuer passes before its first MMA
      const uint32_t ones[8] = {0x3F803F80u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
      tmem_st8(tbase + ((uint32_t)(q4 * 32) << 16) + ONES_COL, ones);
    }

    // First half of an item's prologue: raw x tile -> XR[t] (TMEM, bf16) and lrelu(x) in place in XA.  It touches
    // neither D1 nor D2, so in kernels with a fused tail it runs for item it+1 while item it's tail MMAs execute
    // (its x tile was prefetched a whole item earlier); only the accumulators' bias rows and the publish remain
    // between the tail epilogue and the next item's first dilated conv.
    auto stage_item = [&](int it_) {
      const int item_ = blockIdx.x + it_ * gridDim.x;
      const int b_ = item_ / p.n_chains, chain_ = item_ - b_ * p.n_chains;
      const int g0_ = chain_ * V - SD;
      uint8_t* xa_ = sXA + (size_t)xbuf(it_) * xa_bytes + (size_t)(c0 / 8) * plane_bytes;
      const int reach_ = G > p.out.halo ? G : p.out.halo;
      const bool edge_ = g0_ <= reach_ || g0_ + TR + reach_ >= L - 1;
      mbar_wait_relaxed(&bars.x_full[xbuf(it_)], xpar(it_));
      uint32_t w[CW / 2];
      const uint8_t* src = xa_ + (size_t)(G + t * 128 + m) * 16;
#pragma unroll
      for (int pl = 0; pl < CW / 8; ++pl) {
        const uint4 u = *reinterpret_cast<const uint4*>(src + (size_t)pl * plane_bytes);
        w[4 * pl] = u.x; w[4 * pl + 1] = u.y; w[4 * pl + 2] = u.z; w[4 * pl + 3] = u.w;
      }
      tmem_st<CW / 2>(tlane + XR_OFF + c0 / 2, w);
#pragma unroll
      for (int i = 0; i < CW / 2; ++i) w[i] = lrelu_bf16x2(w[i], slope2);
      auto rows = [&](int r) {
        uint8_t* dst = xa_ + (size_t)(G + r) * 16;
#pragma unroll
        for (int pl = 0; pl < CW / 8; ++pl)
          *reinterpret_cast<uint4*>(dst + (size_t)pl * plane_bytes) =
              make_uint4(w[4 * pl], w[4 * pl + 1], w[4 * pl + 2], w[4 * pl + 3]);
      };
      if (!edge_) {
        rows(t * 128 + m);
      } else {
        const int tau = g0_ + t * 128 + m;
        if (tau >= 0 && tau < L) {
          rows(tau - g0_);
          if (tau >= 1 && tau <= G && -tau - g0_ >= -G) rows(-tau - g0_);
          const int back = L - 1 - tau;
          if (back >= 1 && back <= G && L - 1 + back - g0_ < TR + G) rows(L - 1 + back - g0_);
        }
      }
    };
    float head_bias[4] = {0.f, 0.f, 0.f, 0.f};   // fused head: its bias, read once (an L2 round trip per item otherwise)
    if (head) {
#pragma unroll
      for (int c = 0; c < 4; ++c)
        if (c < p.head_ch) head_bias[c] = __ldg(p.head_b + c);
    }
    bool staged = false;   // the current item's first prologue half already ran (during the previous item's tail)

    int gb = 0;
    for (int it = 0; blockIdx.x + it * gridDim.x < n_items; ++it) {
      const int item = blockIdx.x + it * gridDim.x;
      const int buf = xbuf(it);
      const int b = item / p.n_chains, chain = item - b * p.n_chains;
      const int g0 = chain * V - SD;            // sequence row of staged row 0
      uint8_t* xa = sXA + (size_t)buf * xa_bytes + (size_t)(c0 / 8) * plane_bytes;   // this thread's first plane

      // Items whose staged rows stay clear of both sequence ends take the unconditional store path.
      const int reach = G > p.out.halo ? G : p.out.halo;   // mirror rows: `guard` in XA, `halo` in the output
      const bool edge = g0 <= reach || g0 + TR + reach >= L - 1;
      // Writes lrelu'd channels [c0, c0+CW) of this thread's row (packed bf16 pairs) into XA.  Near a sequence
      // end: rows outside [0, L) are skipped and rows within `guard` of the end are also written to their mirror
      // position (what ReflectionPad1d of the next dilated conv reads).
      auto row_store = [&](int r, const uint32_t* w) {      // r = staged row (0 = sequence row g0)
        if (DBG(1)) return;
        uint8_t* dst = xa + (size_t)(G + r) * 16;
#pragma unroll
        for (int pl = 0; pl < CW / 8; ++pl)
          *reinterpret_cast<uint4*>(dst + (size_t)pl * plane_bytes) =
              make_uint4(w[4 * pl], w[4 * pl + 1], w[4 * pl + 2], w[4 * pl + 3]);
      };
      auto put_xa = [&](const uint32_t* w) {
        if (!edge) {
          row_store(t * 128 + m, w);
          return;
        }
        const int tau = g0 + t * 128 + m;
        if (tau < 0 || tau >= L) return;
        row_store(tau - g0, w);
        if (tau >= 1 && tau <= G && -tau - g0 >= -G) row_store(-tau - g0, w);
        const int back = L - 1 - tau;
        if (back >= 1 && back <= G && L - 1 + back - g0 < TR + G) row_store(L - 1 + back - g0, w);
      };
      // Tail transposed conv: no mirrors — the rows either side of the sequence are its zero padding.
      auto put_xa_up = [&](const uint32_t* w) {
        if (!edge) {
          row_store(t * 128 + m, w);
          return;
        }
        const int tau = g0 + t * 128 + m;
        if (tau < 0 || tau >= L) return;
        row_store(tau - g0, w);
        uint32_t z[CW / 2];
#pragma unroll
        for (int i = 0; i < CW / 2; ++i) z[i] = 0;
        if (tau == 0) row_store(-1 - g0, z);
        if (tau == L - 1) row_store(L - g0, z);
      };

      // Ring mode only: accumulators start from the bias row (tcgen05.st), then every MMA accumulates.
      auto init_acc = [&](uint32_t taddr, const float* bias_row) {
        uint32_t bv[CW];
        const float4* src = reinterpret_cast<const float4*>(bias_row + c0);
#pragma unroll
        for (int i = 0; i < CW / 4; ++i) {
          const float4 f = src[i];
          bv[4 * i] = __float_as_uint(f.x); bv[4 * i + 1] = __float_as_uint(f.y);
          bv[4 * i + 2] = __float_as_uint(f.z); bv[4 * i + 3] = __float_as_uint(f.w);
        }
        tmem_st<CW>(taddr + c0, bv);
      };
      auto init_acc16 = [&](uint32_t taddr, const float* bias_row, int c) {
        uint32_t bv[16];
        const float4* src = reinterpret_cast<const float4*>(bias_row + c0 + c);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 f = src[i];
          bv[4 * i] = __float_as_uint(f.x); bv[4 * i + 1] = __float_as_uint(f.y);
          bv[4 * i + 2] = __float_as_uint(f.z); bv[4 * i + 3] = __float_as_uint(f.w);
        }
        tmem_st<16>(taddr + c0 + c, bv);
      };

      // Where this thread's row goes in XA for the whole item (byte offsets within a plane, -1 = nowhere): its own
      // staged row, the mirror rows ReflectionPad1d reads near a sequence end, and for the fused transposed conv
      // the zero rows either side of the sequence.  Computed once per item: epilogue 2 stores chunk by chunk.
      int o_main = (G + t * 128 + m) * 16, o_mlo = -1, o_mhi = -1, o_zlo = -1, o_zhi = -1;
      if (edge) {
        const int tau = g0 + t * 128 + m;
        if (tau < 0 || tau >= L) o_main = -1;
        else {
          if (tau >= 1 && tau <= G && -tau - g0 >= -G) o_mlo = (G - tau - g0) * 16;
          const int back = L - 1 - tau;
          if (back >= 1 && back <= G && L - 1 + back - g0 < TR + G) o_mhi = (G + L - 1 + back - g0) * 16;
          if (tau == 0) o_zlo = (G - 1 - g0) * 16;
          if (tau == L - 1) o_zhi = (G + L - g0) * 16;
        }
      }
      // 16 channels (two planes) of this thread's row -> XA.  `zero_pad`: the round feeds the fused transposed conv
      auto xa_store16 = [&](int c, const uint32_t* a8, bool zero_pad) {
        uint8_t* base = xa + (size_t)(c / 8) * plane_bytes;
        const uint4 lo = make_uint4(a8[0], a8[1], a8[2], a8[3]), hi = make_uint4(a8[4], a8[5], a8[6], a8[7]);
        if (!edge) {
          *reinterpret_cast<uint4*>(base + o_main) = lo;
          *reinterpret_cast<uint4*>(base + plane_bytes + o_main) = hi;
          return;
        }
        if (o_main < 0) return;
        *reinterpret_cast<uint4*>(base + o_main) = lo;
        *reinterpret_cast<uint4*>(base + plane_bytes + o_main) = hi;
        if (!zero_pad) {
          if (o_mlo >= 0) {
            *reinterpret_cast<uint4*>(base + o_mlo) = lo;
            *reinterpret_cast<uint4*>(base + plane_bytes + o_mlo) = hi;
          }
          if (o_mhi >= 0) {
            *reinterpret_cast<uint4*>(base + o_mhi) = lo;
            *reinterpret_cast<uint4*>(base + plane_bytes + o_mhi) = hi;
          }
        } else {
          const uint4 z = make_uint4(0, 0, 0, 0);
          if (o_zlo >= 0) {
            *reinterpret_cast<uint4*>(base + o_zlo) = z;
            *reinterpret_cast<uint4*>(base + plane_bytes + o_zlo) = z;
          }
          if (o_zhi >= 0) {
            *reinterpret_cast<uint4*>(base + o_zhi) = z;
            *reinterpret_cast<uint4*>(base + plane_bytes + o_zhi) = z;
          }
        }
      };

      // ---- prologue: raw x tile -> XR[t] (TMEM, bf16) and lrelu(x) in place; accumulators <- block 0 biases
      if (EARLY_STAGE) {
        if (!staged) stage_item(it);
        tr(200);
        tmem_wait_st();          // raw x in XR[t] (and, first item, the ones operand)
        fence_proxy_async();
        tc_fence_before();
        mbar_arrive(&bars.xa_ready[t]);
        if (SPLIT_E2) mbar_arrive(&bars.xr_ready[t]);
        tr(201);
      } else {
      mbar_wait_relaxed(&bars.x_full[buf], xpar(it));
      tr(200);
      {
        uint32_t w[CW / 2];
        const uint8_t* src = xa + (size_t)(G + t * 128 + m) * 16;
#pragma unroll
        for (int pl = 0; pl < CW / 8; ++pl) {
          const uint4 u = *reinterpret_cast<const uint4*>(src + (size_t)pl * plane_bytes);
          w[4 * pl] = u.x; w[4 * pl + 1] = u.y; w[4 * pl + 2] = u.z; w[4 * pl + 3] = u.w;
        }
        // critical half first (what the dilated conv needs), then raw x for the 1x1 shortcut
        if (!BIAS_MMA && !OVERLAY) init_acc(tlane, s_bias[0][0]);
        if (!SPLIT_E2) tmem_st<CW / 2>(tlane + XR_OFF + c0 / 2, w);
        uint32_t a[CW / 2];
#pragma unroll
        for (int i = 0; i < CW / 2; ++i) a[i] = lrelu_bf16x2(w[i], slope2);
        put_xa(a);
        tmem_wait_st();
        fence_proxy_async();
        tc_fence_before();
        mbar_arrive(&bars.xa_ready[t]);
        if (SPLIT_E2) {
          tmem_st<CW / 2>(tlane + XR_OFF + c0 / 2, w);
          if (!BIAS_MMA) init_acc(tlane + CP, s_bias[0][1]);
          tmem_wait_st();
          tc_fence_before();
          mbar_arrive(&bars.xr_ready[t]);
        }
        tr(201);
      }
      }

      for (int blk = 0; blk < nb; ++blk, ++gb) {
        const uint32_t ph = gb & 1;
        const bool last = blk == nb - 1;
        // ---- epilogue 1: h = lrelu(D1) -> bf16 over the low columns of D1 (A operand of the 1x1 conv)
        {
          mbar_wait(&bars.d1_full[t], n_d1++ & 1);
          tc_fence_after();
          tr(210);
          trw(310);
          uint32_t w[CW / 2];
          if (OVERLAY) {
            // D1 carries no bias row (raw x lived in its upper half): add b1 here; 16 columns at a time with the next
            // TMEM load in flight, so that 64 channels per thread do not need 64 + 32 live registers
            uint32_t v[2][16];
            tmem_ld16_nw(tlane + c0, v[0]);
#pragma unroll
            for (int c = 0; c < CW; c += 16) {
              uint32_t* vc = v[(c / 16) & 1];
              tmem_wait_ld();
              if (c + 16 < CW) tmem_ld16_nw(tlane + c0 + c + 16, v[((c / 16) & 1) ^ 1]);
              const float4* bb = reinterpret_cast<const float4*>(&s_bias[RING ? blk : 0][0][RING ? c0 + c : 0]);
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const float4 b4 = bb[i];
                w[c / 2 + 2 * i] = pack_bf16x2(lrelu_f(__uint_as_float(vc[4 * i]) + b4.x, slope),
                                               lrelu_f(__uint_as_float(vc[4 * i + 1]) + b4.y, slope));
                w[c / 2 + 2 * i + 1] = pack_bf16x2(lrelu_f(__uint_as_float(vc[4 * i + 2]) + b4.z, slope),
                                                   lrelu_f(__uint_as_float(vc[4 * i + 3]) + b4.w, slope));
              }
            }
          } else {
          uint32_t v[CW];
          tmem_ld_nw<CW>(tlane + c0, v);
          tmem_wait_ld();
          tr(212);
          if (DBG(64)) {
#pragma unroll
            for (int i = 0; i < CW / 2; ++i) w[i] = v[i];
          } else {
#pragma unroll
            for (int i = 0; i < CW / 2; ++i)
              w[i] = pack_bf16x2(lrelu_f(__uint_as_float(v[2 * i]), slope), lrelu_f(__uint_as_float(v[2 * i + 1]), slope));
          }
          }
          tr(213);
          // every warp of the quarter must have read D1 before any of them overwrites its low columns
          if (CS > 1) asm volatile("bar.sync %0, %1;" ::"r"(pair_bar), "n"(CS * 32) : "memory");
          if (!DBG(2)) tmem_st<CW / 2>(tlane + c0 / 2, w);
          tmem_wait_st();
          tr(214);
          tc_fence_before();
          mbar_arrive(&bars.h_ready[t]);
          tr(211);
          trw(311);
        }
        // ---- epilogue 2: x_new = D2 -> XR[t] (raw) and XA (lrelu), or the stage output
        {
          mbar_wait(&bars.d2_full[t], ph);
          // This tile's XA rows (and, at a sequence end, mirror rows inside the next tile) are about to be rewritten,
          // and the NEXT tile's dilated-conv MMAs of this block read the last d of them (tap 0).  Nothing this warp has
          // waited for covers those MMAs: d1_full[t] was committed before they were issued.  They are issued right
          // behind it, a whole epilogue 1 + 1x1 round before this point, so the wait never spins — but a late issue
          // (instruction-cache miss with several kernels resident, a starved warp) did let the rewrite win: one
          // inference in a few hundred had d rows of one block computed from the block's OUTPUT.
          if (KT > 1 && t + 1 < KT) mbar_wait(&bars.d1_full[t + 1], (n_d1 - 1) & 1);
          tc_fence_after();
          tr(220);
          trw(320);
          if (CHUNKED_E2 && (!last || tail)) {
            // What the next MMA round needs: lrelu(x_new) in XA, raw x_new in XR[t] and the accumulators' bias rows
            // (the 1x1 MMAs of this tile have retired, so D1 | D2 are free).  Done in chunks of 16 columns — load,
            // convert, store, with the next chunk's TMEM load in flight — so that the live set stays ~50 registers
            // (the all-at-once form held 2 * CW values and spilled in this loop) and the first stores start early.
            static_assert(CW % 16 == 0, "epilogue 2 works in 16-column chunks");
            uint32_t v[2][16];
            tmem_ld16_nw(tlane + CP + c0, v[0]);
#pragma unroll
            for (int c = 0; c < CW; c += 16) {
              uint32_t* vc = v[(c / 16) & 1];
              tmem_wait_ld();
              if (c + 16 < CW) tmem_ld16_nw(tlane + CP + c0 + c + 16, v[((c / 16) & 1) ^ 1]);
              uint32_t w8[8], a8[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) w8[i] = pack_bf16x2(__uint_as_float(vc[2 * i]), __uint_as_float(vc[2 * i + 1]));
#pragma unroll
              for (int i = 0; i < 8; ++i) a8[i] = lrelu_bf16x2(w8[i], slope2);
              xa_store16(c, a8, last && up);
              if (!last) {
                tmem_st8(tlane + XR_OFF + (c0 + c) / 2, w8);
                if (!BIAS_MMA && !OVERLAY) init_acc16(tlane, s_bias[RING ? blk + 1 : 0][0], c);
              }
            }
            tr(225);
            tmem_wait_st();
            fence_proxy_async();
            tr(226);
            tc_fence_before();
            mbar_arrive(&bars.xa_ready[t]);
            tr(221);
            trw(321);
            if (SPLIT_E2 && !last) {
              if (!BIAS_MMA) {   // off the critical path: the D2 bias row, needed only after the next dilated conv
                init_acc(tlane + CP, s_bias[RING ? blk + 1 : 0][1]);
                tmem_wait_st();
                tc_fence_before();
              }
              mbar_arrive(&bars.xr_ready[t]);   // raw x went out with the chunks above
            }
          } else if (!last || tail) {
            // Narrow stages (two CTAs per SM, instruction-issue bound): everything at once — one TMEM load and one
            // wait; the chunked form's extra waits cost more than its shorter live ranges gain (measured: +10 % at 48).
            uint32_t v[CW], w[CW / 2];
            tmem_ld_nw<CW>(tlane + CP + c0, v);
            tmem_wait_ld();
            tr(223);
#pragma unroll
            for (int i = 0; i < CW / 2; ++i) w[i] = pack_bf16x2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]));
            if (!SPLIT_E2 && !last && !DBG(2)) tmem_st<CW / 2>(tlane + XR_OFF + c0 / 2, w);
            uint32_t a[CW / 2];
#pragma unroll
            for (int i = 0; i < CW / 2; ++i) a[i] = DBG(32) ? w[i] : lrelu_bf16x2(w[i], slope2);
            tr(224);
            if (last && head) tmem_st<CW / 2>(tlane + XR_OFF + c0 / 2, a);   // A operand of the fused head (TMEM)
            else if (last && up) put_xa_up(a);
            else put_xa(a);
            tr(225);
            tmem_wait_st();
            if (!DBG(8) && !(last && head)) fence_proxy_async();
            tr(226);
            tc_fence_before();
            mbar_arrive(&bars.xa_ready[t]);
            tr(221);
            trw(321);
            if (SPLIT_E2 && !last) {
              if (!DBG(2)) tmem_st<CW / 2>(tlane + XR_OFF + c0 / 2, w);
              tmem_wait_st();
              tc_fence_before();
              mbar_arrive(&bars.xr_ready[t]);
            }
          } else {
            uint32_t v[CW], w[CW / 2];
            tmem_ld_nw<CW>(tlane + CP + c0, v);
            tmem_wait_ld();
            const int tau = g0 + t * 128 + m;
            const bool own = tau >= chain * V && tau < (chain + 1) * V && tau < L;
            if (own) {
              const TcCanonOut& o = p.out;
              if (o.lrelu) {
#pragma unroll
                for (int i = 0; i < CW / 2; ++i)
                  w[i] = pack_bf16x2(lrelu_f(__uint_as_float(v[2 * i]), slope), lrelu_f(__uint_as_float(v[2 * i + 1]), slope));
              } else {
#pragma unroll
                for (int i = 0; i < CW / 2; ++i) w[i] = pack_bf16x2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]));
              }
              __nv_bfloat16* ob = o.base + (int64_t)b * o.batch_stride + (int64_t)(c0 / 8) * o.plane_rows * 8;
              auto put = [&](int row, bool zero) {
#pragma unroll
                for (int pl = 0; pl < CW / 8; ++pl)
                  *reinterpret_cast<uint4*>(ob + ((int64_t)pl * o.plane_rows + row) * 8) =
                      zero ? make_uint4(0, 0, 0, 0) : make_uint4(w[4 * pl], w[4 * pl + 1], w[4 * pl + 2], w[4 * pl + 3]);
              };
              put(o.halo + tau, false);
              if (edge) {
                if (o.margin == 1) {          // halo rows mirror the interior (ReflectionPad1d of the consumer)
                  if (tau >= 1 && tau <= o.halo) put(o.halo - tau, false);
                  const int back = o.L - 1 - tau;
                  if (back >= 1 && back <= o.halo) put(o.halo + o.L - 1 + back, false);
                } else if (o.margin == 2) {   // zero rows either side (transposed conv of the consumer)
                  if (tau == 0) put(o.halo - 1, true);
                  if (tau == o.L - 1) put(o.halo + o.L, true);
                }
              }
            }
            tr(222);
          }
        }
      }
      if (EARLY_STAGE) {
        staged = blockIdx.x + (it + 1) * gridDim.x < n_items;
        if (staged) stage_item(it + 1);
      }
      if (head) {
        // ---- epilogue 3: y[t][c] = bias[c] + sum_j P[reflect(t + j - 3)][j * ch + c], tanh -> fp32 [batch][ch][L].
        // Thread = row: the shifted rows come from other lanes by shuffle; rows that belong to a neighbouring warp
        // (first / last three lanes) come from that warp's edge rows in shared memory.  ReflectionPad1d of the head
        // (melgan_generator.py:66) is the index map reflect(): its sources are never more than three rows away.
        auto head_epi = [&](auto ch_c) {
          constexpr int CH = decltype(ch_c)::value;
          constexpr int HT = 7, HN = CH == 4 ? 32 : 16;
          constexpr int LAST_W = 4 * KT - 1;
          const int wic = t * 4 + q4;                       // this warp's position in the chain (32 rows each)
          float* eb = sEdge + wic * (6 * 32);
          mbar_wait(&bars.d1_full[t], n_d1++ & 1);
          tc_fence_after();
          tr(240);
          uint32_t v[HN];
          tmem_ld_nw<HN>(tlane, v);
          tmem_wait_ld();
          tr(241);
          if (lane < 3 || lane >= 29) {
            float4* dst = reinterpret_cast<float4*>(eb + (lane < 3 ? lane : lane - 26) * 32);
#pragma unroll
            for (int i = 0; i < (HT * CH + 3) / 4; ++i)
              dst[i] = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]), __uint_as_float(v[4 * i + 2]),
                                   __uint_as_float(v[4 * i + 3]));
          }
          mbar_arrive(&bars.edge[wic]);                     // release: the rows above are visible to the waiters
          const int tau = g0 + t * 128 + m;
          const bool own = tau >= chain * V && tau < (chain + 1) * V && tau < L;
          float y[CH];
#pragma unroll
          for (int c = 0; c < CH; ++c) y[c] = head_bias[c];
          int sl[HT];                                       // source lane of tap j (outside [0, 32): a neighbour's row)
#pragma unroll
          for (int j = 0; j < HT; ++j) {
            int sr = tau + j - 3;
            sr = sr < 0 ? -sr : sr;
            sr = sr >= L ? 2 * (L - 1) - sr : sr;
            sl[j] = lane + (sr - tau);
#pragma unroll
            for (int c = 0; c < CH; ++c) {
              const float sv = __shfl_sync(0xffffffffu, __uint_as_float(v[j * CH + c]), sl[j] & 31);
              if ((unsigned)sl[j] < 32u) y[c] += sv;
            }
          }
          tr(243);
          if (wic > 0) mbar_wait(&bars.edge[wic - 1], it & 1);
          if (wic < LAST_W) mbar_wait(&bars.edge[wic + 1], it & 1);
          tr(244);
          if (own) {
#pragma unroll
            for (int j = 0; j < HT; ++j) {
              const float* nb = nullptr;
              if (sl[j] < 0 && wic > 0) nb = eb - 6 * 32 + (sl[j] + 6) * 32 + j * CH;          // rows 29..31 of warp wic - 1
              else if (sl[j] >= 32 && wic < LAST_W) nb = eb + 6 * 32 + (sl[j] - 32) * 32 + j * CH;   // rows 0..2 of wic + 1
              if (nb) {
#pragma unroll
                for (int c = 0; c < CH; ++c) y[c] += nb[c];
              }
            }
            float* of = p.head_out + (int64_t)b * CH * L + tau;
#pragma unroll
            for (int c = 0; c < CH; ++c) of[(int64_t)c * L] = tanh_fast(y[c]);
          }
        };
        if (p.head_ch == 4) head_epi(std::integral_constant<int, 4>{});
        else head_epi(std::integral_constant<int, 1>{});
        tr(245);
      }
      auto signal_item = [&]() {   // this warp's stores of the item are out: count it for utterance b
        if (p.done_out) {
          __syncwarp();
          if (lane == 0) smem_count_add(&bars.items_done[it & 1]);
        }
      };
      if (up) {
        // ---- epilogue 3: transposed-conv accumulator -> next stage's input; GEMM row q holds output rows
        // q f - P .. q f - P + f - 1, columns ordered [plane][phase][8]
        mbar_wait(&bars.d1_full[t], n_d1++ & 1);
        tc_fence_after();
        tr(230);
        const int tau = g0 + t * 128 + m;
        const bool own = tau >= chain * V && tau < (chain + 1) * V && tau <= L;   // row L: x[L-1] w[r+f] tail
        const TcCanonOut& o = p.up_out;
        const int f = p.up_f;
        const int n0 = tau * f - p.up_P;
        // column unit u (8 channels) = (plane u / f, phase u % f): walk (plane, phase) and the store pointer together
        int pl = (cs * UW / 8) / f, phi = (cs * UW / 8) - pl * f;
        __nv_bfloat16* optr = o.base + (int64_t)b * o.batch_stride + ((int64_t)pl * o.plane_rows + o.halo + n0 + phi) * 8;
        const bool interior = own && n0 >= 0 && n0 + f <= o.L;   // every phase of this row lands inside the output
        // Even f with (halo - P) even (the host shifts the tensor's halo by a row if needed): the two units of a
        // 16-column batch are rows n, n+1 of one plane at a 32-byte aligned address -> one 256-bit store, and a
        // warp's stores tile a contiguous span (16-byte stores at a 32-byte stride crawl through the LSU).
        const bool paired = (f & 1) == 0 && ((o.halo - p.up_P) & 1) == 0;
        const int64_t wrap = ((int64_t)o.plane_rows - (f - 1)) * 8;
        for (int c = cs * UW; c < (cs + 1) * UW; c += 16) {
          uint32_t v[16];
          tmem_ld16_nw(tlane + c, v);
          tmem_wait_ld();
          uint32_t w[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) w[i] = pack_bf16x2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]));
          if (paired && interior) {
            if (!DBG(256)) st_global_v8(optr, w);
            phi += 2;
            optr += phi == f ? wrap + 8 : 16;
            if (phi == f) phi = 0;
            continue;
          }
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            if (own && n0 + phi >= 0 && n0 + phi < o.L)
              *reinterpret_cast<uint4*>(optr) = make_uint4(w[4 * h], w[4 * h + 1], w[4 * h + 2], w[4 * h + 3]);
            if (++phi == f) {
              phi = 0;
              optr += wrap;
            } else {
              optr += 8;
            }
          }
        }
        tr(231);
      }
      signal_item();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == PRODUCER_WARP) tmem_dealloc(tbase, TMEM_ALLOC);
}

constexpr size_t SMEM_MAX = 227 * 1024 - 8 * 1024;   // static: barriers, TMEM base, ring mode's fp32 bias rows (6 KB)

// Bias operands of the bias MMAs: {b1, b2 + bs} per block and the fused tail's bias as [n x 16 B] planes, one zero plane.
// Fused head: packed weights [cp/8][head_n][8] + per epilogue warp six accumulator rows of 32 fp32 (edge exchange).
size_t head_smem_bytes(const TcStackParams& p) {
  return p.head_w ? (size_t)p.cp * p.head_n * 2 + (size_t)4 * tc_stack_kt(p.cp) * 6 * 32 * 4 : 0;
}

size_t bias_smem_bytes(const TcStackParams& p) {
  if (p.cp > 96) return 0;   // ring mode pre-loads its accumulators instead
  if (p.k24) return ((size_t)p.nblocks * 2 * p.cp + (p.up_w ? (size_t)p.up_n : 0) +
                     std::max<size_t>(p.cp, p.up_w ? (size_t)p.up_n : 0)) * 16 + 128 * 16;   // + the ones plane
  const size_t tail_n = p.up_w ? (size_t)p.up_n : 0;
  return ((size_t)p.nblocks * 2 * p.cp + tail_n + std::max<size_t>(p.cp, tail_n)) * 16;
}

// cudaFuncSetAttribute is per device and costs a driver call: done once per (instantiation, device).
template <int CP, int KT, int CS, int TAIL>
cudaError_t launch(const TcStackParams& p, size_t smem, int grid, int device, cudaStream_t stream) {
  auto kern = tc_stack_kernel<CP, KT, CS, TAIL>;
  static std::atomic<uint64_t> attr_done{0};
  const uint64_t bit = 1ull << (device & 63);
  if (!(attr_done.load(std::memory_order_acquire) & bit)) {
    // dynamic + static shared memory share the 227 KB of a CTA: the C = 256 instantiation carries 8 KB of fp32 bias rows
    cudaFuncAttributes fa{};
    cudaError_t e = cudaFuncGetAttributes(&fa, kern);
    if (e != cudaSuccess) return e;
    const int dyn_max = std::min<int>((int)SMEM_MAX, 227 * 1024 - (int)fa.sharedSizeBytes);
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_max);
    if (e != cudaSuccess) return e;
    attr_done.fetch_or(bit, std::memory_order_release);
  }
  return launch_pdl(kern, dim3(grid), dim3(SERVICE_THREADS + KT * CS * 128), smem, stream, p);
}

}  // namespace

int tc_stack_ctas_per_sm(int cp) { return tc_stack_kt(cp) * 5 * cp / 2 + 8 <= 256 ? 2 : 1; }

int tc_stack_cs(int cp) { return cp >= 128 ? 4 : cp == 96 ? 2 : 1; }

// 128-row tiles per work item.  Narrow stages (cp <= 48) run two CTAs per SM with half the TMEM each; one CTA per
// SM with twice the tiles measured slower (cp = 32: KT 6 -> 0.099 ms vs KT 3 x 2 CTAs -> 0.087 ms): a single conv
// issuer then has to feed every tile.  TTSB200_STACK_KT_<cp>=<kt> overrides the choice for experiments (the value
// must have an instantiation in dispatch()).
int tc_stack_kt(int cp) {
  static int cache[4][2] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}};   // cp -> kt, filled on first use per cp
  int dflt;
  switch (cp) {
    case 256: return 1;
    case 192: return 1;
    case 128: return 1;
    case 96: return 2;
    case 64: return 3;
    case 48: dflt = 2; break;
    case 32: dflt = 3; break;
    default: return 0;
  }
  const int slot = cp == 48 ? 0 : 1;
  if (cache[slot][0] != cp) {
    char name[32];
    snprintf(name, sizeof name, "TTSB200_STACK_KT_%d", cp);
    const char* e = getenv(name);
    cache[slot][1] = e ? atoi(e) : dflt;
    cache[slot][0] = cp;
  }
  return cache[slot][1];
}

// Fills kt / guard / w_sets for a stage (cp, dilations, nblocks set by the caller); returns false when the
// stage cannot run on the fused kernel.
bool tc_stack_configure(TcStackParams& p) {
  p.kt = tc_stack_kt(p.cp);
  if (p.kt == 0 || p.res_taps != TAPS || p.nblocks < 1 || p.nblocks > TC_STACK_MAX_BLOCKS) return false;
  int dmax = 1;
  p.sum_d = 0;
  for (int i = 0; i < p.nblocks; ++i) {
    p.sum_d += p.dil[i];
    dmax = dmax > p.dil[i] ? dmax : p.dil[i];
  }
  p.guard = (dmax + 7) / 8 * 8;
  const bool head = p.head_w != nullptr;
  // fused head: 7 taps x {1, 4} channels packed on N (the epilogue's shuffle sum is written for those shapes)
  if (head && (p.cp > 64 || p.head_taps != 7 || (p.head_ch != 1 && p.head_ch != 4) ||
               p.head_n != (p.head_ch == 4 ? 32 : 16)))
    return false;
  const bool up = p.up_w != nullptr;
  if (up) {
    // the tail conv's accumulator overlays D1|D2, its weights take the streamed W1 slot, and with a channel
    // split it must stay clear of the partner warp's D2 columns
    const int cs = tc_stack_cs(p.cp), n = p.up_n;
    if (head || p.cp > 96 || n < 16 || n > 256 || n % (16 * cs) || n > 2 * p.cp || (cs > 1 && n > p.cp) ||
        p.up_f < 1 || n % (8 * p.up_f))
      return false;
  }
  const int hh = head ? (p.head_taps - 1) / 2 : up ? 1 : 0;
  if (hh > p.guard) p.guard = (hh + 7) / 8 * 8;
  if (p.kt * 128 - 2 * (p.sum_d + hh) < 32 || p.guard + p.sum_d + hh > 127) return false;
  const size_t xa = (p.cp > 192 ? 1 : 2) * (size_t)(p.cp / 8) * (p.kt * 128 + 2 * p.guard) * 16;   // C = 256: one x tile
  const size_t wset = (size_t)(TAPS + 2) * p.cp * p.cp * 2;
  const size_t wh = head_smem_bytes(p) + bias_smem_bytes(p);
  // two CTAs per SM when the TMEM footprint allows it: each then gets half of the shared memory
  const size_t budget = tc_stack_ctas_per_sm(p.cp) == 2 ? SMEM_MAX / 2 - 1024 : SMEM_MAX;
  if (p.cp > 96) {   // ring mode: w_sets = number of [RING_KB x cp] chunk slots
    if (head) return false;
    const size_t slot = (size_t)RING_KB * p.cp * 2;
    if (xa + wh + 4 * slot > budget) return false;
    const size_t n = (budget - xa - wh) / slot;
    p.w_sets = (int)(n > RING_MAX_SLOTS ? RING_MAX_SLOTS : n);
    return true;
  }
  if (xa + wset + wh > budget) return false;
  p.w_sets = (!up && xa + p.nblocks * wset + wh <= budget) ? p.nblocks : 1;
  if (up) {   // tail weights: resident beside the block weights when they fit, else sharing the streamed W2S slot
    const size_t wu = (size_t)2 * p.cp * p.up_n * 2;
    if (xa + p.nblocks * wset + wu <= budget) {
      p.w_sets = p.nblocks;
      p.up_mode = 1;
    } else if (xa + wset + wu <= budget) p.up_mode = 1;
    else if (p.up_n <= p.cp) p.up_mode = 2;
    else return false;
  }
  return true;
}

cudaError_t dispatch(const TcStackParams& p, size_t smem, int grid, int device, cudaStream_t stream);

size_t tc_stack_smem_bytes(const TcStackParams& p) {
  if (p.cp > 96)
    return (p.cp > 192 ? 1 : 2) * (size_t)(p.cp / 8) * (p.kt * 128 + 2 * p.guard) * 16 +
           (size_t)p.w_sets * RING_KB * p.cp * 2 + bias_smem_bytes(p);
  return bias_smem_bytes(p) + 2 * (size_t)(p.cp / 8) * (p.kt * 128 + 2 * p.guard) * 16 +
         (size_t)p.w_sets * (TAPS + 2) * p.cp * p.cp * 2 +
         head_smem_bytes(p) +
         (p.up_w && p.up_mode == 1 ? (size_t)2 * p.cp * p.up_n * 2 : 0);
}

void tc_stack_count_info(const TcStackParams& p, int* n_units, int* unit_rows, int* row_offset) {
  const int V = p.kt * 128 - 2 * (p.sum_d + (p.head_w ? (p.head_taps - 1) / 2 : p.up_w ? 1 : 0));
  *n_units = ceil_div(p.L + (p.up_w ? 1 : 0), V);   // chains per utterance
  // chain c owns GEMM rows [c V, (c + 1) V); with the fused transposed conv they become output rows
  // [c V f - P, (c + 1) V f - P)
  *unit_rows = V * (p.up_w ? p.up_f : 1);
  *row_offset = p.up_w ? p.up_P : 0;
}

cudaError_t tc_stack_launch(const TcStackParams& p_in, int batch, const TcLaunchCtx& ctx, cudaStream_t stream) {
  TcStackParams p = p_in;
  if (p.kt != tc_stack_kt(p.cp) || p.kt == 0) return cudaErrorInvalidConfiguration;
  const size_t smem = tc_stack_smem_bytes(p);
  if (smem > SMEM_MAX) return cudaErrorInvalidConfiguration;
  const int V = p.kt * 128 - 2 * (p.sum_d + (p.head_w ? (p.head_taps - 1) / 2 : p.up_w ? 1 : 0));
  p.batch = batch;
  p.n_chains = ceil_div(p.L + (p.up_w ? 1 : 0), V);   // the transposed conv also has GEMM row L
  const int items = p.n_chains * batch;
  const int slots = ctx.sms * tc_stack_ctas_per_sm(p.cp);
  int grid = items < slots ? items : slots;
  // test hook (read at plan creation): few CTAs, so that each walks many items on small inputs
  if (ctx.stack_max_grid > 0) grid = std::max(1, std::min(grid, ctx.stack_max_grid));
  p.debug = ctx.stack_debug;
  p.trace = trace_begin(ctx.trace_file, stream);
  const cudaError_t rc = dispatch(p, smem, grid, ctx.device, stream);
  trace_end(ctx.trace_file, p.trace, p.cp, stream);
  return rc;
}

cudaError_t dispatch(const TcStackParams& p, size_t smem, int grid, int device, cudaStream_t stream) {
  const int tail = p.head_w ? 1 : p.up_w ? 2 : 0;
#define TTSB_STACK_CASE(CP, KT, CS, TAIL) \
  if (p.cp == CP && p.kt == KT && tail == TAIL) return launch<CP, KT, CS, TAIL>(p, smem, grid, device, stream)
  TTSB_STACK_CASE(256, 1, 4, 0);
  TTSB_STACK_CASE(192, 1, 4, 0);
  TTSB_STACK_CASE(128, 1, 4, 0);
  TTSB_STACK_CASE(96, 2, 2, 0);
  TTSB_STACK_CASE(96, 2, 2, 2);
  TTSB_STACK_CASE(64, 3, 1, 0);
  TTSB_STACK_CASE(64, 3, 1, 1);
  TTSB_STACK_CASE(64, 3, 1, 2);
  TTSB_STACK_CASE(48, 4, 1, 0);
  TTSB_STACK_CASE(48, 4, 1, 1);
  TTSB_STACK_CASE(48, 4, 1, 2);
  TTSB_STACK_CASE(32, 6, 1, 0);
  TTSB_STACK_CASE(32, 6, 1, 1);
  TTSB_STACK_CASE(32, 6, 1, 2);
  // experiment variants (TTSB200_STACK_KT_<cp>)
  TTSB_STACK_CASE(48, 2, 1, 0);
  TTSB_STACK_CASE(48, 2, 1, 1);
  TTSB_STACK_CASE(48, 2, 1, 2);
  TTSB_STACK_CASE(32, 3, 1, 0);
  TTSB_STACK_CASE(32, 3, 1, 1);
  TTSB_STACK_CASE(32, 3, 1, 2);
  TTSB_STACK_CASE(32, 5, 1, 1);
#undef TTSB_STACK_CASE
  return cudaErrorInvalidConfiguration;
}

}  // namespace ttsb
