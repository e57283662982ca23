// BF16 implicit-GEMM 1-D convolution on tcgen05 / TMEM for the vocoder's dense contractions:
// k7 projection and head, ConvTranspose1d upsampling (as two phase classes of 2-tap convs), the
// dilated k3 residual conv and the fused (1x1 conv + 1x1 shortcut) of ResidualStack
// (TTS/vocoder/models/melgan_generator.py:29-78, TTS/vocoder/layers/melgan.py:13-38).
//
// GEMM view per CTA: D[128 time rows, ncols <= 256 out-channels] += A[128, K] * W[ncols, K]^T with
// K = sources x taps x C_in.  The activation tile (128 + tap span rows, all channels) is brought in
// ONCE by cp.async.bulk straight into the canonical UMMA layout; every tap of a (dilated) conv is
// the same tile read at a different start row, i.e. only the A descriptor's start address
// changes (verified on hardware by tests/cuda/umma_probe.cu).  Weights stream through a 3-stage
// bulk-copy ring.  fp32 accumulators live in TMEM; the epilogue warps read them back with
// tcgen05.ld, add bias, apply LeakyReLU / tanh and write the next layer's operand layout.
//
// Warp roles (192 threads): warp 0 = copy producer + TMEM owner, warp 1 = MMA issuer (one
// thread), warps 2..5 = LeakyReLU prologue in shared memory + epilogue (one TMEM lane quarter each).
#include <atomic>
#include <cstdlib>

#include "common.cuh"
#include "tc_common.cuh"
#include "tc_conv.cuh"
#include "trace.cuh"

namespace ttsb {

using namespace tc;

namespace {

constexpr int NT = 320;            // producer, MMA issuer, 8 epilogue warps (two per TMEM lane quarter)
constexpr int EPI_THREADS = 256;

// Accumulator columns [c, c + nc) of this warp's lanes <- bias row (the MMAs then always accumulate).
__device__ __forceinline__ void bias_to_acc(uint32_t trow, int c, int nc, int ncols, const float* s_bias) {
  uint32_t bv[32];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float4 f4 = *reinterpret_cast<const float4*>(&s_bias[min(c + 4 * i, ncols - 4)]);
    bv[4 * i] = __float_as_uint(f4.x); bv[4 * i + 1] = __float_as_uint(f4.y);
    bv[4 * i + 2] = __float_as_uint(f4.z); bv[4 * i + 3] = __float_as_uint(f4.w);
  }
  tmem_st<16>(trow + c, bv);
  if (nc > 16) tmem_st<16>(trow + c + 16, bv + 16);
}

// Accumulator tile -> output tensor, for one epilogue warp: TMEM lane quarter q4 (rows t0 + q4*32 + lane), the
// 32-column batches hf, hf+2, ... of group G.  REINIT: each batch is overwritten with the bias row as soon as it
// has been read, ready for the next tile's MMAs (persistent kernel).
template <bool REINIT>
__device__ __forceinline__ void conv_epilogue(const TcConvParams& p, const TcGroup& G, uint32_t trow, int q4, int hf,
                                              int lane, int t0, int b, const float* s_bias) {
  const int ncols = G.ncols;
  const int q = t0 + q4 * 32 + lane;       // GEMM row of this thread
  const bool row_ok = q < p.Lt;
  const int f = p.up_f;
  const int row0 = q * f - p.up_P;         // output row of phase 0
  __nv_bfloat16* ob = p.out_c + (int64_t)b * p.out_batch_stride;
  // (plane, phase) of the first 8-column unit; advanced incrementally (no division in the loop)
  int plane = ((G.col_first >> 3) + hf * 4) / f, r = ((G.col_first >> 3) + hf * 4) - plane * f;
  const int out_L = p.out_L, halo = p.out_halo, margin = p.margin;
  const bool to_canon = p.out_mode == 0, act = p.out_lrelu != 0;
  const float slope = p.slope;
  // rows that also own halo rows of the output (mirror / zero margins) are rare: test once per thread
  const int reach = margin == 1 ? halo : 0;
  const bool near_edge = row_ok && margin != 0 && (row0 <= reach || row0 + f - 1 >= out_L - 1 - reach);
  const bool whole = row_ok && row0 >= 0 && row0 + f <= out_L;   // every phase of this row lands inside the output
  // Even-stride transposed conv: units (plane, r), (plane, r + 1) are adjacent output rows, so a pair goes out as
  // one 32-byte store (the host keeps halo - P even, which makes the address 32-byte aligned).  16-byte stores
  // at a stride of f rows leave every sector half written per instruction and crawl.
  const bool paired = (f & 1) == 0 && ((halo - p.up_P) & 1) == 0;
  // one output row of one plane, plus the halo rows it owns near a sequence end
  auto put_row = [&](int pl, int tau, uint4 w4) {
    if (tau < 0 || tau >= out_L) return;
    __nv_bfloat16* pb = ob + (int64_t)pl * p.out_plane_rows * 8;
    *reinterpret_cast<uint4*>(pb + (int64_t)(halo + tau) * 8) = w4;
  };
  auto put_margins = [&](int pl, int tau, uint4 w4) {
    if (tau < 0 || tau >= out_L) return;
    __nv_bfloat16* pb = ob + (int64_t)pl * p.out_plane_rows * 8;
    if (margin == 1) {         // halo rows mirror the interior (ReflectionPad1d of the consumer)
      if (tau >= 1 && tau <= halo) *reinterpret_cast<uint4*>(pb + (int64_t)(halo - tau) * 8) = w4;
      const int back = out_L - 1 - tau;
      if (back >= 1 && back <= halo) *reinterpret_cast<uint4*>(pb + (int64_t)(halo + out_L - 1 + back) * 8) = w4;
    } else if (margin == 2) {  // zero rows either side (transposed conv of the consumer)
      const uint4 z = make_uint4(0, 0, 0, 0);
      if (tau == 0) *reinterpret_cast<uint4*>(pb + (int64_t)(halo - 1) * 8) = z;
      if (tau == out_L - 1) *reinterpret_cast<uint4*>(pb + (int64_t)(halo + out_L) * 8) = z;
    }
  };
  for (int c = hf * 32; c < ncols; c += 64) {
    const int nc = min(32, ncols - c);     // 16 or 32
    uint32_t v[32];
#ifdef TTSB_STACK_TRACE
    if (p.debug & 4) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = i + lane;
    } else
#endif
    {
    tmem_ld16_nw(trow + c, v);
    if (nc > 16) tmem_ld16_nw(trow + c + 16, v + 16);
    tmem_wait_ld();
    }
#ifdef TTSB_STACK_TRACE
    if (!(p.debug & 2))
#endif
    if (REINIT) bias_to_acc(trow, c, nc, ncols, s_bias);
#ifdef TTSB_STACK_TRACE
    if (p.debug & 8) {   // knock-out: no global stores
      for (int u = 0; u < 8; ++u)
        if (++r == f) {
          r = 0;
          ++plane;
        }
      continue;
    }
#endif
    if (to_canon) {
      uint32_t w[16];
      if (act) {
#pragma unroll
        for (int i = 0; i < 16; ++i)
          w[i] = pack_bf16x2(lrelu_f(__uint_as_float(v[2 * i]), slope), lrelu_f(__uint_as_float(v[2 * i + 1]), slope));
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) w[i] = pack_bf16x2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]));
      }
      if (paired) {
#pragma unroll
        for (int u = 0; u < 4; u += 2) {
          if (u * 8 < nc) {
            const int tau = row0 + r;
            if (whole) {
              st_global_v8(ob + ((int64_t)plane * p.out_plane_rows + halo + tau) * 8, w + 4 * u);
            } else if (row_ok) {
              put_row(plane, tau, make_uint4(w[4 * u], w[4 * u + 1], w[4 * u + 2], w[4 * u + 3]));
              put_row(plane, tau + 1, make_uint4(w[4 * u + 4], w[4 * u + 5], w[4 * u + 6], w[4 * u + 7]));
            }
            if (near_edge) {
              put_margins(plane, tau, make_uint4(w[4 * u], w[4 * u + 1], w[4 * u + 2], w[4 * u + 3]));
              put_margins(plane, tau + 1, make_uint4(w[4 * u + 4], w[4 * u + 5], w[4 * u + 6], w[4 * u + 7]));
            }
            r += 2;
            if (r == f) {
              r = 0;
              ++plane;
            }
          }
        }
      } else {
#pragma unroll
        for (int u = 0; u < 4; ++u) {          // 8-column units: one (plane, phase) each
          if (u * 8 < nc) {
            const uint4 w4 = make_uint4(w[4 * u], w[4 * u + 1], w[4 * u + 2], w[4 * u + 3]);
            if (row_ok) put_row(plane, row0 + r, w4);
            if (near_edge) put_margins(plane, row0 + r, w4);
            if (++r == f) {
              r = 0;
              ++plane;
            }
          }
        }
      }
    } else {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (u * 8 < nc) {
          const int tau = row0 + r;
          if (row_ok && tau >= 0 && tau < out_L) {
            float* of = p.out_f + ((int64_t)b * p.cout_real + plane * 8) * out_L + tau;
#pragma unroll
            for (int i = 0; i < 8; ++i)
              if (plane * 8 + i < p.cout_real) of[(int64_t)i * out_L] = tanh_fast(__uint_as_float(v[u * 8 + i]));
          }
          if (++r == f) {
            r = 0;
            ++plane;
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)              // skip the partner warp's batch
      if (++r == f) {
        r = 0;
        ++plane;
      }
    __syncwarp();   // edge lanes took other branches: reconverge before the next warp-wide tcgen05.ld
  }
}

// Interior tiles of the activation-resident kernel: every GEMM row of the tile is valid and all its output rows lie
// clear of the sequence ends (no mirror / zero margins to write), so the epilogue is load -> convert -> store with the
// addresses advanced incrementally: ~45 instructions per 32-column batch instead of the general routine's ~120 with
// its per-row edge predicates.  The accumulator already holds the bias (bias MMA) and is not re-initialised.
template <bool ACT>
__device__ __forceinline__ void conv_epilogue_fast(const TcConvParams& p, const TcGroup& G, uint32_t trow, int q4, int hf,
                                                   int lane, int t0, int b, float slope) {
  const int ncols = G.ncols, f = p.up_f;
  const int q = t0 + q4 * 32 + lane;
  const int u0 = (G.col_first >> 3) + hf * 4;          // first 8-column unit of this warp's first batch
  int plane = u0 / f, r = u0 - plane * f;
  const int64_t prow = (int64_t)p.out_plane_rows * 8;  // elements per plane
  __nv_bfloat16* ptr = p.out_c + (int64_t)b * p.out_batch_stride + (int64_t)plane * prow +
                       (int64_t)(p.out_halo + q * f - p.up_P + r) * 8;
  const bool paired = (f & 1) == 0 && ((p.out_halo - p.up_P) & 1) == 0;
  const int64_t wrap = prow - (int64_t)f * 8;          // pointer step when the phase wraps into the next plane
  auto step = [&](int units) {                         // advance (phase, pointer) by `units` 8-column units
    r += units;
    ptr += units * 8;
    if (r >= f) {                                      // units <= 2 and f >= units, or f == 1 handled by the loop below
      r -= f;
      ptr += wrap;
    }
  };
  for (int c = hf * 32; c < ncols; c += 64) {
    const int nc = min(32, ncols - c);
    uint32_t v[32], w[16];
    tmem_ld16_nw(trow + c, v);
    if (nc > 16) tmem_ld16_nw(trow + c + 16, v + 16);
    tmem_wait_ld();
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      float lo = __uint_as_float(v[2 * i]), hi = __uint_as_float(v[2 * i + 1]);
      if (ACT) {
        lo = lrelu_f(lo, slope);
        hi = lrelu_f(hi, slope);
      }
      w[i] = pack_bf16x2(lo, hi);
    }
    const int nu = nc >> 3;                            // 2 or 4 units in this batch
    if (paired) {
#pragma unroll
      for (int u = 0; u < 4; u += 2) {
        if (u < nu) st_global_v8(ptr, w + 4 * u);
        step(2);
      }
      step(2);                                         // the partner warp's batch: 4 units
      step(2);
    } else {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (u < nu) *reinterpret_cast<uint4*>(ptr) = make_uint4(w[4 * u], w[4 * u + 1], w[4 * u + 2], w[4 * u + 3]);
        step(1);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) step(1);
    }
  }
}

__global__ void __launch_bounds__(NT) tc_conv_kernel(const __grid_constant__ TcConvParams p, const int kb_stage,
                                                      const int nst) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar_a, bar_a_ready, bar_d, bar_full[TC_MAX_STAGES], bar_empty[TC_MAX_STAGES];
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float s_bias[256];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int t0 = blockIdx.x * TC_TILE_M;
  const TcGroup& G = p.grp[blockIdx.y];
  const int b = blockIdx.z;
  const int ncols = G.ncols;

  const uint32_t a_bytes0 = (uint32_t)p.src[0].planes * p.src[0].rows * 16;
  const uint32_t a_bytes1 = p.nsrc > 1 ? (uint32_t)p.src[1].planes * p.src[1].rows * 16 : 0;
  const uint32_t w_stage_bytes = (uint32_t)(kb_stage / 8) * ncols * 16;
  uint8_t* sA0 = smem;
  uint8_t* sA1 = smem + a_bytes0;
  uint8_t* sW = smem + a_bytes0 + a_bytes1;

  uint32_t tcols = 32;
  while (tcols < (uint32_t)ncols) tcols <<= 1;
  const bool need_lrelu = p.src[0].lrelu || (p.nsrc > 1 && p.src[1].lrelu);

  if (tid == 0) {
    mbar_init(&bar_a, 1);
    mbar_init(&bar_a_ready, EPI_THREADS);
    mbar_init(&bar_d, 1);
    for (int s = 0; s < nst; ++s) {
      mbar_init(&bar_full[s], 1);
      mbar_init(&bar_empty[s], 1);
    }
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(&tmem_base_s, tcols);
  for (int c = tid; c < ncols; c += NT) s_bias[c] = G.bias[c];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_base_s;
  pdl_launch_dependents();
  pdl_wait();

  if (warp == 0) {
    if (elect_one()) {
      Tracer tr = make_tracer(p.trace, 0, true);
      tr(10);
      // ---- activation tile(s): one bulk copy per 8-channel plane
      mbar_expect_tx(&bar_a, a_bytes0 + a_bytes1);
      for (int s = 0; s < p.nsrc; ++s) {
        const TcSrc& S = p.src[s];
        const __nv_bfloat16* gx = S.x + (int64_t)b * S.batch_stride + (int64_t)(S.row_base + t0) * 8;
        uint8_t* dst = s ? sA1 : sA0;
        const uint32_t plane_bytes = (uint32_t)S.rows * 16;
        for (int pl = 0; pl < S.planes; ++pl)
          bulk_g2s(dst + (size_t)pl * plane_bytes, gx + (int64_t)pl * S.plane_rows * 8, plane_bytes, &bar_a);
      }
      // ---- weight ring
      int it = 0;
      const __nv_bfloat16* gw = G.w;
      for (int s = 0; s < p.nsrc; ++s) {
        const int cp = p.src[s].planes * 8;
        for (int j = 0; j < p.src[s].taps; ++j)
          for (int c0 = 0; c0 < cp; c0 += kb_stage, ++it) {
            const int kb = min(kb_stage, cp - c0);
            const int st = it % nst;
            const uint32_t ph = (it / nst) & 1;
            mbar_wait(&bar_empty[st], ph ^ 1);
            tr(11);
            const uint32_t bytes = (uint32_t)(kb / 8) * ncols * 16;
            mbar_expect_tx(&bar_full[st], bytes);
            bulk_g2s(sW + (size_t)st * w_stage_bytes, gw, bytes, &bar_full[st]);
            gw += (int64_t)kb * ncols;
          }
      }
    }
  } else if (warp == 1) {
    // ---- MMA issuer: the whole warp walks the loops (addresses stay in uniform registers), one lane issues
    const bool leader = elect_one() != 0;
    Tracer tr = make_tracer(p.trace, 1, leader);
    tr(20);
    mbar_wait(&bar_a, 0);          // activation tile(s) landed
    mbar_wait(&bar_a_ready, 0);    // ... LeakyReLU'd in place if asked, and the accumulator holds the bias
    tc_fence_after();
    tr(21);
    const uint32_t idesc = make_idesc_bf16(TC_TILE_M, ncols);
    const uint32_t lbo_b = (uint32_t)ncols * 16;
    const uint64_t b_tmpl = make_desc(0, lbo_b, 128);
    int it = 0;
    for (int s = 0; s < p.nsrc; ++s) {
      const TcSrc& S = p.src[s];
      const int cp = S.planes * 8;
      const uint32_t lbo_a = (uint32_t)S.rows * 16;
      const uint64_t a_tmpl = make_desc(0, lbo_a, 128);
      const uint32_t a_base = smem_u32(s ? sA1 : sA0);
      for (int j = 0; j < S.taps; ++j) {
        const uint32_t a_tap = a_base + (uint32_t)S.tap_row[j] * 16;
        for (int c0 = 0; c0 < cp; c0 += kb_stage, ++it) {
          const int kb = min(kb_stage, cp - c0);
          const int st = it % nst;
          const uint32_t ph = (it / nst) & 1;
          mbar_wait(&bar_full[st], ph);
          tc_fence_after();
          tr(22);
          if (leader) {
            uint32_t a_lo = (a_tap + (uint32_t)(c0 >> 3) * lbo_a) >> 4;
            uint32_t b_lo = smem_u32(sW + (size_t)st * w_stage_bytes) >> 4;
            for (int kk = 0; kk < kb / 16; ++kk) {
              umma_bf16(tbase, a_tmpl | a_lo, b_tmpl | b_lo, idesc, 1);
              a_lo += lbo_a >> 3;   // two planes further, in 16-byte units
              b_lo += lbo_b >> 3;
            }
            umma_commit(&bar_empty[st]);  // frees the stage once these MMAs have read it
          }
          __syncwarp();
        }
      }
    }
    if (leader) umma_commit(&bar_d);
    __syncwarp();
    tr(23);
  } else {
    // ---- warps 2..9: bias -> accumulator, optional LeakyReLU over the staged tile, then the epilogue.
    // Two warps share a TMEM lane quarter and take alternate 32-column batches.
    const int et = tid - 64;
    const int q4 = warp & 3;                 // TMEM lane quarter this warp may touch
    const int hf = (warp - 2) >> 2;          // which of the alternating batches
    const uint32_t trow = tbase + ((uint32_t)(q4 * 32) << 16);
    Tracer tr = make_tracer(p.trace, 2, et == 0);
    tr(30);
    // the accumulator starts from the bias row, so the MMAs always accumulate and the epilogue has no add
    for (int c = hf * 32; c < ncols; c += 64) bias_to_acc(trow, c, min(32, ncols - c), ncols, s_bias);
    tmem_wait_st();
    if (need_lrelu) {
      const Slope2 slope2 = make_slope2(p.slope);
      mbar_wait(&bar_a, 0);
      for (int s = 0; s < p.nsrc; ++s) {
        if (!p.src[s].lrelu) continue;
        uint4* v = reinterpret_cast<uint4*>(s ? sA1 : sA0);
        const int nvec = (s ? a_bytes1 : a_bytes0) / 16;
        for (int e = et; e < nvec; e += EPI_THREADS) {
          uint4 u = v[e];
          u.x = lrelu_bf16x2(u.x, slope2); u.y = lrelu_bf16x2(u.y, slope2);
          u.z = lrelu_bf16x2(u.z, slope2); u.w = lrelu_bf16x2(u.w, slope2);
          v[e] = u;
        }
      }
      fence_proxy_async();
    }
    tc_fence_before();
    mbar_arrive(&bar_a_ready);
    tr(31);
    mbar_wait_relaxed(&bar_d, 0);
    tc_fence_after();
    tr(32);
    conv_epilogue<false>(p, G, trow, q4, hf, lane, t0, b, s_bias);
  }
  if (warp >= 2) {
    Tracer tr = make_tracer(p.trace, 3, tid == 64);
    tr(33);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, tcols);
}

// ---- persistent, weights-resident variant ---------------------------------------------------------------
// For layers whose output is wide and whose K is deep (projection, transposed convs) the kernel above spends
// most of its life streaming the same weights into every CTA and serialising load -> MMA -> epilogue.  Here a
// CTA keeps ONE column group's weights in shared memory for its whole life and walks over (utterance, time
// tile) units: activations stream through a ring of 64-channel chunks, two TMEM accumulators alternate, so the
// loads of unit i+1, the MMAs of unit i and the epilogue of unit i-1 overlap.
constexpr int CP_MAX_SLOTS = 6;
constexpr int CP_CHUNK_PLANES = 8;   // 64 channels per activation chunk
// A chunk is 8 small plane copies; their issue cost (~450 + 90 per copy cycles of the issuing warp, see
// tests/cuda/ingest_probe.cu) exceeds the chunk's MMA time, so two warps issue alternate chunks.
// (Measured alternative that lost: ONE 4-D tensor-map copy per chunk, box {8 channels, rows, 8 planes, 1} — the
// layout's 16-byte inner extent makes the TMA unit fetch 16 bytes per request: convT@6 27.1 -> 29.5 us.)
constexpr int NTP = NT + 64;   // + second activation producer, + the warp that counts finished units (done_out)

__global__ void __launch_bounds__(NTP) tc_convp_kernel(const __grid_constant__ TcConvParams p, const int nslots,
                                                       const int n_tiles, const int batch) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar_w, a_full[CP_MAX_SLOTS], a_empty[CP_MAX_SLOTS], acc_full[2], acc_free[2];
  __shared__ uint32_t units_done[2];   // epilogue warps that have issued their stores, summed over the even / odd units
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float s_bias[256];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool follow = p.follow.done != nullptr;   // the previous kernel is followed segment by segment (follow.cuh)
  const int ngroups = p.ngroups;
  const TcGroup& G = p.grp[blockIdx.x % ngroups];
  const int ncols = G.ncols;
  const TcSrc& S = p.src[0];
  const int planes = S.planes, taps = S.taps;
  const uint32_t lbo_a = (uint32_t)S.rows * 16;
  const uint32_t slot_bytes = CP_CHUNK_PLANES * lbo_a;
  const uint32_t tap_bytes = (uint32_t)planes * ncols * 16;
  uint8_t* sW = smem;                               // [tap][plane][ncols][8]
  uint8_t* sA = smem + (size_t)taps * tap_bytes;    // ring of [<= 8 planes][rows][8] chunks
  const int nchunks = (planes + CP_CHUNK_PLANES - 1) / CP_CHUNK_PLANES;
  const uint32_t acc_stride = ncols <= 32 ? 32 : ncols <= 64 ? 64 : ncols <= 128 ? 128 : 256;
  const int first = blockIdx.x / ngroups, stride = gridDim.x / ngroups;
  const int n_units = n_tiles * batch;

  if (tid == 0) {
    mbar_init(&bar_w, 1);
    for (int s = 0; s < nslots; ++s) {
      mbar_init(&a_full[s], 1);
      mbar_init(&a_empty[s], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_free[i], EPI_THREADS);
    }
    units_done[0] = units_done[1] = 0;
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(&tmem_base_s, 2 * acc_stride);
  for (int c = tid; c < ncols; c += NTP) s_bias[c] = G.bias[c];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_base_s;
  if (follow && !p.follow.early) {   // late form: the check precedes this CTA's trigger (follow.cuh follow_armed)
    if (tid == 0) follow_armed(p.follow);
    __syncthreads();
  }
  pdl_launch_dependents();   // every thread: the CTA counts as triggered once all of its threads have
  if (warp != 0 && !follow) pdl_wait();   // warp 0 first starts the weight load (weights are written by no kernel)

  constexpr int PRODUCER2 = NT / 32;   // the extra warp: odd activation chunks
  constexpr int SIGNAL_WARP = PRODUCER2 + 1;
  if (warp == SIGNAL_WARP) {
    // unit by unit: every epilogue warp has issued its stores -> make them visible device-wide, count the unit for its
    // utterance (the fence stalls until the stores have drained, so no epilogue warp executes it)
    if (p.done_out && elect_one()) {
      int i = 0;
      for (int u = first; u < n_units; u += stride, ++i) {
        const int b = u / n_tiles, tile = u - b * n_tiles;
        smem_count_wait(&units_done[i & 1], (uint32_t)((i >> 1) + 1) * (EPI_THREADS / 32));
        __threadfence();
        atomicAdd(p.done_out + b * FOLLOW_SEGS + tile / p.count_units_per_seg, 1);
      }
    }
  } else if (warp == 0 || warp == PRODUCER2) {
    if (elect_one()) {
      const uint32_t mine = warp == 0 ? 0 : 1;
      if (warp == 0) {
        mbar_expect_tx(&bar_w, (uint32_t)taps * tap_bytes);
        for (int j = 0; j < taps; ++j)
          bulk_g2s(sW + (size_t)j * tap_bytes, G.w + (size_t)j * (tap_bytes / 2), tap_bytes, &bar_w);
        if (!follow) pdl_wait();
      }
      uint32_t gc = 0;
      FollowState fst;
      for (int u = first; u < n_units; u += stride) {
        const int b = u / n_tiles, t0 = (u - b * n_tiles) * TC_TILE_M;
        if (follow) follow_wait(p.follow, b, t0 + p.follow_in_lo, t0 + p.follow_in_lo + S.rows, fst);
        const __nv_bfloat16* gx = S.x + (int64_t)b * S.batch_stride + (int64_t)(S.row_base + t0) * 8;
        for (int c = 0; c < nchunks; ++c, ++gc) {
          if ((gc & 1) != mine) continue;
          const int slot = gc % nslots;
          mbar_wait(&a_empty[slot], ((gc / nslots) & 1) ^ 1);
          const int npl = min(CP_CHUNK_PLANES, planes - c * CP_CHUNK_PLANES);
          mbar_expect_tx(&a_full[slot], (uint32_t)npl * lbo_a);
          for (int pl = 0; pl < npl; ++pl)
            bulk_g2s(sA + (size_t)slot * slot_bytes + (size_t)pl * lbo_a,
                     gx + (int64_t)(c * CP_CHUNK_PLANES + pl) * S.plane_rows * 8, lbo_a, &a_full[slot]);
        }
      }
    }
  } else if (warp == 1) {
    const bool leader = elect_one() != 0;
    const uint32_t idesc = make_idesc_bf16(TC_TILE_M, ncols);
    const uint32_t lbo_b = (uint32_t)ncols * 16;
    const uint64_t a_tmpl = make_desc(0, lbo_a, 128), b_tmpl = make_desc(0, lbo_b, 128);
    const uint32_t w_base = smem_u32(sW), a_base = smem_u32(sA);
    Tracer tr = make_tracer(p.trace, 1, leader);
    tr(19);
    mbar_wait(&bar_w, 0);
    tr(18);
    uint32_t gc = 0;
    int i = 0;
    for (int u = first; u < n_units; u += stride, ++i) {
      const int buf = i & 1;
      mbar_wait(&acc_free[buf], (i >> 1) & 1);   // epilogue has drained this accumulator and reloaded the bias
      tc_fence_after();
      tr(20);
      const uint32_t d = tbase + buf * acc_stride;
      for (int c = 0; c < nchunks; ++c, ++gc) {
        const int slot = gc % nslots;
        mbar_wait(&a_full[slot], (gc / nslots) & 1);
        tc_fence_after();
        tr(22);
        if (leader) {
          const int npl = min(CP_CHUNK_PLANES, planes - c * CP_CHUNK_PLANES);
          for (int j = 0; j < taps; ++j) {
            uint32_t a_lo = (a_base + slot * slot_bytes + (uint32_t)S.tap_row[j] * 16) >> 4;
            uint32_t b_lo = (w_base + j * tap_bytes + (uint32_t)(c * CP_CHUNK_PLANES) * lbo_b) >> 4;
            for (int kk = 0; kk < npl / 2; ++kk) {
              umma_bf16(d, a_tmpl | a_lo, b_tmpl | b_lo, idesc, 1);
              a_lo += lbo_a >> 3;   // two planes further, in 16-byte units
              b_lo += lbo_b >> 3;
            }
          }
          umma_commit(&a_empty[slot]);
        }
        __syncwarp();
      }
      if (leader) umma_commit(&acc_full[buf]);
      __syncwarp();
      tr(21);
    }
  } else {
    Tracer tr = make_tracer(p.trace, warp, lane == 0);
    const int q4 = warp & 3, hf = (warp - 2) >> 2;
    const uint32_t trow = tbase + ((uint32_t)(q4 * 32) << 16);
    for (int buf = 0; buf < 2; ++buf)
      for (int c = hf * 32; c < ncols; c += 64) bias_to_acc(trow + buf * acc_stride, c, min(32, ncols - c), ncols, s_bias);
    tmem_wait_st();
    tc_fence_before();
    mbar_arrive(&acc_free[0]);
    mbar_arrive(&acc_free[1]);
    int i = 0;
    for (int u = first; u < n_units; u += stride, ++i) {
      const int buf = i & 1;
      mbar_wait_relaxed(&acc_full[buf], (i >> 1) & 1);
      tc_fence_after();
      tr(30);
      const int b = u / n_tiles, t0 = (u - b * n_tiles) * TC_TILE_M;
      conv_epilogue<true>(p, G, trow + buf * acc_stride, q4, hf, lane, t0, b, s_bias);
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(&acc_free[buf]);
      if (p.done_out) {
        __syncwarp();
        if (lane == 0) smem_count_add(&units_done[i & 1]);
      }
      tr(31);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 2 * acc_stride);
}

// ---- persistent, activation-resident variant (weights streamed) -------------------------------------------
// The weights-resident kernel above must keep its column groups narrow (64-128 columns: a whole group's weights have
// to fit in shared memory), and an SS MMA with M = 128 fetches 4 KB of A per K-step however narrow N is, so at
// N = 64 the tensor pipe idles behind the shared-memory reads (measured 50 cycles per MMA against a 32-cycle floor;
// convT@3 ran at 22 % of the BF16 peak).  Here the roles are swapped: a CTA keeps one (utterance, time tile) unit's
// ACTIVATION tile resident, walks ALL column groups of the layer over it with N up to 256 per MMA (A and B reads
// balance, the pipe runs at its N/2-cycle floor), and the weights stream from L2 through a ring of >= 16 KB stages
// fed by two producer warps.  Two TMEM accumulators alternate between groups, so the epilogue of group g overlaps
// the MMAs of group g + 1.  The bias is the first MMA of every group (accumulate off): A = a constant [128 x 16]
// operand in shared memory with K0 = K1 = 1, B = the group's bias as hi / lo bf16 K-rows, second K-plane of both
// pointed at one zero plane — so the epilogue neither adds a bias nor re-loads accumulators through tcgen05.st
// (128 B/cycle/SM).  Weight traffic is units x layer weights from L2 (44-132 MB per launch at cfg2).
constexpr int CS_MAX_STAGES = 8;
constexpr int CS_MAX_COLS = 2048;          // all columns of a layer (bias table)
constexpr int NTS = NT + 32;               // + second weight producer

__global__ void __launch_bounds__(NTS) tc_convs_kernel(const __grid_constant__ TcConvParams p, const int kb_stage,
                                                       const int nst, const int na, const int n_tiles, const int batch) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t a_full[2], a_free[2], w_full[CS_MAX_STAGES], w_empty[CS_MAX_STAGES], acc_full[2],
      acc_free[2];
  __shared__ uint32_t tmem_base_s;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ngroups = p.ngroups;
  const int ncols = p.grp[0].ncols;          // all groups of a layer have the same width (pack_tc_layer)
  const TcSrc& S = p.src[0];
  const int planes = S.planes, taps = S.taps, cp = planes * 8;
  const uint32_t lbo_a = (uint32_t)S.rows * 16;
  const uint32_t a_bytes = (uint32_t)planes * lbo_a;
  const uint32_t w_stage_bytes = (uint32_t)(kb_stage / 8) * ncols * 16;
  uint8_t* sA = smem;                                   // [na][planes][rows][8]
  uint8_t* sW = smem + (size_t)na * a_bytes;            // ring of [kb/8 planes][ncols][8] stages
  uint8_t* sOnes = sW + (size_t)nst * w_stage_bytes;    // [128 rows][8]: K0 = K1 = 1 (A of the bias MMA)
  uint8_t* sBias = sOnes + 128 * 16;                    // [group][ncols][8]: K0 / K1 = bias hi / lo (B of the bias MMA)
  uint8_t* sZero = sBias + (size_t)ngroups * ncols * 16;   // max(128, ncols) zero rows: second K-plane of both
  const uint32_t acc_stride = ncols <= 32 ? 32 : ncols <= 64 ? 64 : ncols <= 128 ? 128 : 256;
  const int n_units = n_tiles * batch;
  const int kchunks = (cp + kb_stage - 1) / kb_stage;
  const int stages_per_group = taps * kchunks;

  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&a_full[i], 1);
      mbar_init(&a_free[i], 1);
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_free[i], EPI_THREADS);
    }
    for (int s = 0; s < nst; ++s) {
      mbar_init(&w_full[s], 1);
      mbar_init(&w_empty[s], 1);
    }
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(&tmem_base_s, 2 * acc_stride);
  for (int e = tid; e < 128; e += NTS) *reinterpret_cast<uint4*>(sOnes + e * 16) = make_uint4(0x3F803F80u, 0u, 0u, 0u);
  for (int g = 0; g < ngroups; ++g)
    for (int c = tid; c < ncols; c += NTS) {
      const float bv = p.grp[g].bias[c];
      const __nv_bfloat16 hi = __float2bfloat16(bv), lo = __float2bfloat16(bv - __bfloat162float(hi));
      *reinterpret_cast<uint4*>(sBias + ((size_t)g * ncols + c) * 16) =
          make_uint4((uint32_t)__bfloat16_as_ushort(hi) | ((uint32_t)__bfloat16_as_ushort(lo) << 16), 0u, 0u, 0u);
    }
  for (int e = tid; e < (ncols > 128 ? ncols : 128); e += NTS)
    *reinterpret_cast<uint4*>(sZero + e * 16) = make_uint4(0u, 0u, 0u, 0u);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_base_s;
  pdl_launch_dependents();

  constexpr int PRODUCER2 = NT / 32;   // the extra warp: odd weight stages
  if (warp == 0 || warp == PRODUCER2) {
    if (elect_one()) {
      const uint32_t mine = warp == 0 ? 0 : 1;
      Tracer tr = make_tracer(p.trace, warp == 0 ? 0 : 11, true);
      tr(9);
      // Weights are written by no kernel: the ring starts filling before the previous kernel has drained.
      uint32_t it = 0;
      int i = 0;
      bool waited = false;
      for (int u = blockIdx.x; u < n_units; u += gridDim.x, ++i) {
        if (warp == 0) {   // this unit's activation tile: one bulk copy per 8-channel plane
          if (!waited) {
            pdl_wait();    // the previous kernel's output is complete from here on
            waited = true;
          }
          const int ab = i % na;
          if (i >= na) mbar_wait(&a_free[ab], ((i / na) - 1) & 1);
          const int b = u / n_tiles, t0 = (u - b * n_tiles) * TC_TILE_M;
          const __nv_bfloat16* gx = S.x + (int64_t)b * S.batch_stride + (int64_t)(S.row_base + t0) * 8;
          mbar_expect_tx(&a_full[ab], a_bytes);
          for (int pl = 0; pl < planes; ++pl)
            bulk_g2s(sA + (size_t)ab * a_bytes + (size_t)pl * lbo_a, gx + (int64_t)pl * S.plane_rows * 8, lbo_a,
                     &a_full[ab]);
          tr(10);
        }
        for (int g = 0; g < ngroups; ++g) {
          const __nv_bfloat16* gw = p.grp[g].w;
          for (int j = 0; j < taps; ++j)
            for (int c0 = 0; c0 < cp; c0 += kb_stage, ++it) {
              const int kb = min(kb_stage, cp - c0);
              if ((it & 1) == mine) {
                const int st = it % nst;
                mbar_wait(&w_empty[st], ((it / nst) & 1) ^ 1);
                tr(11);
                const uint32_t bytes = (uint32_t)(kb / 8) * ncols * 16;
                mbar_expect_tx(&w_full[st], bytes);
                bulk_g2s(sW + (size_t)st * w_stage_bytes, gw, bytes, &w_full[st]);
                tr(12);
              }
              gw += (int64_t)kb * ncols;
            }
        }
      }
    }
  } else if (warp == 1) {
    pdl_wait();
    const bool leader = elect_one() != 0;
    const uint32_t idesc = make_idesc_bf16(TC_TILE_M, ncols);
    const uint32_t lbo_b = (uint32_t)ncols * 16;
    const uint64_t a_tmpl = make_desc(0, lbo_a, 128), b_tmpl = make_desc(0, lbo_b, 128);
    const uint32_t w_base = smem_u32(sW), a_base0 = smem_u32(sA);
    const uint64_t ones_desc = make_desc(smem_u32(sOnes), smem_u32(sZero) - smem_u32(sOnes), 128);
    uint32_t it = 0, gg = 0;
    int i = 0;
    Tracer tr = make_tracer(p.trace, 1, leader);
    tr(19);
    for (int u = blockIdx.x; u < n_units; u += gridDim.x, ++i) {
      const int ab = i % na;
      mbar_wait(&a_full[ab], (i / na) & 1);
      tc_fence_after();
      tr(18);
      const uint32_t a_base = a_base0 + (uint32_t)ab * a_bytes;
      for (int g = 0; g < ngroups; ++g, ++gg) {
        const int buf = gg & 1;
        mbar_wait(&acc_free[buf], (gg >> 1) & 1);   // drained by the epilogue
        tc_fence_after();
        tr(20);
        const uint32_t d = tbase + buf * acc_stride;
        if (leader) {   // D <- bias (overwrites)
          const uint32_t ba = smem_u32(sBias + (size_t)g * ncols * 16);
          umma_bf16(d, ones_desc, make_desc(ba, smem_u32(sZero) - ba, 128), idesc, 0);
        }
        __syncwarp();
        for (int j = 0; j < taps; ++j) {
          const uint32_t a_tap = a_base + (uint32_t)S.tap_row[j] * 16;
          for (int c0 = 0; c0 < cp; c0 += kb_stage, ++it) {
            const int kb = min(kb_stage, cp - c0);
            const int st = it % nst;
            mbar_wait(&w_full[st], (it / nst) & 1);
            tc_fence_after();
            tr(22);
            if (leader) {
              uint32_t a_lo = (a_tap + (uint32_t)(c0 >> 3) * lbo_a) >> 4;
              uint32_t b_lo = (w_base + (uint32_t)st * w_stage_bytes) >> 4;
              for (int kk = 0; kk < kb / 16; ++kk) {
                umma_bf16(d, a_tmpl | a_lo, b_tmpl | b_lo, idesc, 1);
                a_lo += lbo_a >> 3;   // two planes further, in 16-byte units
                b_lo += lbo_b >> 3;
              }
              umma_commit(&w_empty[st]);
            }
            __syncwarp();
          }
        }
        if (leader) {
          umma_commit(&acc_full[buf]);
          if (g == ngroups - 1) umma_commit(&a_free[ab]);   // the tile is not read again
        }
        __syncwarp();
        tr(21);
      }
    }
    (void)stages_per_group;
  } else {
    pdl_wait();
    const int q4 = warp & 3, hf = (warp - 2) >> 2;
    const uint32_t trow = tbase + ((uint32_t)(q4 * 32) << 16);
    const int reach = p.margin == 1 ? p.out_halo : 0;
    mbar_arrive(&acc_free[0]);
    mbar_arrive(&acc_free[1]);
    uint32_t gg = 0;
    Tracer tr = make_tracer(p.trace, warp, lane == 0);
    tr(29);
    for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
      const int b = u / n_tiles, t0 = (u - b * n_tiles) * TC_TILE_M;
      for (int g = 0; g < ngroups; ++g, ++gg) {
        const int buf = gg & 1;
        mbar_wait_relaxed(&acc_full[buf], (gg >> 1) & 1);
        tc_fence_after();
        tr(30);
        // interior tile (warp-uniform): every row valid, no output row within reach of a sequence end
        const bool interior = p.out_mode == 0 && t0 + TC_TILE_M <= p.Lt && t0 * p.up_f - p.up_P > reach &&
                              (t0 + TC_TILE_M - 1) * p.up_f - p.up_P + p.up_f - 1 < p.out_L - 1 - reach;
        if (interior) {
          if (p.out_lrelu) conv_epilogue_fast<true>(p, p.grp[g], trow + buf * acc_stride, q4, hf, lane, t0, b, p.slope);
          else conv_epilogue_fast<false>(p, p.grp[g], trow + buf * acc_stride, q4, hf, lane, t0, b, p.slope);
        } else {
          conv_epilogue<false>(p, p.grp[g], trow + buf * acc_stride, q4, hf, lane, t0, b, nullptr);
        }
        tc_fence_before();
        mbar_arrive(&acc_free[buf]);
        tr(31);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 2 * acc_stride);
}

// ---- layout helpers --------------------------------------------------------------------------------
__global__ void prep_mel_kernel(const float* __restrict__ mel, int C, int T, int pad, __nv_bfloat16* __restrict__ out,
                                int planes, int plane_rows, int halo, int* __restrict__ counters, int n_counters,
                                int armed_id, int early) {
  // one thread per (plane row, plane); row index covers halo margins too (reflect of the padded signal)
  //
  // Hand-over counters (follow.cuh): block 0 zeroes them once the previous inference has completed; the armed flag
  // behind them keeps the kernels that follow (resident early: dependent launches cascade) from polling before that.
  const bool zeroes = blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0;
  if (zeroes && !early) {   // late form: armed = 0 before ANY dependent can launch (follow.cuh follow_armed)
    if (threadIdx.x == 0)
      asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(counters + n_counters), "r"(0) : "memory");
    __syncthreads();
  }
  pdl_launch_dependents();
  pdl_wait();   // the previous inference on this stream has completed
  if (zeroes) {
    for (int i = threadIdx.x; i < n_counters; i += blockDim.x) counters[i] = 0;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0)
      asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(counters + n_counters), "r"(armed_id) : "memory");
  }
  const int L = T + 2 * pad;
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  const int pl = blockIdx.y, b = blockIdx.z;
  if (row >= plane_rows) return;
  int t = row - halo;                       // position in the replicate-padded signal
  bool ok = t > -L && t < 2 * L - 1;        // mirror only defined within one reflection
  if (t < 0) t = -t;
  if (t >= L) t = 2 * (L - 1) - t;
  int src = t - pad;
  src = src < 0 ? 0 : (src >= T ? T - 1 : src);
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float a = 0.f, c = 0.f;
    const int c0 = pl * 8 + 2 * i;
    if (ok && c0 < C) a = mel[((int64_t)b * C + c0) * T + src];
    if (ok && c0 + 1 < C) c = mel[((int64_t)b * C + c0 + 1) * T + src];
    w[i] = pack_bf16x2(a, c);
  }
  *reinterpret_cast<uint4*>(out + (((int64_t)b * planes + pl) * plane_rows + row) * 8) = make_uint4(w[0], w[1], w[2], w[3]);
}

__global__ void untile_kernel(const __nv_bfloat16* __restrict__ x, int64_t batch_stride, int C, int64_t L, int plane_rows,
                              int halo, int undo_lrelu, float slope, float* __restrict__ out) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int c = blockIdx.y, b = blockIdx.z;
  if (t >= L) return;
  float v = __bfloat162float(x[(int64_t)b * batch_stride + ((int64_t)(c >> 3) * plane_rows + halo + t) * 8 + (c & 7)]);
  if (undo_lrelu && v < 0.f) v /= slope;
  out[((int64_t)b * C + c) * L + t] = v;
}

}  // namespace

namespace {
constexpr size_t SMEM_MAX = 227 * 1024 - 3 * 1024;     // static barriers + bias table live in the same 227 KB
constexpr size_t SMEM_HALF = 113 * 1024 - 3 * 1024;    // budget that lets two CTAs share an SM
struct SmemPlan { size_t a_bytes, total; int kb_stage, nst; };
// Weight ring geometry: the deepest / widest ring that still lets two CTAs share an SM (their loads, MMAs and
// epilogues then overlap); if the activation tile alone is too big for that, the best ring that fits at all.
SmemPlan smem_plan(const TcConvParams& p) {
  SmemPlan s{};
  s.a_bytes = (size_t)p.src[0].planes * p.src[0].rows * 16;
  if (p.nsrc > 1) s.a_bytes += (size_t)p.src[1].planes * p.src[1].rows * 16;
  int nmax = 0;
  for (int g = 0; g < p.ngroups; ++g) nmax = nmax > p.grp[g].ncols ? nmax : p.grp[g].ncols;
  // {channels per stage, stages}.  Sharing the SM is only worth it with a ring deep enough to cover the L2
  // latency (>= 3 stages of >= 32 channels); otherwise one CTA per SM with the deepest ring that fits.
  static const int half_opts[][2] = {{64, 4}, {64, 3}, {32, 4}, {32, 3}};
  static const int full_opts[][2] = {{64, 4}, {64, 3}, {32, 4}, {32, 3}, {32, 2}, {16, 3}, {16, 2}};
  for (auto& o : half_opts) {
    const size_t tot = s.a_bytes + (size_t)o[1] * (o[0] / 8) * nmax * 16;
    if (tot <= SMEM_HALF) {
      s.kb_stage = o[0]; s.nst = o[1]; s.total = tot;
      return s;
    }
  }
  for (auto& o : full_opts) {
    const size_t tot = s.a_bytes + (size_t)o[1] * (o[0] / 8) * nmax * 16;
    if (tot <= SMEM_MAX) {
      s.kb_stage = o[0]; s.nst = o[1]; s.total = tot;
      return s;
    }
  }
  s.total = (size_t)-1;
  return s;
}
}  // namespace

// bytes of dynamic shared memory of the launch
size_t tc_conv_smem_bytes(const TcConvParams& p) { return smem_plan(p).total; }

namespace {
// Ring depth of the persistent variant for this launch; 0 = not eligible (two sources, in-kernel LeakyReLU,
// fp32 output, or weights that do not fit beside a useful ring).
int convp_slots(const TcConvParams& p, size_t* smem_bytes) {
  if (p.nsrc != 1 || p.src[0].lrelu || p.out_mode != 0 || (p.src[0].planes & 1)) return 0;
  size_t w = 0;
  for (int g = 0; g < p.ngroups; ++g) {
    if (p.grp[g].ncols > 256) return 0;
    const size_t wg = (size_t)p.src[0].taps * p.src[0].planes * p.grp[g].ncols * 16;
    w = w > wg ? w : wg;
  }
  const size_t slot = (size_t)CP_CHUNK_PLANES * p.src[0].rows * 16;
  if (w + 2 * slot > SMEM_MAX) return 0;
  size_t n = (SMEM_MAX - w) / slot;
  const size_t nchunks = (p.src[0].planes + CP_CHUNK_PLANES - 1) / CP_CHUNK_PLANES;
  if (n > CP_MAX_SLOTS) n = CP_MAX_SLOTS;
  if (n > 2 * nchunks) n = 2 * nchunks;   // two whole tiles in flight is as deep as the pipeline gets
  if (n < 2) n = 2;
  *smem_bytes = w + n * slot;
  return (int)n;
}
}  // namespace

TcLaunchCtx tc_launch_ctx(int device) {
  TcLaunchCtx c;
  c.device = device;
  cudaDeviceGetAttribute(&c.sms, cudaDevAttrMultiProcessorCount, device);
  auto env_int = [](const char* name, int dflt) {
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
  };
  c.stack_max_grid = env_int("TTSB200_STACK_MAX_GRID", 0);
  c.stack_debug = env_int("TTSB200_STACK_DEBUG", 0);
  c.conv_debug = env_int("TTSB200_CONV_DEBUG", 0);
  c.conv_persistent = env_int("TTSB200_CONV_PERSISTENT", 2);
  c.pqmf_direct = getenv("TTSB200_PQMF_DIRECT") ? 1 : 0;
  c.no_k24 = getenv("TTSB200_NO_K24") ? 1 : 0;
  c.no_follow = getenv("TTSB200_NO_FOLLOW") ? 1 : 0;
  c.trace_file = getenv("TTSB200_STACK_TRACE");
  return c;
}

namespace {
// Geometry of the activation-resident kernel for this launch; nst == 0 = not eligible (two sources, in-kernel
// LeakyReLU, fp32 output, uneven groups, or a tile that leaves no room for a useful weight ring).
struct ConvsPlan { int kb, nst, na; size_t smem; };
constexpr size_t CS_SMEM_MAX = 227 * 1024 - 1024;   // static: barriers
ConvsPlan convs_plan(const TcConvParams& p) {
  ConvsPlan r{0, 0, 0, 0};
  if (p.nsrc != 1 || p.src[0].lrelu || p.out_mode != 0 || (p.src[0].planes & 1)) return r;
  const int ncols = p.grp[0].ncols;
  if (ncols > 256 || (size_t)ncols * p.ngroups > CS_MAX_COLS) return r;
  for (int g = 1; g < p.ngroups; ++g)
    if (p.grp[g].ncols != ncols) return r;
  const size_t a = (size_t)p.src[0].planes * p.src[0].rows * 16;
  // bias-MMA operands: ones plane, every group's bias plane, one zero plane
  const size_t extra = 128 * 16 + (size_t)p.ngroups * ncols * 16 + (size_t)(ncols > 128 ? ncols : 128) * 16;
  // widest stage first: a bulk copy costs its issuing warp ~500 cycles whatever its size (ingest_probe.cu), so the
  // ring wants few large stages; at least three of them, and a second activation buffer when it still fits
  for (int kb : {64, 32, 16}) {
    const size_t stage = (size_t)(kb / 8) * ncols * 16;
    if (a + extra + 3 * stage > CS_SMEM_MAX) continue;
    r.kb = kb;
    r.na = (2 * a + extra + 4 * stage <= CS_SMEM_MAX) ? 2 : 1;
    size_t n = (CS_SMEM_MAX - extra - r.na * a) / stage;
    r.nst = (int)(n > CS_MAX_STAGES ? CS_MAX_STAGES : n);
    r.smem = r.na * a + r.nst * stage + extra;
    return r;
  }
  return r;
}
}  // namespace

bool tc_conv_count_info(const TcConvParams& p, int batch, const TcLaunchCtx& ctx, int* n_units, int* unit_rows,
                        int* row_offset, int* mult) {
  size_t smem = 0;
  if (ctx.conv_persistent != 2 || convp_slots(p, &smem) <= 0 || p.ngroups > ctx.sms || batch <= 0) return false;
  *n_units = ceil_div(p.Lt, TC_TILE_M);   // time tiles per utterance, each finished once per column group
  *unit_rows = TC_TILE_M * p.up_f;        // GEMM row q -> output rows q f - P .. q f - P + f - 1
  *row_offset = p.up_P;
  *mult = p.ngroups;
  return true;
}

cudaError_t tc_conv_launch(const TcConvParams& p_in, int batch, const TcLaunchCtx& ctx, cudaStream_t stream) {
  TcConvParams p = p_in;
  // cudaFuncSetAttribute is per device and costs a driver call: once per device, not per launch
  static std::atomic<uint64_t> attr_done{0};
  const uint64_t bit = 1ull << (ctx.device & 63);
  if (!(attr_done.load(std::memory_order_acquire) & bit)) {
    cudaError_t e = cudaFuncSetAttribute(tc_convp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_MAX);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(tc_convs_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CS_SMEM_MAX);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(tc_conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_MAX);
    if (e != cudaSuccess) return e;
    attr_done.fetch_or(bit, std::memory_order_release);
  }
  p.debug = ctx.conv_debug;
  if (ctx.conv_persistent == 1) {
    const ConvsPlan cs = convs_plan(p);
    if (cs.nst >= 3) {
      const int n_tiles = ceil_div(p.Lt, TC_TILE_M);
      const int units = n_tiles * batch;
      p.trace = trace_begin(ctx.trace_file, stream);
      const cudaError_t rc = launch_pdl(tc_convs_kernel, dim3(units < ctx.sms ? units : ctx.sms), dim3(NTS), cs.smem,
                                        stream, p, cs.kb, cs.nst, cs.na, n_tiles, batch);
      trace_end(ctx.trace_file, p.trace, 1000 + p.trace_id, stream);
      return rc;
    }
  }
  {
    size_t smem = 0;
    const int sms = ctx.sms;
    const int nslots = ctx.conv_persistent == 2 ? convp_slots(p, &smem) : 0;
    if (nslots > 0 && p.ngroups <= sms) {
      const int n_tiles = ceil_div(p.Lt, TC_TILE_M);
      int per_group = sms / p.ngroups;
      if (per_group > n_tiles * batch) per_group = n_tiles * batch;
      p.trace = trace_begin(ctx.trace_file, stream);
      const cudaError_t rc =
          launch_pdl(tc_convp_kernel, dim3(per_group * p.ngroups), dim3(NTP), smem, stream, p, nslots, n_tiles, batch);
      trace_end(ctx.trace_file, p.trace, 1000 + p.trace_id, stream);
      return rc;
    }
  }
  const SmemPlan sp = smem_plan(p);
  if (sp.total > SMEM_MAX) return cudaErrorInvalidConfiguration;
  const dim3 grid(ceil_div(p.Lt, TC_TILE_M), p.ngroups, batch);
  p.trace = trace_begin(ctx.trace_file, stream);
  const cudaError_t rc = launch_pdl(tc_conv_kernel, grid, dim3(NT), sp.total, stream, p, sp.kb_stage, sp.nst);
  trace_end(ctx.trace_file, p.trace, 1000 + p.trace_id, stream);
  return rc;
}

cudaError_t tc_prep_mel_launch(const float* mel, int batch, int C, int T, int pad, __nv_bfloat16* out, int planes,
                               int plane_rows, int halo, int* counters, int n_counters, int armed_id, int early,
                               cudaStream_t stream) {
  const dim3 grid(ceil_div(plane_rows, 128), planes, batch);
  // Launched behind the previous kernel of the stream like every other kernel of the chain: its CTAs are resident
  // when that kernel (the last one of the previous inference) drains, and griddepcontrol.wait inside holds them
  // until it has completed — which is also when its counters may be zeroed.
  return launch_pdl(prep_mel_kernel, grid, dim3(128), 0, stream, mel, C, T, pad, out, planes, plane_rows, halo, counters,
                    n_counters, armed_id, early);
}

cudaError_t tc_untile_launch(const __nv_bfloat16* x, int64_t batch_stride, int batch, int C, int64_t L, int plane_rows,
                             int halo, int undo_lrelu, float slope, float* out, cudaStream_t stream) {
  const dim3 grid(ceil_div(L, 256), C, batch);
  untile_kernel<<<grid, 256, 0, stream>>>(x, batch_stride, C, L, plane_rows, halo, undo_lrelu, slope, out);
  return cudaGetLastError();
}

}  // namespace ttsb
