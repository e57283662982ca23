// Fused ResidualStack on tcgen05 / TMEM (TTS/vocoder/layers/melgan.py:13-38): all n blocks
//     x <- S_b x + W2_b lrelu( W1_b (*) lrelu(x)[reflect, dilation 3^b] + b1_b ) + (b2_b + bs_b)
// of one upsampling stage run in ONE kernel on a time tile of kt*128 rows.  The stage input is read
// from HBM once, every intermediate (x, lrelu(x), h) lives in shared memory in the canonical UMMA
// layout, accumulators live in TMEM, and only the stack output goes back to HBM.  Per tile the
// valid region shrinks by the dilation reach of each block (halo recompute, sum_d rows per side).
//
// Shared-memory tensors ([cp/8 planes][RS rows][8 ch] bf16, RS = kt*128 + 2*guard):
//   XR  raw x           — A operand of the 1x1 shortcut, overwritten in place by each block's output
//   XA  lrelu(x)        — A operand of the dilated conv (taps = start-row offsets); after that
//                          conv has been issued for the whole tile the same buffer holds h = lrelu(.)
//                          (A operand of the 1x1 conv) and is then refilled with lrelu(x_new)
//   W1 / W2S            — weights of the current block; the next block's are fetched by the bulk-copy
//                          engine as soon as the MMAs that read the buffer have retired
// TMEM: D1[t] (dilated conv) and D2[t] (1x1 + shortcut) for each of the kt row tiles.
//
// Sequence edges: ReflectionPad1d precedes every dilated conv, so before each block the rows just
// outside [0, L) are rewritten with the mirror of the rows just inside (only in CTAs that see an edge).
//
// Warps: 0 = bulk-copy producer + TMEM owner, 1 = MMA issuer, 2..9 = prologue/epilogue (two warps per
// TMEM lane quarter, splitting the 16-column chunks).
#include "common.cuh"
#include "tc_common.cuh"
#include "tc_conv.cuh"
#include "tc_store.cuh"

namespace ttsb {

using namespace tc;

namespace {

constexpr int NT = 320;
constexpr int EPI_THREADS = 256;

__device__ __forceinline__ float lrelu_f(float v, float s) { return v >= 0.f ? v : v * s; }
__device__ __forceinline__ uint32_t lrelu_bf16x2(uint32_t u, float slope) {
  return pack_bf16x2(lrelu_f(__uint_as_float(u << 16), slope), lrelu_f(__uint_as_float(u & 0xFFFF0000u), slope));
}
__device__ __forceinline__ void epi_barrier() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

struct Bars {
  uint64_t x, xready, w1, w2, w1_free, w2_free;
  uint64_t d1[TC_STACK_MAX_KT], h[TC_STACK_MAX_KT], d2[TC_STACK_MAX_KT];
};

__global__ void __launch_bounds__(NT) tc_stack_kernel(const __grid_constant__ TcStackParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) Bars bars;
  __shared__ uint32_t tmem_base_s;
  __shared__ float s_bias[TC_STACK_MAX_BLOCKS][2][96];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int cp = p.cp, kt = p.kt, planes = cp >> 3;
  const int TR = kt * 128, RS = TR + 2 * p.guard, G = p.guard;
  const int V = TR - 2 * p.sum_d;                  // rows of final output this CTA owns
  const int g0 = blockIdx.x * V - p.sum_d;         // sequence row of staged row 0
  const int b = blockIdx.y;
  const int half = (p.res_taps - 1) / 2;

  const uint32_t plane_bytes = (uint32_t)RS * 16;
  const uint32_t act_bytes = (uint32_t)planes * plane_bytes;
  uint8_t* sXR = smem;
  uint8_t* sXA = smem + act_bytes;
  uint8_t* sW1 = smem + 2 * act_bytes;
  const uint32_t w1_bytes = (uint32_t)p.res_taps * cp * cp * 2;
  const uint32_t w2_bytes = 2u * cp * cp * 2;
  uint8_t* sW2 = sW1 + w1_bytes;

  uint32_t tcols = 32;
  while (tcols < (uint32_t)(2 * kt * cp)) tcols <<= 1;

  if (tid == 0) {
    mbar_init(&bars.x, 1);
    mbar_init(&bars.xready, EPI_THREADS);
    mbar_init(&bars.w1, 1);
    mbar_init(&bars.w2, 1);
    mbar_init(&bars.w1_free, 1);
    mbar_init(&bars.w2_free, 1);
    for (int t = 0; t < kt; ++t) {
      mbar_init(&bars.d1[t], 1);
      mbar_init(&bars.h[t], EPI_THREADS);
      mbar_init(&bars.d2[t], 1);
    }
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(&tmem_base_s, tcols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_base_s;

  if (warp == 0) {
    if (elect_one()) {
      // stage input tile: rows g0 .. g0+TR of every plane, into XR at staged row G
      mbar_expect_tx(&bars.x, (uint32_t)planes * TR * 16);
      const __nv_bfloat16* gx = p.x + (int64_t)b * p.x_batch_stride + (int64_t)(p.x_halo + g0) * 8;
      for (int pl = 0; pl < planes; ++pl)
        bulk_g2s(sXR + (size_t)pl * plane_bytes + (size_t)G * 16, gx + (int64_t)pl * p.x_plane_rows * 8,
                 (uint32_t)TR * 16, &bars.x);
      for (int blk = 0; blk < p.nblocks; ++blk) {
        if (blk > 0) mbar_wait(&bars.w1_free, (blk - 1) & 1);
        mbar_expect_tx(&bars.w1, w1_bytes);
        bulk_g2s(sW1, p.w1[blk], w1_bytes, &bars.w1);
        if (blk > 0) mbar_wait(&bars.w2_free, (blk - 1) & 1);
        mbar_expect_tx(&bars.w2, w2_bytes);
        bulk_g2s(sW2, p.w2s[blk], w2_bytes, &bars.w2);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      const uint32_t idesc = make_idesc_bf16(128, cp);
      const uint32_t lbo_a = plane_bytes, lbo_b = (uint32_t)cp * 16;
      const uint32_t xa = smem_u32(sXA), xr = smem_u32(sXR), w1 = smem_u32(sW1), w2 = smem_u32(sW2);
      const int ksteps = cp >> 4;
      for (int blk = 0; blk < p.nblocks; ++blk) {
        const uint32_t ph = blk & 1;
        const int d = p.dil[blk];
        mbar_wait(&bars.xready, ph);
        mbar_wait(&bars.w1, ph);
        tc_fence_after();
        for (int t = 0; t < kt; ++t) {                 // dilated conv: taps are start-row offsets into XA
          uint32_t acc = 0;
          for (int j = 0; j < p.res_taps; ++j) {
            const uint32_t a_row = xa + (uint32_t)(G + t * 128 + (j - half) * d) * 16;
            const uint32_t w_tap = w1 + (uint32_t)j * cp * cp * 2;
            for (int kk = 0; kk < ksteps; ++kk) {
              umma_bf16(tbase + t * cp, make_desc(a_row + 2 * kk * lbo_a, lbo_a, 128),
                        make_desc(w_tap + 2 * kk * lbo_b, lbo_b, 128), idesc, acc);
              acc = 1;
            }
          }
          umma_commit(&bars.d1[t]);
        }
        umma_commit(&bars.w1_free);
        mbar_wait(&bars.w2, ph);
        for (int t = 0; t < kt; ++t) {                 // 1x1 on h (in the XA buffer) + 1x1 shortcut on raw x
          mbar_wait(&bars.h[t], ph);
          tc_fence_after();
          const uint32_t row = (uint32_t)(G + t * 128) * 16;
          uint32_t acc = 0;
          for (int s = 0; s < 2; ++s) {
            const uint32_t a_row = (s ? xr : xa) + row;
            const uint32_t w_s = w2 + (uint32_t)s * cp * cp * 2;
            for (int kk = 0; kk < ksteps; ++kk) {
              umma_bf16(tbase + (kt + t) * cp, make_desc(a_row + 2 * kk * lbo_a, lbo_a, 128),
                        make_desc(w_s + 2 * kk * lbo_b, lbo_b, 128), idesc, acc);
              acc = 1;
            }
          }
          umma_commit(&bars.d2[t]);
        }
        umma_commit(&bars.w2_free);
      }
    }
  } else {
    // ---------------- warps 2..9: prologue + epilogues ----------------
    const int et = tid - 64;                    // 0..255
    const int q4 = warp & 3;                    // TMEM lane quarter
    const int hf = (warp - 2) >> 2;             // which half of the 16-column chunks
    const int m = q4 * 32 + lane;               // row within a 128-row tile
    const uint32_t tlane = tbase + ((uint32_t)(q4 * 32) << 16);
    const int nchunks = cp >> 4;

    for (int e = et; e < p.nblocks * 2 * cp; e += EPI_THREADS) {
      const int blk = e / (2 * cp), r = e - blk * 2 * cp;
      s_bias[blk][r / cp][r % cp] = r < cp ? p.b1[blk][r] : p.b2s[blk][r - cp];
    }

    // rows just outside the sequence <- mirror of the rows just inside (ReflectionPad1d), for dilation d
    auto mirror = [&](int d) {
      const bool left = g0 < 0, right = g0 + TR > p.L;
      if (!left && !right) return;
      const int per_side = half * d * planes * 2;
      for (int e = et; e < 2 * per_side; e += EPI_THREADS) {
        const int side = e / per_side, r = e - side * per_side;
        if (side == 0 ? !left : !right) continue;
        const int buf = r & 1, q = r >> 1, pl = q % planes, i = q / planes + 1;
        const int dst = (side == 0 ? -i : p.L - 1 + i) - g0, src = (side == 0 ? i : p.L - 1 - i) - g0;
        if (dst < -G || dst >= TR + G || src < 0 || src >= TR) continue;
        uint8_t* base = (buf ? sXA : sXR) + (size_t)pl * plane_bytes;
        *reinterpret_cast<uint4*>(base + (size_t)(G + dst) * 16) = *reinterpret_cast<const uint4*>(base + (size_t)(G + src) * 16);
      }
    };

    // prologue: XA = lrelu(XR) over the staged rows
    mbar_wait(&bars.x, 0);
    for (int e = et; e < planes * TR; e += EPI_THREADS) {
      const int pl = e / TR, r = e - pl * TR;
      const size_t off = (size_t)pl * plane_bytes + (size_t)(G + r) * 16;
      uint4 u = *reinterpret_cast<const uint4*>(sXR + off);
      u.x = lrelu_bf16x2(u.x, p.slope); u.y = lrelu_bf16x2(u.y, p.slope);
      u.z = lrelu_bf16x2(u.z, p.slope); u.w = lrelu_bf16x2(u.w, p.slope);
      *reinterpret_cast<uint4*>(sXA + off) = u;
    }
    epi_barrier();
    mirror(p.dil[0]);
    fence_proxy_async();
    mbar_arrive(&bars.xready);

    for (int blk = 0; blk < p.nblocks; ++blk) {
      const uint32_t ph = blk & 1;
      const bool last = blk == p.nblocks - 1;
      // ---- epilogue 1: h = lrelu(D1 + b1) -> XA buffer (all dilated-conv MMAs of the tile have retired)
      mbar_wait(&bars.d1[kt - 1], ph);
      tc_fence_after();
      for (int t = 0; t < kt; ++t) {
        for (int c = hf; c < nchunks; c += 2) {
          float v[16];
          tmem_ld16(tlane + t * cp + c * 16, v);
          uint32_t w[8];
#pragma unroll
          for (int i = 0; i < 8; ++i)
            w[i] = pack_bf16x2(lrelu_f(v[2 * i] + s_bias[blk][0][c * 16 + 2 * i], p.slope),
                               lrelu_f(v[2 * i + 1] + s_bias[blk][0][c * 16 + 2 * i + 1], p.slope));
          uint8_t* dst = sXA + (size_t)(2 * c) * plane_bytes + (size_t)(G + t * 128 + m) * 16;
          *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
          *reinterpret_cast<uint4*>(dst + plane_bytes) = make_uint4(w[4], w[5], w[6], w[7]);
        }
        fence_proxy_async();
        mbar_arrive(&bars.h[t]);
      }
      // ---- epilogue 2: x_new = D2 + (b2 + bs)
      for (int t = 0; t < kt; ++t) {
        mbar_wait(&bars.d2[t], ph);
        tc_fence_after();
        const int tau = g0 + t * 128 + m;       // sequence row of this thread
        const bool own = tau >= (int)blockIdx.x * V && tau < ((int)blockIdx.x + 1) * V && tau < p.L;
        for (int c = hf; c < nchunks; c += 2) {
          float v[16];
          tmem_ld16(tlane + (kt + t) * cp + c * 16, v);
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] += s_bias[blk][1][c * 16 + i];
          if (!last) {
            uint32_t w[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) w[i] = pack_bf16x2(v[2 * i], v[2 * i + 1]);
            const size_t off = (size_t)(2 * c) * plane_bytes + (size_t)(G + t * 128 + m) * 16;
            *reinterpret_cast<uint4*>(sXR + off) = make_uint4(w[0], w[1], w[2], w[3]);
            *reinterpret_cast<uint4*>(sXR + off + plane_bytes) = make_uint4(w[4], w[5], w[6], w[7]);
#pragma unroll
            for (int i = 0; i < 8; ++i) w[i] = lrelu_bf16x2(w[i], p.slope);
            *reinterpret_cast<uint4*>(sXA + off) = make_uint4(w[0], w[1], w[2], w[3]);
            *reinterpret_cast<uint4*>(sXA + off + plane_bytes) = make_uint4(w[4], w[5], w[6], w[7]);
          } else if (own) {
            canon_store16(p.out, b, c * 16, tau, v, p.slope);
          }
        }
      }
      if (!last) {
        epi_barrier();
        mirror(p.dil[blk + 1]);
        fence_proxy_async();
        mbar_arrive(&bars.xready);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, tcols);
}

constexpr size_t SMEM_MAX = 227 * 1024 - 8 * 1024;   // static: barriers + bias table

}  // namespace

size_t tc_stack_smem_bytes(const TcStackParams& p) {
  const size_t RS = (size_t)p.kt * 128 + 2 * p.guard;
  return 2 * (size_t)(p.cp / 8) * RS * 16 + (size_t)(p.res_taps + 2) * p.cp * p.cp * 2;
}

cudaError_t tc_stack_launch(const TcStackParams& p, int batch, cudaStream_t stream) {
  const size_t smem = tc_stack_smem_bytes(p);
  if (smem > SMEM_MAX || p.cp > 96 || p.cp % 16 || p.kt < 1 || p.kt > TC_STACK_MAX_KT ||
      p.nblocks < 1 || p.nblocks > TC_STACK_MAX_BLOCKS || 2 * p.kt * p.cp > 512)
    return cudaErrorInvalidConfiguration;
  cudaError_t e = cudaFuncSetAttribute(tc_stack_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_MAX);
  if (e != cudaSuccess) return e;
  const int V = p.kt * 128 - 2 * p.sum_d;
  const dim3 grid(ceil_div(p.L, V), batch);
  tc_stack_kernel<<<grid, NT, smem, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace ttsb
