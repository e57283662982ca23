// fp32 implicit-GEMM 1-D convolution for the exact (max-abs <= 1e-5) vocoder path.
//
// One kernel serves every parametrised layer of the reference generator
// (TTS/vocoder/models/melgan_generator.py:29-78, TTS/vocoder/layers/melgan.py:13-38):
//   * k7 projection / head:  reflect pad fused into the index map, replicate inference padding
//     folded in as a clamp, LeakyReLU prologue and tanh epilogue optional;
//   * ConvTranspose1d(k=2f, stride f): f phases, each a 2-tap conv over the input with
//     zero-fill edges, written with stride f;
//   * residual block: dilated k3 with LeakyReLU prologue, then ONE launch for
//     conv1x1(lrelu(h)) + shortcut1x1(x) + both biases (two GEMM sources, one accumulator).
//
// GEMM view: rows = output channels (tile 16*CPT), columns = time (tile 128), K = taps * C_in
// walked in slices of 16 channels.  256 threads, each 8 time positions x CPT channels in
// registers; operands staged through double-buffered shared memory with register prefetch.
// fp32 FFMA only — TF32 cannot hold 1e-5 through ~30 stacked convs.
#include "common.cuh"

namespace ttsb {

namespace {

constexpr int TN = 128;  // time positions per CTA
constexpr int BK = 16;   // input channels per K slice
constexpr int NT = 256;  // threads per CTA

// 0 <= slope <= 1 (validated at plan creation), so lrelu(v) = max(v, slope v): no predicate chain
__device__ __forceinline__ float lrelu_f(float v, float slope) { return fmaxf(v, v * slope); }

// Resolve a virtual position to a stored column; returns -1 for "contributes zero".
__device__ __forceinline__ int resolve_pos(const ConvSrc& s, int pos) {
  if (s.edge_zero) return (pos >= 0 && pos < s.L) ? pos : -1;
  if (pos < 0) pos = -pos;                       // ReflectionPad1d: -i -> i
  if (pos >= s.L) pos = 2 * (s.L - 1) - pos;     //                  L-1+i -> L-1-i
  if (s.clamp_pad) {                             // F.pad(..., 'replicate') folded in
    pos -= s.clamp_pad;
    pos = pos < 0 ? 0 : (pos >= s.Lreal ? s.Lreal - 1 : pos);
  }
  return pos;
}

template <int CPT>
__global__ void __launch_bounds__(NT, CPT >= 6 ? 2 : (CPT >= 3 ? 3 : 4)) conv_f32_kernel(const ConvParams p) {
  constexpr int TM = 16 * CPT;
  __shared__ __align__(16) float sA[2][BK][TN];
  __shared__ __align__(16) float sW[2][BK][TM];

  const int tid = threadIdx.x;
  const int tx = tid & 15;   // time group: positions tx*4..+3 and 64+tx*4..+3
  const int ty = tid >> 4;   // channel group: channels ty*CPT..+CPT-1
  const int t0 = blockIdx.x * TN;
  const int co0 = blockIdx.y * TM;
  const int b = blockIdx.z / p.n_phase;
  const int ph = blockIdx.z - b * p.n_phase;
  const int phase_shift = p.n_phase > 1 ? (ph + p.up_P) / p.up_f : 0;
  const int out_off = p.n_phase > 1 ? ph : 0;

  const int lt = tid & (TN - 1);  // loader: this thread's time column
  const int lk = tid >> 7;        // loader: K rows lk, lk+2, ..., lk+14
  const bool t_ok = (t0 + lt) < p.Lt;

  const int nck0 = (p.src[0].C + BK - 1) / BK;
  const int its0 = p.src[0].taps * nck0;
  const int nck1 = p.nsrc > 1 ? (p.src[1].C + BK - 1) / BK : 0;
  const int n_it = its0 + (p.nsrc > 1 ? p.src[1].taps * nck1 : 0);

  float acc[CPT][8];
#pragma unroll
  for (int c = 0; c < CPT; ++c)
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[c][i] = 0.f;

  float ra[BK / 2];
  float rw[CPT];
  bool ra_lrelu = false;   // the prefetched slice still wants its LeakyReLU (applied in stash(), see fetch())

  // loader constants of this thread: the K row and tile column of each of its CPT weight elements (e = tid + NT * i
  // over a [BK][TM] slice), as one offset into the [k][CoutPad] weight rows
  int wk[CPT], woff[CPT];
#pragma unroll
  for (int i = 0; i < CPT; ++i) {
    const int e = tid + NT * i;
    wk[i] = e / TM;
    woff[i] = wk[i] * p.CoutPad + (e - wk[i] * TM);
  }

  auto fetch = [&](int it) {
    const int si = it >= its0 ? 1 : 0;
    const ConvSrc& s = p.src[si];
    const int r = it - (si ? its0 : 0);
    const int nck = si ? nck1 : nck0;
    const int j = r / nck;
    const int c0 = (r - j * nck) * BK;
    const int col = t_ok ? resolve_pos(s, t0 + lt + j * s.dil + s.shift + phase_shift) : -1;
    const bool full = c0 + BK <= s.C;   // uniform over the CTA: every channel of the slice exists (all but a last slice)
    // the raw value only: touching it here (the LeakyReLU prologue used to) makes the warp wait for the load
    // BEFORE the FMAs of the current slice — half of all warp time in the kernel (ncu: long-scoreboard stalls on
    // exactly that multiply); the activation is applied when the slice is stashed, after the FMAs
    const float* xp = s.x + (int64_t)b * s.batch_stride + (int64_t)(c0 + lk) * s.Lreal + (col >= 0 ? col : 0);
    const int64_t xstep = 2 * (int64_t)s.Lreal;
    if (full && col >= 0) {
#pragma unroll
      for (int i = 0; i < BK / 2; ++i) ra[i] = __ldg(xp + i * xstep);
    } else {
#pragma unroll
      for (int i = 0; i < BK / 2; ++i) ra[i] = (col >= 0 && c0 + lk + 2 * i < s.C) ? __ldg(xp + i * xstep) : 0.f;
    }
    ra_lrelu = s.lrelu != 0;
    const float* wp = s.w + ((int64_t)(ph * s.taps + j) * s.C + c0) * p.CoutPad + co0;
    if (full) {
#pragma unroll
      for (int i = 0; i < CPT; ++i) rw[i] = __ldg(wp + woff[i]);
    } else {
#pragma unroll
      for (int i = 0; i < CPT; ++i) rw[i] = (c0 + wk[i] < s.C) ? __ldg(wp + woff[i]) : 0.f;
    }
  };
  auto stash = [&](int buf) {
#pragma unroll
    for (int i = 0; i < BK / 2; ++i) sA[buf][lk + 2 * i][lt] = ra_lrelu ? lrelu_f(ra[i], p.slope) : ra[i];
#pragma unroll
    for (int i = 0; i < CPT; ++i) (&sW[buf][0][0])[tid + NT * i] = rw[i];   // [k][m] is linear in e = tid + NT * i
  };

  fetch(0);
  stash(0);
  __syncthreads();
  for (int it = 0; it < n_it; ++it) {
    const int buf = it & 1;
    if (it + 1 < n_it) fetch(it + 1);
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      const float4 a0 = *reinterpret_cast<const float4*>(&sA[buf][k][tx * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&sA[buf][k][64 + tx * 4]);
      float w[CPT];
#pragma unroll
      for (int c = 0; c < CPT; ++c) w[c] = sW[buf][k][ty * CPT + c];
#pragma unroll
      for (int c = 0; c < CPT; ++c) {
        acc[c][0] = fmaf(w[c], a0.x, acc[c][0]);
        acc[c][1] = fmaf(w[c], a0.y, acc[c][1]);
        acc[c][2] = fmaf(w[c], a0.z, acc[c][2]);
        acc[c][3] = fmaf(w[c], a0.w, acc[c][3]);
        acc[c][4] = fmaf(w[c], a1.x, acc[c][4]);
        acc[c][5] = fmaf(w[c], a1.y, acc[c][5]);
        acc[c][6] = fmaf(w[c], a1.z, acc[c][6]);
        acc[c][7] = fmaf(w[c], a1.w, acc[c][7]);
      }
    }
    if (it + 1 < n_it) stash(buf ^ 1);
    __syncthreads();
  }

  // epilogue: bias (+tanh), NCL store
  const bool vec_ok = p.out_stride == 1 && (p.out_L & 3) == 0 &&
                      ((reinterpret_cast<uintptr_t>(p.out) & 15) == 0) && ((p.out_batch_stride & 3) == 0);
#pragma unroll
  for (int c = 0; c < CPT; ++c) {
    const int co = co0 + ty * CPT + c;
    if (co >= p.Cout) continue;
    const float bv = p.bias[co];
    float* orow = p.out + (int64_t)b * p.out_batch_stride + (int64_t)co * p.out_L;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int t = t0 + h * 64 + tx * 4;
      float v[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        v[i] = acc[c][h * 4 + i] + bv;
        if (p.tanh_out) v[i] = tanhf(v[i]);
      }
      if (vec_ok && t + 3 < p.Lt) {
        *reinterpret_cast<float4*>(orow + t) = make_float4(v[0], v[1], v[2], v[3]);
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (t + i < p.Lt) orow[(int64_t)(t + i) * p.out_stride + out_off] = v[i];
      }
    }
  }
}

}  // namespace

int conv_f32_pick_cpt(int Cout) {
  // candidates 1,2,3,4,6,8: least padded channels first, then the widest tile
  const int cand[6] = {8, 6, 4, 3, 2, 1};
  int best = 1;
  int64_t best_pad = -1;
  for (int i = 0; i < 6; ++i) {
    const int tm = 16 * cand[i];
    const int64_t padded = round_up(Cout, tm);
    if (best_pad < 0 || padded < best_pad) {
      best_pad = padded;
      best = cand[i];
    }
  }
  return best;
}

cudaError_t conv_f32_launch(const ConvParams& p, int batch, int cpt, cudaStream_t stream) {
  const dim3 grid(ceil_div(p.Lt, TN), p.CoutPad / (16 * cpt), batch * p.n_phase);
  const dim3 block(NT);
  switch (cpt) {
    case 1: conv_f32_kernel<1><<<grid, block, 0, stream>>>(p); break;
    case 2: conv_f32_kernel<2><<<grid, block, 0, stream>>>(p); break;
    case 3: conv_f32_kernel<3><<<grid, block, 0, stream>>>(p); break;
    case 4: conv_f32_kernel<4><<<grid, block, 0, stream>>>(p); break;
    case 6: conv_f32_kernel<6><<<grid, block, 0, stream>>>(p); break;
    case 8: conv_f32_kernel<8><<<grid, block, 0, stream>>>(p); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

}  // namespace ttsb
