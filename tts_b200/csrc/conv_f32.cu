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

// CPT output channels x 8 time positions per thread; YG channel groups per CTA, so the tile is YG * CPT channels x
// 2048 / YG time positions; K walked in slices of BK input channels.  Wide layers take YG = 16 (128 positions); layers of
// 24 / 32 / 48 channels take fewer, wider groups so that the thread tile stays 6-8 channels deep (FMAs per shared-memory
// read) without padding the channel tile, and BK = 8 so that 24 input channels are three whole slices.
template <int CPT, int YG, int BK>
__global__ void __launch_bounds__(NT, CPT >= 6 ? 2 : (CPT >= 3 ? 3 : 4)) conv_f32_kernel(const ConvParams p) {
  constexpr int TM = YG * CPT;
  constexpr int XG = NT / YG;                     // time groups
  constexpr int TN = 8 * XG;                      // time positions per CTA
  constexpr int NA = BK * TN / NT;                // staged input values per thread and slice
  constexpr int NCOL = TN > NT ? TN / NT : 1;     // distinct time columns among them
  constexpr int NW = (BK * TM + NT - 1) / NT;     // staged weights per thread and slice
  static_assert(BK * TN % NT == 0 && (TN >= NT ? TN % NT == 0 : NT % TN == 0), "loader mapping");
  __shared__ __align__(16) float sA[2][BK][TN];
  __shared__ __align__(16) float sW[2][BK][TM];

  const int tid = threadIdx.x;
  const int tx = tid % XG;   // time group: positions tx*4..+3 and TN/2+tx*4..+3
  const int ty = tid / XG;   // channel group: channels ty*CPT..+CPT-1
  const int t0 = blockIdx.x * TN;
  const int co0 = blockIdx.y * TM;
  const int b = blockIdx.z / p.n_phase;
  const int ph = blockIdx.z - b * p.n_phase;
  const int phase_shift = p.n_phase > 1 ? (ph + p.up_P) / p.up_f : 0;
  const int out_off = p.n_phase > 1 ? ph : 0;

  // loader: element e = tid + NT * i of a [BK][TN] slice sits in row e / TN, column e % TN — NCOL columns per thread
  int lt[NCOL];
  bool t_ok[NCOL];
#pragma unroll
  for (int c = 0; c < NCOL; ++c) {
    lt[c] = (tid + NT * c) % TN;
    t_ok[c] = (t0 + lt[c]) < p.Lt;
  }

  const int nck0 = (p.src[0].C + BK - 1) / BK;
  const int its0 = p.src[0].taps * nck0;
  const int nck1 = p.nsrc > 1 ? (p.src[1].C + BK - 1) / BK : 0;
  const int n_it = its0 + (p.nsrc > 1 ? p.src[1].taps * nck1 : 0);

  float acc[CPT][8];
#pragma unroll
  for (int c = 0; c < CPT; ++c)
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[c][i] = 0.f;

  float ra[NA];
  float rw[NW];
  bool ra_lrelu = false;   // the prefetched slice still wants its LeakyReLU (applied in stash(), see fetch())

  // loader constants of this thread: the K row and tile column of each of its weight elements (e = tid + NT * i
  // over a [BK][TM] slice), as one offset into the [k][CoutPad] weight rows
  int wk[NW], woff[NW];
#pragma unroll
  for (int i = 0; i < NW; ++i) {
    const int e = tid + NT * i;
    wk[i] = e < BK * TM ? e / TM : BK;   // BK: no such element (a slice smaller than the CTA)
    woff[i] = wk[i] * p.CoutPad + (e - wk[i] * TM);
  }

  auto fetch = [&](int it) {
    const int si = it >= its0 ? 1 : 0;
    const ConvSrc& s = p.src[si];
    const int r = it - (si ? its0 : 0);
    const int nck = si ? nck1 : nck0;
    const int j = r / nck;
    const int c0 = (r - j * nck) * BK;
    const bool full = c0 + BK <= s.C;   // uniform over the CTA: every channel of the slice exists (all but a last slice)
    // the raw value only: touching it here (the LeakyReLU prologue used to) makes the warp wait for the load
    // BEFORE the FMAs of the current slice — half of all warp time in the kernel (ncu: long-scoreboard stalls on
    // exactly that multiply); the activation is applied when the slice is stashed, after the FMAs
    const float* xb = s.x + (int64_t)b * s.batch_stride + (int64_t)c0 * s.Lreal;
#pragma unroll
    for (int c = 0; c < NCOL; ++c) {
      const int col = t_ok[c] ? resolve_pos(s, t0 + lt[c] + j * s.dil + s.shift + phase_shift) : -1;
      const float* xp = xb + (col >= 0 ? col : 0);
      if (full && col >= 0) {
#pragma unroll
        for (int i = c; i < NA; i += NCOL) ra[i] = __ldg(xp + (int64_t)((tid + NT * i) / TN) * s.Lreal);
      } else {
#pragma unroll
        for (int i = c; i < NA; i += NCOL) {
          const int row = (tid + NT * i) / TN;
          ra[i] = (col >= 0 && c0 + row < s.C) ? __ldg(xp + (int64_t)row * s.Lreal) : 0.f;
        }
      }
    }
    ra_lrelu = s.lrelu != 0;
    const float* wp = s.w + ((int64_t)(ph * s.taps + j) * s.C + c0) * p.CoutPad + co0;
    if (full && BK * TM % NT == 0) {
#pragma unroll
      for (int i = 0; i < NW; ++i) rw[i] = __ldg(wp + woff[i]);
    } else {
#pragma unroll
      for (int i = 0; i < NW; ++i) rw[i] = (wk[i] < BK && c0 + wk[i] < s.C) ? __ldg(wp + woff[i]) : 0.f;
    }
  };
  auto stash = [&](int buf) {
#pragma unroll
    for (int i = 0; i < NA; ++i)
      sA[buf][(tid + NT * i) / TN][(tid + NT * i) % TN] = ra_lrelu ? lrelu_f(ra[i], p.slope) : ra[i];
#pragma unroll
    for (int i = 0; i < NW; ++i)
      if (BK * TM % NT == 0 || wk[i] < BK) (&sW[buf][0][0])[tid + NT * i] = rw[i];   // [k][m] is linear in e
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
      const float4 a1 = *reinterpret_cast<const float4*>(&sA[buf][k][TN / 2 + tx * 4]);
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
      const int t = t0 + h * (TN / 2) + tx * 4;
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

// ---- 7-tap conv to a few output channels (the output head: Conv1d(C, 1 or 4, k7) + tanh) --------------------------
// The implicit GEMM above gives every thread 8 time positions x CPT channels and pads the channel tile to 16: with 4
// (or 1) real channels each staged input value feeds ONE FMA per thread and 12 of 16 channel groups idle — the layer ran
// at 3.5 % of the FFMA peak, bound by shared-memory reads, and re-read its input once per tap.  Here a thread owns 4
// consecutive time positions of EVERY output channel: per input channel it reads its 10-value window once (three
// float4) and each tap's CH weights as one broadcast load, 28 * CH FMAs per 3 + 7 shared-memory instructions; the input
// is read once.  Index map (reflect, LeakyReLU prologue) and epilogue (bias, tanh) are conv_f32_kernel's.
constexpr int HEAD_T = 1024;     // output positions per CTA
constexpr int HEAD_K = 8;        // input channels per staged slice
constexpr int HEAD_ROW = HEAD_T + 8;   // staged positions t0 - 4 .. t0 + HEAD_T + 3
constexpr int HEAD_TAPS = 7;
constexpr int HEAD_MAX_C = 64;

template <int CH>
__global__ void __launch_bounds__(NT) head_f32_kernel(const ConvParams p) {
  __shared__ __align__(16) float sx[HEAD_K][HEAD_ROW];
  __shared__ __align__(16) float sw[HEAD_MAX_C * HEAD_TAPS * CH];
  const ConvSrc& s = p.src[0];
  const int tid = threadIdx.x;
  const int t0 = blockIdx.x * HEAD_T;
  const int b = blockIdx.z;
  constexpr int HH = (HEAD_TAPS - 1) / 2;

  for (int e = tid; e < s.C * HEAD_TAPS * CH; e += NT) {   // [ci][tap][ch] <- [tap][ci][CoutPad]
    const int ch = e % CH, j = (e / CH) % HEAD_TAPS, ci = e / (CH * HEAD_TAPS);
    sw[e] = ch < p.Cout ? __ldg(s.w + ((int64_t)j * s.C + ci) * p.CoutPad + ch) : 0.f;
  }
  // stored column of each staged position this thread fills (-1: not needed or zero), the same for every channel
  constexpr int NI = (HEAD_ROW + NT - 1) / NT;
  int col[NI];
#pragma unroll
  for (int i = 0; i < NI; ++i) {
    const int idx = tid + i * NT, pos = t0 - 4 + idx;
    col[i] = (idx < HEAD_ROW && pos >= -HH && pos < p.Lt + HH) ? resolve_pos(s, pos) : -1;
  }

  float acc[CH][4];
#pragma unroll
  for (int c = 0; c < CH; ++c)
#pragma unroll
    for (int r = 0; r < 4; ++r) acc[c][r] = 0.f;

  for (int c0 = 0; c0 < s.C; c0 += HEAD_K) {
    const int nk = min(HEAD_K, s.C - c0);
    if (c0 > 0) __syncthreads();   // the previous slice has been consumed
    for (int k = 0; k < nk; ++k) {
      const float* xr = s.x + (int64_t)b * s.batch_stride + (int64_t)(c0 + k) * s.Lreal;
#pragma unroll
      for (int i = 0; i < NI; ++i) {
        const int idx = tid + i * NT;
        if (idx < HEAD_ROW) {
          float v = col[i] >= 0 ? __ldg(xr + col[i]) : 0.f;
          if (s.lrelu) v = lrelu_f(v, p.slope);
          sx[k][idx] = v;
        }
      }
    }
    __syncthreads();
    for (int k = 0; k < nk; ++k) {
      float v[12];
#pragma unroll
      for (int q = 0; q < 3; ++q) {
        const float4 f = *reinterpret_cast<const float4*>(&sx[k][4 * tid + 4 * q]);
        v[4 * q] = f.x; v[4 * q + 1] = f.y; v[4 * q + 2] = f.z; v[4 * q + 3] = f.w;
      }
      const float* wr = sw + (c0 + k) * HEAD_TAPS * CH;
#pragma unroll
      for (int j = 0; j < HEAD_TAPS; ++j) {
        float w[CH];
        if (CH == 4) {
          const float4 f = *reinterpret_cast<const float4*>(wr + j * 4);
          w[0] = f.x; w[1 % CH] = f.y; w[2 % CH] = f.z; w[3 % CH] = f.w;
        } else {
#pragma unroll
          for (int c = 0; c < CH; ++c) w[c] = wr[j * CH + c];
        }
#pragma unroll
        for (int c = 0; c < CH; ++c)
#pragma unroll
          for (int r = 0; r < 4; ++r) acc[c][r] = fmaf(w[c], v[r + j + 4 - HH], acc[c][r]);   // position t + j - HH
      }
    }
  }

  const int t = t0 + 4 * tid;
  if (t >= p.Lt) return;
  const bool vec_ok = (p.out_L & 3) == 0 && ((reinterpret_cast<uintptr_t>(p.out) & 15) == 0) &&
                      ((p.out_batch_stride & 3) == 0) && t + 3 < p.Lt;
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    if (c >= p.Cout) continue;
    const float bv = p.bias[c];
    float* orow = p.out + (int64_t)b * p.out_batch_stride + (int64_t)c * p.out_L;
    float y[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      y[r] = acc[c][r] + bv;
      if (p.tanh_out) y[r] = tanhf(y[r]);
    }
    if (vec_ok) {
      *reinterpret_cast<float4*>(orow + t) = make_float4(y[0], y[1], y[2], y[3]);
    } else {
#pragma unroll
      for (int r = 0; r < 4; ++r)
        if (t + r < p.Lt) orow[t + r] = y[r];
    }
  }
}

}  // namespace

// The narrow-output kernel serves a plain (one source, one phase, unit stride and dilation) centred 7-tap conv to at most
// 4 channels from at most HEAD_MAX_C: every shipped generator's output head.
bool head_f32_usable(const ConvParams& p) {
  const ConvSrc& s = p.src[0];
  return p.nsrc == 1 && p.n_phase == 1 && p.out_stride == 1 && p.Cout <= 4 && s.C <= HEAD_MAX_C && s.taps == HEAD_TAPS &&
         s.dil == 1 && s.shift == -(HEAD_TAPS - 1) / 2 && s.L > HEAD_TAPS && p.Lt == s.L;
}

cudaError_t head_f32_launch(const ConvParams& p, int batch, cudaStream_t stream) {
  const dim3 grid(ceil_div(p.Lt, HEAD_T), 1, batch);
  if (p.Cout == 1) head_f32_kernel<1><<<grid, NT, 0, stream>>>(p);
  else head_f32_kernel<4><<<grid, NT, 0, stream>>>(p);
  return cudaGetLastError();
}

// Tile codes: 1, 2, 3, 4, 6, 8 = that many channels per thread in 16 groups (tile 16 * code channels x 128 positions, K
// slices of 16); 24 / 32 / 48 = a tile of exactly that many channels in 4 / 4 / 8 groups of 6 / 8 / 6 (512 / 512 / 256
// positions, K slices of 8 / 8 / 16).
int conv_f32_tile_channels(int code) { return code >= 24 ? code : 16 * code; }

int conv_f32_pick_cpt(int Cout) {
  if (Cout == 24 || Cout == 32 || Cout == 48) return Cout;   // narrow layers: a tile of their own (no padding)
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

namespace {
template <int CPT, int YG, int BK>
cudaError_t launch_tile(const ConvParams& p, int batch, cudaStream_t stream) {
  const dim3 grid(ceil_div(p.Lt, 8 * (NT / YG)), p.CoutPad / (YG * CPT), batch * p.n_phase);
  conv_f32_kernel<CPT, YG, BK><<<grid, dim3(NT), 0, stream>>>(p);
  return cudaGetLastError();
}
}  // namespace

cudaError_t conv_f32_launch(const ConvParams& p, int batch, int cpt, cudaStream_t stream) {
  if (p.CoutPad % conv_f32_tile_channels(cpt) != 0) return cudaErrorInvalidValue;
  switch (cpt) {
    case 1: return launch_tile<1, 16, 16>(p, batch, stream);
    case 2: return launch_tile<2, 16, 16>(p, batch, stream);
    case 3: return launch_tile<3, 16, 16>(p, batch, stream);
    case 4: return launch_tile<4, 16, 16>(p, batch, stream);
    case 6: return launch_tile<6, 16, 16>(p, batch, stream);
    case 8: return launch_tile<8, 16, 16>(p, batch, stream);
    case 24: return launch_tile<6, 4, 8>(p, batch, stream);
    case 32: return launch_tile<8, 4, 8>(p, batch, stream);
    case 48: return launch_tile<6, 8, 16>(p, batch, stream);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace ttsb
