// PQMF synthesis / analysis (TTS/vocoder/layers/pqmf.py:48-56) as polyphase FIR kernels.
//
// Reference synthesis: zero-stuff x N with gain N (conv_transpose1d with updown_filter*N,
// stride N), then conv1d with G [1,N,K], K = taps+1, zero padding taps/2.  Only every N-th tap
// meets a non-zero sample, so for output t = N q + r
//     y[t] = N * sum_k sum_{j : (r + j - pad) % N == 0} G[k][j] * x[k][q + (r + j - pad)/N].
// For N=4, K=63 that is 4 bands x 16 positions x[k][q-7 .. q+8] per group of 4 outputs:
// 63 MACs/sample instead of the reference's 252, and no zero-stuffed intermediate in HBM.
//
// Roofline: 8 B/sample (4 B sub-band in + 4 B out) vs 126 FLOP/sample -> 15.8 FLOP/B, which is
// at the fp32-FFMA ridge of a B200 (~11 FLOP/B), so the specialised kernel keeps both the FFMA
// pipe and the LSU lean: x windows come from shared memory as aligned float4, coefficients as
// one broadcast float4 per (band, offset), and stores are staged for full-line writes.
#include "common.cuh"

namespace ttsb {

namespace {

// ---- specialised N=4, K=63 (the only PQMF the reference instantiates: multiband_melgan_generator.py:24)
constexpr int SQ_THREADS = 128;
constexpr int SQ_QPT = 4;                        // q (groups of 4 outputs) per thread
constexpr int SQ_TILE = SQ_THREADS * SQ_QPT;     // 512 q per CTA -> 2048 output samples
constexpr int SQ_HALO_L = 8;                     // window is q-7..q+8; 8 keeps float4 alignment
constexpr int SQ_SMEM_Q = SQ_TILE + 16;          // 8 left + 8 right

__global__ void __launch_bounds__(SQ_THREADS) pqmf_synth_4x63_kernel(const float* __restrict__ x,
                                                                      const float* __restrict__ G,
                                                                      float* __restrict__ y, int Ts) {
  __shared__ __align__(16) float sx[4][SQ_SMEM_Q];
  __shared__ __align__(16) float4 sc[4][16];              // sc[k][off+7] = 4*G[k][4*off+31-r], r=0..3
  __shared__ __align__(16) float so[SQ_THREADS * 16 + 16];

  const int tid = threadIdx.x;
  const int b = blockIdx.y;
  const int q0 = blockIdx.x * SQ_TILE;
  const float* xb = x + (int64_t)b * 4 * Ts;

  pdl_launch_dependents();
  if (tid < 64) {
    const int k = tid >> 4, o = (tid & 15) - 7;
    float c[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int j = 4 * o + 31 - r;
      c[r] = (j >= 0 && j < 63) ? 4.f * G[k * 63 + j] : 0.f;
    }
    sc[k][tid & 15] = make_float4(c[0], c[1], c[2], c[3]);
  }
  pdl_wait();   // the head's sub-band signals are complete from here on
  for (int e = tid; e < 4 * SQ_SMEM_Q; e += SQ_THREADS) {
    const int k = e / SQ_SMEM_Q, s = e - k * SQ_SMEM_Q;
    const int q = q0 - SQ_HALO_L + s;
    sx[k][s] = (q >= 0 && q < Ts) ? __ldg(xb + (int64_t)k * Ts + q) : 0.f;
  }
  __syncthreads();

  float acc[SQ_QPT][4];
#pragma unroll
  for (int i = 0; i < SQ_QPT; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;

#pragma unroll
  for (int k = 0; k < 4; ++k) {
    float w[20];  // x[k][q_first-8 .. q_first+11], q_first = q0 + 4*tid
#pragma unroll
    for (int v = 0; v < 5; ++v) {
      const float4 t = *reinterpret_cast<const float4*>(&sx[k][4 * tid + 4 * v]);
      w[4 * v + 0] = t.x; w[4 * v + 1] = t.y; w[4 * v + 2] = t.z; w[4 * v + 3] = t.w;
    }
#pragma unroll
    for (int o = 0; o < 16; ++o) {       // offset o-7
      const float4 c = sc[k][o];
#pragma unroll
      for (int i = 0; i < SQ_QPT; ++i) {
        const float xv = w[i + o + 1];   // x[k][q_first + i + (o-7)] -> window index i + o - 7 + 8
        acc[i][0] = fmaf(c.x, xv, acc[i][0]);
        acc[i][1] = fmaf(c.y, xv, acc[i][1]);
        acc[i][2] = fmaf(c.z, xv, acc[i][2]);
        acc[i][3] = fmaf(c.w, xv, acc[i][3]);
      }
    }
  }

  // stage 16 consecutive outputs per thread, then write full lines
#pragma unroll
  for (int i = 0; i < SQ_QPT; ++i)
    *reinterpret_cast<float4*>(&so[tid * 16 + 4 * i]) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
  __syncthreads();
  const int64_t Tout = 4 * (int64_t)Ts;
  float* yb = y + (int64_t)b * Tout + (int64_t)q0 * 4;
  const int64_t remain = Tout - (int64_t)q0 * 4;  // multiple of 4
  const bool aligned = (reinterpret_cast<uintptr_t>(y) & 15) == 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int e = (tid + SQ_THREADS * i) * 4;
    if (e < remain) {
      const float4 v = *reinterpret_cast<const float4*>(&so[e]);
      if (aligned) {
        *reinterpret_cast<float4*>(yb + e) = v;
      } else {
        yb[e] = v.x; yb[e + 1] = v.y; yb[e + 2] = v.z; yb[e + 3] = v.w;
      }
    }
  }
}

// ---- generic (any N, any taps): one thread per output sample ------------------------------------
__global__ void pqmf_synth_generic_kernel(const float* __restrict__ x, const float* __restrict__ G,
                                          float* __restrict__ y, int N, int K, int pad, int64_t Ts,
                                          int64_t Tout) {
  const int b = blockIdx.y;
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= Tout) return;
  // j must satisfy (t + j - pad) % N == 0
  int64_t base = t - pad;
  int j0 = (int)(((-base) % N + N) % N);
  float acc = 0.f;
  for (int j = j0; j < K; j += N) {
    const int64_t m = (base + j) / N;  // exact
    if (m < 0 || m >= Ts) continue;
    for (int k = 0; k < N; ++k) acc = fmaf(G[k * K + j], x[((int64_t)b * N + k) * Ts + m], acc);
  }
  y[(int64_t)b * Tout + t] = acc * (float)N;
}

// analysis: y[b][k][m] = sum_j H[k][j] * x[b][N m + j - pad]   (zero outside)
__global__ void pqmf_analysis_kernel(const float* __restrict__ x, const float* __restrict__ H,
                                     float* __restrict__ y, int N, int K, int pad, int64_t T,
                                     int64_t Tout) {
  extern __shared__ float sh[];  // [N][K] filter, then the x window of the CTA
  float* sH = sh;
  float* sx = sh + N * K;
  const int b = blockIdx.y;
  const int64_t m0 = (int64_t)blockIdx.x * blockDim.x;
  const int win = blockDim.x * N + K;
  for (int e = threadIdx.x; e < N * K; e += blockDim.x) sH[e] = H[e];
  for (int e = threadIdx.x; e < win; e += blockDim.x) {
    const int64_t p = m0 * N - pad + e;
    sx[e] = (p >= 0 && p < T) ? x[(int64_t)b * T + p] : 0.f;
  }
  __syncthreads();
  const int64_t m = m0 + threadIdx.x;
  if (m >= Tout) return;
  for (int k = 0; k < N; ++k) {
    float acc = 0.f;
    for (int j = 0; j < K; ++j) acc = fmaf(sH[k * K + j], sx[threadIdx.x * N + j], acc);
    y[((int64_t)b * N + k) * Tout + m] = acc;
  }
}

}  // namespace

cudaError_t pqmf_synthesis_launch(const float* x, const float* G, int batch, int bands, int taps,
                                  int64_t Ts, float* y, cudaStream_t stream) {
  const int K = taps + 1, pad = taps / 2;
  const int64_t Tout = (int64_t)bands * Ts + 2 * pad - taps;
  if (batch <= 0 || Ts <= 0 || Tout <= 0) return cudaSuccess;
  if (bands == 4 && taps == 62 && Ts < (1 << 30)) {
    const dim3 grid(ceil_div(Ts, SQ_TILE), batch);
    return launch_pdl(pqmf_synth_4x63_kernel, grid, dim3(SQ_THREADS), 0, stream, x, G, y, (int)Ts);
  } else {
    const dim3 grid(ceil_div(Tout, 256), batch);
    pqmf_synth_generic_kernel<<<grid, 256, 0, stream>>>(x, G, y, bands, K, pad, Ts, Tout);
  }
  return cudaGetLastError();
}

cudaError_t pqmf_analysis_launch(const float* x, const float* H, int batch, int bands, int taps,
                                 int64_t T, float* y, cudaStream_t stream) {
  const int K = taps + 1, pad = taps / 2;
  const int64_t Tout = (T + 2 * pad - K) / bands + 1;
  if (batch <= 0 || T <= 0 || Tout <= 0) return cudaSuccess;
  const int threads = 128;
  const size_t smem = (size_t)(bands * K + threads * bands + K) * sizeof(float);
  const dim3 grid(ceil_div(Tout, threads), batch);
  pqmf_analysis_kernel<<<grid, threads, smem, stream>>>(x, H, y, bands, K, pad, T, Tout);
  return cudaGetLastError();
}

}  // namespace ttsb
