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
#include <algorithm>
#include <cmath>
#include <cstdlib>

#include <cuda_fp16.h>

#include "common.cuh"
#include "follow.cuh"

namespace ttsb {

namespace {

// Loads of data the PREVIOUS kernel of a programmatic-dependent-launch chain wrote.  They must stay behind
// griddepcontrol.wait: a read-only-path load (__ldg / const __restrict__ -> ld.global.nc) is assumed by the
// compiler to read immutable memory and may be hoisted above the wait (seen: the factorised kernel's first loads
// moved up and read the sub-bands before the head kernel had written them).  volatile asm pins them.
__device__ __forceinline__ float4 ld_after_wait4(const float* p) {
  float4 v;
  asm volatile("ld.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float ld_after_wait(const float* p) {
  float v;
  asm volatile("ld.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}

// Where the samples go: row b starts at y + b * batch_stride and holds samples [lo, hi) of the synthesis output.
struct PqmfOut {
  int64_t batch_stride, lo, hi;
  int coalesce;   // fp32 stores as 512-byte runs per warp instruction (windowed output: may be peer memory)
};

// 16 consecutive output samples of one thread -> global memory as full 32-byte sectors: two 32-byte stores of fp32,
// or (fp16 waveform: TTSB200_OUT_F16, half the device->host bytes of the result) one 32-byte store of packed halves.
// `y0` points at sample 0 of this batch item's row; only samples t in [lo, hi) are stored (the whole row, or the
// kept range of a long-form / streaming window whose halo samples are computed but never written).
__device__ __forceinline__ void store16(float* y0, const float (&acc)[4][4], int64_t t_first, int64_t lo, int64_t hi) {
  float* yt = y0 + t_first;
  if (t_first >= lo && t_first + 16 <= hi && (reinterpret_cast<uintptr_t>(yt) & 31) == 0) {
#pragma unroll
    for (int h = 0; h < 2; ++h)
      asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(yt + 8 * h), "f"(acc[2 * h][0]),
                   "f"(acc[2 * h][1]), "f"(acc[2 * h][2]), "f"(acc[2 * h][3]), "f"(acc[2 * h + 1][0]),
                   "f"(acc[2 * h + 1][1]), "f"(acc[2 * h + 1][2]), "f"(acc[2 * h + 1][3])
                   : "memory");
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int r = 0; r < 4; ++r)
        if (t_first + 4 * i + r >= lo && t_first + 4 * i + r < hi) yt[4 * i + r] = acc[i][r];
  }
}
__device__ __forceinline__ void store16(__half* y0, const float (&acc)[4][4], int64_t t_first, int64_t lo, int64_t hi) {
  __half* yt = y0 + t_first;
  if (t_first >= lo && t_first + 16 <= hi && (reinterpret_cast<uintptr_t>(yt) & 31) == 0) {
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __half2 lo = __floats2half2_rn(acc[i][0], acc[i][1]), hi = __floats2half2_rn(acc[i][2], acc[i][3]);
      w[2 * i] = *reinterpret_cast<const uint32_t*>(&lo);
      w[2 * i + 1] = *reinterpret_cast<const uint32_t*>(&hi);
    }
    asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(yt), "r"(w[0]), "r"(w[1]), "r"(w[2]),
                 "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7])
                 : "memory");
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int r = 0; r < 4; ++r)
        if (t_first + 4 * i + r >= lo && t_first + 4 * i + r < hi) yt[4 * i + r] = __float2half_rn(acc[i][r]);
  }
}
__device__ __forceinline__ void store1(float* y, float v) { *y = v; }
__device__ __forceinline__ void store1(__half* y, float v) { *y = __float2half_rn(v); }

// ---- specialised N=4, K=63 (the only PQMF the reference instantiates: multiband_melgan_generator.py:24)
constexpr int SQ_THREADS = 64;
constexpr int SQ_QPT = 4;                        // q (groups of 4 outputs) per thread
constexpr int SQ_TILE = SQ_THREADS * SQ_QPT;     // 256 q per CTA -> 1024 output samples: small CTAs, several waves,
                                                 // so one CTA's loads and stores hide behind another's FMAs
constexpr int SQ_HALO_L = 8;                     // window is q-7..q+8; 8 keeps float4 alignment
constexpr int SQ_SMEM_Q = SQ_TILE + 16;          // 8 left + 8 right

template <typename OutT>
__global__ void __launch_bounds__(SQ_THREADS) pqmf_synth_4x63_kernel(const float* x,
                                                                      const float* __restrict__ G,
                                                                      OutT* __restrict__ y, int Ts, PqmfOut o) {
  __shared__ __align__(16) float sx[4][SQ_SMEM_Q];
  __shared__ __align__(16) float4 sc[4][16];              // sc[k][off+7] = 4*G[k][4*off+31-r], r=0..3

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
  // window rows q0-8 .. q0+TILE+8 of the four bands; whole float4s when the rows allow it
  const bool vec = (Ts & 3) == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0;
  if (vec) {
    for (int e = tid; e < SQ_SMEM_Q; e += SQ_THREADS) {     // e = k * (SMEM_Q/4) + s4
      const int k = e / (SQ_SMEM_Q / 4), s4 = e - k * (SQ_SMEM_Q / 4);
      const int q = q0 - SQ_HALO_L + 4 * s4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (q >= 0 && q + 3 < Ts) v = ld_after_wait4(xb + (int64_t)k * Ts + q);
      *reinterpret_cast<float4*>(&sx[k][4 * s4]) = v;       // q < 0 or q >= Ts: zero padding (whole vectors)
    }
  } else {
    for (int e = tid; e < 4 * SQ_SMEM_Q; e += SQ_THREADS) {
      const int k = e / SQ_SMEM_Q, s = e - k * SQ_SMEM_Q;
      const int q = q0 - SQ_HALO_L + s;
      sx[k][s] = (q >= 0 && q < Ts) ? ld_after_wait(xb + (int64_t)k * Ts + q) : 0.f;
    }
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

  // each thread owns 4*QPT consecutive outputs, written as 32-byte stores (full sectors, no staging pass)
  const int64_t t_first = (int64_t)(q0 + SQ_QPT * tid) * 4;
  static_assert(SQ_QPT == 4, "store16 writes 16 samples per thread");
  store16(y + (int64_t)b * o.batch_stride - o.lo, acc, t_first, o.lo, o.hi);
}

// ---- factorised N=4, K=63 synthesis ------------------------------------------------------------------------
// The reference's synthesis bank is cosine-modulated (pqmf.py:22-32): G[k][j] = q[j] * C[k][j] with
// C[k][j] = 2 cos((2k+1) pi/8 (j - 30.5) - (-1)^k pi/4), which has C[k][j+8] = -C[k][j].  So
//     y[4q+r] = sum_o w[r][o] * v[(4o+31-r) mod 8][q+o],   v[p][m] = sum_k C[k][p] x[k][m],
//     w[r][o] = 4 q[j] (-1)^(j div 8), j = 4o+31-r, o = -7..8:
// 8 modulated signals (32 MACs per sub-band sample = 8 per output) + a 16-tap FIR per output = 24 MACs per
// output sample instead of 63.  Output r uses v[7-r] at even o and v[3-r] at odd o.  The table (C then w,
// pqmf_factor_4x63 below) is built at plan creation and only if G really has this structure.
constexpr int SF_THREADS = 128;
constexpr int SF_TILE = SF_THREADS * 4;          // 512 q per CTA -> 2048 output samples
constexpr int SF_WIN = SF_TILE + 16;             // v window: q0-8 .. q0+TILE+7

template <typename OutT>
__global__ void __launch_bounds__(SF_THREADS) pqmf_synth_4x63_fact_kernel(const float* x,
                                                                           const float* __restrict__ tbl,
                                                                           OutT* __restrict__ y, int Ts, PqmfOut o,
                                                                           const TcFollow follow) {
  __shared__ __align__(16) float sv[8][SF_WIN];
  __shared__ __align__(16) float st[PQMF_FACT_TABLE];   // [p][k] modulation, then [p][8] FIR taps

  const int tid = threadIdx.x;
  const int b = blockIdx.y;
  const int q0 = blockIdx.x * SF_TILE;
  const float* xb = x + (int64_t)b * 4 * Ts;

  if (follow.done && !follow.early) {   // late form: the check precedes this CTA's trigger (follow.cuh follow_armed)
    if (tid == 0) follow_armed(follow);
    __syncthreads();
  }
  pdl_launch_dependents();   // every thread: the CTA counts as triggered once all of its threads have
  if (tid < PQMF_FACT_TABLE) st[tid] = tbl[tid];
  // The head kernel counts finished chains per time segment (follow.cuh): follow it segment by segment when it
  // does, else wait for its whole grid.
  if (follow.done) {
    if (tid == 0) {
      FollowState fst;
      follow_wait(follow, b, q0 - 8, q0 + SF_TILE + 8, fst);
    }
  } else {
    pdl_wait();   // the head's sub-band signals are complete from here on
  }
  __syncthreads();

  // ---- phase 1: v[p][m] for the window.  Thread tid: 4 consecutive m (one float4 per band) ...
  {
    const int mw = 4 * tid;                       // window index of the first m
    const int m = q0 - 8 + mw;
    float4 xv[4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
      xv[k] = (m >= 0 && m + 3 < Ts) ? ld_after_wait4(xb + (int64_t)k * Ts + m)
                                     : make_float4(0.f, 0.f, 0.f, 0.f);   // Ts % 4 == 0: whole vectors in or out
#pragma unroll
    for (int p = 0; p < 8; ++p) {
      const float4 c = *reinterpret_cast<const float4*>(&st[4 * p]);
      float4 v;
      v.x = fmaf(c.w, xv[3].x, fmaf(c.z, xv[2].x, fmaf(c.y, xv[1].x, c.x * xv[0].x)));
      v.y = fmaf(c.w, xv[3].y, fmaf(c.z, xv[2].y, fmaf(c.y, xv[1].y, c.x * xv[0].y)));
      v.z = fmaf(c.w, xv[3].z, fmaf(c.z, xv[2].z, fmaf(c.y, xv[1].z, c.x * xv[0].z)));
      v.w = fmaf(c.w, xv[3].w, fmaf(c.z, xv[2].w, fmaf(c.y, xv[1].w, c.x * xv[0].w)));
      *reinterpret_cast<float4*>(&sv[p][mw]) = v;
    }
    // ... plus one (m, p) pair of the 16-sample right halo
    const int mh = SF_TILE + (tid >> 3), ph = tid & 7;
    const int mg = q0 - 8 + mh;
    float acc = 0.f;
    if (mg >= 0 && mg < Ts) {
#pragma unroll
      for (int k = 0; k < 4; ++k) acc = fmaf(st[4 * ph + k], ld_after_wait(xb + (int64_t)k * Ts + mg), acc);
    }
    sv[ph][mh] = acc;
  }
  __syncthreads();

  // ---- phase 2: 4 q x 4 phases per thread
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;
#pragma unroll
  for (int p = 0; p < 8; ++p) {
    float w[20];   // v[p][q_first-8 .. q_first+11]
#pragma unroll
    for (int u = 0; u < 5; ++u) {
      const float4 t = *reinterpret_cast<const float4*>(&sv[p][4 * tid + 4 * u]);
      w[4 * u + 0] = t.x; w[4 * u + 1] = t.y; w[4 * u + 2] = t.z; w[4 * u + 3] = t.w;
    }
    const float4 c0 = *reinterpret_cast<const float4*>(&st[32 + 8 * p]);
    const float4 c1 = *reinterpret_cast<const float4*>(&st[32 + 8 * p + 4]);
    const float c[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
    constexpr int dummy = 0;
    (void)dummy;
    const int r = p >= 4 ? 7 - p : 3 - p;        // the output phase this signal feeds
    const int base = p >= 4 ? 2 : 1;             // window index of (i = 0, first o): o = -6 (even) or -7 (odd)
#pragma unroll
    for (int e = 0; e < 8; ++e)
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[i][r] = fmaf(c[e], w[i + base + 2 * e], acc[i][r]);
  }

  OutT* yb = y + (int64_t)b * o.batch_stride - o.lo;
  if (sizeof(OutT) == 4 && o.coalesce) {
    // fp32 into a window of a larger buffer, possibly a peer GPU's memory over NVLink (long-form ranks): a thread's 16
    // samples as two 32-byte stores make a warp instruction 32 separate sectors 64 bytes apart — fine for the local
    // L2, but over NVLink every sector travels as a packet of its own (measured: ~300 GB/s into the consumer, and the
    // fp16 waveform, whose stores are contiguous, ran 28 % faster at 8 ranks).  Transpose through the (now idle) v
    // tile so that every warp instruction writes 512 contiguous bytes.  Row stride 130 float4: conflict-free both ways.
    __syncthreads();                                     // every thread is done reading sv
    float4* stg = reinterpret_cast<float4*>(&sv[0][0]);
#pragma unroll
    for (int j = 0; j < 4; ++j) stg[j * 130 + tid] = make_float4(acc[j][0], acc[j][1], acc[j][2], acc[j][3]);
    __syncthreads();
    float* yf = reinterpret_cast<float*>(yb);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int g = i * SF_THREADS + tid;                // float4 index within the CTA's 2,048 samples
      const float4 v = stg[(g & 3) * 130 + (g >> 2)];
      const int64_t t = (int64_t)q0 * 4 + 4 * g;
      float* p = yf + t;
      if (t >= o.lo && t + 4 <= o.hi && (reinterpret_cast<uintptr_t>(p) & 15) == 0) {
        *reinterpret_cast<float4*>(p) = v;
      } else {
        const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int r = 0; r < 4; ++r)
          if (t + r >= o.lo && t + r < o.hi) p[r] = e[r];
      }
    }
    return;
  }
  const int64_t t_first = (int64_t)(q0 + 4 * tid) * 4;
  if (t_first >= o.hi) return;
  store16(yb, acc, t_first, o.lo, o.hi);
}

// ---- generic (any N, any taps): one thread per output sample ------------------------------------
template <typename OutT>
__global__ void pqmf_synth_generic_kernel(const float* __restrict__ x, const float* __restrict__ G,
                                          OutT* __restrict__ y, int N, int K, int pad, int64_t Ts,
                                          int64_t Tout, PqmfOut o) {
  const int b = blockIdx.y;
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= Tout || t < o.lo || t >= o.hi) return;
  // j must satisfy (t + j - pad) % N == 0
  int64_t base = t - pad;
  int j0 = (int)(((-base) % N + N) % N);
  float acc = 0.f;
  for (int j = j0; j < K; j += N) {
    const int64_t m = (base + j) / N;  // exact
    if (m < 0 || m >= Ts) continue;
    for (int k = 0; k < N; ++k) acc = fmaf(G[k * K + j], x[((int64_t)b * N + k) * Ts + m], acc);
  }
  store1(y + (int64_t)b * o.batch_stride + (t - o.lo), acc * (float)N);
}

__global__ void f32_to_f16_kernel(const float* __restrict__ x, __half* __restrict__ y, int64_t n) {
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i + 4 <= n && ((reinterpret_cast<uintptr_t>(x + i) & 15) == 0) && ((reinterpret_cast<uintptr_t>(y + i) & 7) == 0)) {
    const float4 v = *reinterpret_cast<const float4*>(x + i);
    const __half2 lo = __floats2half2_rn(v.x, v.y), hi = __floats2half2_rn(v.z, v.w);
    *reinterpret_cast<uint2*>(y + i) = make_uint2(*reinterpret_cast<const uint32_t*>(&lo), *reinterpret_cast<const uint32_t*>(&hi));
  } else {
    for (int64_t j = i; j < n && j < i + 4; ++j) y[j] = __float2half_rn(x[j]);
  }
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

// Host side, at plan creation: G [4][63] -> the factorised kernel's table, or false when G is not the
// cosine-modulated bank of pqmf.py:22-32 (a checkpoint may carry any buffer; the direct kernel then runs).
bool pqmf_factor_4x63(const float* G, float* tbl) {
  const double PI = 3.14159265358979323846;
  double C[4][63], gmax = 0.0;
  for (int k = 0; k < 4; ++k)
    for (int j = 0; j < 63; ++j) {
      C[k][j] = 2.0 * std::cos((2 * k + 1) * (PI / 8.0) * (j - 30.5) - ((k & 1) ? -1.0 : 1.0) * PI / 4.0);
      gmax = std::max(gmax, std::fabs((double)G[k * 63 + j]));
    }
  double q[63];
  for (int j = 0; j < 63; ++j) {
    int kb = 0;
    for (int k = 1; k < 4; ++k)
      if (std::fabs(C[k][j]) > std::fabs(C[kb][j])) kb = k;
    q[j] = (double)G[kb * 63 + j] / C[kb][j];
    for (int k = 0; k < 4; ++k)
      if (std::fabs((double)G[k * 63 + j] - q[j] * C[k][j]) > 1e-6 * gmax) return false;
  }
  for (int p = 0; p < 8; ++p)
    for (int k = 0; k < 4; ++k) tbl[4 * p + k] = (float)C[k][p];
  for (int p = 0; p < 8; ++p) {
    const int r = p >= 4 ? 7 - p : 3 - p, o0 = p >= 4 ? -6 : -7;
    for (int e = 0; e < 8; ++e) {
      const int j = 4 * (o0 + 2 * e) + 31 - r;
      tbl[32 + 8 * p + e] = (j >= 0 && j < 63) ? (float)(4.0 * q[j] * (((j / 8) & 1) ? -1.0 : 1.0)) : 0.f;
    }
  }
  return true;
}

namespace {
template <typename OutT>
cudaError_t fact_launch(const float* x, const float* tbl, int batch, int64_t Ts, OutT* y, PqmfOut o, const TcFollow& follow,
                        cudaStream_t stream) {
  // shared-memory carve-out preference: a hint, set once per process (first device that launches)
  static const cudaError_t carve = cudaFuncSetAttribute(pqmf_synth_4x63_fact_kernel<OutT>,
                                                        cudaFuncAttributePreferredSharedMemoryCarveout, 100);
  if (carve != cudaSuccess) return carve;
  const dim3 grid(ceil_div(Ts, SF_TILE), batch);
  return launch_pdl(pqmf_synth_4x63_fact_kernel<OutT>, grid, dim3(SF_THREADS), 0, stream, x, tbl, y, (int)Ts, o, follow);
}
template <typename OutT>
cudaError_t direct_launch(const float* x, const float* G, int batch, int bands, int taps, int64_t Ts, OutT* y,
                          PqmfOut o, cudaStream_t stream) {
  const int K = taps + 1, pad = taps / 2;
  const int64_t Tout = (int64_t)bands * Ts + 2 * pad - taps;
  if (bands == 4 && taps == 62 && Ts < (1 << 30)) {
    const dim3 grid(ceil_div(Ts, SQ_TILE), batch);
    // 5.4 KB of static shared memory per 64-thread CTA: with the default carve-out only 21 CTAs fit an SM and cfg2's
    // 23.5 CTAs per SM spill into a second, nearly empty wave; ask for the full carve-out (32 CTAs per SM).
    static const cudaError_t carve = cudaFuncSetAttribute(pqmf_synth_4x63_kernel<OutT>,
                                                          cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    if (carve != cudaSuccess) return carve;
    return launch_pdl(pqmf_synth_4x63_kernel<OutT>, grid, dim3(SQ_THREADS), 0, stream, x, G, y, (int)Ts, o);
  }
  const dim3 grid(ceil_div(Tout, 256), batch);
  pqmf_synth_generic_kernel<OutT><<<grid, 256, 0, stream>>>(x, G, y, bands, K, pad, Ts, Tout, o);
  return cudaGetLastError();
}
}  // namespace

namespace {
PqmfOut make_out(const PqmfCrop* crop, int64_t Tout) {
  if (crop && crop->len > 0) return PqmfOut{crop->batch_stride, crop->lo, crop->lo + crop->len, 1};
  return PqmfOut{Tout, 0, Tout, 0};
}
}  // namespace

cudaError_t pqmf_synthesis_fact_launch(const float* x, const float* tbl, int batch, int64_t Ts, void* y, int out_f16,
                                       const PqmfCrop* crop, const TcFollow* follow, cudaStream_t stream) {
  if (batch <= 0 || Ts <= 0) return cudaSuccess;
  const PqmfOut o = make_out(crop, 4 * Ts);
  const TcFollow f = follow ? *follow : TcFollow{};
  return out_f16 ? fact_launch(x, tbl, batch, Ts, static_cast<__half*>(y), o, f, stream)
                 : fact_launch(x, tbl, batch, Ts, static_cast<float*>(y), o, f, stream);
}
// usable when whole float4s of every band are in or out of range and aligned
bool pqmf_fact_usable(const float* x, int64_t Ts) {
  return (Ts & 3) == 0 && Ts < (1 << 29) && (reinterpret_cast<uintptr_t>(x) & 15) == 0;
}

cudaError_t pqmf_synthesis_launch(const float* x, const float* G, int batch, int bands, int taps,
                                  int64_t Ts, void* y, int out_f16, const PqmfCrop* crop, cudaStream_t stream) {
  const int64_t Tout = (int64_t)bands * Ts + 2 * (taps / 2) - taps;
  if (batch <= 0 || Ts <= 0 || Tout <= 0) return cudaSuccess;
  const PqmfOut o = make_out(crop, Tout);
  return out_f16 ? direct_launch(x, G, batch, bands, taps, Ts, static_cast<__half*>(y), o, stream)
                 : direct_launch(x, G, batch, bands, taps, Ts, static_cast<float*>(y), o, stream);
}

cudaError_t f32_to_f16_launch(const float* x, void* y, int64_t n, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  f32_to_f16_kernel<<<ceil_div(ceil_div(n, 4), 256), 256, 0, stream>>>(x, static_cast<__half*>(y), n);
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
