// Shared host/device helpers for the tts_b200 CUDA library (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "../../include/tts_b200.h"

namespace ttsb {

// ---- error plumbing (thread-local message behind ttsb200_last_error) -------------------------
void set_error(const std::string& msg);
int fail(int code, const std::string& msg);

#define TTSB_CUDA(expr)                                                                        \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess)                                                                     \
      return ::ttsb::fail(TTSB200_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
  } while (0)

// ---- programmatic dependent launch -------------------------------------------------------------------
// The kernels of one inference run back to back on one stream.  Launched with the programmatic-serialization
// attribute, kernel N+1 may start its preamble (barrier init, TMEM allocation, weight prefetch) while kernel N
// drains; it calls pdl_wait() before touching anything kernel N wrote (or still reads).  pdl_wait() returns when
// every prerequisite grid has completed and flushed, so chains of such kernels stay correctly ordered.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <class... KArgs, class... Args>
cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
#endif

static inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }
static inline int ceil_div(int64_t x, int64_t m) { return (int)((x + m - 1) / m); }

// ---- fp32 implicit-GEMM convolution (conv_f32.cu) ---------------------------------------------
// One launch computes, for every batch item b, GEMM-time index t in [0, Lt) and channel o:
//   out[b][o][t*out_stride + out_offset] =
//       epi( bias[o] + sum_{s<nsrc} sum_{j<taps_s} sum_{c<C_s} w_s[ph][j][c][o] * pro_s( x_s[b][c][pos_s(t,j)] ) )
// pos_s(t,j) = t + j*dil_s + shift_s (+ phase-dependent shift for transposed-conv phases) and
// is resolved by the edge rule: reflect (ReflectionPad1d), then optional replicate clamp
// (inference_padding), or zero-fill (transposed conv).
struct ConvSrc {
  const float* x;         // [B][C][Lreal]
  const float* w;         // [phases][taps][C][CoutPad]   (o contiguous)
  int64_t batch_stride;   // C * Lreal
  int C;
  int L;                  // virtual length seen by the reflect rule (= Lreal + 2*clamp_pad)
  int Lreal;              // stored row length
  int clamp_pad;          // replicate padding folded into the index map (0 = none)
  int taps;
  int dil;
  int shift;
  int edge_zero;          // 0: reflect, 1: zero-fill outside [0, L)
  int lrelu;              // apply LeakyReLU(slope) to the loaded value
};

struct ConvParams {
  ConvSrc src[2];
  int nsrc;
  const float* bias;      // [CoutPad]
  float* out;             // [B][Cout][out_L]
  int64_t out_batch_stride;
  int Cout;
  int CoutPad;
  int Lt;                 // GEMM-time extent
  int out_L;              // stored row length of out
  int out_stride;         // 1, or f for transposed-conv phases
  int n_phase;            // 1, or f
  int up_f, up_P;         // transposed conv: phase r -> shift (r+P)/f, out_offset r
  int tanh_out;
  float slope;
};

// cpt = tile code (conv_f32.cu): 1..8 = output channels per thread of a 16*cpt-channel x 128-position tile, 24 / 32 / 48 = a
// tile of exactly that many channels x 512 / 512 / 256 positions; 256 threads either way.
int conv_f32_pick_cpt(int Cout);
int conv_f32_tile_channels(int code);
cudaError_t conv_f32_launch(const ConvParams& p, int batch, int cpt, cudaStream_t stream);
// A centred 7-tap conv to <= 4 output channels (the output head) has a kernel of its own: conv_f32.cu head_f32_kernel.
bool head_f32_usable(const ConvParams& p);
cudaError_t head_f32_launch(const ConvParams& p, int batch, cudaStream_t stream);

// ---- PQMF (pqmf.cu) ----------------------------------------------------------------------------
// y: fp32, or fp16 when out_f16 (ttsb200_out_dtype).  crop (optional): only samples [lo, lo + len) of every row are
// stored, row b at y + b * batch_stride (elements) — the kept range of a long-form / streaming window goes straight
// to its place in the final buffer, which may be peer memory.
struct PqmfCrop {
  int64_t lo, len, batch_stride;
};
cudaError_t pqmf_synthesis_launch(const float* x, const float* G, int batch, int bands, int taps,
                                  int64_t Ts, void* y, int out_f16, const PqmfCrop* crop, cudaStream_t stream);
cudaError_t f32_to_f16_launch(const float* x, void* y, int64_t n, cudaStream_t stream);
// Factorised 4-band / 63-tap synthesis (24 instead of 63 MACs per sample): table built by pqmf_factor_4x63 at
// plan creation; false = G is not the reference's cosine-modulated bank, keep the direct kernel.
constexpr int PQMF_FACT_TABLE = 96;
bool pqmf_factor_4x63(const float* G_host, float* tbl_host);
bool pqmf_fact_usable(const float* x, int64_t Ts);
// follow (optional): how the kernel that writes x counts its finished work; the launch then follows that kernel
// segment by segment instead of waiting for its whole grid (follow.cuh).
struct TcFollow;
cudaError_t pqmf_synthesis_fact_launch(const float* x, const float* tbl, int batch, int64_t Ts, void* y, int out_f16,
                                       const PqmfCrop* crop, const TcFollow* follow, cudaStream_t stream);
cudaError_t pqmf_analysis_launch(const float* x, const float* H, int batch, int bands, int taps,
                                 int64_t T, float* y, cudaStream_t stream);

}  // namespace ttsb
