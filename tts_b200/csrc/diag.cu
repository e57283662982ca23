// Diagnostics behind the bench's roofline denominators that MEASURED_PEAKS.json does not carry.
//
// ttsb200_diag_ffma_tflops: the fp32 path (conv_f32.cu) computes on the CUDA cores, so its roofline is the FFMA
// rate, measured here with a register-resident FMA loop on every SM (16 independent chains per thread, 1,024
// threads per SM) and CUDA events around it.  ttsb200_diag_hostlink_gbs: the pinned host<->device copy rate of
// this box, the ceiling of the end-to-end figure (the waveform has to reach host memory).
#include <algorithm>
#include <vector>

#include "common.cuh"

namespace ttsb {
namespace {

constexpr int FFMA_CHAINS = 16;

__global__ void __launch_bounds__(256) ffma_kernel(float* out, int iters, float b, float c) {
  float a[FFMA_CHAINS];
#pragma unroll
  for (int i = 0; i < FFMA_CHAINS; ++i) a[i] = (float)(threadIdx.x + i) * 1e-3f;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
#pragma unroll
      for (int i = 0; i < FFMA_CHAINS; ++i) a[i] = fmaf(a[i], b, c);
    }
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < FFMA_CHAINS; ++i) s += a[i];
  if (s == 123.456f) out[0] = s;   // keeps the chains alive
}

}  // namespace
}  // namespace ttsb

using namespace ttsb;

extern "C" {

double ttsb200_diag_ffma_tflops(int32_t device) {
  int prev = 0;
  if (cudaGetDevice(&prev) != cudaSuccess || cudaSetDevice(device) != cudaSuccess) return fail(TTSB200_ERR_CUDA, "cudaSetDevice");
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  float* d = nullptr;
  cudaMalloc(&d, 64);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int iters = 4096, grid = sms * 8, threads = 256;
  double best = 0.0;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(e0);
    ffma_kernel<<<grid, threads>>>(d, iters, 0.999f, 1e-3f);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) break;
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double flops = 2.0 * grid * threads * (double)iters * 4 * FFMA_CHAINS;
    if (rep > 0 && ms > 0.f) best = std::max(best, flops / (ms * 1e-3) / 1e12);
  }
  const cudaError_t err = cudaGetLastError();
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d);
  cudaSetDevice(prev);
  if (err != cudaSuccess) return fail(TTSB200_ERR_CUDA, std::string("ffma probe: ") + cudaGetErrorString(err));
  return best;
}

// direction 0: host -> device, 1: device -> host.  `streams` concurrent copies of `bytes` each from / into pinned
// host buffers; returns the aggregate GB/s (best of 5).
double ttsb200_diag_hostlink_gbs(int32_t device, int32_t direction, int64_t bytes, int32_t streams) {
  if (bytes <= 0 || streams <= 0 || streams > 16) return fail(TTSB200_ERR_INVALID, "bytes/streams out of range");
  int prev = 0;
  if (cudaGetDevice(&prev) != cudaSuccess || cudaSetDevice(device) != cudaSuccess) return fail(TTSB200_ERR_CUDA, "cudaSetDevice");
  std::vector<void*> h(streams, nullptr), d(streams, nullptr);
  std::vector<cudaStream_t> st(streams, nullptr);
  cudaError_t err = cudaSuccess;
  for (int i = 0; i < streams && err == cudaSuccess; ++i) {
    err = cudaMallocHost(&h[i], bytes);
    if (err == cudaSuccess) err = cudaMalloc(&d[i], bytes);
    if (err == cudaSuccess) err = cudaStreamCreateWithFlags(&st[i], cudaStreamNonBlocking);
  }
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  double best = 0.0;
  for (int rep = 0; rep < 6 && err == cudaSuccess; ++rep) {
    cudaDeviceSynchronize();
    cudaEventRecord(e0, st[0]);
    for (int i = 1; i < streams; ++i) cudaStreamWaitEvent(st[i], e0, 0);
    for (int i = 0; i < streams; ++i)
      cudaMemcpyAsync(direction ? h[i] : d[i], direction ? d[i] : h[i], bytes,
                      direction ? cudaMemcpyDeviceToHost : cudaMemcpyHostToDevice, st[i]);
    for (int i = 1; i < streams; ++i) {
      cudaEvent_t done;
      cudaEventCreateWithFlags(&done, cudaEventDisableTiming);
      cudaEventRecord(done, st[i]);
      cudaStreamWaitEvent(st[0], done, 0);
      cudaEventDestroy(done);
    }
    cudaEventRecord(e1, st[0]);
    err = cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep > 0 && ms > 0.f) best = std::max(best, (double)bytes * streams / (ms * 1e-3) / 1e9);
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  for (int i = 0; i < streams; ++i) {
    if (st[i]) cudaStreamDestroy(st[i]);
    if (d[i]) cudaFree(d[i]);
    if (h[i]) cudaFreeHost(h[i]);
  }
  cudaSetDevice(prev);
  if (err != cudaSuccess) return fail(TTSB200_ERR_CUDA, std::string("host link probe: ") + cudaGetErrorString(err));
  return best;
}

}  // extern "C"
