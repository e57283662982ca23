// Timeline capture for CTA 0 (debug aid).  Build with TTSB200_NVCC_DEFS=-DTTSB_STACK_TRACE and run with
// TTSB200_STACK_TRACE=<file>: role r appends {tag, clock64} pairs; the launcher dumps "id role tag clock" lines.
#pragma once

#include <cstdio>
#include <cstdlib>
#include <vector>

#include <cuda_runtime.h>

namespace ttsb {

#ifdef TTSB_STACK_TRACE
struct Tracer {
  long long* buf;
  int n;
  __device__ __forceinline__ void operator()(int tag) {
    if (buf && n < 254) {
      buf[n] = tag;
      buf[n + 1] = clock64();
      n += 2;
    }
  }
};
__device__ __forceinline__ Tracer make_tracer(long long* trace, int role, bool leader) {
  return Tracer{(trace && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && leader) ? trace + role * 256 : nullptr, 0};
}
#else   // compiled out: the trace points cost instruction-cache space
struct Tracer {
  __device__ __forceinline__ void operator()(int) {}
};
__device__ __forceinline__ Tracer make_tracer(long long*, int, bool) { return Tracer{}; }
#endif

constexpr int TRACE_ROLES = 40;

// Host side: allocate before the launch, dump + free after it.
inline long long* trace_begin(const char* file, cudaStream_t stream) {
  if (!file) return nullptr;
  long long* d = nullptr;
  cudaMalloc(&d, TRACE_ROLES * 256 * sizeof(long long));
  cudaMemsetAsync(d, 0, TRACE_ROLES * 256 * sizeof(long long), stream);
  return d;
}
inline void trace_end(const char* file, long long* d, int id, cudaStream_t stream) {
  if (!d) return;
  std::vector<long long> h(TRACE_ROLES * 256);
  cudaStreamSynchronize(stream);
  cudaMemcpy(h.data(), d, h.size() * sizeof(long long), cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (FILE* f = fopen(file, "a")) {
    for (int r = 0; r < TRACE_ROLES; ++r)
      for (int i = 0; i < 256 && h[r * 256 + i]; i += 2)
        fprintf(f, "%d %d %d %lld\n", id, r, (int)h[r * 256 + i], h[r * 256 + i + 1]);
    fclose(f);
  }
}

}  // namespace ttsb
