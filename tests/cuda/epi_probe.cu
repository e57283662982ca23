// Probe for the costs that bound the fused-stack epilogues on the narrow stages (run on a B200):
//   1. tcgen05.ld / tcgen05.st throughput per SM with 4 / 8 / 16 warps (one or two CTAs per SM): is a tile's
//      accumulator read-out (128 rows x C fp32 columns, twice per block) TMEM-port bound?
//   2. single-warp latency of ld(x32)+wait, st(x16)+wait, fence.proxy.async after 4 STS.128, mbarrier arrive->wake;
//   3. the hand-off round trip: 128 threads arrive -> issuer wakes -> one MMA (N = 32) -> commit -> waiters wake.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o epi_probe tests/cuda/epi_probe.cu
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../tts_b200/csrc/tc_common.cuh"

using namespace ttsb::tc;

__device__ __forceinline__ void tmem_ld32_nw(uint32_t taddr, uint32_t* r) {
  tmem_ld16_nw(taddr, r);
  tmem_ld16_nw(taddr + 16, r + 16);
}

// mode 0: ld x16 (wait every `batch` loads); 1: st x8 ; 2: ld x32 as one instruction
__global__ void __launch_bounds__(512, 1) tmem_rate_kernel(int mode, int iters, int batch, long long* cycles, uint32_t* sink) {
  __shared__ uint32_t tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (warp == 0) tmem_alloc(&tmem_base, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_base;
  const uint32_t tlane = tbase + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 32;
  uint32_t acc = 0;
  uint32_t v[32];
  for (int i = 0; i < 32; ++i) v[i] = tid + i;
  tmem_st<32>(tlane, v);
  tmem_wait_st();
  __syncthreads();
  const long long t0 = clock64();
  if (mode == 0) {
    for (int i = 0; i < iters; i += batch) {
      for (int j = 0; j < batch; ++j) {
        tmem_ld16_nw(tlane + (j & 1) * 16, v + (j & 1) * 16);
      }
      tmem_wait_ld();
      acc += v[0] + v[16];
    }
  } else if (mode == 1) {
    for (int i = 0; i < iters; i += batch) {
      for (int j = 0; j < batch; ++j) tmem_st8(tlane + (j & 3) * 8, v + (j & 3) * 8);
      tmem_wait_st();
    }
  } else {
    for (int i = 0; i < iters; i += batch) {
      for (int j = 0; j < batch; ++j) {
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
            "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
              "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
              "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
              "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
            : "r"(tlane));
      }
      tmem_wait_ld();
      acc += v[0] + v[31];
    }
  }
  const long long t1 = clock64();
  __syncthreads();
  if (tid == 0) cycles[blockIdx.x] = t1 - t0;
  if (acc == 0x12345678u) sink[0] = acc;
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 256);
}

// Single-warp latencies, written as cycle deltas: [0] ld x32 + wait, [1] st x16 + wait, [2] 4 STS.128 + fence.proxy.async,
// [3] fence.proxy.async alone, [4] tcgen05 fence pair, [5] mbarrier arrive (count 1) + try_wait already complete
__global__ void latency_kernel(long long* out, uint32_t* sink) {
  __shared__ uint32_t tmem_base;
  __shared__ __align__(16) uint4 buf[4 * 128];
  __shared__ uint64_t bar;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (tid == 0) {
    mbar_init(&bar, blockDim.x);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(&tmem_base, 128);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tlane = tmem_base + ((uint32_t)(warp * 32) << 16);
  uint32_t v[32];
  for (int i = 0; i < 32; ++i) v[i] = tid * 3 + i;
  tmem_st<32>(tlane, v);
  tmem_wait_st();
  long long t[8];
  uint32_t acc = 0;
  for (int rep = 0; rep < 3; ++rep) {
    __syncwarp();
    t[0] = clock64();
    tmem_ld32_nw(tlane, v);
    tmem_wait_ld();
    acc += v[5];
    t[1] = clock64();
    tmem_st<16>(tlane + 32, v);
    tmem_wait_st();
    t[2] = clock64();
    for (int i = 0; i < 4; ++i) buf[i * 128 + tid] = make_uint4(v[i], v[i + 1], v[i + 2], v[i + 3]);
    fence_proxy_async();
    t[3] = clock64();
    fence_proxy_async();
    t[4] = clock64();
    tc_fence_before();
    tc_fence_after();
    t[5] = clock64();
    mbar_arrive(&bar);
    mbar_wait(&bar, rep & 1);
    t[6] = clock64();
  }
  if (tid == 0)
    for (int i = 0; i < 6; ++i) out[i] = t[i + 1] - t[i];
  if (acc == 0x12345678u) sink[0] = acc;
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, 128);
}

// Hand-off round trip: 4 epilogue warps arrive on `go`; warp 4 (issuer) waits, issues `nmma` MMAs of N = 32 and commits
// to `done`; the epilogue warps wait on `done`.  Reports the mean cycles per round seen by thread 0 (arrive -> wake).
__global__ void roundtrip_kernel(int rounds, int nmma, int ts, long long* out) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t go, done;
  __shared__ uint32_t tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < (8 * 160 * 16 + 4 * 32 * 16) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (tid == 0) {
    mbar_init(&go, 128);
    mbar_init(&done, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(&tmem_base, 128);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_base;
  if (warp == 4) {
    const bool leader = elect_one() != 0;
    const uint32_t idesc = make_idesc_bf16(128, 32);
    const uint32_t sA = smem_u32(smem), sB = sA + 8 * 160 * 16;
    const uint64_t ad = make_desc(sA, 160 * 16, 128), bd = make_desc(sB, 32 * 16, 128);
    for (int r = 0; r < rounds; ++r) {
      mbar_wait(&go, r & 1);
      tc_fence_after();
      if (leader) {
        for (int k = 0; k < nmma; ++k) {
          if (ts) umma_bf16_ts(tbase, tbase + 64 + 8 * (k & 3), bd, idesc, 1);
          else umma_bf16(tbase, ad + (k & 3), bd, idesc, 1);
        }
        umma_commit(&done);
      }
      __syncwarp();
    }
  } else {
    const long long t0 = clock64();
    for (int r = 0; r < rounds; ++r) {
      tc_fence_before();
      mbar_arrive(&go);
      mbar_wait(&done, r & 1);
      tc_fence_after();
    }
    const long long t1 = clock64();
    if (tid == 0) out[blockIdx.x] = (t1 - t0) / rounds;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 128);
}

int main() {
  setvbuf(stdout, NULL, _IOLBF, 0);
  long long* dC;
  uint32_t* dS;
  cudaMalloc(&dC, 8 * 1024);
  cudaMalloc(&dS, 64);
  auto run_rate = [&](int mode, int warps, int grid, int batch) {
    const int iters = 4096;
    tmem_rate_kernel<<<grid, warps * 32>>>(mode, iters, batch, dC, dS);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
      printf("CUDA error: %s\n", cudaGetErrorString(e));
      exit(1);
    }
    std::vector<long long> h(grid);
    cudaMemcpy(h.data(), dC, 8 * grid, cudaMemcpyDeviceToHost);
    long long mx = 0;
    for (auto c : h) mx = c > mx ? c : mx;
    const double per_cta_bytes = (double)warps * iters * 32 * (mode == 1 ? 8 : mode == 2 ? 32 : 16) * 4;
    const int ctas_per_sm = grid > 148 ? 2 : 1;
    printf("TMEM %s warps/CTA=%2d CTAs/SM=%d batch=%d: %.1f cycles/instr/warp, %.0f B/cycle/SM\n",
           mode == 0 ? "ld.x16" : mode == 1 ? "st.x8 " : "ld.x32", warps, ctas_per_sm, batch, (double)mx / iters,
           per_cta_bytes * ctas_per_sm / mx);
  };
  for (int mode : {0, 2, 1})
    for (int warps : {1, 4, 8, 16})
      for (int batch : {1, 4}) run_rate(mode, warps, 148, batch);
  run_rate(0, 8, 296, 2);
  run_rate(0, 16, 296, 2);
  run_rate(1, 8, 296, 4);

  latency_kernel<<<1, 32>>>(dC, dS);
  cudaDeviceSynchronize();
  long long h[8];
  cudaMemcpy(h, dC, 8 * 6, cudaMemcpyDeviceToHost);
  printf("LAT 1 warp : ld.x32+wait %lld | st.x16+wait %lld | 4xSTS.128+fence.proxy.async %lld | fence.proxy.async %lld | "
         "tcgen05 fence pair %lld | arrive+try_wait(done) %lld\n", h[0], h[1], h[2], h[3], h[4], h[5]);
  latency_kernel<<<148, 128>>>(dC, dS);
  cudaDeviceSynchronize();
  cudaMemcpy(h, dC, 8 * 6, cudaMemcpyDeviceToHost);
  printf("LAT 4 warps: ld.x32+wait %lld | st.x16+wait %lld | 4xSTS.128+fence.proxy.async %lld | fence.proxy.async %lld | "
         "tcgen05 fence pair %lld\n", h[0], h[1], h[2], h[3], h[4]);

  cudaFuncSetAttribute(roundtrip_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  for (int ts : {0, 1})
    for (int nmma : {0, 1, 2, 6}) {
      roundtrip_kernel<<<148, 160, 8 * 160 * 16 + 4 * 32 * 16>>>(2000, nmma, ts, dC);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) {
        printf("CUDA error: %s\n", cudaGetErrorString(e));
        return 1;
      }
      cudaMemcpy(h, dC, 8, cudaMemcpyDeviceToHost);
      printf("ROUNDTRIP %s nmma=%d (N=32): %lld cycles arrive -> issuer -> MMA -> commit -> wake\n", ts ? "TS" : "SS", nmma,
             h[0]);
    }
  printf("PROBE DONE\n");
  return 0;
}
