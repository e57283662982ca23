// Probe: how fast can one SM pull L2-resident bytes into shared memory with cp.async.bulk while every SM does
// the same?  This is what bounds the kernels that stream weights per tile (ring mode of the fused stack kernel).
//   mode 0: every CTA streams the SAME region (weights shared by all tiles)
//   mode 1: every CTA streams its own region (activations)
//   mode 2: clusters of 2, each CTA loads half of every chunk and multicasts it to both (same region)
// Sweeps chunk size and ring depth.  Reports bytes / cycle / SM from the device-wide elapsed time.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o ingest_probe tests/cuda/ingest_probe.cu
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../tts_b200/csrc/tc_common.cuh"

using namespace ttsb::tc;

constexpr int MAX_SLOTS = 16;

// wait flavours: 0 = try_wait with a suspend-time hint (tc_common.cuh mbar_wait), 1 = try_wait without a hint,
// 2 = test_wait spin
__device__ __forceinline__ void wait_nohint(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tW1_LOOP:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra W1_DONE;\n\t"
      "bra W1_LOOP;\n\tW1_DONE:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void wait_test(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tW2_LOOP:\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra W2_DONE;\n\t"
      "bra W2_LOOP;\n\tW2_DONE:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void wait_any(int flavour, uint64_t* bar, uint32_t parity) {
  if (flavour == 0) mbar_wait(bar, parity);
  else if (flavour == 1) wait_nohint(bar, parity);
  else wait_test(bar, parity);
}

// cycles per wait on a barrier whose phase has ALREADY completed
__global__ void wait_cost_kernel(int flavour, int iters, long long* cycles) {
  __shared__ __align__(8) uint64_t bar;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
    mbar_arrive(&bar);
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) wait_any(flavour, &bar, 0);
    *cycles = clock64() - t0;
  }
}
// ping-pong between two warps through two barriers: cycles per round trip (two hand-offs)
__global__ void pingpong_kernel(int flavour, int iters, long long* cycles) {
  __shared__ __align__(8) uint64_t a, b;
  if (threadIdx.x == 0) {
    mbar_init(&a, 1);
    mbar_init(&b, 1);
    fence_mbar_init();
  }
  __syncthreads();
  const long long t0 = clock64();
  if (threadIdx.x == 0) {
    for (int i = 0; i < iters; ++i) {
      mbar_arrive(&a);
      wait_any(flavour, &b, i & 1);
    }
    *cycles = clock64() - t0;
  } else if (threadIdx.x == 32) {
    for (int i = 0; i < iters; ++i) {
      wait_any(flavour, &a, i & 1);
      mbar_arrive(&b);
    }
  }
}

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_remote(uint64_t* bar, uint32_t cta) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(bar)), "r"(cta));
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}
__device__ __forceinline__ void bulk_g2s_mc(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;" ::"r"(
          smem_u32(dst_smem)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar)), "h"(mask)
      : "memory");
}

// region_bytes: bytes of the streamed region (wraps); n_chunks: chunks each CTA pulls
__global__ void ingest_kernel(const uint8_t* src, size_t region_bytes, size_t cta_stride, uint32_t chunk, int slots,
                              int n_chunks, int split, int flavour, long long* cycles) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t full[MAX_SLOTS], empty[MAX_SLOTS];
  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int i = 0; i < MAX_SLOTS; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    fence_mbar_init();
  }
  __syncthreads();
  const uint8_t* base = src + (size_t)blockIdx.x * cta_stride;
  const long long t0 = clock64();
  if (tid == 0) {
    size_t off = 0;
    const uint32_t piece = chunk / split;
    for (int i = 0; i < n_chunks; ++i) {
      const int s = i % slots;
      if (i >= slots) wait_any(flavour, &empty[s], ((i / slots) - 1) & 1);
      mbar_expect_tx(&full[s], chunk);
      for (int k = 0; k < split; ++k) bulk_g2s(smem + (size_t)s * chunk + k * piece, base + off + k * piece, piece, &full[s]);
      off += chunk;
      if (off + chunk > region_bytes) off = 0;
    }
  } else if (tid == 32) {
    for (int i = 0; i < n_chunks; ++i) {
      const int s = i % slots;
      wait_any(flavour, &full[s], (i / slots) & 1);
      mbar_arrive(&empty[s]);
    }
  }
  __syncthreads();
  if (tid == 0 && blockIdx.x == 0) *cycles = clock64() - t0;
}

// P producer warps (chunk i is issued by warp i % P), or one warp whose first `lanes` lanes each issue a piece.
__global__ void ingest_par_kernel(const uint8_t* src, size_t region_bytes, size_t cta_stride, uint32_t chunk, int slots,
                                  int n_chunks, int P, int lanes, long long* cycles) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t full[MAX_SLOTS], empty[MAX_SLOTS];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int i = 0; i < MAX_SLOTS; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    fence_mbar_init();
  }
  __syncthreads();
  const uint8_t* base = src + (size_t)blockIdx.x * cta_stride;
  const long long t0 = clock64();
  const int chunks_in_region = (int)(region_bytes / chunk);
  if (warp < P) {
    const uint32_t piece = chunk / lanes;
    for (int i = warp; i < n_chunks; i += P) {
      const int s = i % slots;
      if (i >= slots) wait_nohint(&empty[s], ((i / slots) - 1) & 1);
      const size_t off = (size_t)(i % chunks_in_region) * chunk;
      if (lane == 0) mbar_expect_tx(&full[s], chunk);
      __syncwarp();
      if (lane < lanes) bulk_g2s(smem + (size_t)s * chunk + lane * piece, base + off + lane * piece, piece, &full[s]);
      __syncwarp();
    }
  } else if (tid == P * 32) {
    for (int i = 0; i < n_chunks; ++i) {
      const int s = i % slots;
      wait_nohint(&full[s], (i / slots) & 1);
      mbar_arrive(&empty[s]);
    }
  }
  __syncthreads();
  if (tid == 0 && blockIdx.x == 0) *cycles = clock64() - t0;
}

int main() {
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  int khz = 0;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const size_t pool = (size_t)sms * 2 * (1 << 20);
  uint8_t* src;
  cudaMalloc(&src, pool);
  cudaMemset(src, 1, pool);
  long long* cyc;
  cudaMalloc(&cyc, 8);
  cudaFuncSetAttribute(ingest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  printf("sms %d clock %d kHz\n", sms, khz);
  printf("%-10s %8s %6s %6s %5s %10s %10s %10s\n", "mode", "chunk", "slots", "split", "cta/sm", "us", "B/cyc/SM", "TB/s");
  const size_t total_per_cta = 24u << 20;
  auto run = [&](int mode, uint32_t chunk, int slots, int split, int ctas_per_sm, int flavour) {
    const int n_chunks = (int)(total_per_cta / ctas_per_sm / chunk);
    const size_t smem = (size_t)chunk * slots;
    if (smem * ctas_per_sm > 200 * 1024 || slots > MAX_SLOTS) return;
    const int grid = sms * ctas_per_sm;
    const size_t region = mode == 1 ? (1u << 20) : (1179648u);   // 1.1 MB of weights, or 1 MB per CTA
    const size_t stride = mode == 1 ? (1u << 20) : 0;
    float best = 1e9f;
    long long hc = 0;
    for (int rep = 0; rep < 3; ++rep) {
      cudaEventRecord(e0);
      ingest_kernel<<<grid, 64, smem>>>(src, region, stride, chunk, slots, n_chunks, split, flavour, cyc);
      cudaEventRecord(e1);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) {
        printf("error %s\n", cudaGetErrorString(e));
        exit(1);
      }
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      if (ms < best) {
        best = ms;
        cudaMemcpy(&hc, cyc, 8, cudaMemcpyDeviceToHost);
      }
    }
    const double bytes_per_sm = (double)n_chunks * chunk * ctas_per_sm;
    printf("%-10s w%d %8u %6d %6d %5d %10.1f %10.2f %10.2f\n", mode == 0 ? "same" : "distinct", flavour, chunk,
           slots, split, ctas_per_sm, best * 1e3, bytes_per_sm / (double)hc, bytes_per_sm * sms / (best * 1e-3) / 1e12);
    fflush(stdout);
  };
  cudaFuncSetAttribute(ingest_par_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  // region: 0 = the same 1.1 MB for all CTAs, 1 = 1 MB per CTA (148 MB, HBM), 2 = 128 KB per CTA (19 MB, L2-resident)
  auto runp = [&](int mode, uint32_t chunk, int slots, int P, int lanes) {
    const int n_chunks = (int)(total_per_cta / chunk);
    const size_t smem = (size_t)chunk * slots;
    if (smem > 200 * 1024 || slots > MAX_SLOTS) return;
    const size_t region = mode == 0 ? 1179648u : mode == 1 ? (1u << 20) : (1u << 17);
    const size_t stride = mode == 0 ? 0 : region;
    float best = 1e9f;
    long long hc = 0;
    for (int rep = 0; rep < 3; ++rep) {
      cudaEventRecord(e0);
      ingest_par_kernel<<<sms, (P + 1) * 32, smem>>>(src, region, stride, chunk, slots, n_chunks, P, lanes, cyc);
      cudaEventRecord(e1);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) {
        printf("error %s\n", cudaGetErrorString(e));
        exit(1);
      }
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      if (ms < best) {
        best = ms;
        cudaMemcpy(&hc, cyc, 8, cudaMemcpyDeviceToHost);
      }
    }
    const double bytes_per_sm = (double)n_chunks * chunk;
    printf("par %-8s chunk %6u slots %2d warps %d lanes %2d  %8.1f us  %7.2f B/cyc/SM  %6.2f TB/s\n",
           mode == 0 ? "same" : mode == 1 ? "hbm" : "l2", chunk, slots, P, lanes, best * 1e3, bytes_per_sm / (double)hc,
           bytes_per_sm * sms / (best * 1e-3) / 1e12);
    fflush(stdout);
  };
  for (int mode : {0, 2}) {
    for (uint32_t chunk : {4096u, 12288u, 36864u})
      for (int P : {1, 2, 4}) runp(mode, chunk, chunk > 20000 ? 4 : 8, P, 1);
    for (int lanes : {2, 4, 8, 16}) runp(mode, 12288, 8, 1, lanes);
    for (int lanes : {4, 12}) runp(mode, 36864, 4, 1, lanes);
    runp(mode, 73728, 2, 1, 1);
    runp(mode, 98304, 2, 1, 1);
    runp(mode, 98304, 2, 2, 1);
    runp(mode, 49152, 4, 1, 12);   // an activation tile as 12 plane copies issued by 12 lanes
    runp(mode, 49152, 4, 2, 12);
  }
  return 0;
}
