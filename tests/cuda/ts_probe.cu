// Probe for the "A operand in TMEM" form of tcgen05.mma used by the fused ResidualStack v2 kernel
// (run on a B200):
//   1. correctness: A [128 x K] bf16 written to TMEM with tcgen05.st 32x32b (lane = row, one 32-bit column =
//      two consecutive K elements), B [N x K] in shared memory (canonical K-major), D = A B^T;
//   2. in-place reuse: the bf16 A operand is written over the low columns of a consumed fp32 accumulator;
//   3. throughput: cycles per MMA for SS (A in smem) vs TS (A in TMEM) at N = 32..256, back to back.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o ts_probe tests/cuda/ts_probe.cu
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../tts_b200/csrc/tc_common.cuh"

using namespace ttsb::tc;

// A: [128][K] bf16 row-major (global), B: [K/8][N][8] bf16, D: [128][N] fp32.
// a_col: TMEM column where the bf16 A operand is placed; d_col: accumulator column.
__global__ void ts_kernel(const __nv_bfloat16* A, const __nv_bfloat16* B, float* D, int N, int K, int a_col,
                          int d_col) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar_load, bar_mma;
  __shared__ uint32_t tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t bytesB = (K / 8) * N * 16;
  if (tid == 0) {
    mbar_init(&bar_load, 1);
    mbar_init(&bar_mma, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(&tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_base;
  const uint32_t tlane = tbase + ((uint32_t)(warp * 32) << 16);
  // every thread stores its row of A: K/2 columns, 8 at a time
  const int row = warp * 32 + lane;
  for (int c = 0; c < K / 2; c += 8) {
    uint32_t r[8];
    for (int i = 0; i < 8; ++i) r[i] = *reinterpret_cast<const uint32_t*>(A + (size_t)row * K + 2 * (c + i));
    tmem_st8(tlane + a_col + c, r);
  }
  tmem_wait_st();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (tid == 0) {
    mbar_expect_tx(&bar_load, bytesB);
    bulk_g2s(smem, B, bytesB, &bar_load);
    mbar_wait(&bar_load, 0);
    tc_fence_after();
    const uint32_t idesc = make_idesc_bf16(128, N);
    const uint32_t lboB = N * 16;
    for (int k = 0; k < K / 16; ++k)
      umma_bf16_ts(tbase + d_col, tbase + a_col + 8 * k, make_desc(smem_u32(smem) + 2 * k * lboB, lboB, 128), idesc,
                   k > 0);
    umma_commit(&bar_mma);
  }
  __syncwarp();
  mbar_wait(&bar_mma, 0);
  tc_fence_after();
  for (int c = 0; c < N; c += 16) {
    float v[16];
    tmem_ld16(tlane + d_col + c, v);
    for (int i = 0; i < 16; ++i)
      if (c + i < N) D[(size_t)row * N + c + i] = v[i];   // N = 24: M = 128 with N a multiple of 8, not 16
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 512);
}

// Throughput: `iters` MMAs of [128 x N x 16] issued back to back by one thread; A from smem (mode 0) or TMEM (1).
__global__ void rate_kernel(int N, int mode, int iters, int rowsA, long long* cycles, int nacc) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar_mma;
  __shared__ uint32_t tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < (2 * rowsA * 16 + 2 * 256 * 16) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (tid == 0) {
    mbar_init(&bar_mma, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(&tmem_base, 512);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_base;
  if (warp == 0 && elect_one()) {
    const uint32_t idesc = make_idesc_bf16(128, N);
    const uint32_t lboA = rowsA * 16, lboB = N * 16;
    const uint32_t sA = smem_u32(smem), sB = sA + 2 * rowsA * 16;
    const uint64_t ad = make_desc(sA, lboA, 128), bd = make_desc(sB, lboB, 128);
    const long long t0 = clock64();
    // 4 MMAs per asm block, descriptors precomputed: the issue loop is a handful of instructions per MMA
    const uint32_t d0 = tbase, d1 = tbase + (nacc > 1 ? N : 0);
    const uint64_t a0 = ad, a1 = ad + 1, a2 = ad + 2, a3 = ad + 3;
    const uint32_t t0a = tbase + 448, t1a = tbase + 456, t2a = tbase + 464, t3a = tbase + 472;
    for (int i = 0; i < iters; i += 4) {
      if (mode == 0)
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %8, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], %2, %6, %7, p;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%1], %3, %6, %7, p;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], %4, %6, %7, p;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%1], %5, %6, %7, p;\n\t}" ::"r"(d0), "r"(d1), "l"(a0), "l"(a1),
            "l"(a2), "l"(a3), "l"(bd), "r"(idesc), "r"(1)
            : "memory");
      else
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %8, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], [%2], %6, %7, p;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%1], [%3], %6, %7, p;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], [%4], %6, %7, p;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%1], [%5], %6, %7, p;\n\t}" ::"r"(d0), "r"(d1), "r"(t0a), "r"(t1a),
            "r"(t2a), "r"(t3a), "l"(bd), "r"(idesc), "r"(1)
            : "memory");
    }
    umma_commit(&bar_mma);
    mbar_wait(&bar_mma, 0);
    const long long t1 = clock64();
    cycles[blockIdx.x] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 512);
}

static float bf16r(float x) { return __bfloat162float(__float2bfloat16(x)); }

int main() {
  int fails = 0;
  const int cfgs[][4] = {{96, 96, 0, 256}, {96, 96, 96, 0}, {48, 48, 200, 0}, {32, 32, 0, 480}, {192, 192, 256, 0},
                         {96, 16, 96, 0}, {256, 64, 256, 0},
                         {24, 32, 64, 0}, {40, 48, 200, 0}};
  for (auto& c : cfgs) {
    const int N = c[0], K = c[1], a_col = c[2], d_col = c[3];
    std::vector<__nv_bfloat16> hA(128 * K), hB(K * N);
    std::vector<float> fA(128 * K), fB(N * K);
    srand(N * 131 + K);
    for (int m = 0; m < 128; ++m)
      for (int k = 0; k < K; ++k) {
        float v = bf16r((rand() % 2001 - 1000) / 1000.f);
        fA[m * K + k] = v;
        hA[m * K + k] = __float2bfloat16(v);
      }
    for (int n = 0; n < N; ++n)
      for (int k = 0; k < K; ++k) {
        float v = bf16r((rand() % 2001 - 1000) / 1000.f);
        fB[n * K + k] = v;
        hB[((size_t)(k / 8) * N + n) * 8 + k % 8] = __float2bfloat16(v);
      }
    __nv_bfloat16 *dA, *dB;
    float* dD;
    cudaMalloc(&dA, hA.size() * 2);
    cudaMalloc(&dB, hB.size() * 2);
    cudaMalloc(&dD, 128 * N * 4);
    cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(ts_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    ts_kernel<<<1, 128, (K / 8) * N * 16>>>(dA, dB, dD, N, K, a_col, d_col);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
      printf("CUDA error: %s\n", cudaGetErrorString(e));
      return 1;
    }
    std::vector<float> hD(128 * N);
    cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost);
    double mx = 0;
    for (int m = 0; m < 128; ++m)
      for (int n = 0; n < N; ++n) {
        double ref = 0;
        for (int k = 0; k < K; ++k) ref += (double)fA[m * K + k] * fB[n * K + k];
        mx = fmax(mx, fabs(ref - hD[m * N + n]));
      }
    const bool ok = mx < 1e-3;
    printf("TS N=%3d K=%3d a_col=%3d d_col=%3d: max abs err %.3e %s\n", N, K, a_col, d_col, mx, ok ? "OK" : "FAIL");
    fails += !ok;
    cudaFree(dA); cudaFree(dB); cudaFree(dD);
  }
  long long* dC;
  cudaMalloc(&dC, 8 * 148);
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int Ns[] = {16, 24, 32, 48, 96, 128, 192, 256};
  for (int mode = 0; mode < 2; ++mode)
    for (int N : Ns)
      for (int nacc : {1, 2}) {
        const int grid = 148;
        if (nacc * N > 448 && nacc > 1) continue;
        const int iters = 2000, rowsA = 160;
        rate_kernel<<<grid, 128, 2 * rowsA * 16 + 2 * 256 * 16>>>(N, mode, iters, rowsA, dC, nacc);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) {
          printf("CUDA error: %s\n", cudaGetErrorString(e));
          return 1;
        }
        long long h[148];
        cudaMemcpy(h, dC, 8 * grid, cudaMemcpyDeviceToHost);
        long long mxc = 0;
        for (int i = 0; i < grid; ++i) mxc = h[i] > mxc ? h[i] : mxc;
        printf("RATE %s N=%3d nacc=%d: %.1f cycles/MMA (floor %d)\n", mode ? "TS" : "SS", N, nacc,
               (double)mxc / iters, N / 2);
      }
  printf(fails ? "PROBE FAILED\n" : "PROBE PASSED\n");
  return fails != 0;
}
