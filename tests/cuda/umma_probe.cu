// Standalone probe that pins the tcgen05 conventions the BF16 path relies on (run on a B200):
//   * canonical K-major no-swizzle operand layout [K/8][rows][8] with LBO = plane stride, SBO = 128 B
//   * a row-shifted A start address (dilated-conv tap) gives D[m] = sum_k A[m + shift][k] B[n][k]
//   * TMEM accumulator readback with tcgen05.ld 32x32b (lane = row, column = n)
//   * several N (16..256) and K sizes, accumulate flag across K steps
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o umma_probe tests/cuda/umma_probe.cu
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../tts_b200/csrc/tc_common.cuh"

using namespace ttsb::tc;

// A: [K/8][rowsA][8] bf16, B: [K/8][N][8] bf16, D: [128][N] fp32.  swap=1 swaps the LBO/SBO fields.
__global__ void probe_kernel(const __nv_bfloat16* A, const __nv_bfloat16* B, float* D, int rowsA, int N, int K,
                             int shift, int swap) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar_load, bar_mma;
  __shared__ uint32_t tmem_base;
  uint8_t* sA = smem;
  uint8_t* sB = smem + (size_t)(K / 8) * rowsA * 16;
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t bytesA = (K / 8) * rowsA * 16, bytesB = (K / 8) * N * 16;
  if (tid == 0) {
    mbar_init(&bar_load, 1);
    mbar_init(&bar_mma, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(&tmem_base, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_base;
  if (tid == 0) {
    mbar_expect_tx(&bar_load, bytesA + bytesB);
    bulk_g2s(sA, A, bytesA, &bar_load);
    bulk_g2s(sB, B, bytesB, &bar_load);
    mbar_wait(&bar_load, 0);
    tc_fence_after();
    const uint32_t idesc = make_idesc_bf16(128, N);
    const uint32_t lboA = rowsA * 16, lboB = N * 16, sbo = 128;
    for (int k = 0; k < K / 16; ++k) {
      const uint32_t a_addr = smem_u32(sA) + 2 * k * lboA + shift * 16;
      const uint32_t b_addr = smem_u32(sB) + 2 * k * lboB;
      const uint64_t ad = swap ? make_desc(a_addr, sbo, lboA) : make_desc(a_addr, lboA, sbo);
      const uint64_t bd = swap ? make_desc(b_addr, sbo, lboB) : make_desc(b_addr, lboB, sbo);
      umma_bf16(tbase, ad, bd, idesc, k > 0);
    }
    umma_commit(&bar_mma);
  }
  __syncwarp();
  mbar_wait(&bar_mma, 0);
  tc_fence_after();
  // 4 warps read the 4 lane quarters
  const int row = warp * 32 + (tid & 31);
  for (int c = 0; c < N; c += 16) {
    float v[16];
    tmem_ld16(tbase + ((uint32_t)(warp * 32) << 16) + c, v);
    for (int i = 0; i < 16; ++i) D[(size_t)row * N + c + i] = v[i];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 256);
}

static float bf(float x) { return __bfloat162float(__float2bfloat16(x)); }

int run(int N, int K, int shift, int swap) {
  const int rowsA = 128 + 32;
  std::vector<float> a((size_t)rowsA * K), b((size_t)N * K);
  for (auto& v : a) v = bf((rand() % 2001 - 1000) / 1000.f);
  for (auto& v : b) v = bf((rand() % 2001 - 1000) / 1000.f);
  std::vector<__nv_bfloat16> ha((size_t)rowsA * K), hb((size_t)N * K);
  for (int r = 0; r < rowsA; ++r)
    for (int k = 0; k < K; ++k) ha[((size_t)(k / 8) * rowsA + r) * 8 + k % 8] = __float2bfloat16(a[(size_t)r * K + k]);
  for (int n = 0; n < N; ++n)
    for (int k = 0; k < K; ++k) hb[((size_t)(k / 8) * N + n) * 8 + k % 8] = __float2bfloat16(b[(size_t)n * K + k]);
  __nv_bfloat16 *dA, *dB;
  float* dD;
  cudaMalloc(&dA, ha.size() * 2);
  cudaMalloc(&dB, hb.size() * 2);
  cudaMalloc(&dD, (size_t)128 * N * 4);
  cudaMemcpy(dA, ha.data(), ha.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hb.data(), hb.size() * 2, cudaMemcpyHostToDevice);
  cudaMemset(dD, 0xff, (size_t)128 * N * 4);
  const size_t smem = ha.size() * 2 + hb.size() * 2;
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  probe_kernel<<<1, 128, smem>>>(dA, dB, dD, rowsA, N, K, shift, swap);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("N=%d K=%d shift=%d swap=%d: CUDA error %s\n", N, K, shift, swap, cudaGetErrorString(e));
    return 2;
  }
  std::vector<float> d((size_t)128 * N);
  cudaMemcpy(d.data(), dD, d.size() * 4, cudaMemcpyDeviceToHost);
  double maxerr = 0;
  for (int m = 0; m < 128; ++m)
    for (int n = 0; n < N; ++n) {
      double ref = 0;
      for (int k = 0; k < K; ++k) ref += (double)a[(size_t)(m + shift) * K + k] * b[(size_t)n * K + k];
      double err = fabs(ref - d[(size_t)m * N + n]);
      if (!(err <= maxerr)) maxerr = err;
    }
  printf("N=%3d K=%3d shift=%2d swap=%d: max abs err %.3e %s\n", N, K, shift, swap, maxerr,
         maxerr < 1e-3 ? "OK" : "MISMATCH");
  cudaFree(dA); cudaFree(dB); cudaFree(dD);
  return maxerr < 1e-3 ? 0 : 1;
}

int main() {
  int bad = 0;
  bad += run(64, 64, 0, 0) != 0;
  const int Ns[] = {16, 32, 48, 96, 192, 256};
  for (int N : Ns) bad += run(N, 96, 0, 0) != 0;
  const int shifts[] = {1, 2, 3, 8, 9, 18, 27};
  for (int s : shifts) bad += run(96, 48, s, 0) != 0;
  bad += run(96, 384, 5, 0) != 0;
  printf(bad ? "PROBE FAILED (%d)\n" : "PROBE PASSED\n", bad);
  return bad ? 1 : 0;
}
