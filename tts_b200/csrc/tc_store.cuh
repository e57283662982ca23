// Epilogue store of 16 consecutive channels of one output row into a canonical bf16 tensor, plus the
// halo rows the consumer's padding rule needs (shared by the conv and fused-stack kernels).
#pragma once

#include "tc_common.cuh"
#include "tc_conv.cuh"

namespace ttsb {
namespace tc {

__device__ __forceinline__ void canon_store16(const TcCanonOut& o, int b, int ch, int tau, float* v, float slope) {
  if (o.lrelu) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = v[i] >= 0.f ? v[i] : v[i] * slope;
  }
  uint4 lo, hi;
  lo.x = pack_bf16x2(v[0], v[1]);   lo.y = pack_bf16x2(v[2], v[3]);
  lo.z = pack_bf16x2(v[4], v[5]);   lo.w = pack_bf16x2(v[6], v[7]);
  hi.x = pack_bf16x2(v[8], v[9]);   hi.y = pack_bf16x2(v[10], v[11]);
  hi.z = pack_bf16x2(v[12], v[13]); hi.w = pack_bf16x2(v[14], v[15]);
  __nv_bfloat16* ob = o.base + (int64_t)b * o.batch_stride;
  const int64_t pl0 = (int64_t)(ch >> 3) * o.plane_rows;
  const int64_t pl1 = pl0 + o.plane_rows;
  auto put = [&](int row, const uint4& a, const uint4& c) {
    *reinterpret_cast<uint4*>(ob + (pl0 + row) * 8) = a;
    *reinterpret_cast<uint4*>(ob + (pl1 + row) * 8) = c;
  };
  put(o.halo + tau, lo, hi);
  if (o.margin == 1) {          // halo rows mirror the interior (ReflectionPad1d of the consumer)
    if (tau >= 1 && tau <= o.halo) put(o.halo - tau, lo, hi);
    const int back = o.L - 1 - tau;
    if (back >= 1 && back <= o.halo) put(o.halo + o.L - 1 + back, lo, hi);
  } else if (o.margin == 2) {   // zero rows either side (transposed conv of the consumer)
    const uint4 z = make_uint4(0, 0, 0, 0);
    if (tau == 0) put(o.halo - 1, z, z);
    if (tau == o.L - 1) put(o.halo + o.L, z, z);
  }
}

}  // namespace tc
}  // namespace ttsb
