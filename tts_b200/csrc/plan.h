// Plan = immutable, device-resident, kernel-ready form of one generator's weights.
#pragma once

#include <mutex>
#include <vector>

#include "common.cuh"
#include "tc_conv.cuh"

namespace ttsb {

// fp32 weights of one conv, laid out [phases][taps][Cin][CoutPad] (CoutPad contiguous).
struct ConvF32 {
  int Cin = 0, Cout = 0, CoutPad = 0, taps = 0, phases = 1, cpt = 1;
  size_t w_off = 0;  // float offset into the plan blob
  size_t b_off = 0;  // [CoutPad]
};

// BF16 weights of one conv for the tcgen05 kernel: per column group, [src][tap][Cin_pad/8][ncols][8].
struct TcGroupPlan {
  size_t w_off = 0;     // bf16-element offset into the plan's bf16 blob
  size_t bias_off = 0;  // float offset into the fp32 blob, [ncols]
  int ncols = 0, col_first = 0;
};
struct TcLayer {
  int cin_pad = 0, cout_pad = 0, cout = 0, up_f = 1;
  std::vector<TcGroupPlan> groups;
};

struct BlockPlan {
  TcLayer tc_conv1, tc_conv2s;  // BF16: dilated k3; fused (1x1 on lrelu(h)) + (1x1 shortcut on x)
  int dilation = 1;
  bool has_k24 = false;          // 24-channel stage: dilated conv also packed as five K = 16 steps with the bias inside
  size_t w1_k24_off = 0;         // bf16-element offset of that packing (stack_bf16_tc.cu issue_k24)
  ConvF32 conv1;     // dilated k3                                   melgan.py:20-24
  ConvF32 conv2;     // 1x1 after the second LeakyReLU               melgan.py:26-27
  ConvF32 shortcut;  // 1x1 on the un-activated input                melgan.py:31-34
  size_t b2s_off = 0;  // conv2.bias + shortcut.bias, [CoutPad]
};

struct StagePlan {
  int f = 1, P = 0, Cin = 0, Cout = 0;
  int layer_up = 0, layer_stack = 0;  // reference layers[] indices
  ConvF32 up;                         // ConvTranspose1d as f two-tap phases
  TcLayer tc_up;                      // BF16: two phase classes (q = 0 / 1), 2 taps each
  std::vector<BlockPlan> blocks;
};

struct HostStaging {
  std::mutex mu;
  cudaStream_t stream = nullptr;
  float* d_in = nullptr;
  float* d_out = nullptr;
  void* d_ws = nullptr;
  size_t in_bytes = 0, out_bytes = 0, ws_bytes = 0;
};

}  // namespace ttsb

struct ttsb200_plan {
  ttsb200_generator_config cfg{};
  int device = 0;
  float* blob = nullptr;  // all fp32 weights
  size_t blob_floats = 0;
  ttsb::ConvF32 proj, head;
  ttsb::TcLayer tc_proj, tc_head;
  size_t head_ts_off = 0;           // fused-head operand [cp/8][head_ts_n][8] (taps packed on N), bf16-element offset
  int head_ts_n = 0;                // its GEMM columns; 0 = this head shape has no fused form
  __nv_bfloat16* blob16 = nullptr;  // all BF16 weights (UMMA canonical K-major layout)
  size_t blob16_elems = 0;
  int halo = 8;                     // margin rows either side of every canonical activation plane
  std::vector<ttsb::StagePlan> stages;
  size_t pqmf_G_off = 0;  // [N][taps+1]
  size_t pqmf_fact_off = 0;   // factorised-synthesis table (pqmf.cu), valid when pqmf_fact
  bool pqmf_fact = false;
  int hop_generator = 1;  // prod(upsample_factors)
  ttsb::TcLaunchCtx ctx;  // device facts + debug knobs, read once at creation (nothing on the launch path calls getenv)
  ttsb::HostStaging host;
};

struct ttsb200_profile {
  struct Rec {
    const char* name;
    int layer;
    double flops, bytes;
    cudaEvent_t start, stop;
  };
  std::vector<Rec> recs;
  std::vector<cudaEvent_t> pool;  // recycled events
};
