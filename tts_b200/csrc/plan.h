// Plan = immutable, device-resident, kernel-ready form of one generator's weights.
#pragma once

#include <mutex>
#include <vector>

#include "common.cuh"

namespace ttsb {

// fp32 weights of one conv, laid out [phases][taps][Cin][CoutPad] (CoutPad contiguous).
struct ConvF32 {
  int Cin = 0, Cout = 0, CoutPad = 0, taps = 0, phases = 1, cpt = 1;
  size_t w_off = 0;  // float offset into the plan blob
  size_t b_off = 0;  // [CoutPad]
};

struct BlockPlan {
  int dilation = 1;
  ConvF32 conv1;     // dilated k3                                   melgan.py:20-24
  ConvF32 conv2;     // 1x1 after the second LeakyReLU               melgan.py:26-27
  ConvF32 shortcut;  // 1x1 on the un-activated input                melgan.py:31-34
  size_t b2s_off = 0;  // conv2.bias + shortcut.bias, [CoutPad]
};

struct StagePlan {
  int f = 1, P = 0, Cin = 0, Cout = 0;
  int layer_up = 0, layer_stack = 0;  // reference layers[] indices
  ConvF32 up;                         // ConvTranspose1d as f two-tap phases
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
  std::vector<ttsb::StagePlan> stages;
  size_t pqmf_G_off = 0;  // [N][taps+1]
  int hop_generator = 1;  // prod(upsample_factors)
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
