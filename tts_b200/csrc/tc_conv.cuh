// Host-visible description of one launch of the tcgen05 implicit-GEMM convolution (conv_bf16_tc.cu).
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "follow.cuh"

namespace ttsb {

// Device facts and debug knobs of a plan, read ONCE at plan creation: nothing on the launch path queries the
// driver or the environment (the streaming path, BASELINE config 5, is latency-bound).
struct TcLaunchCtx {
  int device = 0;
  int sms = 148;
  int stack_max_grid = 0;      // TTSB200_STACK_MAX_GRID: test hook, few CTAs so each walks many items
  int stack_debug = 0;         // TTSB200_STACK_DEBUG / TTSB200_CONV_DEBUG: knock-out bits (trace builds only)
  int conv_debug = 0;
  int conv_persistent = 2;     // TTSB200_CONV_PERSISTENT: 2 weights-resident persistent kernel (narrow column groups,
                               // the default: fastest at cfg2), 1 activation-resident persistent kernel (weights
                               // streamed, N = 192-256 per MMA; measured 0.321 vs 0.309 ms per step), 0 generic only
  int no_follow = 0;           // TTSB200_NO_FOLLOW: every kernel waits for the whole previous grid (A/B experiments)
  int no_k24 = 0;              // TTSB200_NO_K24: keep the six-K-step dilated conv on 24-channel stages (A/B experiments)
  int pqmf_direct = 0;         // TTSB200_PQMF_DIRECT: direct 63-tap synthesis instead of the factorised form
  const char* trace_file = nullptr;   // TTSB200_STACK_TRACE (trace builds only)
};
TcLaunchCtx tc_launch_ctx(int device);

constexpr int TC_MAX_TAPS = 8;
constexpr int TC_TILE_M = 128;   // GEMM rows (time positions) per CTA = UMMA M
constexpr int TC_MAX_STAGES = 4; // weight ring depth

// Activations are "canonical": [batch][C_pad/8 planes][plane_rows][8] bf16, plane row = halo + t.
struct TcSrc {
  const __nv_bfloat16* x;
  int64_t batch_stride;   // elements
  int planes;             // C_pad / 8
  int plane_rows;         // rows per plane in global memory
  int row_base;           // global plane row of smem row 0 when t0 == 0  (= halo + min tap shift)
  int rows;               // rows staged in shared memory (TC_TILE_M + tap span)
  int taps;
  int tap_row[TC_MAX_TAPS];  // smem row of GEMM row 0 for tap j
  int lrelu;              // LeakyReLU applied in shared memory before the MMAs read it
};

// One CTA column group: <= 256 accumulator columns.  The layer's columns are ordered [plane][phase][8]:
// 8 output channels x all `up_f` phases of a transposed conv are adjacent, so one thread's values for a
// plane form `up_f` consecutive 16-byte rows of the output tensor.
struct TcGroup {
  const __nv_bfloat16* w;  // [src][tap][C_pad/8][ncols][8]
  const float* bias;       // [ncols]
  int ncols;               // multiple of 16 and of 8 * up_f
  int col_first;           // first layer column of the group
};

constexpr int TC_MAX_GROUPS = 32;

struct TcConvParams {
  TcSrc src[2];
  int nsrc;
  TcGroup grp[TC_MAX_GROUPS];
  int ngroups;
  int Lt;                 // GEMM rows: output positions, or input positions + 1 for a transposed conv
  int up_f, up_P;         // output row of GEMM row q, phase r:  q * up_f - up_P + r   (plain conv: 1, 0)
  // output
  int out_mode;           // 0: canonical bf16   1: fp32 NCL + tanh
  __nv_bfloat16* out_c;   // canonical out
  int64_t out_batch_stride;
  int out_plane_rows;
  int out_halo;
  int out_L;              // valid output rows
  int out_lrelu;          // store LeakyReLU(y) (the only consumer wants the activated tensor)
  int margin;             // 0 none, 1 reflect (halo rows mirror the interior), 2 zero (rows -1 and L)
  float* out_f;           // fp32 NCL out [B][cout_real][out_L]
  int cout_real;
  float slope;
  // Hand-over per time segment (follow.cuh; persistent kernel only — tc_conv_count_info()): every CTA counts its
  // finished time tiles in done_out[b][tile / count_units_per_seg]; with follow.done the loaders poll the previous
  // kernel's counters instead of the kernel waiting for that whole grid.  follow_in_lo: sequence row of the first
  // staged input row of GEMM row 0 (min tap shift).
  int* done_out;
  int count_units_per_seg;
  TcFollow follow;
  int follow_in_lo;
  long long* trace;       // optional timeline buffer (trace.cuh)
  int trace_id;
  int debug;              // knock-out bits for timeline experiments (trace builds only)
};

// Where a kernel's result goes: canonical bf16 for the next tcgen05 kernel (halo rows filled per the
// consumer's padding rule), shared by the conv and the fused-stack kernels.
struct TcCanonOut {
  __nv_bfloat16* base;     // batch 0
  int64_t batch_stride;    // elements
  int plane_rows;
  int halo;
  int L;                   // valid rows
  int lrelu;               // store LeakyReLU(y)
  int margin;              // 0 none, 1 reflect, 2 zero rows at -1 and L
};

// Fused ResidualStack (melgan.py:36-39): x <- S_b x + W2_b lrelu(W1_b * lrelu(x)) for b = 0..n-1 in ONE
// persistent kernel: lrelu(x) in shared memory, raw x / h / accumulators in TMEM (stack_bf16_tc.cu).
constexpr int TC_STACK_MAX_BLOCKS = 4;
constexpr int TC_STACK_MAX_KT = 6;
struct TcStackParams {
  const __nv_bfloat16* x;  // canonical raw input (no margins needed)
  int64_t x_batch_stride;
  int x_plane_rows, x_halo;
  int cp;                  // padded channels (multiple of 16, <= 96)
  int L;                   // sequence length
  int nblocks;
  int res_taps;            // res_kernel (odd)
  int dil[TC_STACK_MAX_BLOCKS];
  const __nv_bfloat16* w1[TC_STACK_MAX_BLOCKS];   // [tap][cp/8][cp][8]
  const __nv_bfloat16* w2s[TC_STACK_MAX_BLOCKS];  // [2][cp/8][cp][8]  (conv 1x1, shortcut 1x1)
  const float* b1[TC_STACK_MAX_BLOCKS];           // [cp]
  const float* b2s[TC_STACK_MAX_BLOCKS];          // [cp]  conv2.bias + shortcut.bias
  int k24;                 // cp == 32 carrying 24 real channels: w1[] is the five-K-step packing (bias inside), see
                           // stack_bf16_tc.cu issue_k24 / schedule_bf16.inl pack_w1_k24
  int kt;                  // 128-row M tiles per work item (fixed per cp: tc_stack_kt)
  int sum_d;               // one-sided receptive field of the stack = sum over blocks of dil
  int guard;               // spare rows either side of the staged tile (>= max dilation reach)
  int w_sets;              // nblocks: all weights resident in shared memory; 1: streamed per block
  int batch, n_chains;     // filled by tc_stack_launch: work items = batch x n_chains
  // Optional fused output head (last stage only, cp <= 64): LeakyReLU -> ReflectionPad1d -> Conv1d(k) -> Tanh
  // (melgan_generator.py:65-77) computed from the stack output while it is still in shared memory; the stack
  // output itself is then never written to HBM.
  // Taps are packed on N: column j * head_ch + c is tap j of output channel c, so one K = cp MMA round with
  // A = lrelu(x_out) in TMEM gives every tap's partial product and the epilogue sums the shifted rows.
  const __nv_bfloat16* head_w;   // [cp/8][head_n][8]; nullptr = no head, write the canonical stack output
  const float* head_b;           // [head_ch]
  int head_taps, head_ch;        // kernel size (7), real output channels (1 or 4)
  int head_n;                    // GEMM columns: head_taps * head_ch rounded up to 16
  float* head_out;               // fp32 [batch][head_ch][L]
  // Optional fused ConvTranspose1d of the NEXT stage (melgan_generator.py:47-56): the stack output (activated,
  // still in shared memory) is contracted with the two-tap polyphase weights, y[q f - P + r] = a[q] w[r] +
  // a[q-1] w[r+f], and the kernel writes the next stage's input instead of its own output.
  const __nv_bfloat16* up_w;     // [2 taps][cp/8][up_n][8], columns ordered [plane][phase][8]; nullptr = none
  const float* up_b;             // [up_n]
  int up_f, up_P, up_n;          // stride, output offset (padding), GEMM columns = f * C_next_pad
  int up_mode;                   // filled by tc_stack_configure: 1 weights resident, 2 streamed through the W2S slot
  TcCanonOut up_out;             // raw input of the next stage's stack (no margins), L = next length
  // Hand-over between consecutive kernels per time segment instead of per grid (follow.cuh): the signal warp counts
  // finished chains in done_out[b][chain / count_units_per_seg]; with follow.done the kernel skips
  // griddepcontrol.wait and its loader waits for the segments that cover the rows of the item it is about to load.
  int* done_out;
  int count_units_per_seg;
  TcFollow follow;
  int debug;               // knock-out bits for timeline experiments (trace builds only)
  long long* trace;        // optional (TTSB200_STACK_TRACE): [role][256] {tag, clock} pairs written by CTA 0
  TcCanonOut out;
  float slope;
};
int tc_stack_ctas_per_sm(int cp);                 // 2 when two CTAs' accumulators fit the 512 TMEM columns
int tc_stack_kt(int cp);                          // 0: this channel count has no fused kernel
int tc_stack_cs(int cp);                          // epilogue warps per TMEM lane quarter (channel split)
bool tc_stack_configure(TcStackParams& p);        // fills kt / sum_d / guard / w_sets from cp, dil[], nblocks
size_t tc_stack_smem_bytes(const TcStackParams& p);
cudaError_t tc_stack_launch(const TcStackParams& p, int batch, const TcLaunchCtx& ctx, cudaStream_t stream);
// What a consumer of this launch's output needs to follow it: work units (chains) per utterance, output rows per
// unit and the row offset (follow.cuh TcFollow).
void tc_stack_count_info(const TcStackParams& p, int* n_units, int* unit_rows, int* row_offset);

size_t tc_conv_smem_bytes(const TcConvParams& p);
// True when tc_conv_launch will run this launch on the persistent weights-resident kernel, the one that counts
// finished work (done_out) and can follow a counting producer (follow.done); fills what a consumer needs (TcFollow):
// time tiles per utterance, output rows per tile, row offset, counts per tile (column groups).
bool tc_conv_count_info(const TcConvParams& p, int batch, const TcLaunchCtx& ctx, int* n_units, int* unit_rows,
                        int* row_offset, int* mult);
cudaError_t tc_conv_launch(const TcConvParams& p, int batch, const TcLaunchCtx& ctx, cudaStream_t stream);

// mel fp32 [B][C][T] -> canonical bf16 with replicate padding `pad` folded in and reflect margins
// ... and zeroes `n_counters` hand-over counters (first kernel of a forward, ordered behind everything before it)
// ... and publishes armed_id in counters[n_counters] (follow.cuh follow_armed; early = 0: the late form)
cudaError_t tc_prep_mel_launch(const float* mel, int batch, int C, int T, int pad, __nv_bfloat16* out,
                               int planes, int plane_rows, int halo, int* counters, int n_counters, int armed_id,
                               int early, cudaStream_t stream);
// canonical bf16 -> fp32 NCL (per-layer parity taps); undo_lrelu inverts a stored activation
cudaError_t tc_untile_launch(const __nv_bfloat16* x, int64_t batch_stride, int batch, int C, int64_t L, int plane_rows,
                             int halo, int undo_lrelu, float slope, float* out, cudaStream_t stream);

}  // namespace ttsb
