// Host side of the C ABI (include/tts_b200.h): plan construction, the launch schedule of
// generator.inference(), per-launch profiling and the host-buffer entry point.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "plan.h"

// Scratch of ttsb200_generator_forward_static(): lives in the module image, so that entry point never allocates.
// One array per GPU that loads the library, held until exit.  In host code this name is only a handle for
// cudaGetSymbolAddress — never a device address.
static __device__ __align__(256) unsigned char g_static_workspace[TTSB200_STATIC_WORKSPACE_BYTES];

namespace ttsb {

static thread_local std::string g_last_error;
void set_error(const std::string& msg) { g_last_error = msg; }
int fail(int code, const std::string& msg) {
  g_last_error = msg;
  return code;
}

namespace {

int64_t ipow(int64_t b, int e) {
  int64_t r = 1;
  while (e-- > 0) r *= b;
  return r;
}

struct BlobBuilder {
  std::vector<float> data;
  size_t alloc(size_t n) {  // 64-float (256 B) aligned regions
    size_t off = (data.size() + 63) / 64 * 64;
    data.resize(off + n, 0.f);
    return off;
  }
};

// Conv1d weight [Cout][Cin][k] -> [k][Cin][CoutPad]
void pack_conv(BlobBuilder& bb, ConvF32& c, const float* w, const float* b, int Cout, int Cin, int k) {
  c.Cin = Cin; c.Cout = Cout; c.taps = k; c.phases = 1;
  c.cpt = conv_f32_pick_cpt(Cout);
  c.CoutPad = (int)round_up(Cout, conv_f32_tile_channels(c.cpt));
  c.w_off = bb.alloc((size_t)k * Cin * c.CoutPad);
  c.b_off = bb.alloc(c.CoutPad);
  for (int o = 0; o < Cout; ++o)
    for (int ci = 0; ci < Cin; ++ci)
      for (int j = 0; j < k; ++j)
        bb.data[c.w_off + ((size_t)j * Cin + ci) * c.CoutPad + o] = w[((size_t)o * Cin + ci) * k + j];
  for (int o = 0; o < Cout; ++o) bb.data[c.b_off + o] = b[o];
}

// ConvTranspose1d weight [Cin][Cout][2f] -> [f phases][2 taps][Cin][CoutPad]:
// output n = m f + r reads x[m + q] with tap phi and x[m + q - 1] with tap phi + f,
// q = (r + P)/f, phi = (r + P) % f   (derivation: oracle/numpy_oracle.py conv_transpose1d_melgan).
void pack_convT(BlobBuilder& bb, ConvF32& c, const float* w, const float* b, int Cin, int Cout, int f, int P) {
  c.Cin = Cin; c.Cout = Cout; c.taps = 2; c.phases = f;
  c.cpt = conv_f32_pick_cpt(Cout);
  c.CoutPad = (int)round_up(Cout, conv_f32_tile_channels(c.cpt));
  c.w_off = bb.alloc((size_t)f * 2 * Cin * c.CoutPad);
  c.b_off = bb.alloc(c.CoutPad);
  for (int r = 0; r < f; ++r) {
    const int phi = (r + P) % f;
    for (int s = 0; s < 2; ++s)
      for (int ci = 0; ci < Cin; ++ci)
        for (int o = 0; o < Cout; ++o)
          bb.data[c.w_off + (((size_t)r * 2 + s) * Cin + ci) * c.CoutPad + o] =
              w[((size_t)ci * Cout + o) * (2 * f) + phi + f * s];
  }
  for (int o = 0; o < Cout; ++o) bb.data[c.b_off + o] = b[o];
}

int validate_cfg(const ttsb200_generator_config& c) {
  if (c.in_channels <= 0 || c.out_channels <= 0 || c.base_channels <= 0)
    return fail(TTSB200_ERR_INVALID, "channel counts must be positive");
  if (c.proj_kernel <= 0 || (c.proj_kernel - 1) % 2 != 0)
    return fail(TTSB200_ERR_INVALID, " [!] proj_kernel should be an odd number.");  // melgan_generator.py:20-21
  if (c.res_kernel <= 0 || (c.res_kernel - 1) % 2 != 0)
    return fail(TTSB200_ERR_INVALID, " [!] kernel_size has to be odd.");            // melgan.py:9
  if (c.n_stages <= 0 || c.n_stages > TTSB200_MAX_STAGES)
    return fail(TTSB200_ERR_INVALID, "n_stages out of range");
  if (c.num_res_blocks < 0 || c.num_res_blocks > 8) return fail(TTSB200_ERR_INVALID, "num_res_blocks out of range");
  for (int s = 0; s < c.n_stages; ++s) {
    if (c.upsample_factors[s] <= 0 || c.upsample_factors[s] > 64)
      return fail(TTSB200_ERR_INVALID, "upsample factor out of range");
    if ((c.base_channels >> (s + 1)) <= 0) return fail(TTSB200_ERR_INVALID, "base_channels too small for the stage count");
  }
  if (c.pqmf_bands < 0 || (c.pqmf_bands > 0 && c.pqmf_bands != c.out_channels))
    return fail(TTSB200_ERR_INVALID, "pqmf_bands must equal out_channels");
  if (c.pqmf_bands > 0 && c.pqmf_taps <= 0) return fail(TTSB200_ERR_INVALID, "pqmf_taps must be positive");
  if (!(c.act_slope >= 0.f && c.act_slope <= 1.f))   // kernels use lrelu(v) = max(v, slope v)
    return fail(TTSB200_ERR_INVALID, "act_slope must be in [0, 1]");
  return 0;
}

struct Shape {
  int64_t L0;                  // padded frames
  std::vector<int64_t> L;      // length after each stage
  int64_t max_act;             // max C*L over every intermediate, per item
};

Shape shapes_for(const ttsb200_plan* plan, int frames, int pad) {
  Shape s;
  s.L0 = (int64_t)frames + 2 * (int64_t)pad;
  int64_t L = s.L0;
  s.max_act = (int64_t)plan->cfg.base_channels * L;
  for (const StagePlan& st : plan->stages) {
    L *= st.f;
    s.L.push_back(L);
    s.max_act = std::max<int64_t>(s.max_act, (int64_t)st.Cout * L);
  }
  s.max_act = std::max<int64_t>(s.max_act, (int64_t)plan->cfg.out_channels * L);
  return s;
}

// The reference raises from ReflectionPad1d when pad >= length; mirror that instead of reading
// out of bounds (SURVEY.md §8b "Errors").
int check_lengths(const ttsb200_plan* plan, const Shape& sh) {
  const int pp = (plan->cfg.proj_kernel - 1) / 2;
  const char* msg = "Padding size should be less than the corresponding input dimension";
  if (pp >= sh.L0) return fail(TTSB200_ERR_SHAPE, msg);
  for (size_t s = 0; s < plan->stages.size(); ++s)
    for (const BlockPlan& b : plan->stages[s].blocks)
      if ((int64_t)((plan->cfg.res_kernel - 1) / 2) * b.dilation >= sh.L[s]) return fail(TTSB200_ERR_SHAPE, msg);
  if (pp >= sh.L.back()) return fail(TTSB200_ERR_SHAPE, msg);
  return 0;
}

struct Launcher {
  cudaStream_t stream;
  ttsb200_profile* prof;
  int batch;

  cudaEvent_t take_event() {
    if (!prof->pool.empty()) {
      cudaEvent_t e = prof->pool.back();
      prof->pool.pop_back();
      return e;
    }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
  }
  template <typename F>
  cudaError_t run(const char* name, int layer, double flops, double bytes, F&& f) {
    if (!prof) return f();
    ttsb200_profile::Rec r{name, layer, flops, bytes, take_event(), take_event()};
    cudaEventRecord(r.start, stream);
    cudaError_t e = f();
    cudaEventRecord(r.stop, stream);
    prof->recs.push_back(r);
    return e;
  }
};

ConvSrc make_src(const float* x, const float* w, int C, int64_t Lreal, int taps, int dil, int shift,
                 bool edge_zero, bool lrelu, int clamp_pad = 0) {
  ConvSrc s{};
  s.x = x; s.w = w; s.C = C; s.Lreal = (int)Lreal; s.L = (int)Lreal + 2 * clamp_pad;
  s.clamp_pad = clamp_pad; s.batch_stride = (int64_t)C * Lreal;
  s.taps = taps; s.dil = dil; s.shift = shift; s.edge_zero = edge_zero ? 1 : 0; s.lrelu = lrelu ? 1 : 0;
  return s;
}

int copy_tap(const ttsb200_forward_args* a, int layer, const float* src, int C, int64_t L, cudaStream_t st) {
  for (int i = 0; i < a->n_taps; ++i)
    if (a->tap_layers[i] == layer)
      TTSB_CUDA(cudaMemcpyAsync(a->tap_out[i], src, (size_t)a->batch * C * L * sizeof(float),
                                cudaMemcpyDeviceToDevice, st));
  return 0;
}

// Last step of both schedules: `sub` holds the head output [B][out_channels][Lin] (fp32).  Either it already is the
// result (no PQMF / skip_pqmf, fp32 out), or it is converted to fp16, or PQMF synthesis turns it into the waveform
// (the synthesis kernels store fp32 or fp16 themselves).
int finish_output(const ttsb200_plan* plan, const ttsb200_forward_args* a, Launcher& ln, const float* sub, bool to_out,
                  int64_t Lin, const TcFollow* follow = nullptr) {
  const ttsb200_generator_config& cfg = plan->cfg;
  const int B = a->batch;
  const int f16 = a->out_dtype == TTSB200_OUT_F16;
  cudaStream_t st = static_cast<cudaStream_t>(a->stream);
  const bool windowed = a->out_keep_len > 0;
  if (to_out) {
    // `sub` is a workspace buffer whenever a conversion or a window is asked for (see the schedules)
    const int rows = B * cfg.out_channels;
    const size_t es = f16 ? 2 : 4;
    if (f16 && !windowed) {
      const int64_t n = (int64_t)rows * Lin;
      TTSB_CUDA(ln.run("out.f32_to_f16", -1, 0.0, 6.0 * n, [&] { return f32_to_f16_launch(sub, a->out, n, st); }));
    } else if (windowed) {
      const void* src = sub;
      if (f16)   // only the PQMF synthesis kernels store fp16 into a window themselves
        return fail(TTSB200_ERR_INVALID, "an output window with fp16 output needs a generator that ends in PQMF synthesis");
      const size_t dpitch = (size_t)(cfg.out_channels == 1 ? a->out_batch_stride : a->out_keep_len) * es;
      TTSB_CUDA(cudaMemcpy2DAsync(a->out, dpitch, static_cast<const char*>(src) + (size_t)a->out_keep_lo * es,
                                  (size_t)Lin * es, (size_t)a->out_keep_len * es, rows, cudaMemcpyDeviceToDevice, st));
    }
    return 0;
  }
  const int N = cfg.pqmf_bands;
  const PqmfCrop crop{a->out_keep_lo, a->out_keep_len, a->out_batch_stride};
  const double out_samples = windowed ? (double)a->out_keep_len * B : (double)N * Lin * B;
  TTSB_CUDA(ln.run("pqmf.synthesis", -1, 2.0 * (cfg.pqmf_taps + 1) * Lin * B,
                   4.0 * N * Lin * B + (f16 ? 2.0 : 4.0) * out_samples, [&] {
    if (plan->pqmf_fact && !plan->ctx.pqmf_direct && pqmf_fact_usable(sub, Lin))
      return pqmf_synthesis_fact_launch(sub, plan->blob + plan->pqmf_fact_off, B, Lin, a->out, f16,
                                        windowed ? &crop : nullptr, follow, st);
    return pqmf_synthesis_launch(sub, plan->blob + plan->pqmf_G_off, B, N, cfg.pqmf_taps, Lin, a->out, f16,
                                 windowed ? &crop : nullptr, st);
  }));
  return 0;
}

// fp32 schedule: 1 + S*(1 + 2n) + 1 (+1 PQMF) launches, every intermediate NCL fp32 in `workspace`.
int forward_f32(const ttsb200_plan* plan, const ttsb200_forward_args* a) {
  const ttsb200_generator_config& cfg = plan->cfg;
  const Shape sh = shapes_for(plan, a->frames, a->pad);
  if (int rc = check_lengths(plan, sh)) return rc;
  const int B = a->batch;
  const size_t buf_floats = (size_t)round_up((int64_t)B * sh.max_act, 64);
  if (!a->workspace || a->workspace_bytes < 3 * buf_floats * sizeof(float))
    return fail(TTSB200_ERR_WORKSPACE, "workspace smaller than ttsb200_workspace_bytes()");
  float* cur = static_cast<float*>(a->workspace);
  float* spare0 = cur + buf_floats;
  float* spare1 = spare0 + buf_floats;
  cudaStream_t st = static_cast<cudaStream_t>(a->stream);
  Launcher ln{st, a->profile, B};
  const float* W = plan->blob;
  const float slope = cfg.act_slope;
  const int pp = (cfg.proj_kernel - 1) / 2;

  auto base_params = [&](const ConvF32& c, float* out, int64_t Lt, int64_t out_L) {
    ConvParams p{};
    p.nsrc = 1; p.bias = W + c.b_off; p.out = out; p.Cout = c.Cout; p.CoutPad = c.CoutPad;
    p.Lt = (int)Lt; p.out_L = (int)out_L; p.out_stride = 1; p.n_phase = 1; p.up_f = 1; p.up_P = 0;
    p.out_batch_stride = (int64_t)c.Cout * out_L; p.tanh_out = 0; p.slope = slope;
    return p;
  };
  auto io_bytes = [&](double cin_l, double cout_l, double wfloats) { return 4.0 * (B * (cin_l + cout_l) + wfloats); };

  // layers[0..1]: ReflectionPad1d + Conv1d k7; inference_padding folded into the index map
  {
    ConvParams p = base_params(plan->proj, cur, sh.L0, sh.L0);
    p.src[0] = make_src(a->mel, W + plan->proj.w_off, cfg.in_channels, a->frames, cfg.proj_kernel, 1, -pp,
                        false, false, a->pad);
    const double macs = (double)cfg.proj_kernel * cfg.in_channels * cfg.base_channels * sh.L0 * B;
    TTSB_CUDA(ln.run("f32.proj.k7", 1, 2 * macs,
                     io_bytes((double)cfg.in_channels * a->frames, (double)cfg.base_channels * sh.L0,
                              (double)cfg.proj_kernel * cfg.in_channels * cfg.base_channels),
                     [&] { return conv_f32_launch(p, B, plan->proj.cpt, st); }));
    if (int rc = copy_tap(a, 1, cur, cfg.base_channels, sh.L0, st)) return rc;
  }
  int64_t Lin = sh.L0;
  for (size_t s = 0; s < plan->stages.size(); ++s) {
    const StagePlan& sp = plan->stages[s];
    const int64_t L = sh.L[s];
    {  // LeakyReLU + ConvTranspose1d: f phases, two taps each (x[m+q], x[m+q-1]), zero-fill edges
      ConvParams p = base_params(sp.up, spare0, Lin, L);
      p.src[0] = make_src(cur, W + sp.up.w_off, sp.Cin, Lin, 2, -1, 0, true, true);
      p.out_stride = sp.f; p.n_phase = sp.f; p.up_f = sp.f; p.up_P = sp.P;
      const double macs = 2.0 * sp.f * sp.Cin * sp.Cout * Lin * B;
      TTSB_CUDA(ln.run("f32.upsample.convT", sp.layer_up, 2 * macs,
                       io_bytes((double)sp.Cin * Lin, (double)sp.Cout * L, 2.0 * sp.f * sp.Cin * sp.Cout),
                       [&] { return conv_f32_launch(p, B, sp.up.cpt, st); }));
      std::swap(cur, spare0);
      if (int rc = copy_tap(a, sp.layer_up, cur, sp.Cout, L, st)) return rc;
    }
    const int C = sp.Cout;
    const int half = (cfg.res_kernel - 1) / 2;
    for (const BlockPlan& bp : sp.blocks) {
      {  // h = conv_k3_dilated(reflect_pad(lrelu(x)))
        ConvParams p = base_params(bp.conv1, spare1, L, L);
        p.src[0] = make_src(cur, W + bp.conv1.w_off, C, L, cfg.res_kernel, bp.dilation, -half * bp.dilation,
                            false, true);
        const double macs = (double)cfg.res_kernel * C * C * L * B;
        TTSB_CUDA(ln.run("f32.resblock.conv1_dilated", sp.layer_stack, 2 * macs,
                         io_bytes((double)C * L, (double)C * L, (double)cfg.res_kernel * C * C),
                         [&] { return conv_f32_launch(p, B, bp.conv1.cpt, st); }));
      }
      {  // x <- W2 * lrelu(h) + S * x + (b2 + bs)      one accumulator, two GEMM sources
        ConvParams p = base_params(bp.conv2, spare0, L, L);
        p.nsrc = 2;
        p.bias = W + bp.b2s_off;
        p.src[0] = make_src(spare1, W + bp.conv2.w_off, C, L, 1, 1, 0, false, true);
        p.src[1] = make_src(cur, W + bp.shortcut.w_off, C, L, 1, 1, 0, false, false);
        const double macs = 2.0 * C * C * L * B;
        TTSB_CUDA(ln.run("f32.resblock.conv2+shortcut", sp.layer_stack, 2 * macs,
                         io_bytes(2.0 * C * L, (double)C * L, 2.0 * C * C),
                         [&] { return conv_f32_launch(p, B, bp.conv2.cpt, st); }));
        std::swap(cur, spare0);
      }
    }
    if (int rc = copy_tap(a, sp.layer_stack, cur, C, L, st)) return rc;
    Lin = L;
  }
  {  // LeakyReLU + ReflectionPad1d + Conv1d k7 + Tanh
    const int Cl = plan->head.Cin;
    const bool to_out = cfg.pqmf_bands == 0 || a->skip_pqmf;
    const bool direct = to_out && a->out_dtype == TTSB200_OUT_F32 && a->out_keep_len == 0;
    float* dst = direct ? static_cast<float*>(a->out) : spare0;
    ConvParams p = base_params(plan->head, dst, Lin, Lin);
    p.src[0] = make_src(cur, W + plan->head.w_off, Cl, Lin, cfg.proj_kernel, 1, -pp, false, true);
    p.tanh_out = 1;
    const double macs = (double)cfg.proj_kernel * Cl * cfg.out_channels * Lin * B;
    const int ih = 4 + 3 * cfg.n_stages;
    TTSB_CUDA(ln.run("f32.head.k7+tanh", ih, 2 * macs,
                     io_bytes((double)Cl * Lin, (double)cfg.out_channels * Lin,
                              (double)cfg.proj_kernel * Cl * cfg.out_channels),
                     [&] { return head_f32_usable(p) ? head_f32_launch(p, B, st) : conv_f32_launch(p, B, plan->head.cpt, st); }));
    if (int rc = copy_tap(a, ih + 1, dst, cfg.out_channels, Lin, st)) return rc;
    if (int rc = finish_output(plan, a, ln, dst, to_out, Lin)) return rc;
  }
  return 0;
}

#include "schedule_bf16.inl"

}  // namespace
}  // namespace ttsb

using namespace ttsb;

extern "C" {

int32_t ttsb200_abi_version(void) { return TTSB200_ABI_VERSION; }
const char* ttsb200_last_error(void) { return g_last_error.c_str(); }

int32_t ttsb200_device_ok(int32_t device) {
  cudaDeviceProp prop{};
  TTSB_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return fail(TTSB200_ERR_ARCH, std::string("tts_b200 carries sm_100a code only; device is ") + prop.name +
                                      " (sm_" + std::to_string(prop.major) + std::to_string(prop.minor) + ")");
  return 0;
}

int32_t ttsb200_expected_tensors(const ttsb200_generator_config* cfg) {
  if (!cfg) return fail(TTSB200_ERR_INVALID, "cfg is NULL");
  return 2 * (2 + cfg->n_stages * (1 + 3 * cfg->num_res_blocks)) + (cfg->pqmf_bands > 0 ? 1 : 0);
}

int32_t ttsb200_plan_create(const ttsb200_generator_config* cfg, const float* const* tensors, int32_t n_tensors,
                            int32_t device, ttsb200_plan** out_plan) {
  if (!cfg || !tensors || !out_plan) return fail(TTSB200_ERR_INVALID, "NULL argument");
  if (int rc = validate_cfg(*cfg)) return rc;
  if (n_tensors != ttsb200_expected_tensors(cfg))
    return fail(TTSB200_ERR_INVALID, "expected " + std::to_string(ttsb200_expected_tensors(cfg)) +
                                         " tensors, got " + std::to_string(n_tensors));
  for (int i = 0; i < n_tensors; ++i)
    if (!tensors[i]) return fail(TTSB200_ERR_INVALID, "tensor " + std::to_string(i) + " is NULL");
  if (int rc = ttsb200_device_ok(device)) return rc;

  ttsb200_plan* plan = new ttsb200_plan();
  plan->cfg = *cfg;
  plan->device = device;
  BlobBuilder bb;
  int ti = 0;
  pack_conv(bb, plan->proj, tensors[ti], tensors[ti + 1], cfg->base_channels, cfg->in_channels, cfg->proj_kernel);
  ti += 2;
  int hop = 1;
  int clast = cfg->base_channels;
  for (int s = 0; s < cfg->n_stages; ++s) {
    StagePlan sp;
    sp.f = cfg->upsample_factors[s];
    sp.P = sp.f / 2 + sp.f % 2;  // melgan_generator.py:46-47
    sp.Cin = cfg->base_channels >> s;
    sp.Cout = cfg->base_channels >> (s + 1);
    sp.layer_up = 3 + 3 * s;
    sp.layer_stack = 4 + 3 * s;
    pack_convT(bb, sp.up, tensors[ti], tensors[ti + 1], sp.Cin, sp.Cout, sp.f, sp.P);
    ti += 2;
    const int C = sp.Cout;
    sp.blocks.resize(cfg->num_res_blocks);
    for (int b = 0; b < cfg->num_res_blocks; ++b) {
      BlockPlan& bp = sp.blocks[b];
      bp.dilation = (int)ipow(cfg->res_kernel, b);  // melgan.py:14
      pack_conv(bb, bp.conv1, tensors[ti], tensors[ti + 1], C, C, cfg->res_kernel);
      pack_conv(bb, bp.conv2, tensors[ti + 2], tensors[ti + 3], C, C, 1);
      ti += 4;
    }
    for (int b = 0; b < cfg->num_res_blocks; ++b) {
      BlockPlan& bp = sp.blocks[b];
      pack_conv(bb, bp.shortcut, tensors[ti], tensors[ti + 1], C, C, 1);
      ti += 2;
      bp.b2s_off = bb.alloc(bp.conv2.CoutPad);
      for (int o = 0; o < C; ++o)
        bb.data[bp.b2s_off + o] = bb.data[bp.conv2.b_off + o] + bb.data[bp.shortcut.b_off + o];
    }
    plan->stages.push_back(std::move(sp));
    hop *= cfg->upsample_factors[s];
    clast = C;
  }
  pack_conv(bb, plan->head, tensors[ti], tensors[ti + 1], cfg->out_channels, clast, cfg->proj_kernel);
  ti += 2;
  if (cfg->pqmf_bands > 0) {
    const size_t n = (size_t)cfg->pqmf_bands * (cfg->pqmf_taps + 1);
    plan->pqmf_G_off = bb.alloc(n);
    std::memcpy(&bb.data[plan->pqmf_G_off], tensors[ti], n * sizeof(float));
    if (cfg->pqmf_bands == 4 && cfg->pqmf_taps == 62) {
      float tbl[ttsb::PQMF_FACT_TABLE];
      if (ttsb::pqmf_factor_4x63(tensors[ti], tbl)) {
        plan->pqmf_fact_off = bb.alloc(ttsb::PQMF_FACT_TABLE);
        std::memcpy(&bb.data[plan->pqmf_fact_off], tbl, sizeof tbl);
        plan->pqmf_fact = true;
      }
    }
    ti += 1;
  }
  plan->hop_generator = hop;
  plan->ctx = tc_launch_ctx(device);   // before packing: the conv kernel choice sets the column-group widths
  Bf16Builder b16;
  if (int rc = pack_bf16(plan, tensors, bb, b16)) {
    delete plan;
    return rc;
  }
  plan->blob_floats = bb.data.size();
  plan->blob16_elems = b16.data.size();
  // Say so once when a stage has no fused stack kernel (channel counts other than 32 / 48 / 64 / 96 / 128 / 192 / 256
  // after padding to 16, more than four blocks, dilations too wide for the tile): it then runs 2 launches per block
  // on the generic conv kernel, several times slower — a different speed class, not an error.
  if (cfg->num_res_blocks > 0 && !getenv("TTSB200_QUIET"))
    for (int s = 0; s < cfg->n_stages; ++s)
      if (stack_tiling(plan, s).kt == 0)
        fprintf(stderr, "tts_b200: BF16 path: stage %d (%d channels, %d blocks) has no fused ResidualStack kernel; "
                        "it will run %d generic conv launches (set TTSB200_QUIET=1 to silence)\n",
                s, plan->stages[s].Cout, cfg->num_res_blocks, 2 * cfg->num_res_blocks);

  int prev = 0;
  cudaGetDevice(&prev);
  cudaError_t e = cudaSetDevice(device);
  if (e == cudaSuccess) e = cudaMalloc(&plan->blob, plan->blob_floats * sizeof(float));
  if (e == cudaSuccess)
    e = cudaMemcpy(plan->blob, bb.data.data(), plan->blob_floats * sizeof(float), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMalloc(&plan->blob16, plan->blob16_elems * sizeof(__nv_bfloat16));
  if (e == cudaSuccess)
    e = cudaMemcpy(plan->blob16, b16.data.data(), plan->blob16_elems * sizeof(__nv_bfloat16), cudaMemcpyHostToDevice);
  cudaSetDevice(prev);
  if (e != cudaSuccess) {
    if (plan->blob) cudaFree(plan->blob);
    if (plan->blob16) cudaFree(plan->blob16);
    delete plan;
    return fail(TTSB200_ERR_CUDA, std::string("plan upload: ") + cudaGetErrorString(e));
  }
  *out_plan = plan;
  return 0;
}

void ttsb200_plan_destroy(ttsb200_plan* plan) {
  if (!plan) return;
  if (plan->blob) cudaFree(plan->blob);
  if (plan->blob16) cudaFree(plan->blob16);
  if (plan->host.d_in) cudaFree(plan->host.d_in);
  if (plan->host.d_out) cudaFree(plan->host.d_out);
  if (plan->host.d_ws) cudaFree(plan->host.d_ws);
  if (plan->host.stream) cudaStreamDestroy(plan->host.stream);
  delete plan;
}

int64_t ttsb200_output_length(const ttsb200_plan* plan, int32_t frames, int32_t pad, int32_t skip_pqmf) {
  if (!plan) return fail(TTSB200_ERR_INVALID, "plan is NULL");
  const int64_t L = ((int64_t)frames + 2 * (int64_t)pad) * plan->hop_generator;
  if (plan->cfg.pqmf_bands > 0 && !skip_pqmf) {
    const int taps = plan->cfg.pqmf_taps;
    return plan->cfg.pqmf_bands * L + 2 * (taps / 2) - taps;
  }
  return L;
}

int64_t ttsb200_workspace_bytes(const ttsb200_plan* plan, int32_t batch, int32_t frames, int32_t pad,
                                int32_t precision) {
  if (!plan) return fail(TTSB200_ERR_INVALID, "plan is NULL");
  if (batch <= 0 || frames <= 0 || pad < 0) return fail(TTSB200_ERR_INVALID, "batch/frames must be positive, pad >= 0");
  const Shape sh = shapes_for(plan, frames, pad);
  if (precision == TTSB200_BF16) return (int64_t)bf16_workspace(plan, batch, sh).total_bytes;
  return 3 * round_up((int64_t)batch * sh.max_act, 64) * (int64_t)sizeof(float);
}

int32_t ttsb200_layer_shape(const ttsb200_plan* plan, int32_t layer, int32_t frames, int32_t pad, int32_t* channels,
                            int64_t* length) {
  if (!plan || !channels || !length) return fail(TTSB200_ERR_INVALID, "NULL argument");
  const Shape sh = shapes_for(plan, frames, pad);
  const int S = plan->cfg.n_stages;
  if (layer == 1) {
    *channels = plan->cfg.base_channels;
    *length = sh.L0;
    return 0;
  }
  for (int s = 0; s < S; ++s)
    if (layer == 3 + 3 * s || layer == 4 + 3 * s) {
      *channels = plan->stages[s].Cout;
      *length = sh.L[s];
      return 0;
    }
  if (layer == 5 + 3 * S) {
    *channels = plan->cfg.out_channels;
    *length = sh.L.back();
    return 0;
  }
  return fail(TTSB200_ERR_INVALID, "layer " + std::to_string(layer) + " has no tappable output");
}

int32_t ttsb200_generator_forward(const ttsb200_plan* plan, const ttsb200_forward_args* a) {
  if (!plan || !a) return fail(TTSB200_ERR_INVALID, "NULL argument");
  if (!a->mel || !a->out) return fail(TTSB200_ERR_INVALID, "mel/out is NULL");
  if (a->batch <= 0 || a->frames <= 0 || a->pad < 0)
    return fail(TTSB200_ERR_INVALID, "batch/frames must be positive, pad >= 0");
  if (a->n_taps < 0 || (a->n_taps > 0 && (!a->tap_layers || !a->tap_out)))
    return fail(TTSB200_ERR_INVALID, "tap arrays missing");
  int fmax = 1;
  for (const StagePlan& sp : plan->stages) fmax = std::max(fmax, sp.f);
  if ((int64_t)a->batch * fmax > 65535)
    return fail(TTSB200_ERR_INVALID, "batch too large for one launch (batch * max upsample factor > 65535); split the batch");
  if (a->out_dtype != TTSB200_OUT_F32 && a->out_dtype != TTSB200_OUT_F16)
    return fail(TTSB200_ERR_INVALID, "unknown out_dtype");
  if (a->out_keep_len < 0 || a->out_keep_lo < 0 ||
      (a->out_keep_len > 0 &&
       (a->out_keep_lo + a->out_keep_len > ttsb200_output_length(plan, a->frames, a->pad, a->skip_pqmf) ||
        a->out_batch_stride < a->out_keep_len)))
    return fail(TTSB200_ERR_INVALID, "output window outside the result, or out_batch_stride < out_keep_len");
  if (a->out_keep_len > 0 && (plan->cfg.pqmf_bands == 0 || a->skip_pqmf) && plan->cfg.out_channels > 1 &&
      a->out_batch_stride != (int64_t)plan->cfg.out_channels * a->out_keep_len)
    return fail(TTSB200_ERR_INVALID, "a windowed multi-channel output must be dense: out_batch_stride = channels * out_keep_len");
  // The caller's current device need not be the plan's (a generator on cuda:1 called while cuda:0 is current,
  // which the reference supports): launches, attribute calls and events all go to plan->device.
  int prev = 0;
  TTSB_CUDA(cudaGetDevice(&prev));
  if (prev != plan->device) TTSB_CUDA(cudaSetDevice(plan->device));
  struct Restore {
    int prev, cur;
    ~Restore() { if (prev != cur) cudaSetDevice(prev); }
  } restore{prev, plan->device};
  switch (a->precision) {
    case TTSB200_F32: return forward_f32(plan, a);
    case TTSB200_BF16: return forward_bf16(plan, a);
    default: return fail(TTSB200_ERR_INVALID, "unknown precision");
  }
}

int32_t ttsb200_generator_forward_static(const ttsb200_plan* plan, const ttsb200_forward_args* a) {
  if (!plan || !a) return fail(TTSB200_ERR_INVALID, "NULL argument");
  if (a->profile)   // a profile creates CUDA events on demand
    return fail(TTSB200_ERR_INVALID, "ttsb200_generator_forward_static takes no profile (it would allocate events)");
  if (plan->ctx.trace_file)
    return fail(TTSB200_ERR_INVALID, "ttsb200_generator_forward_static: the plan was created with a trace file (debug builds allocate)");
  if (a->batch <= 0 || a->frames <= 0 || a->pad < 0)
    return fail(TTSB200_ERR_INVALID, "batch/frames must be positive, pad >= 0");
  const int64_t need = ttsb200_workspace_bytes(plan, a->batch, a->frames, a->pad, a->precision);
  if (need < 0) return (int32_t)need;
  if ((uint64_t)need > (uint64_t)TTSB200_STATIC_WORKSPACE_BYTES)
    return fail(TTSB200_ERR_WORKSPACE, "ttsb200_generator_forward_static: this shape needs " + std::to_string(need) +
                                           " bytes of scratch, the static array holds " +
                                           std::to_string((uint64_t)TTSB200_STATIC_WORKSPACE_BYTES) +
                                           " (TTSB200_STATIC_WORKSPACE_BYTES); use ttsb200_generator_forward with a caller workspace");
  // Set-up, once per device, with that device current: the array's device address on the plan's GPU.
  constexpr int MAX_DEVICES = 64;
  static std::mutex mu;
  static void* addr[MAX_DEVICES] = {};
  if (plan->device < 0 || plan->device >= MAX_DEVICES) return fail(TTSB200_ERR_INVALID, "device index out of range");
  void* ws = nullptr;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[plan->device]) {
      int prev = 0;
      TTSB_CUDA(cudaGetDevice(&prev));
      if (prev != plan->device) TTSB_CUDA(cudaSetDevice(plan->device));
      void* p = nullptr;
      const cudaError_t e = cudaGetSymbolAddress(&p, g_static_workspace);
      if (prev != plan->device) cudaSetDevice(prev);
      if (e != cudaSuccess || !p)
        return fail(TTSB200_ERR_CUDA, std::string("cudaGetSymbolAddress(g_static_workspace): ") + cudaGetErrorString(e));
      addr[plan->device] = p;
    }
    ws = addr[plan->device];
  }
  ttsb200_forward_args b = *a;
  b.workspace = ws;
  b.workspace_bytes = (size_t)TTSB200_STATIC_WORKSPACE_BYTES;
  return ttsb200_generator_forward(plan, &b);
}

int32_t ttsb200_generator_forward_host(ttsb200_plan* plan, const float* mel_host, int32_t batch, int32_t frames,
                                       int32_t pad, int32_t precision, float* out_host) {
  if (!plan || !mel_host || !out_host) return fail(TTSB200_ERR_INVALID, "NULL argument");
  if (batch <= 0 || frames <= 0 || pad < 0) return fail(TTSB200_ERR_INVALID, "batch/frames must be positive, pad >= 0");
  HostStaging& h = plan->host;
  std::lock_guard<std::mutex> lock(h.mu);
  int prev = 0;
  cudaGetDevice(&prev);
  TTSB_CUDA(cudaSetDevice(plan->device));
  struct Restore {
    int d;
    ~Restore() { cudaSetDevice(d); }
  } restore{prev};
  if (!h.stream) TTSB_CUDA(cudaStreamCreateWithFlags(&h.stream, cudaStreamNonBlocking));
  const size_t in_b = (size_t)batch * plan->cfg.in_channels * frames * sizeof(float);
  const int64_t out_len = ttsb200_output_length(plan, frames, pad, 0);
  // [batch, 1, samples] with PQMF synthesis, [batch, out_channels, samples] without (MelganGenerator(out_channels=4))
  const size_t out_ch = plan->cfg.pqmf_bands > 0 ? 1 : (size_t)plan->cfg.out_channels;
  const size_t out_b = (size_t)batch * out_ch * (size_t)out_len * sizeof(float);
  const int64_t ws = ttsb200_workspace_bytes(plan, batch, frames, pad, precision);
  if (ws < 0) return (int32_t)ws;
  auto ensure = [](void** p, size_t* have, size_t need) -> cudaError_t {
    if (*have >= need) return cudaSuccess;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *have = 0;
    cudaError_t e = cudaMalloc(p, need);
    if (e == cudaSuccess) *have = need;
    return e;
  };
  TTSB_CUDA(ensure(reinterpret_cast<void**>(&h.d_in), &h.in_bytes, in_b));
  TTSB_CUDA(ensure(reinterpret_cast<void**>(&h.d_out), &h.out_bytes, out_b));
  TTSB_CUDA(ensure(&h.d_ws, &h.ws_bytes, (size_t)ws));
  TTSB_CUDA(cudaMemcpyAsync(h.d_in, mel_host, in_b, cudaMemcpyHostToDevice, h.stream));
  ttsb200_forward_args a{};
  a.mel = h.d_in; a.batch = batch; a.frames = frames; a.pad = pad; a.precision = precision;
  a.out = h.d_out; a.workspace = h.d_ws; a.workspace_bytes = h.ws_bytes; a.stream = h.stream;
  if (int rc = ttsb200_generator_forward(plan, &a)) return rc;
  TTSB_CUDA(cudaMemcpyAsync(out_host, h.d_out, out_b, cudaMemcpyDeviceToHost, h.stream));
  TTSB_CUDA(cudaStreamSynchronize(h.stream));
  return 0;
}

int32_t ttsb200_pqmf_synthesis(const float* x, const float* G, int32_t batch, int32_t bands, int32_t taps, int64_t Ts,
                               float* y, void* stream) {
  if (!x || !G || !y) return fail(TTSB200_ERR_INVALID, "NULL argument");
  if (batch <= 0 || bands <= 0 || taps <= 0 || Ts <= 0 || batch > 65535)
    return fail(TTSB200_ERR_INVALID, "batch/bands/taps/Ts out of range");
  TTSB_CUDA(pqmf_synthesis_launch(x, G, batch, bands, taps, Ts, y, 0, nullptr, static_cast<cudaStream_t>(stream)));
  return 0;
}

int32_t ttsb200_pqmf_analysis(const float* x, const float* H, int32_t batch, int32_t bands, int32_t taps, int64_t T,
                              float* y, void* stream) {
  if (!x || !H || !y) return fail(TTSB200_ERR_INVALID, "NULL argument");
  if (batch <= 0 || bands <= 0 || taps <= 0 || T <= 0 || batch > 65535)
    return fail(TTSB200_ERR_INVALID, "batch/bands/taps/T out of range");
  TTSB_CUDA(pqmf_analysis_launch(x, H, batch, bands, taps, T, y, static_cast<cudaStream_t>(stream)));
  return 0;
}

int32_t ttsb200_pqmf_factor_4x63(const float* G_host, float* table96_host) {
  if (!G_host || !table96_host) return fail(TTSB200_ERR_INVALID, "NULL argument");
  return pqmf_factor_4x63(G_host, table96_host) ? 1 : 0;
}

int32_t ttsb200_profile_create(ttsb200_profile** out_profile) {
  if (!out_profile) return fail(TTSB200_ERR_INVALID, "NULL argument");
  *out_profile = new ttsb200_profile();
  return 0;
}

void ttsb200_profile_reset(ttsb200_profile* p) {
  if (!p) return;
  for (auto& r : p->recs) {
    p->pool.push_back(r.start);
    p->pool.push_back(r.stop);
  }
  p->recs.clear();
}

void ttsb200_profile_destroy(ttsb200_profile* p) {
  if (!p) return;
  ttsb200_profile_reset(p);
  for (cudaEvent_t e : p->pool) cudaEventDestroy(e);
  delete p;
}

int32_t ttsb200_profile_read(ttsb200_profile* p, ttsb200_launch_record* records, int32_t max_records) {
  if (!p) return fail(TTSB200_ERR_INVALID, "profile is NULL");
  const int n = (int)p->recs.size();
  if (!records) return n;
  const int m = n < max_records ? n : max_records;
  for (int i = 0; i < m; ++i) {
    const auto& r = p->recs[i];
    float ms = 0.f;
    TTSB_CUDA(cudaEventElapsedTime(&ms, r.start, r.stop));
    records[i].name = r.name;
    records[i].layer = r.layer;
    records[i].ms = ms;
    records[i].flops = r.flops;
    records[i].bytes = r.bytes;
  }
  return m;
}

int32_t ttsb200_launch_count(const ttsb200_plan* plan, int32_t precision, int32_t skip_pqmf) {
  if (!plan) return fail(TTSB200_ERR_INVALID, "plan is NULL");
  const int S = plan->cfg.n_stages, n = plan->cfg.num_res_blocks;
  int launches = 1 + S + 1 + ((plan->cfg.pqmf_bands > 0 && !skip_pqmf) ? 1 : 0);
  for (int s = 0; s < S; ++s)
    launches += (precision == TTSB200_BF16 && n > 0 && stack_tiling(plan, s).kt > 0) ? 1 : 2 * n;
  if (precision == TTSB200_BF16 && n > 0 && head_fusable(plan)) --launches;   // head folded into the last stack
  if (precision == TTSB200_BF16)
    for (int s = 0; s + 1 < S; ++s)
      if (up_fusable(plan, s)) --launches;   // next stage's transposed conv folded into this stack
  return launches + (precision == TTSB200_BF16 ? 1 : 0);
}

}  // extern "C"
