// BF16 / tcgen05 half of vocoder.cu (included inside its anonymous namespace): weight packing into
// the UMMA canonical layout and the launch schedule of generator.inference().
//
// Every intermediate is a "canonical" bf16 tensor [B][C_pad/8][rows][8], rows = roundup(L,128) + 2*halo,
// whose halo rows are filled by the PRODUCER with what its single consumer's padding rule needs
// (mirror rows for ReflectionPad1d, zero rows for the transposed conv), and stored already
// LeakyReLU-activated when the consumer only ever reads lrelu(x).  So every kernel's loads are
// plain contiguous bulk copies and no pad / activation kernel exists.

struct Bf16Builder {
  std::vector<__nv_bfloat16> data;
  size_t alloc(size_t n) {
    size_t off = (data.size() + 63) / 64 * 64;
    data.resize(off + n, __float2bfloat16(0.f));
    return off;
  }
};

inline int pad16(int c) { return (int)round_up(c, 16); }

// Packs one layer for the tcgen05 conv kernel.  Layer columns are ordered [plane][phase][8 channels]
// (tc_conv.cuh), split into groups of <= 256 columns that hold whole planes.
// wfun(src, tap, ci, phase, co) / bfun(co): reference weights; channels beyond the real counts are zero.
template <class WF, class BF>
int pack_tc_layer(Bf16Builder& b16, BlobBuilder& bb, TcLayer& L, int nsrc, const int* taps, const int* cin, int cout,
                  int up_f, WF wfun, BF bfun, bool resident = false) {
  L.cin_pad = pad16(cin[0]);
  L.cout_pad = pad16(cout);
  L.cout = cout;
  L.up_f = up_f;
  L.groups.clear();
  const int total = up_f * L.cout_pad;
  int unit = 8 * up_f;                       // one plane
  while (unit % 16) unit *= 2;
  if (unit > 256) return fail(TTSB200_ERR_INVALID, "upsample factor too large for the BF16 path");
  // `resident`: the layer runs on the persistent conv kernel, which keeps a whole group's weights in shared
  // memory beside a ring of activation chunks (conv_bf16_tc.cu) — size the groups for that.
  int cap = 256;
  if (resident) {
    const size_t w_budget = (size_t)(227 - 3) * 1024 - (size_t)4 * 8 * (128 + 24) * 16;
    const size_t per_col = (size_t)taps[0] * pad16(cin[0]) * 2;
    if (w_budget / per_col >= (size_t)unit) cap = (int)std::min<size_t>(256, w_budget / per_col);
  }
  int ncols = unit;
  for (int n = cap / unit * unit; n >= unit; n -= unit)
    if (total % n == 0) { ncols = n; break; }
  for (int col0 = 0; col0 < total; col0 += ncols) {
    TcGroupPlan g;
    g.ncols = ncols;
    g.col_first = col0;
    size_t tot = 0;
    for (int s = 0; s < nsrc; ++s) tot += (size_t)taps[s] * pad16(cin[s]) * ncols;
    g.w_off = b16.alloc(tot);
    g.bias_off = bb.alloc(ncols);
    size_t base = g.w_off;
    for (int s = 0; s < nsrc; ++s) {
      const int cp = pad16(cin[s]);
      for (int j = 0; j < taps[s]; ++j) {
        for (int ci = 0; ci < cin[s]; ++ci)
          for (int col = 0; col < ncols; ++col) {
            const int fl = col0 + col;
            const int plane = fl / (8 * up_f), ph = (fl / 8) % up_f, co = plane * 8 + fl % 8;
            if (co >= cout) continue;
            b16.data[base + ((size_t)(ci >> 3) * ncols + col) * 8 + (ci & 7)] = __float2bfloat16(wfun(s, j, ci, ph, co));
          }
        base += (size_t)cp * ncols;
      }
    }
    for (int col = 0; col < ncols; ++col) {
      const int fl = col0 + col;
      const int co = fl / (8 * up_f) * 8 + fl % 8;
      bb.data[g.bias_off + col] = co < cout ? bfun(co) : 0.f;
    }
    L.groups.push_back(g);
  }
  if ((int)L.groups.size() > TC_MAX_GROUPS) return fail(TTSB200_ERR_INVALID, "layer needs too many column groups");
  return 0;
}

// tensors[] in ttsb200_plan_create order; fills plan->tc_* and uploads plan->blob16 later.
int pack_bf16(ttsb200_plan* plan, const float* const* tensors, BlobBuilder& bb, Bf16Builder& b16) {
  const ttsb200_generator_config& cfg = plan->cfg;
  int ti = 0;
  int max_d = (cfg.proj_kernel - 1) / 2;
  {
    const float *w = tensors[ti], *b = tensors[ti + 1];
    const int taps[1] = {cfg.proj_kernel}, cin[1] = {cfg.in_channels};
    const int Cin = cfg.in_channels, k = cfg.proj_kernel;
    if (int rc = pack_tc_layer(b16, bb, plan->tc_proj, 1, taps, cin, cfg.base_channels, 1,
                               [&](int, int j, int ci, int, int co) { return w[((size_t)co * Cin + ci) * k + j]; },
                               [&](int co) { return b[co]; }, plan->ctx.conv_persistent == 2))
      return rc;
    ti += 2;
  }
  int clast = cfg.base_channels;
  for (int s = 0; s < cfg.n_stages; ++s) {
    StagePlan& sp = plan->stages[s];
    {
      const float *w = tensors[ti], *b = tensors[ti + 1];
      const int f = sp.f, P = sp.P, Cout = sp.Cout;
      // y[q f - P + r] = x[q] w[r] + x[q-1] w[r + f]   (GEMM row q = 0..Lin, phase r = 0..f-1)
      const int taps[1] = {2}, cin[1] = {sp.Cin};
      (void)P;
      if (int rc = pack_tc_layer(b16, bb, sp.tc_up, 1, taps, cin, Cout, f,
                                 [&](int, int j, int ci, int ph, int co) {
                                   return w[((size_t)ci * Cout + co) * (2 * f) + ph + f * j];
                                 },
                                 [&](int co) { return b[co]; }, plan->ctx.conv_persistent == 2))
        return rc;
      ti += 2;
    }
    const int C = sp.Cout, k = cfg.res_kernel, nb = cfg.num_res_blocks;
    const int ti_blocks = ti, ti_short = ti + 4 * nb;
    for (int bi = 0; bi < nb; ++bi) {
      BlockPlan& bp = sp.blocks[bi];
      max_d = std::max(max_d, (k - 1) / 2 * bp.dilation);
      const float *w1 = tensors[ti_blocks + 4 * bi], *b1 = tensors[ti_blocks + 4 * bi + 1];
      const float *w2 = tensors[ti_blocks + 4 * bi + 2], *b2 = tensors[ti_blocks + 4 * bi + 3];
      const float *ws = tensors[ti_short + 2 * bi], *bs = tensors[ti_short + 2 * bi + 1];
      {
        const int taps[1] = {k}, cin[1] = {C};
        if (int rc = pack_tc_layer(b16, bb, bp.tc_conv1, 1, taps, cin, C, 1,
                                   [&](int, int j, int ci, int, int co) { return w1[((size_t)co * C + ci) * k + j]; },
                                   [&](int co) { return b1[co]; }))
          return rc;
      }
      if (C == 24 && k == 3) {
        // Five-K-step packing of the dilated conv for the fused stack's k24 mode: K-step ks, half h = one 8-channel
        // plane of one tap, [ks][h][32 columns][8]; the last half is the bias row against the kernel's ones plane
        // (K-row 0 = bf16 high part, K-row 1 = remainder, as in the bias MMAs).  Sized like the plain packing (3
        // taps x 32 x 32) so the kernel's loads do not change.
        static const int tap_of[5][2] = {{0, 0}, {1, 0}, {1, 1}, {2, 2}, {2, -1}};
        static const int ch_of[5][2] = {{0, 8}, {0, 16}, {8, 16}, {0, 8}, {16, 0}};
        bp.has_k24 = true;
        bp.w1_k24_off = b16.alloc((size_t)3 * 32 * 32);
        for (int ks = 0; ks < 5; ++ks)
          for (int h = 0; h < 2; ++h)
            for (int n = 0; n < C; ++n) {
              __nv_bfloat16* dst = &b16.data[bp.w1_k24_off + ((size_t)(ks * 2 + h) * 32 + n) * 8];
              if (tap_of[ks][h] < 0) {
                const __nv_bfloat16 hi = __float2bfloat16(b1[n]);
                dst[0] = hi;
                dst[1] = __float2bfloat16(b1[n] - __bfloat162float(hi));
              } else {
                for (int e = 0; e < 8; ++e)
                  dst[e] = __float2bfloat16(w1[((size_t)n * C + ch_of[ks][h] + e) * k + tap_of[ks][h]]);
              }
            }
      }
      {
        const int taps[2] = {1, 1}, cin[2] = {C, C};
        if (int rc = pack_tc_layer(b16, bb, bp.tc_conv2s, 2, taps, cin, C, 1,
                                   [&](int src, int, int ci, int, int co) {
                                     return (src == 0 ? w2 : ws)[(size_t)co * C + ci];
                                   },
                                   [&](int co) { return b2[co] + bs[co]; }))
          return rc;
      }
    }
    ti += 6 * nb;
    clast = C;
  }
  {
    const float *w = tensors[ti], *b = tensors[ti + 1];
    const int taps[1] = {cfg.proj_kernel}, cin[1] = {clast};
    const int k = cfg.proj_kernel;
    if (int rc = pack_tc_layer(b16, bb, plan->tc_head, 1, taps, cin, cfg.out_channels, 1,
                               [&](int, int j, int ci, int, int co) { return w[((size_t)co * clast + ci) * k + j]; },
                               [&](int co) { return b[co]; }))
      return rc;
    // Operand of the head fused into the last stack kernel (stack_bf16_tc.cu, TAIL = 1): taps packed on N, column
    // j * out + c = tap j of output channel c, K-major [cp/8][n][8].
    if (k == 7 && (cfg.out_channels == 1 || cfg.out_channels == 4)) {
      const int cp = pad16(clast), n = pad16(k * cfg.out_channels);
      plan->head_ts_n = n;
      plan->head_ts_off = b16.alloc((size_t)cp * n);
      for (int ci = 0; ci < clast; ++ci)
        for (int j = 0; j < k; ++j)
          for (int c = 0; c < cfg.out_channels; ++c)
            b16.data[plan->head_ts_off + ((size_t)(ci >> 3) * n + j * cfg.out_channels + c) * 8 + (ci & 7)] =
                __float2bfloat16(w[((size_t)c * clast + ci) * k + j]);
    }
  }
  const int half = (cfg.res_kernel - 1) / 2;
  int sum_d = 0;
  for (int bi = 0; bi < cfg.num_res_blocks; ++bi) sum_d += half * (int)ipow(cfg.res_kernel, bi);
  plan->halo = (int)round_up(std::max(std::max(max_d, sum_d), 1), 8);
  return 0;
}

// Fused-stack configuration for stage s; kt == 0 when the stage must use the per-conv kernels.  With
// `with_head` the output head (LeakyReLU, ReflectionPad1d, Conv1d, Tanh) is folded into the last stage's kernel.
TcStackParams stack_tiling(const ttsb200_plan* plan, int s, bool with_head = false, bool with_up = false) {
  const ttsb200_generator_config& cfg = plan->cfg;
  TcStackParams q{};
  q.cp = pad16(plan->stages[s].Cout);
  q.nblocks = cfg.num_res_blocks;
  q.res_taps = cfg.res_kernel;
  if (q.nblocks < 1 || q.nblocks > TC_STACK_MAX_BLOCKS) return q;
  for (int bi = 0; bi < q.nblocks; ++bi) q.dil[bi] = plan->stages[s].blocks[bi].dilation;
  q.k24 = (q.cp == 32 && plan->stages[s].blocks[0].has_k24 && !plan->ctx.no_k24) ? 1 : 0;
  if (with_head) {
    const TcGroupPlan& g = plan->tc_head.groups[0];
    q.head_w = plan->blob16 + plan->head_ts_off;
    q.head_b = plan->blob + g.bias_off;
    q.head_taps = cfg.proj_kernel;
    q.head_ch = cfg.out_channels;
    q.head_n = plan->head_ts_n;
  }
  if (with_up) {   // the next stage's ConvTranspose1d rides on this stack
    const StagePlan& nx = plan->stages[s + 1];
    const TcGroupPlan& g = nx.tc_up.groups[0];
    q.up_w = plan->blob16 + g.w_off;
    q.up_b = plan->blob + g.bias_off;
    q.up_f = nx.f;
    q.up_P = nx.P;
    q.up_n = g.ncols;
  }
  if (!tc_stack_configure(q)) q.kt = 0;
  return q;
}
// Stage s's stack can also run stage s+1's transposed conv when all its columns fit one accumulator.
bool up_fusable(const ttsb200_plan* plan, int s) {
  if (s + 1 >= plan->cfg.n_stages || plan->cfg.num_res_blocks < 1) return false;
  if (plan->stages[s + 1].tc_up.groups.size() != 1) return false;
  // the fused kernel writes a margin-free tensor: its consumer must be the next stage's fused stack
  if (stack_tiling(plan, s + 1).kt == 0) return false;
  return stack_tiling(plan, s, false, true).kt > 0;
}
// The head rides on the last stage's fused kernel when that kernel exists, the head has a packed-tap operand and
// no per-layer taps were asked for (the stack output is then never materialised).
bool head_fusable(const ttsb200_plan* plan) {
  const ttsb200_generator_config& cfg = plan->cfg;
  const int s = cfg.n_stages - 1;
  return plan->head_ts_n > 0 && plan->tc_head.groups.size() == 1 && stack_tiling(plan, s, true).kt > 0;
}

struct CanonDims {
  int planes, rows;
  size_t elems;  // per batch item
};
inline CanonDims canon_dims(const ttsb200_plan* plan, int C, int64_t L) {
  CanonDims d;
  d.planes = pad16(C) / 8;
  d.rows = (int)round_up(L, TC_TILE_M) + 2 * plan->halo + 128 * TC_STACK_MAX_KT;  // slack: whole-tile bulk loads
  d.elems = (size_t)d.planes * d.rows * 8;
  return d;
}

// Elements between consecutive utterances of a canonical [C, L] tensor.
inline size_t act_stride(const ttsb200_plan* plan, int C, int64_t L) {
  return (size_t)round_up((int64_t)canon_dims(plan, C, L).elems, 128);
}

// Workspace: the canonical mel, then ONE region per activation tensor of the schedule (no ping-pong: with the
// hand-over per time segment, follow.cuh, several kernels of the chain are in flight on the same utterance, so a
// kernel must never write where an older one may still read), a scratch tensor for the per-conv fallback, the fp32
// head output and the hand-over counters.  The tensors grow 2x to 8x per stage, so their sum is ~1.3 x the largest
// one times the three ping-pong buffers there used to be.
struct Bf16Workspace {
  size_t mel_elems, act_elems, head_floats, total_bytes;
  int n_counters;   // [counting launch][B][FOLLOW_SEGS], zeroed by the first kernel of a forward
};
Bf16Workspace bf16_workspace(const ttsb200_plan* plan, int B, const Shape& sh) {
  const ttsb200_generator_config& cfg = plan->cfg;
  Bf16Workspace w{};
  w.mel_elems = (size_t)round_up((int64_t)B * canon_dims(plan, cfg.in_channels, sh.L0).elems, 128);
  // projection output; per stage the transposed conv's output and the stack's (or, per-conv fallback, the second
  // ping-pong tensor); one scratch tensor (h of the per-conv fallback) as large as the largest
  size_t sum = act_stride(plan, cfg.base_channels, sh.L0), mx = sum;
  for (size_t s = 0; s < plan->stages.size(); ++s) {
    const size_t e = act_stride(plan, plan->stages[s].Cout, sh.L[s]);
    sum += 2 * e;
    mx = std::max(mx, e);
  }
  w.act_elems = (size_t)B * (sum + mx);
  w.head_floats = (size_t)round_up((int64_t)B * cfg.out_channels * sh.L.back(), 64);
  w.n_counters = (int)round_up((int64_t)(2 * cfg.n_stages + 2) * B * FOLLOW_SEGS, 64);
  w.total_bytes = (w.mel_elems + w.act_elems) * sizeof(__nv_bfloat16) + w.head_floats * sizeof(float) +
                  (size_t)(w.n_counters + 64) * sizeof(int);   // + the armed flag (follow.cuh follow_armed)
  return w;
}

int forward_bf16(const ttsb200_plan* plan, const ttsb200_forward_args* a) {
  const ttsb200_generator_config& cfg = plan->cfg;
  if (cfg.proj_kernel > TC_MAX_TAPS || cfg.res_kernel > TC_MAX_TAPS)
    return fail(TTSB200_ERR_INVALID, "BF16 path supports kernel sizes up to 7");
  const Shape sh = shapes_for(plan, a->frames, a->pad);
  if (int rc = check_lengths(plan, sh)) return rc;
  const int B = a->batch, halo = plan->halo;
  const Bf16Workspace wsz = bf16_workspace(plan, B, sh);
  if (!a->workspace || a->workspace_bytes < wsz.total_bytes)
    return fail(TTSB200_ERR_WORKSPACE, "workspace smaller than ttsb200_workspace_bytes()");
  __nv_bfloat16* melc = static_cast<__nv_bfloat16*>(a->workspace);
  __nv_bfloat16* act = melc + wsz.mel_elems;     // bump-allocated: one region per activation tensor
  size_t act_used = 0;
  auto fresh = [&](int C, int64_t L) {
    __nv_bfloat16* p = act + act_used;
    act_used += (size_t)B * act_stride(plan, C, L);
    return p;
  };
  __nv_bfloat16* cur = nullptr;                  // the tensor the next launch reads
  float* headf = reinterpret_cast<float*>(act + wsz.act_elems);
  int* counters = reinterpret_cast<int*>(headf + wsz.head_floats);
  // Hand-over between consecutive fused kernels segment by segment (follow.cuh follow_wait): the producer's
  // counters and what they reach when an utterance is complete; null when the last kernel does not count.
  TcFollow prod{};   // how the last launch counts its finished work; done == nullptr: it does not
  const bool follow_ok = !plan->ctx.no_follow;
  // follow.cuh follow_armed: a unique id per inference, or — under stream capture, where a replay would carry the same
  // id every time — the 0 / 1 form that holds the dependent launches back
  cudaStreamCaptureStatus capture = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(static_cast<cudaStream_t>(a->stream), &capture);
  const int armed_early = capture == cudaStreamCaptureStatusNone ? 1 : 0;
  int armed_id = 1;
  if (armed_early) {
    static std::atomic<uint32_t> next_id{(uint32_t)std::chrono::steady_clock::now().time_since_epoch().count() * 2654435761u};
    do armed_id = (int)next_id.fetch_add(1, std::memory_order_relaxed); while (armed_id == 0 || armed_id == 1);
  }
  int n_counting = 0;   // counting launches so far: each gets its own [B][FOLLOW_SEGS] block of counters
  auto counting = [&](int n_units, int unit_rows, int row_offset, int mult, int** done_out, int* units_per_seg) {
    *done_out = counters + (size_t)(n_counting++) * B * FOLLOW_SEGS;
    *units_per_seg = ceil_div(n_units, FOLLOW_SEGS);
    return TcFollow{*done_out, unit_rows, row_offset, *units_per_seg, n_units, mult, counters + wsz.n_counters,
                    armed_id, armed_early};
  };
  cudaStream_t st = static_cast<cudaStream_t>(a->stream);
  Launcher ln{st, a->profile, B};
  const float* Wf = plan->blob;
  const __nv_bfloat16* Wh = plan->blob16;
  const float slope = cfg.act_slope;
  const int pp = (cfg.proj_kernel - 1) / 2;
  const int S = cfg.n_stages, nb = cfg.num_res_blocks;

  auto base = [&](const TcLayer& L, int64_t Lt) {
    TcConvParams p{};
    p.nsrc = 1;
    p.ngroups = (int)L.groups.size();
    for (int g = 0; g < p.ngroups; ++g) {
      const TcGroupPlan& gp = L.groups[g];
      p.grp[g] = TcGroup{Wh + gp.w_off, Wf + gp.bias_off, gp.ncols, gp.col_first};
    }
    p.Lt = (int)Lt; p.up_f = L.up_f; p.up_P = 0; p.cout_real = L.cout; p.slope = slope;
    return p;
  };
  auto src = [&](const __nv_bfloat16* x, int C, int64_t L, int taps, int step, int first, bool lrelu) {
    const CanonDims d = canon_dims(plan, C, L);
    TcSrc s{};
    s.x = x; s.batch_stride = x == melc ? (int64_t)d.elems : (int64_t)act_stride(plan, C, L);
    s.planes = d.planes; s.plane_rows = d.rows;
    const int lo = std::min(first, first + (taps - 1) * step);
    const int hi = std::max(first, first + (taps - 1) * step);
    s.row_base = halo + lo; s.rows = TC_TILE_M + (hi - lo); s.taps = taps;
    for (int j = 0; j < taps; ++j) s.tap_row[j] = first + j * step - lo;
    s.lrelu = lrelu ? 1 : 0;
    return s;
  };
  auto canon_out = [&](TcConvParams& p, __nv_bfloat16* out, int C, int64_t L, bool lrelu, int margin) {
    const CanonDims d = canon_dims(plan, C, L);
    p.out_mode = 0; p.out_c = out; p.out_batch_stride = (int64_t)act_stride(plan, C, L); p.out_plane_rows = d.rows;
    p.out_halo = halo; p.out_L = (int)L; p.out_lrelu = lrelu ? 1 : 0; p.margin = margin;
  };
  int trace_seq = 0;
  auto run = [&](const char* name, int layer, TcConvParams p, double macs, double io_elems) -> int {
    p.trace_id = layer * 10 + trace_seq++ % 10;   // label for the debug timeline (trace.cuh)
    if (tc_conv_smem_bytes(p) > 224 * 1024)
      return fail(TTSB200_ERR_INVALID, std::string(name) + ": tile does not fit shared memory on the BF16 path");
    // the persistent kernel counts finished units per (utterance, time segment) and follows a counting producer
    int n_units = 0, unit_rows = 0, row_offset = 0, mult = 0;
    const bool cnt = follow_ok && tc_conv_count_info(p, B, plan->ctx, &n_units, &unit_rows, &row_offset, &mult);
    TcFollow mine{};
    if (cnt) {
      p.follow = prod;                                 // done == nullptr: wait for the whole previous grid
      p.follow_in_lo = p.src[0].row_base - halo;       // sequence row of staged row 0 at t0 = 0
      mine = counting(n_units, unit_rows, row_offset, mult, &p.done_out, &p.count_units_per_seg);
    }
    TTSB_CUDA(ln.run(name, layer, 2 * macs, 2.0 * io_elems, [&] { return tc_conv_launch(p, B, plan->ctx, st); }));
    prod = mine;
    return 0;
  };
  auto tap = [&](int layer, const __nv_bfloat16* x, int C, int64_t L, bool undo, int x_halo) -> int {
    for (int i = 0; i < a->n_taps; ++i)
      if (a->tap_layers[i] == layer) {
        const CanonDims d = canon_dims(plan, C, L);
        TTSB_CUDA(tc_untile_launch(x, (int64_t)act_stride(plan, C, L), B, C, L, d.rows, x_halo, undo, slope, a->tap_out[i], st));
      }
    return 0;
  };

  {  // mel -> canonical bf16 (replicate inference padding + reflect margin folded in)
    const CanonDims d = canon_dims(plan, cfg.in_channels, sh.L0);
    TTSB_CUDA(ln.run("bf16.prep.mel", 0, 0.0, 4.0 * B * cfg.in_channels * a->frames + 2.0 * B * d.elems, [&] {
      return tc_prep_mel_launch(a->mel, B, cfg.in_channels, a->frames, a->pad, melc, d.planes, d.rows, halo, counters,
                                wsz.n_counters, armed_id, armed_early, st);
    }));
  }
  {
    TcConvParams p = base(plan->tc_proj, sh.L0);
    p.src[0] = src(melc, cfg.in_channels, sh.L0, cfg.proj_kernel, 1, -pp, false);
    cur = fresh(cfg.base_channels, sh.L0);
    canon_out(p, cur, cfg.base_channels, sh.L0, true, 2);
    const double macs = (double)cfg.proj_kernel * cfg.in_channels * cfg.base_channels * sh.L0 * B;
    if (int rc = run("bf16.proj.k7", 1, p, macs,
                     (double)B * (cfg.in_channels + cfg.base_channels) * sh.L0 +
                         (double)cfg.proj_kernel * cfg.in_channels * cfg.base_channels))
      return rc;
    if (int rc = tap(1, cur, cfg.base_channels, sh.L0, true, halo)) return rc;
  }
  int64_t Lin = sh.L0;
  bool head_done = false, up_done = false;
  int cur_halo = halo;   // plane row of sequence row 0 in `cur` (a fused transposed conv may shift it by one)
  for (int s = 0; s < S; ++s) {
    const StagePlan& sp = plan->stages[s];
    const int64_t L = sh.L[s];
    const int C = sp.Cout;
    const bool last_stage = s == S - 1;
    if (up_done) {
      up_done = false;   // the previous stage's stack already produced this stage's input in `cur`
      if (int rc = tap(sp.layer_up, cur, C, L, nb == 0, cur_halo)) return rc;
    } else {
      TcConvParams p = base(sp.tc_up, Lin + 1);               // GEMM rows q = 0..Lin
      p.up_P = sp.P;
      p.src[0] = src(cur, sp.Cin, Lin, 2, -1, 0, false);      // taps read x[q], x[q - 1] (zero rows at -1 and Lin)
      // consumer: block 0 (raw x + reflect) or, with no blocks, the next transposed conv / head
      // consumer: the fused stack (mirrors its own edges: no margins), else block 0 (raw x + reflect) or, with
      // no blocks, the next transposed conv / head
      __nv_bfloat16* nxt = fresh(C, L);
      if (nb > 0) canon_out(p, nxt, C, L, false, stack_tiling(plan, s).kt > 0 ? 0 : 1);
      else canon_out(p, nxt, C, L, true, last_stage ? 1 : 2);
      // a fused stack reads its input at any row offset: shift by a row where that makes the transposed conv's
      // output-row pairs 32-byte aligned (conv_bf16_tc.cu epilogue)
      if (nb > 0 && sp.f % 2 == 0 && stack_tiling(plan, s).kt > 0) p.out_halo = halo + ((halo - sp.P) & 1);
      cur_halo = p.out_halo;
      const double macs = 2.0 * sp.f * sp.Cin * C * Lin * B;
      if (int rc = run("bf16.upsample.convT", sp.layer_up, p, macs,
                       (double)B * (sp.Cin * Lin + C * L) + 2.0 * sp.f * sp.Cin * C))
        return rc;
      cur = nxt;
      if (int rc = tap(sp.layer_up, cur, C, L, nb == 0, cur_halo)) return rc;
    }
    const int half = (cfg.res_kernel - 1) / 2;
    // A tap of this stage's stack output needs that tensor in HBM, which the fused tails never write; every other
    // tap (projection, transposed-conv outputs, head output) reads a tensor the fused schedule materialises anyway.
    bool want_stack = false;
    for (int i = 0; i < a->n_taps; ++i) want_stack |= a->tap_layers[i] == sp.layer_stack;
    const bool fuse_head = last_stage && !want_stack && head_fusable(plan);
    const bool fuse_up = !last_stage && !want_stack && up_fusable(plan, s);
    TcStackParams q = stack_tiling(plan, s, fuse_head, fuse_up);
    if (q.kt > 0) {  // whole ResidualStack in one kernel
      const CanonDims din = canon_dims(plan, C, L);
      q.x = cur; q.x_batch_stride = (int64_t)act_stride(plan, C, L); q.x_plane_rows = din.rows; q.x_halo = cur_halo;
      cur_halo = halo;
      q.L = (int)L; q.slope = slope;
      for (int bi = 0; bi < nb; ++bi) {
        const BlockPlan& bp = sp.blocks[bi];
        q.w1[bi] = Wh + (q.k24 ? bp.w1_k24_off : bp.tc_conv1.groups[0].w_off);
        q.w2s[bi] = Wh + bp.tc_conv2s.groups[0].w_off;
        q.b1[bi] = Wf + bp.tc_conv1.groups[0].bias_off;
        q.b2s[bi] = Wf + bp.tc_conv2s.groups[0].bias_off;
      }
      // its output: the next stage's input when the transposed conv rides along, else this stage's stack output
      // (never written when the head rides along); always a region of its own (see Bf16Workspace)
      __nv_bfloat16* nxt = fuse_head ? nullptr : fuse_up ? fresh(plan->stages[s + 1].Cout, sh.L[s + 1]) : fresh(C, L);
      q.out = TcCanonOut{nxt, (int64_t)act_stride(plan, C, L), din.rows, halo, (int)L, 1, last_stage ? 1 : 2};
      if (follow_ok) q.follow = prod;   // follow the previous kernel segment by segment when it counted its work
      TcFollow mine{};
      if (follow_ok) {
        int n_units = 0, unit_rows = 0, row_offset = 0;
        tc_stack_count_info(q, &n_units, &unit_rows, &row_offset);
        mine = counting(n_units, unit_rows, row_offset, 1, &q.done_out, &q.count_units_per_seg);
      }
      double macs = (double)(cfg.res_kernel + 2) * nb * C * C * L * B;
      double bytes = 2.0 * (2.0 * B * C * L + (double)(cfg.res_kernel + 2) * nb * C * C);
      if (fuse_head) {
        const bool to_out = cfg.pqmf_bands == 0 || a->skip_pqmf;
        q.head_out = (to_out && a->out_dtype == TTSB200_OUT_F32 && a->out_keep_len == 0) ? static_cast<float*>(a->out) : headf;
        macs += (double)cfg.proj_kernel * C * cfg.out_channels * L * B;
        bytes = 2.0 * B * C * L + 4.0 * B * cfg.out_channels * L;
        head_done = true;
      }
      if (fuse_up) {
        const StagePlan& nx = plan->stages[s + 1];
        const CanonDims dn = canon_dims(plan, nx.Cout, sh.L[s + 1]);
        // rows q f - P .. of one GEMM row go out as 32-byte stores: keep (halo - P) even
        cur_halo = halo + ((halo - nx.P) & 1);
        q.up_out = TcCanonOut{nxt, (int64_t)act_stride(plan, nx.Cout, sh.L[s + 1]), dn.rows, cur_halo, (int)sh.L[s + 1], 0, 0};
        macs += 2.0 * nx.f * C * nx.Cout * L * B;
        bytes = 2.0 * B * (C * L + (double)nx.Cout * sh.L[s + 1]);
        up_done = true;
      }
      TTSB_CUDA(ln.run(fuse_head ? "bf16.resstack.fused+head" : fuse_up ? "bf16.resstack.fused+convT" : "bf16.resstack.fused",
                       sp.layer_stack, 2 * macs, bytes, [&] { return tc_stack_launch(q, B, plan->ctx, st); }));
      prod = mine;
      cur = nxt;
    } else {
    // per-conv fallback: every kernel waits for the whole previous grid, so ping-pong between the stage's input and
    // one more tensor of the same shape is safe; h goes to a scratch tensor
    __nv_bfloat16* nxt = nb > 0 ? fresh(C, L) : nullptr;
    __nv_bfloat16* hbuf = act + wsz.act_elems - (size_t)B * act_stride(plan, C, L);   // the scratch region is the largest
    for (int bi = 0; bi < nb; ++bi) {
      const BlockPlan& bp = sp.blocks[bi];
      const int d = bp.dilation;
      {
        TcConvParams p = base(bp.tc_conv1, L);
        p.src[0] = src(cur, C, L, cfg.res_kernel, d, -half * d, true);
        canon_out(p, hbuf, C, L, true, 0);
        if (int rc = run("bf16.resblock.conv1_dilated", sp.layer_stack, p, (double)cfg.res_kernel * C * C * L * B,
                         2.0 * B * C * L + (double)cfg.res_kernel * C * C))
          return rc;
      }
      {
        TcConvParams p = base(bp.tc_conv2s, L);
        p.nsrc = 2;
        p.src[0] = src(hbuf, C, L, 1, 1, 0, false);
        p.src[1] = src(cur, C, L, 1, 1, 0, false);
        const bool last_block = bi == nb - 1;
        if (!last_block) canon_out(p, nxt, C, L, false, 1);
        else canon_out(p, nxt, C, L, true, last_stage ? 1 : 2);
        if (int rc = run("bf16.resblock.conv2+shortcut", sp.layer_stack, p, 2.0 * C * C * L * B,
                         3.0 * B * C * L + 2.0 * C * C))
          return rc;
        std::swap(cur, nxt);
      }
    }
    }
    if (nb > 0)
      if (int rc = tap(sp.layer_stack, cur, C, L, true, halo)) return rc;
    Lin = L;
  }
  {
    const int Cl = plan->stages.back().Cout;
    const bool to_out = cfg.pqmf_bands == 0 || a->skip_pqmf;
    float* dst = (to_out && a->out_dtype == TTSB200_OUT_F32 && a->out_keep_len == 0) ? static_cast<float*>(a->out) : headf;
    const int ih = 4 + 3 * S;
    if (!head_done) {
      TcConvParams p = base(plan->tc_head, Lin);
      p.src[0] = src(cur, Cl, Lin, cfg.proj_kernel, 1, -pp, false);
      p.out_mode = 1; p.out_f = dst; p.out_L = (int)Lin;
      if (int rc = run("bf16.head.k7+tanh", ih, p, (double)cfg.proj_kernel * Cl * cfg.out_channels * Lin * B,
                       (double)B * (Cl + 2.0 * cfg.out_channels) * Lin))
        return rc;
    }
    if (int rc = copy_tap(a, ih + 1, dst, cfg.out_channels, Lin, st)) return rc;
    // the PQMF kernel follows a counting head kernel segment by segment
    if (int rc = finish_output(plan, a, ln, dst, to_out, Lin, head_done && prod.done ? &prod : nullptr)) return rc;
  }
  return 0;
}
