#!/usr/bin/env python
"""Headline benchmark: MB-MelGAN + PQMF batched inference, audio samples/s (x real-time).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--dtype fp32|bf16] [--impl reference]
                    [--workload batched|melgan_cfg1|longform|streaming]

A step = generator.inference() over one batch of 16 x 10 s synthetic mels (BASELINE.json config 2:
MultibandMelganGenerator() defaults, 22,050 Hz, hop 256, 862 frames, inference_padding 2) per GPU.
N > 1 (torchrun): every rank runs the same per-GPU batch on its own utterances — independent
units, no data-path collective ("weak" scaling).  Prints ONE JSON line on rank 0.

What the line says and against which peak (MEASURED_PEAKS.json, driver-written):
  value / ms_per_step   CUDA events on the launch stream around exactly K steps, barrier + synchronize on both sides,
                        max over ranks.  Inputs rotate through a pool larger than L2.  K steps of 0.3 ms are a BURST
                        measurement (the SM clock never leaves its maximum), so `roofline` and `roofline_by_kernel`
                        divide by the BURST peaks (bf16_tflops, hbm_gbs).
  sustained             the same step looped for >= 3 s with no host sync inside, SM clock sampled every few ms during
                        the loop through NVML; its fraction divides by bf16_tflops_sustained.
  roofline_by_kernel    every launch of a step: live CUDA-event time, algorithmic FLOPs / bytes, bound, fraction —
                        tensor kernels against the BF16 peak, PQMF / mel staging as GB/s against the HBM peak.
  fp32                  the reference-precision path (true fp32 FFMA): value, max-abs error against the CPU oracle,
                        and its fraction of the FFMA rate measured on this GPU (ttsb200_diag_ffma_tflops).
  e2e                   public API with pinned HOST buffers, H2D + D2H inside the timed region (fp32 waveform), plus
                        the fp16-waveform variant, the reference callers' blocking pattern, and the measured host-link
                        ceiling of the box.
  cpu_baseline          the UNMODIFIED reference modules (oracle/_ref, staged by oracle/make_ref.py) on the host cores:
                        all cores and one thread, CPU model string, plus cfg1 and a cfg5 chunk.
`--impl reference` prints the reference arm's line (same metric, config label and unit).
"""
import argparse
import json
import os
import platform
import subprocess
import sys
import threading
import time
import warnings

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
warnings.filterwarnings("ignore")

import torch  # noqa: E402

SAMPLE_RATE = 22050
FRAMES = 862          # 10 s at hop 256
BATCH = 16
PAD = 2
WEIGHT_SEED = 54321   # TTS/utils/training.py:8
MEL_SEED = 1234
L2_BYTES = 126 * 1024 * 1024
WORKLOAD_CFG2 = ("BASELINE config 2: MultibandMelganGenerator() defaults (C0 384, factors 2,8,2,2, 3 res blocks) + PQMF "
                 "62-tap synthesis, 22.05 kHz hop 256, batch 16 x 862 frames (10 s) per GPU, inference_padding 2, "
                 "seeded random-init weights")


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "src": "measured"}
    # B200_PROFILING.md fallback figures
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "src": "fallback"}


class ClockSampler:
    """SM clock / power / throttle reasons DURING a timed region, sampled through NVML every `period_s` (a 7 ms region
    needs millisecond sampling; nvidia-smi -lms cannot go below ~100 ms).  Falls back to nvidia-smi without NVML."""
    REASONS = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40,
               "hw_power_brake_slowdown": 0x80, "sw_power_cap": 0x4}

    def __init__(self, index, period_s=0.002):
        self.index, self.period, self.rows, self.stop_flag, self.thread = index, period_s, [], False, None
        self.nv = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(self._nvml_index(index))
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None

    @staticmethod
    def _nvml_index(index):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            try:
                return int(vis.split(",")[index])
            except Exception:
                pass
        return index

    def _loop(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                mhz = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                watts = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
                self.rows.append((time.perf_counter(), mhz, mask, watts))
            except Exception:
                pass
            time.sleep(self.period)

    def start(self):
        self.rows, self.stop_flag = [], False
        self.t0 = time.perf_counter()
        if self.nv is not None:
            self.thread = threading.Thread(target=self._loop, daemon=True)
            self.thread.start()
        return self

    def stop(self):
        self.t1 = time.perf_counter()
        if self.nv is None:
            return self._smi_once()
        self.stop_flag = True
        self.thread.join(timeout=1.0)
        rows = [r for r in self.rows if self.t0 <= r[0] <= self.t1]
        if not rows:                                  # region shorter than one sampling period: sample right now
            try:
                nv = self.nv
                rows = [(self.t1, float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)),
                         int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)),
                         nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)]
            except Exception:
                return self._smi_once()
        mhz = sorted(r[1] for r in rows)
        mask = 0
        for r in rows:
            mask |= r[2]
        return {"sm_mhz": mhz[len(mhz) // 2], "sm_mhz_min": mhz[0], "sm_max_mhz": self.max_mhz,
                "reasons": sorted(n for n, b in self.REASONS.items() if mask & b),
                "power_w_max": round(max(r[3] for r in rows), 1), "samples": len(rows),
                "period_ms": self.period * 1e3, "source": "nvml"}

    def _smi_once(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}",
                                  "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=10).stdout
            r = [c.strip() for c in out.strip().splitlines()[0].split(",")]
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            return {"sm_mhz": float(r[0]), "sm_max_mhz": float(r[1]), "power_w_max": float(r[2]),
                    "reasons": sorted(n for n, v in zip(names, r[3:7]) if v.lower().startswith("active")),
                    "samples": 1, "source": "nvidia-smi (single sample after the region)"}
        except Exception:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["clock query unavailable"], "samples": 0}


def macs_per_item(L0, in_ch=80, out_ch=4, c0=384, factors=(2, 8, 2, 2), n_blocks=3, k=7, rk=3):
    """7*80*C0*L0 + sum_i [2 f_i C_{i-1} C_i L_{i-1} + 5 n C_i^2 L_i] + 7*C_last*out*L_last."""
    L, c, total = L0, c0, k * in_ch * c0 * L0
    for f in factors:
        total += 2 * f * c * (c // 2) * L
        L, c = L * f, c // 2
        total += (rk + 2) * n_blocks * c * c * L
    return total + k * c * out_ch * L


def make_model(dtype, cls="MultibandMelganGenerator", **kw):
    import importlib
    mod = {"MultibandMelganGenerator": "multiband_melgan_generator", "MelganGenerator": "melgan_generator",
           "FullbandMelganGenerator": "fullband_melgan_generator"}[cls]
    C = getattr(importlib.import_module(f"tts_b200.vocoder.models.{mod}"), cls)
    torch.manual_seed(WEIGHT_SEED)
    m = C(**kw)
    m.remove_weight_norm()
    m.inference_padding = PAD
    m.precision = dtype
    return m


# ---------------------------------------------------------------------------------------------------------------
# CPU side: the reference's own modules (oracle/_ref) when staged, else the oracle port.  Checker / baseline only.
# ---------------------------------------------------------------------------------------------------------------
def cpu_model_string():
    try:
        for line in open("/proc/cpuinfo"):
            if line.lower().startswith("model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return platform.processor() or "unknown"


def cpu_generator(cls="MultibandMelganGenerator", **kw):
    """(run(mel) -> waveform, kind): the reference class from oracle/_ref seeded like the GPU arm, or the port."""
    from oracle import make_ref
    classes = make_ref.load()
    if classes is not None:
        torch.manual_seed(WEIGHT_SEED)
        m = classes[cls](**kw)
        m.remove_weight_norm()
        m.eval()
        m.inference_padding = PAD

        def run(mel):
            with torch.no_grad():
                return m.inference(mel)
        return run, "reference"
    from oracle.melgan_oracle import GeneratorSpec, generator_inference
    spec = {"MultibandMelganGenerator": GeneratorSpec.multiband, "MelganGenerator": GeneratorSpec.melgan,
            "FullbandMelganGenerator": GeneratorSpec.fullband}[cls](**kw)
    ours = make_model("fp32", cls, **kw)
    sd = {k: v.detach() for k, v in ours.state_dict().items()}
    return (lambda mel: generator_inference(spec, sd, mel, PAD)), "port"


def time_cpu(run, mel, steps, warmup):
    for _ in range(warmup):
        y = run(mel)
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter()
        y = run(mel)
        ts.append(time.perf_counter() - t0)
    return ts, y


def cpu_reference(steps, warmup, budget_s=100.0):
    """Reference arm / cpu_baseline: the reference generator on all host cores, a bounded sample of config 2."""
    run, kind = cpu_generator()
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(MEL_SEED)
    one = torch.randn(1, 80, FRAMES)
    run(one)
    t0 = time.perf_counter(); run(one); t1 = time.perf_counter() - t0
    items = int(max(1, min(BATCH, budget_s / max(1e-3, (steps + warmup) * t1))))
    mel = torch.randn(items, 80, FRAMES)
    ts, y = time_cpu(run, mel, steps, warmup)
    dt = sum(ts)
    samples = steps * y.shape[0] * y.shape[-1]
    what = ("the UNMODIFIED reference modules staged under oracle/_ref (TTS/vocoder/models/multiband_melgan_generator.py "
            ".inference)" if kind == "reference" else "oracle/melgan_oracle.py port (oracle/_ref not staged)")
    return {"value": samples / dt, "unit": "samples/s", "cores": torch.get_num_threads(), "kind": kind,
            "sample": f"{items} of {BATCH} utterances x {FRAMES} frames per step, {steps} steps; {what}, "
                      f"torch {torch.__version__} CPU kernels",
            "ms_per_step": dt / steps * 1e3, "items": items, "cpu_model": cpu_model_string(),
            "os_cpu_count": cores}


def cpu_extras(budget_s=12.0):
    """one-thread line for cfg2, cfg1 at B=1 (all cores / one thread), one cfg5 chunk window."""
    out = {}
    cores = os.cpu_count() or 1
    run2, kind = cpu_generator()
    torch.manual_seed(MEL_SEED)
    one = torch.randn(1, 80, FRAMES)
    torch.set_num_threads(1)
    ts, y = time_cpu(run2, one, 2, 1)
    out["cfg2_one_thread"] = {"value": y.shape[-1] / min(ts), "unit": "samples/s", "cores": 1,
                              "sample": "1 utterance x 862 frames, best of 2", "ms": min(ts) * 1e3}
    run1, _ = cpu_generator("MelganGenerator")
    mel1 = torch.randn(1, 80, 200)
    for n, key in ((cores, "cfg1_all_cores"), (1, "cfg1_one_thread")):
        torch.set_num_threads(n)
        ts, y = time_cpu(run1, mel1, 5 if n > 1 else 2, 1)
        ts.sort()
        out[key] = {"value": y.shape[-1] / ts[0], "unit": "samples/s", "cores": n, "ms_best": ts[0] * 1e3,
                    "ms_median": ts[len(ts) // 2] * 1e3, "sample": "MelganGenerator() on [1,80,200], inference_padding 2"}
    # cfg5: the steady-state streaming window (context + 20 frames + context) of FB-MelGAN (5,5,4,3)
    from tts_b200.longform import receptive_halo_frames
    fb = make_model("fp32", "FullbandMelganGenerator", upsample_factors=(5, 5, 4, 3))
    ctx = receptive_halo_frames(fb)
    run5, _ = cpu_generator("FullbandMelganGenerator", upsample_factors=(5, 5, 4, 3))
    torch.set_num_threads(cores)
    win = torch.randn(1, 80, 20 + 2 * ctx - 2 * PAD)        # run() pads PAD frames per side itself
    ts, y = time_cpu(run5, win, 5, 1)
    ts.sort()
    out["cfg5_chunk_all_cores"] = {"value": ts[len(ts) // 2] * 1e3, "unit": "ms per 20-frame chunk window", "cores": cores,
                                   "window_frames": 20 + 2 * ctx, "ms_best": ts[0] * 1e3,
                                   "sample": f"FullbandMelganGenerator(5,5,4,3) on a {20 + 2 * ctx}-frame window"}
    out["kind"] = kind
    torch.set_num_threads(cores)
    return out


# ---------------------------------------------------------------------------------------------------------------
def dist_setup(n_gpus):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        dist.init_process_group("nccl" if torch.cuda.is_available() else "gloo", rank=rank, world_size=world,
                                device_id=torch.device("cuda", local) if torch.cuda.is_available() else None)
    return rank, world, local


def allreduce_max(x, world, device):
    if world == 1:
        return x
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()


def kernel_table(recs, steps, peaks, dtype, ffma_peak):
    """Per launch family of one step: live time, algorithmic work, the bound and the fraction of ITS peak."""
    fam = {}
    for r in recs:
        k = (r["name"], r["layer"])
        f = fam.setdefault(k, {"ms": 0.0, "flops": 0.0, "bytes": 0.0, "n": 0})
        f["ms"] += r["ms"]; f["flops"] += r["flops"]; f["bytes"] += r["bytes"]; f["n"] += 1
    total_ms = sum(f["ms"] for f in fam.values())
    rows = []
    for (name, layer), f in sorted(fam.items(), key=lambda kv: -kv[1]["ms"]):
        ms = f["ms"] / f["n"]
        hbm = name.startswith("pqmf") or "prep" in name or name.startswith("out.")
        row = {"kernel": f"{name}@{layer}", "ms": round(ms, 5), "share_of_step": round(f["ms"] / total_ms, 4),
               "algorithmic_flops": f["flops"] / f["n"], "algorithmic_bytes": f["bytes"] / f["n"]}
        gbs = f["bytes"] / f["ms"] / 1e6
        tfs = f["flops"] / f["ms"] / 1e9
        if hbm:
            row.update({"bound": "hbm", "achieved": gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                        "frac": gbs / peaks["hbm_gbs"]})
        elif dtype == "bf16":
            row.update({"bound": "tensor", "achieved": tfs, "peak": peaks["bf16_tflops"], "unit": "TFLOP/s",
                        "frac": tfs / peaks["bf16_tflops"], "hbm_gbs": gbs})
        else:
            row.update({"bound": "fp32 FFMA", "achieved": tfs, "peak": ffma_peak, "unit": "TFLOP/s",
                        "frac": tfs / ffma_peak if ffma_peak else None, "hbm_gbs": gbs})
        rows.append(row)
    return rows, fam, total_ms


def gpu_arm(args):
    from tts_b200 import _lib
    from tts_b200.engine import Profile
    rank, world, local = dist_setup(args.gpus)
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    peaks = load_peaks()
    model = make_model(args.dtype).to(dev).eval()
    eng = model._get_engine()
    model.freeze()
    prec = {"fp32": 0, "bf16": 1}[args.dtype]
    lib = _lib.load()
    ffma_peak = float(lib.ttsb200_diag_ffma_tflops(local)) if (args.dtype == "fp32" or world == 1) else None

    # pool of distinct input batches, total > L2, so a step never finds its mel in cache
    per_batch = BATCH * 80 * FRAMES * 4
    n_pool = max(2, -(-int(L2_BYTES * 1.1) // per_batch))
    g = torch.Generator(device="cpu").manual_seed(MEL_SEED + rank)
    pool = [torch.randn(BATCH, 80, FRAMES, generator=g).to(dev) for _ in range(n_pool)]
    n_out = eng.output_length(FRAMES, PAD)
    out = torch.empty((BATCH, 1, n_out), dtype=torch.float32, device=dev)
    samples_per_step = BATCH * n_out
    stream = torch.cuda.current_stream(dev)

    def step(i, profile=None, engine=eng, precision=prec):
        engine.forward(pool[i % n_pool], PAD, precision, out=out, profile=profile)

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize(dev)
    barrier(world)
    sampler = ClockSampler(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    if rank == 0:
        sampler.start()
    e0.record(stream)
    for i in range(args.steps):
        step(args.warmup + i)
    e1.record(stream)
    torch.cuda.synchronize(dev)
    clocks = sampler.stop() if rank == 0 else None
    barrier(world)
    ms = allreduce_max(e0.elapsed_time(e1), world, dev)
    value = world * samples_per_step * args.steps / (ms * 1e-3)
    flops_per_sample = 2.0 * macs_per_item(FRAMES + 2 * PAD) / n_out   # 51,264
    # The kernels of consecutive steps overlap: what the LAST timed step left in `out` must be what a
    # lone call on the same input gives, bit for bit — a step that ran ahead of its producers would show here.
    last_out = out.clone()
    torch.cuda.synchronize(dev)
    step(args.warmup + args.steps - 1)
    torch.cuda.synchronize(dev)
    consistent = bool(torch.equal(last_out, out))
    assert consistent, "back-to-back steps differ from a lone call: kernels ran ahead of their producers"

    # per-launch CUDA-event pass over the same K steps -> every kernel against its own roofline
    prof = Profile()
    for i in range(args.steps):
        step(args.warmup + i, profile=prof)
    torch.cuda.synchronize(dev)
    recs = prof.read()
    table, fam, total_ms = kernel_table(recs, args.steps, peaks, args.dtype, ffma_peak)
    launches_per_step = len(recs) // max(1, args.steps)
    # dominant kernel: the largest time share; launches within 3 % of it are a tie (two of the stack kernels are, and
    # which one is ahead changes from box to box) that goes to the one doing more algorithmic work
    near = [r for r in table if r["ms"] >= 0.97 * table[0]["ms"]]
    top = max(near, key=lambda r: r["algorithmic_flops"])
    roof = {k: top[k] for k in ("bound", "achieved", "peak", "unit", "frac")}
    roof["dominant_rule"] = "largest time share of the step; ties within 3 % go to the launch with more algorithmic FLOPs"
    if roof["bound"] == "fp32 FFMA":
        roof["bound"] = "tensor"      # contract vocabulary; the denominator is the measured FFMA rate (see note)
        roof["note"] = "fp32 exactness path on the CUDA cores: peak = FFMA rate measured on this GPU by ttsb200_diag_ffma_tflops"
    roof.update({"traffic": None, "peak_source": peaks["src"],
                 "peak_kind": "burst (the timed region is a few ms at the maximum SM clock)" if args.dtype == "bf16"
                 else "measured FFMA loop", "kernel": top["kernel"], "avg_launch_ms": top["ms"],
                 "share_of_step": top["share_of_step"], "algorithmic_flops_per_launch": top["algorithmic_flops"],
                 "algorithmic_bytes_per_launch": top["algorithmic_bytes"]})
    tr = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr):
        roof["traffic"] = json.load(open(tr)).get(f"{args.dtype}:{top['kernel']}")
    achieved = value / world * flops_per_sample / 1e12
    step_peak = peaks["bf16_tflops"] if args.dtype == "bf16" else ffma_peak
    whole = {"flops_per_sample": flops_per_sample, "achieved_tflops": achieved, "peak": step_peak,
             "frac_of_peak": achieved / step_peak if step_peak else None,
             "peak_kind": "bf16 burst" if args.dtype == "bf16" else "fp32 FFMA measured"}

    # sustained: the same step back to back for >= 3 s, no host sync inside, clocks sampled during the loop
    sustained = None
    if not args.no_sustained:
        n_sus = max(args.steps, int(args.sustain_s / (ms / args.steps * 1e-3)) + 1)
        torch.cuda.synchronize(dev)
        barrier(world)
        s2 = ClockSampler(local, period_s=0.01)
        if rank == 0:
            s2.start()
        e0.record(stream)
        for i in range(n_sus):
            step(i)
        e1.record(stream)
        torch.cuda.synchronize(dev)
        sc = s2.stop() if rank == 0 else None
        barrier(world)
        ms_s = allreduce_max(e0.elapsed_time(e1), world, dev)
        v_s = world * samples_per_step * n_sus / (ms_s * 1e-3)
        t_s = v_s / world * flops_per_sample / 1e12
        sustained = {"value": v_s, "unit": "samples/s", "steps": n_sus, "seconds": ms_s / 1e3,
                     "ms_per_step": ms_s / n_sus, "achieved_tflops": t_s,
                     "peak": peaks["bf16_tflops_sustained"] if args.dtype == "bf16" else ffma_peak,
                     "frac_of_sustained_peak": (t_s / peaks["bf16_tflops_sustained"]) if args.dtype == "bf16"
                     else (t_s / ffma_peak if ffma_peak else None)}
        if sc:
            sustained.update({"sm_mhz_median": sc["sm_mhz"], "sm_mhz_min": sc.get("sm_mhz_min"),
                              "power_w_max": sc.get("power_w_max"), "reasons": sc["reasons"],
                              "clock_samples": sc["samples"]})

    # e2e: public API with HOST buffers.  Every step uploads its pinned mel batch, runs inference() and reads
    # the waveform back into pinned memory (inference_to_host); steps rotate over NS streams so one
    # step's copies overlap the others' kernels.  Wall clock around K steps, synchronised on both sides.
    ke = max(3, args.steps)
    host_pool = [torch.randn(BATCH, 80, FRAMES, generator=g).pin_memory() for _ in range(4)]
    NS = 3   # in-flight steps: one stream's H2D + D2H (0.09 + 0.27 ms) is longer than a step's kernels, two streams stall
    streams = [torch.cuda.Stream(dev) for _ in range(NS)]

    def e2e_run(dtype):
        host_out = [torch.empty((BATCH, 1, n_out), dtype=dtype).pin_memory() for _ in range(NS)]
        events = [None] * NS

        def e2e_step(i):
            s = i % NS
            with torch.cuda.stream(streams[s]):
                if events[s] is not None:
                    events[s].synchronize()            # the host buffer of this slot has been consumed
                _, events[s] = model.inference_to_host(host_pool[i % 4], out=host_out[s], dtype=dtype)

        for i in range(3):
            e2e_step(i)
        torch.cuda.synchronize(dev)
        barrier(world)
        t0 = time.perf_counter()
        for i in range(ke):
            e2e_step(i)
        torch.cuda.synchronize(dev)
        te = allreduce_max(time.perf_counter() - t0, world, dev)
        return {"value": world * samples_per_step * ke / te, "unit": "samples/s", "h2d_bytes_per_step": per_batch,
                "d2h_bytes_per_step": host_out[0].numel() * host_out[0].element_size(), "steps": ke,
                "ms_per_step": te / ke * 1e3}

    e2e = e2e_run(torch.float32)
    e2e["api"] = f"MultibandMelganGenerator.inference_to_host(pinned cpu mel, out=pinned cpu wav), {NS} streams"
    e2e["f16_waveform"] = e2e_run(torch.float16)
    e2e["f16_waveform"]["api"] = "same call with dtype=torch.float16: the PQMF kernel stores fp16 (<= 2^-11 relative)"
    # the reference callers' exact pattern, for comparison: inference(cpu mel).cpu() (pageable result, blocking)
    kb = max(3, min(args.steps, 10))
    model.inference(host_pool[0]).cpu()
    torch.cuda.synchronize(dev)
    barrier(world)
    t0 = time.perf_counter()
    for i in range(kb):
        model.inference(host_pool[i % 4]).cpu()
    tb = allreduce_max(time.perf_counter() - t0, world, dev)
    e2e["blocking_cpu_call_value"] = world * samples_per_step * kb / tb
    # measured host-link ceiling of this box with every rank copying at once (pinned, one copy per direction per step)
    barrier(world)
    d2h = float(lib.ttsb200_diag_hostlink_gbs(local, 1, BATCH * n_out * 4, 1))
    barrier(world)
    h2d = float(lib.ttsb200_diag_hostlink_gbs(local, 0, per_batch, 1))
    d2h = -allreduce_max(-d2h, world, dev)     # the slowest rank under concurrent copies
    h2d = -allreduce_max(-h2d, world, dev)
    if d2h > 0 and h2d > 0:
        t_link = BATCH * n_out * 4 / (d2h * 1e9) + 0.0       # D2H and H2D run on separate copy engines: D2H binds
        e2e["hostlink"] = {"d2h_gbs_per_gpu_concurrent": d2h, "h2d_gbs_per_gpu_concurrent": h2d,
                           "ceiling_samples_per_s": world * samples_per_step / max(t_link, per_batch / (h2d * 1e9)),
                           "frac_of_ceiling": None,
                           "how": f"ttsb200_diag_hostlink_gbs: {world} rank(s) copying {BATCH * n_out * 4 >> 20} MiB D2H "
                                  "at once, pinned, min over ranks"}
        e2e["hostlink"]["frac_of_ceiling"] = e2e["value"] / e2e["hostlink"]["ceiling_samples_per_s"]

    line = {"metric": "mb_melgan_pqmf_audio_samples_per_s", "value": value, "unit": "samples/s",
            "x_realtime": value / SAMPLE_RATE, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if args.dtype == "fp32" else "bf16", "data": "synthetic",
            "config": {"workload": WORKLOAD_CFG2,
                       "batch_per_gpu": BATCH, "frames": FRAMES, "samples_per_step_per_gpu": samples_per_step,
                       "l2": f"inputs rotate over a pool of {n_pool} batches = {n_pool * per_batch >> 20} MiB > 126 MiB L2",
                       "parallelism": f"{world} independent replicas, utterances sharded, no collective"},
            "roofline": roof, "whole_step": whole, "sustained": sustained, "roofline_by_kernel": table,
            "e2e": e2e, "back_to_back_equals_lone_call": consistent, "gpu_launches": launches_per_step * args.steps,
            "launches_per_step": launches_per_step, "clocks": clocks,
            "kernel_breakdown_ms_per_step": {f"{k[0]}@{k[1]}": round(v["ms"] / args.steps, 4)
                                             for k, v in sorted(fam.items(), key=lambda kv: -kv[1]["ms"])}}

    if world == 1 and args.dtype == "bf16" and not args.no_fp32:
        # the reference-precision path, driver-visible: a few steps of the fp32 schedule + its error against the oracle
        model.precision = "fp32"
        k32 = 5
        for i in range(2):
            step(i, precision=0)
        torch.cuda.synchronize(dev)
        e0.record(stream)
        for i in range(k32):
            step(i, precision=0)
        e1.record(stream)
        torch.cuda.synchronize(dev)
        ms32 = e0.elapsed_time(e1) / k32
        v32 = samples_per_step / (ms32 * 1e-3)
        from oracle.melgan_oracle import GeneratorSpec, generator_inference
        sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
        ref = generator_inference(GeneratorSpec.multiband(), sd, pool[0][:1].cpu(), PAD)
        err32 = float((model.inference(pool[0][:1]).cpu() - ref).abs().max())
        model.precision = "bf16"
        outb = model.inference(pool[0][:1]).cpu().double()
        snr = float(10 * torch.log10((ref.double() ** 2).sum() / ((ref.double() - outb) ** 2).sum()))
        t32 = v32 * flops_per_sample / 1e12
        line["fp32"] = {"value": v32, "unit": "samples/s", "ms_per_step": ms32, "steps": k32,
                        "max_abs_err_vs_oracle": err32, "tolerance": 1e-5, "achieved_tflops": t32,
                        "ffma_peak_tflops_measured": ffma_peak, "frac_of_ffma_peak": t32 / ffma_peak if ffma_peak else None,
                        "note": "true fp32 FFMA on the CUDA cores (north_star: max-abs <= 1e-5); not the bench dtype"}
        line["bf16_snr_db_vs_fp32_oracle"] = snr

    if rank == 0:
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_reference(3, 1, budget_s=15.0)
            try:
                line["cpu_baseline"]["more"] = cpu_extras()
            except Exception as e:   # the extras must never take the headline line down
                line["cpu_baseline"]["more"] = {"error": repr(e)}
        print(json.dumps(line))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def cfg1_arm(args):
    """BASELINE config 1: MelganGenerator() on one [1,80,200] mel, batch 1 — a latency workload."""
    from tts_b200 import _lib
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    peaks = load_peaks()
    lib = _lib.load()
    ffma_peak = float(lib.ttsb200_diag_ffma_tflops(0))
    res = {}
    g = torch.Generator().manual_seed(MEL_SEED)
    mels = [torch.randn(1, 80, 200, generator=g) for _ in range(8)]
    n_iter = max(50, args.steps * 10)
    flops = None
    for dtype in ("fp32", "bf16"):
        m = make_model(dtype, "MelganGenerator").to(dev).eval()
        m.freeze()
        eng = m._get_engine()
        n_out = eng.output_length(200, PAD)
        flops = 2.0 * macs_per_item(204, 80, 1, 512, (8, 8, 2, 2), 3)
        dmels = [x.to(dev) for x in mels]
        pmels = [x.pin_memory() for x in mels]
        for i in range(max(3, args.warmup)):
            m.inference(dmels[i % 8])
        torch.cuda.synchronize(dev)
        # device-resident, back to back (throughput): CUDA events around n_iter calls
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(n_iter):
            m.inference(dmels[i % 8])
        e1.record()
        torch.cuda.synchronize(dev)
        ms = e0.elapsed_time(e1) / n_iter
        # one call at a time, host mel in -> host waveform out (the reference callers' pattern), wall clock
        lat = []
        for i in range(n_iter):
            torch.cuda.synchronize(dev)
            t0 = time.perf_counter()
            m.inference(pmels[i % 8]).cpu()
            lat.append((time.perf_counter() - t0) * 1e3)
        lat.sort()
        tf = flops / (ms * 1e-3) / 1e12
        peak = peaks["bf16_tflops"] if dtype == "bf16" else ffma_peak
        res[dtype] = {"samples_per_s": n_out / (ms * 1e-3), "x_realtime": n_out / (ms * 1e-3) / SAMPLE_RATE,
                      "ms_per_call_device": ms, "latency_ms_p50": lat[len(lat) // 2],
                      "latency_ms_p99": lat[min(len(lat) - 1, int(len(lat) * 0.99))],
                      "launches": eng.launch_count(_lib.BF16 if dtype == "bf16" else _lib.F32, True),
                      "achieved_tflops": tf, "peak": peak, "frac_of_peak": tf / peak if peak else None,
                      "e2e_samples_per_s": n_out / (lat[len(lat) // 2] * 1e-3)}
    main = res[args.dtype]
    line = {"metric": "melgan_audio_samples_per_s", "value": main["samples_per_s"], "unit": "samples/s",
            "x_realtime": main["x_realtime"], "n_gpus": 1, "steps": n_iter, "warmup": max(3, args.warmup),
            "ms_per_step": main["ms_per_call_device"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if args.dtype == "fp32" else "bf16", "data": "synthetic",
            "config": {"workload": "BASELINE config 1: MelganGenerator() (80 mels, C0 512, factors 8,8,2,2, 3 res blocks) "
                                   "on one [1,80,200] mel, batch 1, inference_padding 2",
                       "flops_per_call": flops, "l2": "rotates over 8 inputs; the 8.5 MB of weights are L2-resident "
                                                      "by design (one utterance at a time is the use case)"},
            "fp32": res["fp32"], "bf16": res["bf16"],
            "e2e": {"value": main["e2e_samples_per_s"], "unit": "samples/s", "h2d_bytes_per_step": 80 * 200 * 4,
                    "d2h_bytes_per_step": 52224 * 4, "api": "MelganGenerator.inference(pinned cpu mel).cpu(), one call "
                                                            "at a time, p50 wall clock"}}
    if not args.no_cpu:
        run1, kind = cpu_generator("MelganGenerator")
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        ts, y = time_cpu(run1, mels[0], 10, 2)
        ts.sort()
        line["cpu_baseline"] = {"value": y.shape[-1] / ts[len(ts) // 2], "unit": "samples/s", "cores": cores, "kind": kind,
                                "sample": "the whole workload: MelganGenerator() on [1,80,200], median of 10",
                                "ms_median": ts[len(ts) // 2] * 1e3, "ms_best": ts[0] * 1e3, "cpu_model": cpu_model_string()}
    print(json.dumps(line))


def longform_arm(args):
    """BASELINE config 4: ONE 10-minute utterance split along time across the ranks (tts_b200/longform.py)."""
    from tts_b200 import longform
    rank, world, local = dist_setup(args.gpus)
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    model = make_model(args.dtype).to(dev).eval()
    model.freeze()
    T = 51680
    mel = torch.randn(1, 80, T, generator=torch.Generator().manual_seed(MEL_SEED)).to(dev)   # replicated: 16.5 MB
    halo = longform.receptive_halo_frames(model)
    n_out = (T + 2 * PAD) * longform.hop_length(model)
    stream = torch.cuda.current_stream(dev)
    # one rank too goes through the plan (a plan over this process alone): same call, and it can be replayed from a
    # CUDA graph, which takes the host's launch jitter out of a 1 ms step
    plan = longform.LongformPlan(model, T, batch=1, mode=None if world > 1 else "local")
    run = lambda: plan.run(mel)

    def timed(fn, steps):
        for _ in range(max(3, args.warmup)):
            fn()
        torch.cuda.synchronize(dev)
        barrier(world)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            w = fn()
        e1.record(stream)
        torch.cuda.synchronize(dev)
        barrier(world)
        return allreduce_max(e0.elapsed_time(e1), world, dev) / steps, w

    # one-GPU time of the same utterance, measured in the same run on rank 0 alone (the denominator of the
    # strong-scaling efficiency; other ranks idle behind the barrier)
    ms1 = None
    if world > 1:
        if rank == 0:
            solo = longform.LongformPlan(model, T, batch=1, mode="local")
            replay1 = solo.graphed(mel)
            for _ in range(3):
                replay1()
            torch.cuda.synchronize(dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(args.steps):
                replay1()
            e1.record(stream)
            torch.cuda.synchronize(dev)
            ms1 = e0.elapsed_time(e1) / args.steps
            solo.close()
        barrier(world)
        ms1 = allreduce_max(ms1 or 0.0, world, dev)
    ms, wav = timed(run, args.steps)
    ms_eager = ms
    if plan is not None:
        # the same call replayed from a CUDA graph: with many ranks the host side of run() (a dozen launches, the
        # window slice, the barrier) takes as long as a rank's kernels
        replay = plan.graphed(mel)
        ms_g, wav_g = timed(replay, args.steps)
        if plan.describe()["launch"] == "cuda graph" and ms_g < ms:
            ms, wav = ms_g, wav_g
    ms_f16 = None
    if world > 1:
        # the waveform as fp16 (every reference caller ends in a 16-bit WAV): half the bytes over NVLink into rank 0
        plan16 = longform.LongformPlan(model, T, batch=1, dtype=torch.float16)
        ms_f16, _ = timed(plan16.graphed(mel), args.steps)
        plan16.close()
    if rank == 0 or world == 1:
        assert wav is not None and wav.shape[-1] == n_out, (None if wav is None else wav.shape, n_out)
    value = n_out / (ms / 1e3)
    line = {"metric": "mb_melgan_pqmf_longform_audio_samples_per_s", "value": value, "unit": "samples/s",
            "x_realtime": value / SAMPLE_RATE, "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32" if args.dtype == "fp32" else "bf16", "data": "synthetic",
            "config": {"workload": "BASELINE config 4: one 600 s utterance [1,80,51680] split along time over the "
                                   f"ranks, {halo}-frame halo sliced from the replicated mel (zero-byte exchange), every "
                                   "rank's last kernel writes its samples straight into rank 0's waveform buffer over "
                                   "NVLink peer memory (no all_gather, no concatenation)", "frames": T, "halo_frames": halo,
                       "parallelism": f"time split x{world}"}}
    line["ms_per_step_eager_launches"] = ms_eager
    if world > 1:
        line["one_gpu_ms_same_run"] = ms1
        line["f16_waveform"] = {"ms_per_step": ms_f16, "value": n_out / (ms_f16 / 1e3),
                                "strong_efficiency": (ms1 / ms_f16) / world if ms1 else None}
        line["strong_efficiency"] = (ms1 / ms) / world if ms1 else None
    line["placement"] = plan.describe()
    if rank == 0:
        print(json.dumps(line))
    if plan is not None:
        plan.close()
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def streaming_arm(args):
    """BASELINE config 5: FB-MelGAN with hop 300 fed 250 ms chunks; per-chunk latency through the public API."""
    from tts_b200.streaming import StreamingVocoder
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    m = make_model(args.dtype, "FullbandMelganGenerator", upsample_factors=(5, 5, 4, 3)).to(dev).eval()
    m.freeze()
    chunk, rate = 20, 24000
    use_graph = not os.environ.get("TTSB200_STREAM_NO_GRAPH")
    sv = StreamingVocoder(m, chunk_frames=chunk, use_graph=use_graph)
    n_chunks = max(50, args.steps * 10)
    mel = torch.randn(1, 80, chunk * (n_chunks + args.warmup + 4), generator=torch.Generator().manual_seed(MEL_SEED))
    host = mel.pin_memory()
    lat, samples = [], 0
    for i in range(n_chunks + args.warmup + 4):
        piece = host[..., i * chunk:(i + 1) * chunk]
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        outs = sv.feed(piece.to(dev, non_blocking=True))      # H2D of the chunk, window inference on the CUDA path
        got = [o.cpu() for o in outs]                         # D2H of the emitted samples (blocking, like the server)
        dt = time.perf_counter() - t0
        if got and i >= args.warmup + 4:
            lat.append(dt)
            samples += sum(g.shape[-1] for g in got)
    lat.sort()
    p50, p99 = lat[len(lat) // 2], lat[min(len(lat) - 1, int(len(lat) * 0.99))]
    line = {"metric": "fb_melgan_stream_chunk_latency_ms", "value": p50 * 1e3, "unit": "ms", "p99_ms": p99 * 1e3,
            "x_realtime": (samples / rate) / sum(lat), "n_gpus": 1, "steps": len(lat), "warmup": args.warmup,
            "ms_per_step": sum(lat) / len(lat) * 1e3, "higher_is_better": False, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32" if args.dtype == "fp32" else "bf16", "data": "synthetic",
            "config": {"workload": "BASELINE config 5: FullbandMelganGenerator(upsample_factors=(5,5,4,3)) hop 300, "
                                   f"24 kHz, stream of [1,80,{chunk}] chunks (250 ms), context {sv.ctx} frames per side, "
                                   "latency = host chunk in -> 6000 samples back on the host",
                       "chunk_frames": chunk, "context_frames": sv.ctx, "samples_per_chunk": chunk * sv.hop,
                       "launch": "CUDA graph per window shape" if use_graph else "kernel by kernel"}}
    if not args.no_cpu:
        run5, kind = cpu_generator("FullbandMelganGenerator", upsample_factors=(5, 5, 4, 3))
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        win = torch.randn(1, 80, chunk + 2 * sv.ctx - 2 * PAD)
        ts, _ = time_cpu(run5, win, 10, 2)
        ts.sort()
        line["cpu_baseline"] = {"value": ts[len(ts) // 2] * 1e3, "unit": "ms", "cores": cores, "kind": kind,
                                "sample": f"one steady-state window ({chunk + 2 * sv.ctx} frames) of the same model on the "
                                          "host cores, median of 10 (compute only, no copies)",
                                "cpu_model": cpu_model_string()}
    print(json.dumps(line))


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = cpu_reference(args.steps, args.warmup)
    line = {"impl": "reference", "metric": "mb_melgan_pqmf_audio_samples_per_s", "value": r["value"],
            "unit": "samples/s", "x_realtime": r["value"] / SAMPLE_RATE, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD_CFG2, "batch_per_gpu": BATCH, "frames": FRAMES,
                       "sample": f"host cores: {r['items']} x 862 frames per step (bounded sample of the 16-utterance "
                                 "batch; throughput is per sample, so the sample size does not change the metric)"},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "cpu_model")},
            "e2e": {"value": r["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--dtype", default=os.environ.get("TTSB200_BENCH_DTYPE", "bf16"), choices=["fp32", "bf16"])
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-fp32", action="store_true", help="skip the fp32 sub-record")
    ap.add_argument("--no-sustained", action="store_true", help="skip the >= 3 s sustained leg")
    ap.add_argument("--sustain-s", type=float, default=3.0)
    ap.add_argument("--workload", default="batched", choices=["batched", "melgan_cfg1", "longform", "streaming"],
                    help="batched = BASELINE config 2 (the headline line); melgan_cfg1 = config 1; longform = config 4; "
                         "streaming = config 5")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        reference_arm(args)
    elif args.workload == "longform":
        longform_arm(args)
    elif args.workload == "streaming":
        streaming_arm(args)
    elif args.workload == "melgan_cfg1":
        cfg1_arm(args)
    else:
        gpu_arm(args)


if __name__ == "__main__":
    main()
