#!/usr/bin/env python
"""Headline benchmark: MB-MelGAN + PQMF batched inference, audio samples/s (x real-time).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--dtype fp32|bf16] [--impl reference]

A step = generator.inference() over one batch of 16 x 10 s synthetic mels (BASELINE.json config 2:
MultibandMelganGenerator() defaults, 22,050 Hz, hop 256, 862 frames, inference_padding 2) per GPU.
N > 1 (torchrun): every rank runs the same per-GPU batch on its own utterances — independent
units, no data-path collective ("weak" scaling).  Prints ONE JSON line on rank 0.

Timing: CUDA events on the launch stream around exactly K steps, barrier + synchronize on both
sides, max over ranks.  Inputs rotate through a pool larger than L2 so no step re-reads its mel
from cache.  `e2e` goes through the public API with pinned HOST buffers (H2D + D2H in the timed
region).  `--impl reference` times the CPU restatement of the reference generator (oracle/, the
same ATen kernels the reference's nn.Sequential dispatches to) on the host cores.
"""
import argparse
import json
import os
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


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "src": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "src": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        sm.sort()
        # under-load median: upper half of the samples (idle samples precede the first launch)
        load = sm[len(sm) // 2:] if sm else []
        med = load[len(load) // 2] if load else None
        return {"sm_mhz": med, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def macs_per_item(L0, in_ch=80, out_ch=4, c0=384, factors=(2, 8, 2, 2), n_blocks=3, k=7, rk=3):
    """7*80*C0*L0 + sum_i [2 f_i C_{i-1} C_i L_{i-1} + 5 n C_i^2 L_i] + 7*C_last*out*L_last."""
    L, c, total = L0, c0, k * in_ch * c0 * L0
    for f in factors:
        total += 2 * f * c * (c // 2) * L
        L, c = L * f, c // 2
        total += (rk + 2) * n_blocks * c * c * L
    return total + k * c * out_ch * L


def make_model(dtype):
    from tts_b200.vocoder.models.multiband_melgan_generator import MultibandMelganGenerator
    torch.manual_seed(WEIGHT_SEED)
    m = MultibandMelganGenerator()
    m.remove_weight_norm()
    m.inference_padding = PAD
    m.precision = dtype
    return m


def cpu_reference(steps, warmup, budget_s=100.0):
    """Reference-arm / cpu_baseline: oracle port of the generator on all host cores."""
    from oracle.melgan_oracle import GeneratorSpec, generator_layers, pqmf_filters, pqmf_synthesis
    import torch.nn.functional as F
    spec = GeneratorSpec.multiband()
    m = make_model("fp32")
    sd = {k: v.detach() for k, v in m.state_dict().items()}
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    H, G, U = pqmf_filters()

    def run(mel):
        with torch.no_grad():
            c = F.pad(mel, (PAD, PAD), mode="replicate")
            return pqmf_synthesis(generator_layers(spec, sd, c), G, U)

    torch.manual_seed(MEL_SEED)
    one = torch.randn(1, 80, FRAMES)
    run(one)
    t0 = time.perf_counter(); run(one); t1 = time.perf_counter() - t0
    items = int(max(1, min(BATCH, budget_s / max(1e-3, (steps + warmup) * t1))))
    mel = torch.randn(items, 80, FRAMES)
    for _ in range(warmup):
        run(mel)
    t0 = time.perf_counter()
    for _ in range(steps):
        y = run(mel)
    dt = time.perf_counter() - t0
    samples = steps * y.shape[0] * y.shape[-1]
    return {"value": samples / dt, "unit": "samples/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{items} of {BATCH} utterances x {FRAMES} frames per step, {steps} steps, "
                      f"oracle/melgan_oracle.py (torch {torch.__version__} CPU kernels, as the reference dispatches)",
            "ms_per_step": dt / steps * 1e3, "items": items}


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


def gpu_arm(args):
    from tts_b200.engine import Profile
    rank, world, local = dist_setup(args.gpus)
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    peaks = load_peaks()
    model = make_model(args.dtype).to(dev).eval()
    eng = model._get_engine()
    model.freeze()
    prec = {"fp32": 0, "bf16": 1}[args.dtype]

    # pool of distinct input batches, total > L2, so a step never finds its mel in cache
    per_batch = BATCH * 80 * FRAMES * 4
    n_pool = max(2, -(-int(L2_BYTES * 1.1) // per_batch))
    g = torch.Generator(device="cpu").manual_seed(MEL_SEED + rank)
    pool = [torch.randn(BATCH, 80, FRAMES, generator=g).to(dev) for _ in range(n_pool)]
    n_out = eng.output_length(FRAMES, PAD)
    out = torch.empty((BATCH, 1, n_out), dtype=torch.float32, device=dev)
    samples_per_step = BATCH * n_out
    stream = torch.cuda.current_stream(dev)

    def step(i, profile=None):
        eng.forward(pool[i % n_pool], PAD, prec, out=out, profile=profile)

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize(dev)
    barrier(world)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    e0.record(stream)
    for i in range(args.steps):
        step(args.warmup + i)
    e1.record(stream)
    torch.cuda.synchronize(dev)
    barrier(world)
    ms = allreduce_max(e0.elapsed_time(e1), world, dev)
    clocks = sampler.stop() if rank == 0 else None
    value = world * samples_per_step * args.steps / (ms * 1e-3)

    # per-launch CUDA-event pass over the same K steps -> dominant kernel family for the roofline
    prof = Profile()
    for i in range(args.steps):
        step(args.warmup + i, profile=prof)
    torch.cuda.synchronize(dev)
    recs = prof.read()
    fam = {}
    for r in recs:
        k = (r["name"], r["layer"])
        f = fam.setdefault(k, {"ms": 0.0, "flops": 0.0, "bytes": 0.0, "n": 0})
        f["ms"] += r["ms"]; f["flops"] += r["flops"]; f["bytes"] += r["bytes"]; f["n"] += 1
    total_ms = sum(f["ms"] for f in fam.values())
    (dname, dlayer), d = max(fam.items(), key=lambda kv: kv[1]["ms"])
    launches_per_step = len(recs) // max(1, args.steps)
    if "pqmf" in dname:
        roof = {"bound": "hbm", "achieved": d["bytes"] / d["ms"] / 1e6, "peak": peaks["hbm_gbs"], "unit": "GB/s"}
    elif args.dtype == "bf16":
        roof = {"bound": "tensor", "achieved": d["flops"] / d["ms"] / 1e9, "peak": peaks["bf16_tflops_sustained"],
                "unit": "TFLOP/s"}
    else:
        sm_max = (clocks or {}).get("sm_max_mhz") or 1965.0
        roof = {"bound": "tensor", "achieved": d["flops"] / d["ms"] / 1e9, "peak": peaks["bf16_tflops_sustained"],
                "unit": "TFLOP/s",
                "note": "fp32 FFMA kernel (exactness path): its own ceiling is the fp32 FFMA peak "
                        f"{148 * 128 * 2 * sm_max / 1e6:.1f} TFLOP/s nominal at {sm_max:.0f} MHz, "
                        "not in MEASURED_PEAKS.json; frac is quoted against the measured bf16 tensor peak"}
    roof.update({"frac": roof["achieved"] / roof["peak"], "traffic": None, "peak_source": peaks["src"],
                 "kernel": f"{dname}@layers[{dlayer}]", "launches_timed": d["n"],
                 "avg_launch_ms": d["ms"] / d["n"], "share_of_step": d["ms"] / total_ms,
                 "algorithmic_flops_per_launch": d["flops"] / d["n"],
                 "algorithmic_bytes_per_launch": d["bytes"] / d["n"]})
    tr = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr):
        roof["traffic"] = json.load(open(tr)).get(f"{args.dtype}:{dname}@{dlayer}")
    flops_per_sample = 2.0 * macs_per_item(FRAMES + 2 * PAD) / n_out   # 51,264
    whole = {"flops_per_sample": flops_per_sample, "achieved_tflops": value / world * flops_per_sample / 1e12,
             "frac_of_bf16_peak": value / world * flops_per_sample / 1e12 / peaks["bf16_tflops_sustained"]}

    # e2e: public API with HOST buffers.  Every step uploads its pinned mel batch, runs inference() and reads
    # the waveform back into pinned memory (inference_to_host); steps rotate over NS streams so one
    # step's copies overlap the others' kernels.  Wall clock around K steps, synchronised on both sides.
    ke = max(3, args.steps)
    host_pool = [torch.randn(BATCH, 80, FRAMES, generator=g).pin_memory() for _ in range(4)]
    NS = 3   # in-flight steps: one stream's H2D + D2H (0.09 + 0.27 ms) is longer than a step's kernels, two streams stall
    host_out = [torch.empty((BATCH, 1, n_out), dtype=torch.float32).pin_memory() for _ in range(NS)]
    streams = [torch.cuda.Stream(dev) for _ in range(NS)]
    events = [None] * NS

    def e2e_step(i):
        s = i % NS
        with torch.cuda.stream(streams[s]):
            if events[s] is not None:
                events[s].synchronize()            # the host buffer of this slot has been consumed
            _, events[s] = model.inference_to_host(host_pool[i % 4], out=host_out[s])

    for i in range(3):
        e2e_step(i)
    torch.cuda.synchronize(dev)
    barrier(world)
    t0 = time.perf_counter()
    for i in range(ke):
        e2e_step(i)
    torch.cuda.synchronize(dev)
    te = allreduce_max(time.perf_counter() - t0, world, dev)
    e2e = {"value": world * samples_per_step * ke / te, "unit": "samples/s", "h2d_bytes_per_step": per_batch,
           "d2h_bytes_per_step": host_out[0].numel() * 4, "steps": ke, "ms_per_step": te / ke * 1e3,
           "api": f"MultibandMelganGenerator.inference_to_host(pinned cpu mel, out=pinned cpu wav), {NS} streams"}
    # the reference callers' exact pattern, for comparison: inference(cpu mel).cpu() (pageable result, blocking)
    kb = max(3, min(args.steps, 10))
    model.inference(host_pool[0]).cpu()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for i in range(kb):
        model.inference(host_pool[i % 4]).cpu()
    e2e["blocking_cpu_call_value"] = world * samples_per_step * kb / (time.perf_counter() - t0)

    line = {"metric": "mb_melgan_pqmf_audio_samples_per_s", "value": value, "unit": "samples/s",
            "x_realtime": value / SAMPLE_RATE, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if args.dtype == "fp32" else "bf16", "data": "synthetic",
            "config": {"workload": "BASELINE config 2: MultibandMelganGenerator() defaults (C0 384, factors 2,8,2,2, "
                                   "3 res blocks) + PQMF 62-tap synthesis, 22.05 kHz hop 256, batch 16 x 862 frames "
                                   "(10 s) per GPU, inference_padding 2, seeded random-init weights",
                       "batch_per_gpu": BATCH, "frames": FRAMES, "samples_per_step_per_gpu": samples_per_step,
                       "l2": f"inputs rotate over a pool of {n_pool} batches = {n_pool * per_batch >> 20} MiB > 126 MiB L2",
                       "parallelism": f"{world} independent replicas, utterances sharded, no collective"},
            "roofline": roof, "whole_step": whole, "e2e": e2e, "gpu_launches": launches_per_step * args.steps,
            "launches_per_step": launches_per_step, "clocks": clocks,
            "kernel_breakdown_ms_per_step": {f"{k[0]}@{k[1]}": round(v["ms"] / args.steps, 4)
                                             for k, v in sorted(fam.items(), key=lambda kv: -kv[1]["ms"])}}
    if rank == 0:
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_reference(3, 1, budget_s=20.0)
        print(json.dumps(line))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


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
    for _ in range(args.warmup):
        wav = longform.longform_inference(model, mel)
    torch.cuda.synchronize(dev)
    barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        wav = longform.longform_inference(model, mel)
    e1.record(stream)
    torch.cuda.synchronize(dev)
    barrier(world)
    ms = allreduce_max(e0.elapsed_time(e1), world, dev)
    assert wav.shape[-1] == n_out
    value = n_out * args.steps / (ms / 1e3)
    line = {"metric": "mb_melgan_pqmf_longform_audio_samples_per_s", "value": value, "unit": "samples/s",
            "x_realtime": value / SAMPLE_RATE, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32" if args.dtype == "fp32" else "bf16", "data": "synthetic",
            "config": {"workload": "BASELINE config 4: one 600 s utterance [1,80,51680] split along time over the "
                                   f"ranks, {halo}-frame halo from the replicated mel (zero-byte exchange), "
                                   "all_gather of the waveform slices", "frames": T, "halo_frames": halo,
                       "parallelism": f"time split x{world}"}}
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def streaming_arm(args):
    """BASELINE config 5: FB-MelGAN with hop 300 fed 250 ms chunks; per-chunk latency through the public API."""
    from tts_b200.streaming import StreamingVocoder
    from tts_b200.vocoder.models.fullband_melgan_generator import FullbandMelganGenerator
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    torch.manual_seed(WEIGHT_SEED)
    m = FullbandMelganGenerator(upsample_factors=(5, 5, 4, 3))
    m.remove_weight_norm()
    m.inference_padding = PAD
    m.precision = args.dtype
    m = m.to(dev).eval()
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
            "config": {"workload": "BASELINE config 2 on the host cores: MultibandMelganGenerator() + PQMF, "
                                   f"{r['items']} x 862 frames per step (bounded sample of the 16-utterance batch)"},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
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
    ap.add_argument("--workload", default="batched", choices=["batched", "longform", "streaming"],
                    help="batched = BASELINE config 2 (the headline line); longform = config 4; streaming = config 5")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        reference_arm(args)
    elif args.workload == "longform":
        longform_arm(args)
    elif args.workload == "streaming":
        streaming_arm(args)
    else:
        gpu_arm(args)


if __name__ == "__main__":
    main()
