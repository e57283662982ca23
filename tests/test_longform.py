"""Long-form time split: host logic on CPU with gloo (world_size 2), exactness on the GPU."""
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp

from oracle.melgan_oracle import GeneratorSpec, generator_inference, random_state_dict, receptive_field_frames
from tts_b200 import longform
from tts_b200.vocoder.models.melgan_generator import MelganGenerator
from tts_b200.vocoder.models.multiband_melgan_generator import MultibandMelganGenerator

SMALL = {"in_channels": 8, "base_channels": 32, "upsample_factors": (2, 3), "num_res_blocks": 2}


def test_partition_and_halo():
    assert longform.frame_partition(10, 4) == [(0, 3), (3, 6), (6, 8), (8, 10)]
    assert longform.frame_partition(2, 4) == [(0, 1), (1, 2), (2, 2), (2, 2)]
    for cls, spec in ((MultibandMelganGenerator, GeneratorSpec.multiband()), (MelganGenerator, GeneratorSpec.melgan())):
        m = cls()
        # tighter than the oracle's conservative bound, never below the measured 12 / 6 frames;
        # sufficiency itself is what the chunked == full tests below establish
        assert longform.receptive_halo_frames(m) <= receptive_field_frames(spec)
        assert longform.receptive_halo_frames(m) >= (12 if spec.pqmf else 6)
    assert longform.hop_length(MultibandMelganGenerator()) == 256


def _oracle_fn(spec, sd):
    return lambda mel: generator_inference(spec, sd, mel, 0)


def test_chunked_equals_full_on_cpu_oracle():
    """The split itself (no CUDA involved): chunks with halo reproduce the full-utterance oracle output."""
    spec = GeneratorSpec.melgan(**SMALL)
    sd = random_state_dict(spec, 3)
    m = MelganGenerator(**SMALL)
    m.inference_padding = 2
    mel = torch.randn(1, 8, 61, generator=torch.Generator().manual_seed(0))
    full = generator_inference(spec, sd, mel, 2)
    for n in (1, 2, 3, 7):
        got = longform.chunked_inference(m, mel, n, infer_fn=_oracle_fn(spec, sd))
        assert got.shape == full.shape
        assert float((got - full).abs().max()) <= 1e-6, n
    # a halo below the receptive field must show up as an error (the test would otherwise prove nothing)
    bad = longform.chunked_inference(m, mel, 3, halo=1, infer_fn=_oracle_fn(spec, sd))
    assert float((bad - full).abs().max()) > 1e-4


def test_chunked_equals_full_mb_default_halo():
    """Default MB-MelGAN + PQMF with the default halo (13 frames): exact on the CPU oracle."""
    spec = GeneratorSpec.multiband()
    sd = random_state_dict(spec, 5)
    m = MultibandMelganGenerator()
    mel = torch.randn(1, 80, 70, generator=torch.Generator().manual_seed(0))
    full = generator_inference(spec, sd, mel, 2)
    got = longform.chunked_inference(m, mel, 3, infer_fn=_oracle_fn(spec, sd))
    assert got.shape == full.shape and float((got - full).abs().max()) <= 1e-6
    short = longform.chunked_inference(m, mel, 3, halo=0, infer_fn=_oracle_fn(spec, sd))
    assert float((short - full).abs().max()) > 1e-3


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_path):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    spec = GeneratorSpec.melgan(**SMALL)
    sd = random_state_dict(spec, 3)
    m = MelganGenerator(**SMALL)
    m.inference_padding = 2
    mel = torch.randn(1, 8, 45, generator=torch.Generator().manual_seed(1))
    wav = longform.longform_inference(m, mel, infer_fn=_oracle_fn(spec, sd))
    mine, rng = longform.longform_inference(m, mel, infer_fn=_oracle_fn(spec, sd), gather=False)
    full = generator_inference(spec, sd, mel, 2)
    ok = wav.shape == full.shape and float((wav - full).abs().max()) <= 1e-6
    ok = ok and float((mine - full[..., rng[0]:rng[1]]).abs().max()) <= 1e-6
    torch.save({"ok": bool(ok), "rank": rank}, f"{out_path}.{rank}")
    dist.destroy_process_group()


def test_longform_two_ranks_gloo(tmp_path):
    port = _free_port()
    out = str(tmp_path / "res")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    for r in range(2):
        assert torch.load(f"{out}.{r}")["ok"], r


@pytest.mark.gpu
@pytest.mark.parametrize("precision,tol", [("fp32", 1e-6), ("bf16", 0.0)])
def test_chunked_equals_full_on_gpu(precision, tol):
    """Every output sample is produced by the same per-element arithmetic whichever chunk it falls in, so the
    split is exact in the interior; fp32 allows last-bit differences from different tile alignment."""
    torch.manual_seed(54321)
    m = MultibandMelganGenerator()
    m.remove_weight_norm()
    m = m.cuda().eval()
    m.precision = precision
    mel = torch.randn(1, 80, 300, generator=torch.Generator().manual_seed(2)).cuda()
    full = m.inference(mel)
    for n in (2, 4):
        got = longform.chunked_inference(m, mel, n)
        assert got.shape == full.shape
        assert float((got - full).abs().max()) <= tol, (n, precision)
    assert m.inference_padding == 2


@pytest.mark.gpu
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_windowed_output_lands_in_place(precision):
    """keep=(lo, n) + out=<view>: the last kernel stores only the kept sample range, at its place inside a larger
    buffer (batch stride > n), and touches nothing around it — what every long-form rank does (LongformPlan.run)."""
    torch.manual_seed(54321)
    m = MultibandMelganGenerator()
    m.remove_weight_norm()
    m = m.cuda().eval()
    m.precision = precision
    mel = torch.randn(2, 80, 60, generator=torch.Generator().manual_seed(4)).cuda()
    full = m.inference_padded(mel, 0)
    hop = 256
    for dt in (torch.float32, torch.float16):
        big = torch.full((2, 1, 70 * hop), 7.0, dtype=dt, device="cuda")
        lo, n, at = 13 * hop, 31 * hop, 5 * hop
        m.inference_padded(mel, 0, out=big[..., at:at + n], keep=(lo, n))
        torch.cuda.synchronize()
        assert torch.equal(big[..., at:at + n], full[..., lo:lo + n].to(dt))
        assert bool((big[..., :at] == 7.0).all()) and bool((big[..., at + n:] == 7.0).all())
    # unaligned window (not a multiple of 16 samples): the per-sample store path
    big = torch.zeros((2, 1, 5000), device="cuda")
    m.inference_padded(mel, 0, out=big[..., 3:3 + 1001], keep=(777, 1001))
    assert torch.equal(big[..., 3:3 + 1001], full[..., 777:777 + 1001]) and float(big[..., 1004:].abs().max()) == 0.0
    # a generator without PQMF places its window with one strided copy
    g = MelganGenerator(**SMALL)
    g.remove_weight_norm()
    g = g.cuda().eval()
    g.precision = precision
    x = torch.randn(2, 8, 40).cuda()
    want = g.inference_padded(x, 0)
    buf = torch.zeros((2, 1, 400), device="cuda")
    g.inference_padded(x, 0, out=buf[..., 100:100 + 60], keep=(30, 60))
    assert torch.equal(buf[..., 100:160], want[..., 30:90]) and float(buf[..., :100].abs().max()) == 0.0
    with pytest.raises(RuntimeError):
        m.inference_padded(mel, 0, out=torch.zeros((2, 1, 10), device="cuda"), keep=(60 * hop - 5, 10))


@pytest.mark.gpu
def test_longform_plan_single_rank_matches_inference():
    """world_size 1: LongformPlan is generator.inference written into the plan's buffer (replicate padding folded
    into the window's index map)."""
    torch.manual_seed(54321)
    m = MultibandMelganGenerator()
    m.remove_weight_norm()
    m = m.cuda().eval()
    mel = torch.randn(1, 80, 90, generator=torch.Generator().manual_seed(6)).cuda()
    plan = longform.LongformPlan(m, 90, batch=1)
    wav = plan.run(mel)
    torch.cuda.synchronize()
    assert plan.mode == "local" and torch.equal(wav, m.inference(mel))


def _nccl_worker(rank, world, port, out_path, precision):
    """One process per GPU, NCCL: LongformPlan (direct placement into rank 0's buffer over peer memory) and the
    portable all_gather form, both against the CPU fp32 oracle of the WHOLE utterance."""
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    res = {"rank": rank}
    try:
        torch.manual_seed(54321)
        m = MultibandMelganGenerator()
        m.remove_weight_norm()
        sd = {k: v.detach().clone() for k, v in m.state_dict().items()}
        m = m.cuda().eval()
        m.precision = precision
        mel = torch.randn(1, 80, 333, generator=torch.Generator().manual_seed(8))
        ref = generator_inference(GeneratorSpec.multiband(), sd, mel, 2)
        plan = longform.LongformPlan(m, 333, batch=1)
        for _ in range(3):                                    # repeated runs reuse the registered destination
            wav = plan.run(mel.cuda())
        if precision == "bf16":                               # ... and so does the same call replayed from a graph
            wav = plan.graphed(mel.cuda())()
        torch.cuda.synchronize()
        dist.barrier()
        res["mode"] = plan.mode
        gathered = longform.longform_inference(m, mel.cuda()).cpu()
        checks = [gathered]
        if rank == 0:
            checks.append(wav.cpu())
        else:
            assert wav is None
        for w in checks:
            assert w.shape == ref.shape
            if precision == "fp32":
                err = float((w - ref).abs().max())
                assert err <= 1e-5, err
            else:
                snr = float(10 * torch.log10(ref.double().pow(2).sum() / (ref.double() - w.double()).pow(2).sum()))
                assert snr >= 40.0, snr
        if rank == 0:
            assert torch.equal(wav.cpu(), gathered)           # same kernels, same tiles: bit-identical placement
        res["ok"] = True
    except Exception as e:  # noqa: BLE001 - reported to the parent
        import traceback
        res["ok"], res["err"] = False, traceback.format_exc()
    torch.save(res, f"{out_path}.{rank}")
    dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_longform_two_ranks_nccl_against_oracle(tmp_path, precision):
    port = _free_port()
    out = str(tmp_path / "res")
    mp.spawn(_nccl_worker, args=(2, port, out, precision), nprocs=2, join=True)
    for r in range(2):
        got = torch.load(f"{out}.{r}")
        assert got["ok"], got.get("err")
    assert torch.load(f"{out}.0")["mode"] in ("symmetric", "ipc"), torch.load(f"{out}.0")["mode"]
