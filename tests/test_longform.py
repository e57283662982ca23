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
