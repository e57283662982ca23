"""Chunked streaming and checkpoint I/O (§8f rank 1): host logic on CPU, exactness on GPU."""
import os

import pytest
import torch

from oracle.melgan_oracle import GeneratorSpec, generator_inference, random_state_dict
from tts_b200.streaming import StreamingVocoder, stream_inference
from tts_b200.vocoder.models.fullband_melgan_generator import FullbandMelganGenerator
from tts_b200.vocoder.models.melgan_generator import MelganGenerator
from tts_b200.vocoder.models.multiband_melgan_generator import MultibandMelganGenerator
from tts_b200.vocoder.utils import io as voc_io

SMALL = {"in_channels": 8, "base_channels": 32, "upsample_factors": (2, 3), "num_res_blocks": 2}


def _oracle_fn(spec, sd):
    return lambda mel: generator_inference(spec, sd, mel, 0)


@pytest.mark.parametrize("pad", [0, 2])
@pytest.mark.parametrize("chunk,feed", [(5, 5), (7, 3), (4, 11), (64, 64)])
def test_stream_equals_full_on_cpu_oracle(pad, chunk, feed):
    """Any chunking / feeding granularity reproduces the whole-utterance output (no CUDA involved)."""
    spec = GeneratorSpec.melgan(**SMALL)
    sd = random_state_dict(spec, 3)
    m = MelganGenerator(**SMALL)
    m.inference_padding = pad
    mel = torch.randn(2, 8, 53, generator=torch.Generator().manual_seed(0))
    full = generator_inference(spec, sd, mel, pad)
    got = stream_inference(m, mel, chunk_frames=chunk, feed_frames=feed, infer_fn=_oracle_fn(spec, sd))
    assert got.shape == full.shape
    assert float((got - full).abs().max()) <= 1e-6


def test_stream_emits_with_bounded_lookahead_and_fixed_windows():
    """Chunk k is emitted as soon as `context` frames of look-ahead exist; interior windows all have one shape."""
    spec = GeneratorSpec.melgan(**SMALL)
    sd = random_state_dict(spec, 3)
    m = MelganGenerator(**SMALL)
    m.inference_padding = 0
    shapes = []

    def infer(mel):
        shapes.append(mel.shape[-1])
        return generator_inference(spec, sd, mel, 0)

    sv = StreamingVocoder(m, chunk_frames=6, infer_fn=infer)
    ctx, hop = sv.ctx, sv.hop
    mel = torch.randn(1, 8, 60, generator=torch.Generator().manual_seed(1))
    assert sv.feed(mel[..., :6 + ctx - 1]) == []                      # one frame short of the look-ahead
    out = sv.feed(mel[..., 6 + ctx - 1:6 + ctx])
    assert len(out) == 1 and out[0].shape[-1] == 6 * hop
    rest = sv.feed(mel[..., 6 + ctx:]) + sv.flush()
    assert sum(o.shape[-1] for o in out + rest) == 60 * hop
    interior = [s for s in shapes[1:] if s == 2 * ctx + 6]
    assert len(interior) >= 5                                        # steady state: context + chunk + context
    assert sv.buf is None and sv.emitted == 0                        # flush() resets the stream
    # too little context must show up as an error
    bad = stream_inference(m, mel, chunk_frames=6, context=0, infer_fn=_oracle_fn(spec, sd))
    assert float((bad - generator_inference(spec, sd, mel, 0)).abs().max()) > 1e-4


def test_stream_mb_default_context():
    spec = GeneratorSpec.multiband()
    sd = random_state_dict(spec, 5)
    m = MultibandMelganGenerator()
    mel = torch.randn(1, 80, 50, generator=torch.Generator().manual_seed(0))
    full = generator_inference(spec, sd, mel, 2)
    got = stream_inference(m, mel, chunk_frames=20, infer_fn=_oracle_fn(spec, sd))
    assert got.shape == full.shape and float((got - full).abs().max()) <= 1e-6


def test_checkpoint_roundtrip_weight_norm_form(tmp_path):
    """TTS/vocoder/utils/io.py: a trainer-style checkpoint (weight_g / weight_v keys, pqmf buffers) loads strictly."""
    torch.manual_seed(7)
    src = MultibandMelganGenerator()
    path = os.path.join(tmp_path, "checkpoint_10.pth.tar")
    # what the reference trainer writes (TTS/vocoder/utils/io.py:21-46); the format is consumed here, not produced
    torch.save({"model": src.state_dict(), "optimizer": None, "scheduler": None, "model_disc": None,
                "optimizer_disc": None, "scheduler_disc": None, "step": 10, "epoch": 3, "date": "November 05, 2026",
                "model_loss": 0.5}, path)
    state = torch.load(path, map_location="cpu", weights_only=False)
    assert any(k.endswith("weight_g") for k in state["model"]) and "pqmf_layer.G" in state["model"]
    dst = MultibandMelganGenerator()
    dst, st = voc_io.load_checkpoint(dst, path)
    assert st["step"] == 10 and st["epoch"] == 3
    for k, v in src.state_dict().items():
        assert torch.equal(v, dst.state_dict()[k]), k
    dst2, _ = voc_io.load_checkpoint(MultibandMelganGenerator(), path, allow_pickle=False)   # safe loader suffices
    assert torch.equal(dst2.state_dict()["pqmf_layer.G"], src.state_dict()["pqmf_layer.G"])
    # a folded model cannot take a weight-norm checkpoint silently
    folded = MultibandMelganGenerator()
    folded.remove_weight_norm()
    with pytest.raises(RuntimeError):
        voc_io.load_checkpoint(folded, path)


@pytest.mark.gpu
@pytest.mark.parametrize("precision,tol", [("fp32", 2e-5), ("bf16", None)])
def test_stream_equals_full_on_gpu(precision, tol):
    """Config-5 model (FB-MelGAN, odd factors 5,5,4,3 = hop 300): streamed chunks == whole utterance on the CUDA path."""
    torch.manual_seed(54321)
    m = FullbandMelganGenerator(upsample_factors=(5, 5, 4, 3))
    m.remove_weight_norm()
    m = m.cuda().eval()
    m.precision = precision
    mel = torch.randn(1, 80, 90, generator=torch.Generator().manual_seed(3)).cuda()
    full = m.inference(mel)
    got = stream_inference(m, mel, chunk_frames=20)
    assert got.shape == full.shape == (1, 1, (90 + 4) * 300)
    err = float((got - full).abs().max())
    if tol is not None:
        assert err <= tol, err
    else:
        # BF16: a sample's rounding pattern depends on which tile it lands in; compare as SNR
        snr = 10 * torch.log10(full.pow(2).sum() / (got - full).pow(2).sum().clamp_min(1e-30))
        assert float(snr) >= 40.0, float(snr)


@pytest.mark.gpu
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_stream_graph_replay_is_bit_identical(precision):
    """use_graph=True replays each window shape from a CUDA graph: same kernels, same launch shapes, so the streamed
    waveform equals the kernel-by-kernel stream bit for bit (first chunks, steady state and flush shapes)."""
    from tts_b200.streaming import StreamingVocoder
    torch.manual_seed(54321)
    m = FullbandMelganGenerator(upsample_factors=(5, 5, 4, 3))
    m.remove_weight_norm()
    m = m.cuda().eval()
    m.precision = precision
    mel = torch.randn(1, 80, 130, generator=torch.Generator().manual_seed(9)).cuda()
    plain = stream_inference(m, mel, chunk_frames=20)
    sv = StreamingVocoder(m, chunk_frames=20, use_graph=True)
    parts = []
    for s in range(0, mel.shape[-1], 20):
        parts += sv.feed(mel[..., s:s + 20])
    parts += sv.flush()
    assert torch.equal(torch.cat(parts, dim=-1), plain)


@pytest.mark.gpu
def test_graph_cache_follows_the_engine():
    """Cached CUDA graphs hold raw pointers into the plan they were captured from: after load_state_dict the
    generator builds a new plan, and the next call must re-capture instead of replaying against the old weights.
    Also: streaming never writes the shared module's inference_padding."""
    from tts_b200.streaming import GraphedInference
    torch.manual_seed(1)
    m = MelganGenerator(**SMALL)
    m.remove_weight_norm()
    m = m.cuda().eval()
    gi = GraphedInference(m)
    mel = torch.randn(1, 8, 40, device="cuda")
    a = gi(mel)
    assert torch.equal(a, m.inference_padded(mel, 0)) and m.inference_padding == 2
    torch.manual_seed(2)
    other = MelganGenerator(**SMALL)
    other.remove_weight_norm()
    m.load_state_dict(other.state_dict())
    b = gi(mel)
    assert torch.equal(b, m.inference_padded(mel, 0))
    assert not torch.equal(a, b)
    assert gi._engine is m._get_engine() and len(gi._graphs) == 1


@pytest.mark.gpu
def test_cfg5_model_bf16_bar_against_the_fp32_oracle():
    """BASELINE config 5 is a BF16 configuration by definition: the streamed BF16 waveform of the (5,5,4,3) FB-MelGAN
    is held to the 40 dB bar against the CPU fp32 oracle of the whole utterance (not only to the GPU's own output)."""
    torch.manual_seed(54321)
    kw = {"upsample_factors": (5, 5, 4, 3)}
    m = FullbandMelganGenerator(**kw)
    m.remove_weight_norm()
    sd = {k: v.detach().clone() for k, v in m.state_dict().items()}
    spec = GeneratorSpec.fullband(**kw)
    m = m.cuda().eval()
    m.precision = "bf16"
    mel = torch.randn(1, 80, 70, generator=torch.Generator().manual_seed(13))
    ref = generator_inference(spec, sd, mel, 2)
    sv = StreamingVocoder(m, chunk_frames=20, use_graph=True)
    parts = []
    for s in range(0, 70, 20):
        parts += sv.feed(mel[..., s:s + 20].cuda())
    got = torch.cat(parts + sv.flush(), dim=-1).cpu()
    assert got.shape == ref.shape
    snr = 10 * torch.log10(ref.double().pow(2).sum() / (ref.double() - got.double()).pow(2).sum())
    assert float(snr) >= 40.0, float(snr)
    m.precision = "fp32"
    got32 = stream_inference(m, mel.cuda(), chunk_frames=20).cpu()
    assert float((got32 - ref).abs().max()) <= 1e-5
