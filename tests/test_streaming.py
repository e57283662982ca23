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
    voc_io.save_checkpoint(src, None, None, None, None, None, 10, 3, str(tmp_path), model_loss=0.5)
    state = torch.load(path, map_location="cpu", weights_only=False)
    assert set(state) >= {"model", "optimizer", "scheduler", "model_disc", "step", "epoch", "date", "model_loss"}
    assert any(k.endswith("weight_g") for k in state["model"]) and "pqmf_layer.G" in state["model"]
    dst = MultibandMelganGenerator()
    dst, st = voc_io.load_checkpoint(dst, path)
    assert st["step"] == 10 and st["epoch"] == 3
    for k, v in src.state_dict().items():
        assert torch.equal(v, dst.state_dict()[k]), k
    # best-model rule: only a lower loss writes
    assert voc_io.save_best_model(0.4, 0.5, src, None, None, None, None, None, 11, 3, str(tmp_path)) == 0.4
    assert os.path.exists(os.path.join(tmp_path, "best_model.pth.tar"))
    assert voc_io.save_best_model(0.6, 0.4, src, None, None, None, None, None, 12, 3, str(tmp_path)) == 0.4
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
