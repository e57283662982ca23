"""Mel front-end: the CPU oracle against fixtures made by independent implementations (torch.stft +
torchaudio's librosa-compatible Slaney filterbank, oracle/gen_golden_mel.py), host-side design data of the product
class, and — on a GPU — the CUDA kernel against the oracle."""
import json

import numpy as np
import pytest
import torch

from oracle import mel_oracle as mo
from tts_b200.utils.audio import AudioProcessor, _slaney_mel_basis

CASES = ["mel_shipped", "mel_default", "mel_preemph_asym"]
# normalised mel units: the shipped configs map 100 dB onto [-4, 4], so 2e-3 is 0.025 dB.  The reference itself computes
# the STFT in complex64 (librosa 0.7.2 default dtype); fp32 FFT rounding on near-silent frames is what the bar allows.
MEL_TOL = 2e-3


def spec_of(g):
    return mo.MelSpec(**json.loads(str(g["spec"])))


def processor(s, **kw):
    return AudioProcessor(sample_rate=s.sample_rate, num_mels=s.num_mels, min_level_db=s.min_level_db, hop_length=s.hop_length,
                          win_length=s.win_length, ref_level_db=s.ref_level_db, fft_size=s.fft_size, preemphasis=s.preemphasis,
                          signal_norm=s.signal_norm, symmetric_norm=s.symmetric_norm, max_norm=s.max_norm, mel_fmin=s.mel_fmin,
                          mel_fmax=s.mel_fmax, spec_gain=s.spec_gain, clip_norm=s.clip_norm, **kw)


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_independent_fixtures(golden, name):
    g = golden(name)
    s = spec_of(g)
    assert np.abs(mo.mel_basis(s) - g["mel_basis"]).max() <= 2e-7
    mel = mo.melspectrogram(g["wav"].astype(np.float64), s)
    assert mel.shape == g["mel"].shape == (s.num_mels, 1 + len(g["wav"]) // s.hop_length)
    assert np.abs(mel - g["mel"]).max() <= 1e-5
    mel32 = mo.melspectrogram(g["wav"].astype(np.float64), s, dtype=np.float32)       # the reference's own precision
    assert np.abs(mel32 - g["mel"]).max() <= 1e-4


@pytest.mark.parametrize("name", CASES)
def test_product_design_data_matches_oracle(golden, name):
    s = spec_of(golden(name))
    ours = _slaney_mel_basis(s.sample_rate, s.fft_size, s.num_mels, s.mel_fmin, s.mel_fmax)
    assert np.abs(ours - mo.mel_basis(s)).max() <= 1e-7                  # two float64 derivations, rounded to float32
    ap = processor(s, device="cpu")
    assert ap.mel_basis.dtype == np.float32 and ap.mel_basis.shape == (s.num_mels, 1 + s.fft_size // 2)
    with pytest.raises(RuntimeError, match="CUDA"):
        ap.melspectrogram(np.zeros(4000, np.float32))                    # no CPU fallback


def test_constructor_contract():
    with pytest.raises(AssertionError):
        AudioProcessor(sample_rate=22050, num_mels=80, min_level_db=0.0, hop_length=256, win_length=1024)
    with pytest.raises(AssertionError):
        AudioProcessor(sample_rate=22050, num_mels=80, min_level_db=-100, hop_length=256, win_length=2048, fft_size=1024)
    ap = AudioProcessor(sample_rate=16000, num_mels=40, min_level_db=-100, frame_shift_ms=12.5, frame_length_ms=50.0,
                        hop_length=None, device="cpu")
    assert (ap.hop_length, ap.win_length) == (200, 800)                  # audio.py:95-101


def test_endpoint_and_trim_oracle_on_constructed_signal():
    s = mo.MelSpec.shipped()
    g = np.random.default_rng(0)
    sr = s.sample_rate
    y = np.concatenate([1e-5 * g.standard_normal(sr // 2), 0.3 * g.standard_normal(sr), 1e-5 * g.standard_normal(2 * sr)])
    cut = mo.trim_silence(y, s, trim_db=60)
    margin = int(sr * 0.01)
    assert abs((len(y) - 2 * margin) - len(cut) - (sr // 2 - margin) - (2 * sr - margin)) <= 4 * s.hop_length
    s20 = mo.MelSpec(spec_gain=20.0)
    assert mo.find_endpoint(y, s20) < len(y)
    assert mo.find_endpoint(0.3 * g.standard_normal(3 * sr), s20) == 3 * sr


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_gpu_mel_matches_oracle_and_fixture(golden, name):
    g = golden(name)
    s = spec_of(g)
    ap = processor(s)
    wav = g["wav"]
    got = ap.melspectrogram(wav)                                         # numpy in (uploaded), CUDA tensor out
    assert got.is_cuda and tuple(got.shape) == g["mel"].shape
    ref = mo.melspectrogram(wav.astype(np.float64), s)
    assert np.abs(got.cpu().numpy() - ref).max() <= MEL_TOL
    assert np.abs(got.cpu().numpy() - g["mel"]).max() <= MEL_TOL
    # batch: rows independent, same numbers
    both = ap.melspectrogram(torch.from_numpy(np.stack([wav, wav[::-1].copy()])).cuda())
    assert torch.equal(both[0], got)
    assert np.abs(both[1].cpu().numpy() - mo.melspectrogram(wav[::-1].astype(np.float64), s)).max() <= MEL_TOL


@pytest.mark.gpu
def test_gpu_mel_edge_cases():
    s = mo.MelSpec(fft_size=512, hop_length=128, win_length=400, num_mels=20, sample_rate=8000)
    ap = processor(s)
    g = np.random.default_rng(3)
    for n in (257, 300, 511, 512, 513, 1151, 1152):                      # shortest legal, ragged tails, odd frame counts
        y = (0.2 * g.standard_normal(n)).astype(np.float32)
        got = ap.melspectrogram(y).cpu().numpy()
        assert got.shape == (20, 1 + n // 128)
        assert np.abs(got - mo.melspectrogram(y.astype(np.float64), s)).max() <= MEL_TOL, n
    with pytest.raises(RuntimeError, match="reflect"):
        ap.melspectrogram(np.zeros(256, np.float32))                     # numpy's reflect pad raises here too
    # mean-var scaling (stats_path configs): StandardScaler.transform per mel row
    mean, std = g.standard_normal(20).astype(np.float32), (0.5 + g.random(20)).astype(np.float32)
    ap.setup_scaler(mean, std)
    y = (0.2 * g.standard_normal(4000)).astype(np.float32)
    ref = mo.melspectrogram(y.astype(np.float64), s, mean=mean, scale=std)
    assert np.abs(ap.melspectrogram(y).cpu().numpy() - ref).max() <= MEL_TOL * 8     # dB units / std, not [-4, 4]
    # all four FFT sizes the kernel instantiates
    for n_fft in (256, 512, 1024, 2048):
        s2 = mo.MelSpec(fft_size=n_fft, hop_length=n_fft // 4, win_length=n_fft, num_mels=32)
        y = (0.1 * g.standard_normal(3 * n_fft + 17)).astype(np.float32)
        assert np.abs(processor(s2).melspectrogram(y).cpu().numpy() - mo.melspectrogram(y.astype(np.float64), s2)).max() <= MEL_TOL


@pytest.mark.gpu
def test_gpu_endpoint_and_trim_match_oracle():
    s = mo.MelSpec(spec_gain=20.0)
    ap = processor(s)
    g = np.random.default_rng(5)
    sr = s.sample_rate
    y = np.concatenate([1e-5 * g.standard_normal(sr // 3), 0.3 * g.standard_normal(sr), 1e-5 * g.standard_normal(2 * sr)]).astype(np.float32)
    assert ap.find_endpoint(y) == mo.find_endpoint(y, s)
    loud = (0.3 * g.standard_normal(2 * sr)).astype(np.float32)
    assert ap.find_endpoint(loud) == mo.find_endpoint(loud, s) == len(loud)
    cut = ap.trim_silence(y)
    want = mo.trim_silence(y, s, trim_db=60)
    assert len(cut) == len(want) and np.array_equal(np.asarray(cut), want.astype(np.float32))


@pytest.mark.gpu
def test_copy_synthesis_stays_on_device():
    """wav -> mel (this front-end) -> MB-MelGAN waveform without leaving the GPU: shapes chain as in the notebooks."""
    from tts_b200.vocoder.models.multiband_melgan_generator import MultibandMelganGenerator
    s = mo.MelSpec.shipped()
    ap = processor(s)
    y = torch.randn(2, 22050, generator=torch.Generator().manual_seed(1)) * 0.1
    mel = ap.melspectrogram(y.cuda())
    assert mel.shape == (2, 80, 87) and mel.is_cuda
    m = MultibandMelganGenerator().cuda().eval()
    m.remove_weight_norm()
    m.inference_padding = 0
    wav = m.inference(mel)
    assert wav.shape == (2, 1, 87 * 256) and bool(torch.isfinite(wav).all())
