"""Writes tests/golden/mel_*.npz from implementations INDEPENDENT of oracle/mel_oracle.py: ``torch.stft`` (the
reference's own torch STFT, TTS/vocoder/layers/losses.py:16-25, documented there as "compatible with audio.py") and
``torchaudio.functional.melscale_fbanks(norm='slaney', mel_scale='slaney')`` (torchaudio's librosa-compatible
filterbank), followed by the reference's dB / normalisation arithmetic (TTS/utils/audio.py:106-134,190-191) in float64.
librosa itself (requirements.txt:6, ==0.7.2) is not installed in this image.

    python oracle/gen_golden_mel.py          # build container only; the fixtures are committed
"""
import json
import os
import sys

import numpy as np
import torch
import torchaudio

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.mel_oracle import MelSpec  # noqa: E402  (configuration container only)

CASES = {
    "mel_shipped": (MelSpec.shipped(), 22050 * 2 + 123, 1),
    "mel_default": (MelSpec(), 9000, 2),
    "mel_preemph_asym": (MelSpec(sample_rate=16000, fft_size=512, hop_length=200, win_length=400, num_mels=40, mel_fmin=95.0,
                                 mel_fmax=7000.0, preemphasis=0.97, ref_level_db=20.0, symmetric_norm=False, max_norm=1.0), 5000, 3),
}


def signal(n, seed):
    g = np.random.default_rng(seed)
    t = np.arange(n) / 22050.0
    y = 0.3 * np.sin(2 * np.pi * 220 * t) + 0.1 * np.sin(2 * np.pi * 3100 * t + 1.0) + 0.05 * g.standard_normal(n)
    y[: n // 10] *= 1e-3          # a near-silent lead-in: exercises the 1e-5 floor and the clip
    return y


def reference_mel(y, s):
    y = torch.from_numpy(np.asarray(y, dtype=np.float64))
    if s.preemphasis != 0:
        y = torch.cat([y[:1], y[1:] - s.preemphasis * y[:-1]])
    win = torch.hann_window(s.win_length, periodic=True, dtype=torch.float64)
    D = torch.stft(y, s.fft_size, s.hop_length, s.win_length, win, center=True, pad_mode="reflect", normalized=False,
                   onesided=True, return_complex=True)
    fb = torchaudio.functional.melscale_fbanks(1 + s.fft_size // 2, s.mel_fmin, s.mel_fmax or s.sample_rate / 2, s.num_mels,
                                               s.sample_rate, norm="slaney", mel_scale="slaney").double().T
    S = s.spec_gain * torch.log10(torch.clamp(fb @ D.abs(), min=1e-5))
    S = S - s.ref_level_db
    Sn = (S - s.min_level_db) / (-s.min_level_db)
    if s.symmetric_norm:
        Sn = 2 * s.max_norm * Sn - s.max_norm
        Sn = Sn.clamp(-s.max_norm, s.max_norm) if s.clip_norm else Sn
    else:
        Sn = s.max_norm * Sn
        Sn = Sn.clamp(0, s.max_norm) if s.clip_norm else Sn
    return Sn.numpy(), fb.numpy()


def main():
    out_dir = os.path.join(ROOT, "tests", "golden")
    for name, (s, n, seed) in CASES.items():
        y = signal(n, seed)
        mel, fb = reference_mel(y, s)
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), wav=y.astype(np.float32), mel=mel.astype(np.float32),
                            mel_basis=fb.astype(np.float32), spec=json.dumps(s.__dict__),
                            made_by=f"torch {torch.__version__} stft + torchaudio {torchaudio.__version__} melscale_fbanks")
        print(name, mel.shape, float(mel.min()), float(mel.max()))


if __name__ == "__main__":
    main()
