"""CPU oracle of the mel front-end either side of the vocoder.  TEST INFRASTRUCTURE ONLY: imported by
tests/, __graft_entry__.smoke() and bench.py's CPU legs — never by the product.

Restates ``AudioProcessor.melspectrogram`` (TTS/utils/audio.py:216-222): optional pre-emphasis (:198-201), ``_stft``
(:258-266) -> ``np.abs`` -> ``_linear_to_mel`` (:203-204, ``mel_basis`` from :85-93) -> ``_amp_to_db`` (:190-191) ->
``_normalize`` (:106-134), plus ``find_endpoint`` (:302-309).

The STFT and the filterbank live in a third-party dependency that is not in the tree: **librosa == 0.7.2**
(requirements.txt:6), absent from this image.  Its published algorithms are restated here:
  * ``librosa.stft(y, n_fft, hop_length, win_length, window='hann', center=True, pad_mode='reflect')``: pad n_fft//2
    samples per side by reflection, frames of n_fft samples every hop, times the periodic Hann window
    (``scipy.signal.get_window('hann', win_length, fftbins=True)``) zero-padded to n_fft about its centre, real FFT;
    1 + len(y) // hop frames; computed in complex64.
  * ``librosa.filters.mel(sr, n_fft, n_mels, fmin, fmax, htk=False, norm=1)``: Slaney mel scale (linear below 1 kHz,
    log above, 200/3 Hz per mel), triangular filters between consecutive mel points, each scaled by
    2 / (f[m+2] - f[m]); float32.
Pinned (tests/test_mel_oracle.py) against independent implementations present here — ``torch.stft`` and
``torchaudio.functional.melscale_fbanks(norm='slaney', mel_scale='slaney')``, which torchaudio documents as
librosa-compatible — through the fixtures oracle/gen_golden_mel.py writes to tests/golden/mel_*.npz.
"""
from dataclasses import dataclass
from typing import Optional

import numpy as np
import scipy.signal


@dataclass
class MelSpec:
    """The AudioProcessor constructor arguments this path reads (TTS/utils/audio.py:15-71)."""
    sample_rate: int = 22050
    fft_size: int = 1024
    hop_length: int = 256
    win_length: int = 1024
    num_mels: int = 80
    mel_fmin: float = 0.0
    mel_fmax: Optional[float] = None
    preemphasis: float = 0.0
    spec_gain: float = 20.0
    ref_level_db: float = 0.0
    min_level_db: float = -100.0
    signal_norm: bool = True
    symmetric_norm: bool = True
    max_norm: float = 4.0
    clip_norm: bool = True

    @staticmethod
    def shipped():
        """TTS/vocoder/configs/multiband_melgan_config.json:6-34 (range normalisation; the shipped stats_path is absent)."""
        return MelSpec(mel_fmin=50.0, mel_fmax=7600.0, spec_gain=1.0)


def hz_to_mel(f):
    f = np.asanyarray(f, dtype=np.float64)
    f_sp = 200.0 / 3
    mels = f / f_sp
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-10) / min_log_hz) / logstep, mels)


def mel_to_hz(m):
    m = np.asanyarray(m, dtype=np.float64)
    f_sp = 200.0 / 3
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)


def mel_basis(spec: MelSpec):
    """librosa.filters.mel as AudioProcessor._build_mel_basis calls it (audio.py:85-93): [num_mels, 1 + fft_size//2] float32."""
    fmax = spec.sample_rate / 2.0 if spec.mel_fmax is None else float(spec.mel_fmax)
    n_freq = 1 + spec.fft_size // 2
    fftfreqs = np.linspace(0, spec.sample_rate / 2.0, n_freq)
    mel_f = mel_to_hz(np.linspace(hz_to_mel(spec.mel_fmin), hz_to_mel(fmax), spec.num_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = np.subtract.outer(mel_f, fftfreqs)
    w = np.zeros((spec.num_mels, n_freq))
    for i in range(spec.num_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        w[i] = np.maximum(0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_f[2:spec.num_mels + 2] - mel_f[:spec.num_mels])
    return (w * enorm[:, None]).astype(np.float32)


def stft_window(spec: MelSpec):
    """Periodic Hann of win_length, zero-padded to fft_size about its centre (librosa.util.pad_center)."""
    w = scipy.signal.get_window("hann", spec.win_length, fftbins=True)
    lpad = (spec.fft_size - spec.win_length) // 2
    out = np.zeros(spec.fft_size)
    out[lpad:lpad + spec.win_length] = w
    return out


def preemphasis(y, coef):
    """scipy.signal.lfilter([1, -coef], [1], y) (audio.py:198-201): y[n] - coef * y[n-1], zero initial state."""
    out = np.array(y, dtype=np.float64, copy=True)
    out[1:] -= coef * np.asarray(y, dtype=np.float64)[:-1]
    return out


def stft_magnitude(y, spec: MelSpec, dtype=np.float64):
    """|librosa.stft(y)|: [1 + fft_size//2, 1 + len(y)//hop]."""
    y = np.asarray(y, dtype=dtype)
    n = spec.fft_size
    if len(y) <= n // 2:
        raise ValueError("reflect padding needs len(y) > fft_size // 2")
    yp = np.pad(y, n // 2, mode="reflect")
    n_frames = 1 + len(y) // spec.hop_length
    idx = np.arange(n)[None, :] + spec.hop_length * np.arange(n_frames)[:, None]
    frames = yp[idx] * stft_window(spec).astype(dtype)[None, :]
    return np.abs(np.fft.rfft(frames, axis=1)).T.astype(dtype)


def amp_to_db(x, spec: MelSpec):
    return spec.spec_gain * np.log10(np.maximum(1e-5, x))                       # audio.py:190-191


def normalize(S, spec: MelSpec, mean=None, scale=None):
    """audio.py:106-134.  mean / scale: the StandardScaler statistics of a stats_path config (TTS/tts/utils/data.py:66-70)."""
    if not spec.signal_norm:
        return S
    if mean is not None:
        return (S - np.asarray(mean)[:, None]) / np.asarray(scale)[:, None]
    S = S - spec.ref_level_db
    Sn = (S - spec.min_level_db) / (-spec.min_level_db)
    if spec.symmetric_norm:
        Sn = 2 * spec.max_norm * Sn - spec.max_norm
        return np.clip(Sn, -spec.max_norm, spec.max_norm) if spec.clip_norm else Sn
    Sn = spec.max_norm * Sn
    return np.clip(Sn, 0, spec.max_norm) if spec.clip_norm else Sn


def melspectrogram(y, spec: MelSpec, mean=None, scale=None, dtype=np.float64):
    """AudioProcessor.melspectrogram (audio.py:216-222): [num_mels, frames]."""
    y = np.asarray(y, dtype=np.float64)
    if spec.preemphasis != 0:
        y = preemphasis(y, spec.preemphasis)
    mag = stft_magnitude(y, spec, dtype)
    S = amp_to_db(mel_basis(spec).astype(dtype) @ mag, spec)
    return normalize(S, spec, mean, scale).astype(dtype)


def find_endpoint(wav, spec: MelSpec, threshold_db=-40, min_silence_sec=0.8):
    """audio.py:302-309 verbatim in behaviour (note: max of the signed samples, not of |wav|, as the reference does)."""
    window_length = int(spec.sample_rate * min_silence_sec)
    hop_length = int(window_length / 4)
    threshold = np.power(10.0, threshold_db / spec.spec_gain)
    for x in range(hop_length, len(wav) - window_length, hop_length):
        if np.max(wav[x:x + window_length]) < threshold:
            return x + hop_length
    return len(wav)


def trim_silence(wav, spec: MelSpec, trim_db=60):
    """AudioProcessor.trim_silence (audio.py:311-316): drop a 0.01 s margin per side, then
    ``librosa.effects.trim(wav, top_db=trim_db, frame_length=win_length, hop_length=hop_length)[0]`` — librosa 0.7.2:
    mean-square per centred frame of the reflect-padded signal (``librosa.feature.rms(center=True)**2``),
    ``power_to_db(mse, ref=np.max, top_db=None)`` (amin 1e-10), frames above ``-top_db`` are non-silent; the signal is
    cut to ``[first * hop, min(len, (last + 1) * hop))``."""
    margin = int(spec.sample_rate * 0.01)
    wav = np.asarray(wav, dtype=np.float64)[margin:-margin]
    n = len(wav)
    half = spec.win_length // 2
    yp = np.pad(wav, half, mode="reflect")
    n_frames = 1 + n // spec.hop_length
    idx = np.arange(spec.win_length)[None, :] + spec.hop_length * np.arange(n_frames)[:, None]
    idx = np.minimum(idx, len(yp) - 1)          # frames are complete for win_length <= 2 * half + hop remainder; guard only
    mse = np.mean(yp[idx] ** 2, axis=1)
    ref = max(1e-10, mse.max())
    db = 10.0 * np.log10(np.maximum(1e-10, mse)) - 10.0 * np.log10(ref)
    keep = np.flatnonzero(db > -trim_db)
    if keep.size == 0:
        return wav[:0]
    start = int(keep[0]) * spec.hop_length
    end = min(n, (int(keep[-1]) + 1) * spec.hop_length)
    return wav[start:end]
