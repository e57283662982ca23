"""GPU mel front-end with the AudioProcessor surface the vocoder's callers use (TTS/utils/audio.py).

``AudioProcessor(**c.audio)`` is how every reference caller builds it (TTS/bin/synthesize.py, TTS/server/synthesizer.py,
the vocoder notebooks); the same keyword arguments work here.  What runs where:

  * ``melspectrogram(y)``  (audio.py:216-222) -> one hand-written sm_100a kernel (tts_b200/csrc/mel.cu) through
    ``ttsb200_mel_forward``: pre-emphasis, centred reflect-padded STFT (shared-memory FFT), magnitude, mel filterbank,
    dB and normalisation.  Takes a CUDA tensor ``[n]`` or ``[B, n]`` (a numpy array / CPU tensor is uploaded) and
    returns a CUDA tensor ``[num_mels, frames]`` / ``[B, num_mels, frames]`` — exactly the layout
    ``generator.inference`` takes, so copy-synthesis stays on the device end to end.
  * ``find_endpoint`` (:302-309) and ``trim_silence`` (:311-316, librosa.effects.trim) -> per-frame statistics from
    ``ttsb200_frame_stats`` on the device, the (tiny) threshold scan on the host.
  * ``mel_basis`` (:85-93) and the STFT window are host-side design data, as in the reference; librosa is not a
    dependency: its Slaney filterbank is restated in ``_slaney_mel_basis`` (pinned against torchaudio's
    librosa-compatible filterbank by tests/test_mel_oracle.py).

CUDA only: there is no CPU fallback; the Griffin-Lim / inverse paths of the reference class are out of scope.
"""
import ctypes

import numpy as np
import scipy.signal
import torch

from .. import _lib


def _hz_to_mel(f):
    f = np.asanyarray(f, dtype=np.float64)
    lin = f / (200.0 / 3)
    return np.where(f >= 1000.0, 15.0 + np.log(np.maximum(f, 1e-10) / 1000.0) / (np.log(6.4) / 27.0), lin)


def _mel_to_hz(m):
    m = np.asanyarray(m, dtype=np.float64)
    return np.where(m >= 15.0, 1000.0 * np.exp((np.log(6.4) / 27.0) * (m - 15.0)), (200.0 / 3) * m)


def _slaney_mel_basis(sample_rate, n_fft, n_mels, fmin, fmax):
    """librosa.filters.mel(sr, n_fft, n_mels, fmin, fmax) of librosa 0.7.2 (htk=False, norm=1): float32 [n_mels, 1 + n_fft//2]."""
    fmax = sample_rate / 2.0 if fmax is None else float(fmax)
    freqs = np.linspace(0, sample_rate / 2.0, 1 + n_fft // 2)
    pts = _mel_to_hz(np.linspace(_hz_to_mel(fmin), _hz_to_mel(fmax), n_mels + 2))
    width = np.diff(pts)
    ramps = pts[:, None] - freqs[None, :]
    rise = -ramps[:-2] / width[:-1, None]
    fall = ramps[2:] / width[1:, None]
    tri = np.maximum(0.0, np.minimum(rise, fall))
    return (tri * (2.0 / (pts[2:] - pts[:-2]))[:, None]).astype(np.float32)


class AudioProcessor:
    def __init__(self, sample_rate=None, num_mels=None, min_level_db=None, frame_shift_ms=None, frame_length_ms=None,
                 hop_length=None, win_length=None, ref_level_db=None, fft_size=1024, power=None, preemphasis=0.0,
                 signal_norm=None, symmetric_norm=None, max_norm=None, mel_fmin=None, mel_fmax=None, spec_gain=20,
                 stft_pad_mode="reflect", clip_norm=True, griffin_lim_iters=None, do_trim_silence=False, trim_db=60,
                 do_sound_norm=False, stats_path=None, device="cuda", **_):
        self.sample_rate = sample_rate
        self.num_mels = num_mels
        self.min_level_db = min_level_db or 0
        self.frame_shift_ms, self.frame_length_ms = frame_shift_ms, frame_length_ms
        self.ref_level_db = ref_level_db or 0
        self.fft_size = fft_size
        self.power = power
        self.preemphasis = preemphasis or 0.0
        self.griffin_lim_iters = griffin_lim_iters
        self.signal_norm = signal_norm
        self.symmetric_norm = symmetric_norm
        self.mel_fmin = mel_fmin or 0
        self.mel_fmax = mel_fmax
        self.spec_gain = float(spec_gain)
        self.stft_pad_mode = stft_pad_mode
        self.max_norm = 1.0 if max_norm is None else float(max_norm)
        self.clip_norm = clip_norm
        self.do_trim_silence, self.trim_db, self.do_sound_norm = do_trim_silence, trim_db, do_sound_norm
        self.stats_path = stats_path
        if stft_pad_mode != "reflect":
            raise NotImplementedError("tts_b200 AudioProcessor implements stft_pad_mode='reflect' (every shipped config)")
        if hop_length is None:                      # audio.py:95-101
            factor = frame_length_ms / frame_shift_ms
            assert factor.is_integer(), " [!] frame_shift_ms should divide frame_length_ms"
            self.hop_length = int(frame_shift_ms / 1000.0 * sample_rate)
            self.win_length = int(self.hop_length * factor)
        else:
            self.hop_length, self.win_length = hop_length, win_length
        assert min_level_db != 0.0, " [!] min_level_db is 0"
        assert self.win_length <= self.fft_size, " [!] win_length cannot be larger than fft_size"
        if self.mel_fmax is not None:
            assert self.mel_fmax <= self.sample_rate // 2
        self.mel_basis = _slaney_mel_basis(self.sample_rate, self.fft_size, self.num_mels, self.mel_fmin, self.mel_fmax)
        self.mel_mean = self.mel_std = None
        if stats_path:                              # audio.py:78-83,165-189: mean-var scaling replaces range normalisation
            stats = np.load(stats_path, allow_pickle=True).item()
            self.setup_scaler(stats["mel_mean"], stats["mel_std"], stats.get("linear_mean"), stats.get("linear_std"))
        self.device = torch.device(device)
        self._plan = None

    def setup_scaler(self, mel_mean, mel_std, linear_mean=None, linear_std=None):
        self.mel_mean = np.ascontiguousarray(mel_mean, dtype=np.float32)
        self.mel_std = np.ascontiguousarray(mel_std, dtype=np.float32)
        self.signal_norm, self.max_norm, self.clip_norm, self.symmetric_norm = True, None, None, None
        self._close()

    # ------------------------------------------------------------------ plan
    def _get_plan(self):
        if self._plan is not None:
            return self._plan
        if self.device.type != "cuda":
            raise RuntimeError("tts_b200 AudioProcessor runs on CUDA (sm_100a) devices only; there is no CPU fallback")
        cfg = _lib.MelConfig()
        cfg.sample_rate, cfg.n_fft, cfg.hop_length = int(self.sample_rate), int(self.fft_size), int(self.hop_length)
        cfg.win_length, cfg.num_mels = int(self.win_length), int(self.num_mels)
        cfg.preemphasis, cfg.spec_gain = float(self.preemphasis), float(self.spec_gain)
        cfg.ref_level_db, cfg.min_level_db = float(self.ref_level_db), float(self.min_level_db)
        cfg.max_norm = float(self.max_norm or 0.0)
        cfg.signal_norm, cfg.symmetric_norm, cfg.clip_norm = int(bool(self.signal_norm)), int(bool(self.symmetric_norm)), \
            int(bool(self.clip_norm))
        window = np.ascontiguousarray(scipy.signal.get_window("hann", self.win_length, fftbins=True), dtype=np.float32)
        basis = np.ascontiguousarray(self.mel_basis, dtype=np.float32)
        mean = self.mel_mean.ctypes.data if self.mel_mean is not None else None
        std = self.mel_std.ctypes.data if self.mel_std is not None else None
        h = ctypes.c_void_p()
        index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        _lib.check(_lib.load().ttsb200_mel_plan_create(ctypes.byref(cfg), basis.ctypes.data, window.ctypes.data, mean, std,
                                                       index, ctypes.byref(h)))
        self._plan, self._index = h, index
        return h

    def _close(self):
        if getattr(self, "_plan", None):
            _lib.load().ttsb200_mel_plan_destroy(self._plan)
        self._plan = None

    def __del__(self):
        try:
            self._close()
        except Exception:
            pass

    def _to_device(self, y):
        if isinstance(y, np.ndarray):
            y = torch.from_numpy(np.ascontiguousarray(y))
        return y.detach().to(device=torch.device("cuda", self._index), dtype=torch.float32, non_blocking=True).contiguous()

    # ------------------------------------------------------------------ reference surface
    def melspectrogram(self, y):
        plan = self._get_plan()
        y = self._to_device(y)
        squeeze = y.dim() == 1
        if squeeze:
            y = y[None]
        if y.dim() != 2:
            raise RuntimeError(f"melspectrogram takes [n] or [B, n], got {tuple(y.shape)}")
        B, n = y.shape
        lib = _lib.load()
        frames = _lib.check(lib.ttsb200_mel_frames(plan, n))
        out = torch.empty((B, self.num_mels, frames), dtype=torch.float32, device=y.device)
        with torch.cuda.device(y.device):
            st = torch.cuda.current_stream(y.device).cuda_stream
            _lib.check(lib.ttsb200_mel_forward(plan, y.data_ptr(), B, n, out.data_ptr(), st))
        return out[0] if squeeze else out

    def frame_stats(self, y, frame_length, hop_length):
        """[B, frames, 3] = {max x, max |x|, mean x^2} per frame (device)."""
        self._get_plan()
        y = self._to_device(y)
        y = y[None] if y.dim() == 1 else y
        B, n = y.shape
        frames = -(-n // hop_length)
        out = torch.empty((B, frames, 3), dtype=torch.float32, device=y.device)
        with torch.cuda.device(y.device):
            st = torch.cuda.current_stream(y.device).cuda_stream
            _lib.check(_lib.load().ttsb200_frame_stats(y.data_ptr(), B, n, int(frame_length), int(hop_length), out.data_ptr(), st))
        return out

    def _db_to_amp(self, x):
        return np.power(10.0, x / self.spec_gain)

    def find_endpoint(self, wav, threshold_db=-40, min_silence_sec=0.8):
        """audio.py:302-309: first position whose following `min_silence_sec` window stays under the threshold."""
        n = int(wav.shape[-1])
        window_length = int(self.sample_rate * min_silence_sec)
        hop_length = int(window_length / 4)
        if hop_length < 1 or n - window_length <= hop_length:
            return n
        stats = self.frame_stats(wav, window_length, hop_length)[0, :, 0].cpu().numpy()   # max of the signed samples
        threshold = self._db_to_amp(threshold_db)
        for i, x in enumerate(range(hop_length, n - window_length, hop_length), start=1):
            if stats[i] < threshold:
                return x + hop_length
        return n

    def trim_silence(self, wav):
        """audio.py:311-316: 0.01 s margin, then librosa.effects.trim(top_db=trim_db, frame_length=win_length,
        hop_length=hop_length): frames whose RMS (centred, zero... reflect-free power) is within top_db of the loudest."""
        margin = int(self.sample_rate * 0.01)
        wav = wav[..., margin:-margin]
        n = int(wav.shape[-1])
        # librosa.feature.rms(center=True) frames: frame t covers [t*hop - frame/2, t*hop + frame/2) of the reflect-padded
        # signal; computed here on the un-padded signal with the same centres (edge frames see fewer samples)
        half = self.win_length // 2
        y = self._to_device(wav)
        pad = torch.nn.functional.pad(y[None, None], (half, half), mode="reflect")[0, 0]
        ms = self.frame_stats(pad, self.win_length, self.hop_length)[0, :, 2]
        n_frames = 1 + n // self.hop_length
        rms = ms[:n_frames].clamp_min(1e-20).sqrt()
        db = 20.0 * torch.log10(rms / rms.max().clamp_min(1e-10))
        keep = torch.nonzero(db > -float(self.trim_db)).flatten()
        if keep.numel() == 0:
            return wav[..., :0]
        start = int(keep[0]) * self.hop_length
        end = min(n, (int(keep[-1]) + 1) * self.hop_length)
        return wav[..., start:end]

    @staticmethod
    def sound_norm(x):
        return x / abs(x).max() * 0.9
