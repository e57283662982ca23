"""MelganGenerator: the reference's Python surface over the B200 CUDA plan.

Surface kept (TTS/vocoder/models/melgan_generator.py:8-97): constructor
arguments, ``layers`` (an ``nn.Sequential`` with the reference's indices and state-dict keys),
``inference_padding``, ``forward``, ``inference``, ``remove_weight_norm``, plus normal
``nn.Module`` behaviour (``load_state_dict``, ``.cuda()``, ``.eval()``, ``.parameters()``).

What runs where:
  * ``inference(c)`` and eval-mode ``forward(c)`` -> hand-written sm_100a kernels through
    include/tts_b200.h (tts_b200/engine.py).  CUDA only; raises on CPU-resident modules.
  * training-mode ``forward(c)`` -> ``self.layers`` with autograd (training is out of scope;
    the trainer needs gradients, train_vocoder.py:111).
New attribute: ``precision`` = "fp32" (max-abs <= 1e-5 vs reference) or "bf16" (tcgen05, SNR >= 40 dB).
"""
import threading

import torch
from torch import nn
from torch.nn.utils import weight_norm

from ..layers.melgan import ResidualStack
from ... import _lib


class MelganGenerator(nn.Module):
    def __init__(self,
                 in_channels=80,
                 out_channels=1,
                 proj_kernel=7,
                 base_channels=512,
                 upsample_factors=(8, 8, 2, 2),
                 res_kernel=3,
                 num_res_blocks=3):
        super().__init__()
        assert (proj_kernel - 1) % 2 == 0, " [!] proj_kernel should be an odd number."
        self.in_channels = in_channels
        self.out_channels = out_channels
        self.proj_kernel = proj_kernel
        self.base_channels = base_channels
        self.upsample_factors = tuple(int(f) for f in upsample_factors)
        self.res_kernel = res_kernel
        self.num_res_blocks = num_res_blocks
        self.inference_padding = 2
        self.precision = "fp32"

        edge = (proj_kernel - 1) // 2
        seq = [nn.ReflectionPad1d(edge),
               weight_norm(nn.Conv1d(in_channels, base_channels, kernel_size=proj_kernel, stride=1, bias=True))]
        width = base_channels
        for f in self.upsample_factors:
            odd = f % 2
            seq.append(nn.LeakyReLU(0.2))
            seq.append(weight_norm(nn.ConvTranspose1d(width, width // 2, 2 * f, stride=f, padding=f // 2 + odd,
                                                      output_padding=odd, bias=True)))
            seq.append(ResidualStack(channels=width // 2, num_res_blocks=num_res_blocks, kernel_size=res_kernel))
            width //= 2
        seq += [nn.LeakyReLU(0.2),
                nn.ReflectionPad1d(edge),
                weight_norm(nn.Conv1d(width, out_channels, proj_kernel, stride=1, bias=True)),
                nn.Tanh()]
        self.layers = nn.Sequential(*seq)

        self._engine = None
        self._engine_sig = None
        self._engine_lock = threading.Lock()
        self._frozen = False

    # ------------------------------------------------------------------ plan management
    def _signature(self):
        return tuple((p.data_ptr(), p._version) for p in self.parameters())

    def _get_engine(self):
        dev = self.layers[1].bias.device
        if dev.type != "cuda":
            raise RuntimeError(
                "tts_b200: this generator runs on CUDA (sm_100a) only — move it with .cuda() first. "
                "There is no CPU fallback for inference.")
        if self._engine is not None and self._frozen:
            return self._engine
        sig = (dev, self._signature())
        with self._engine_lock:
            if self._engine is None or self._engine_sig != sig:
                from ...engine import GeneratorEngine
                # The old engine is released, not closed: another thread may be inside its forward(), and CUDA
                # graphs captured from it (tts_b200/streaming.py) hold a reference.  Its plan is destroyed when the
                # last reference goes (GeneratorEngine.__del__; cudaFree waits for in-flight work).
                self._engine = GeneratorEngine(self, dev)
                self._engine_sig = sig
        return self._engine

    def freeze(self, frozen=True):
        """Skip the per-call weight-change check (latency path). Call after the weights are final."""
        self._frozen = frozen
        return self

    def _apply(self, fn, *a, **kw):  # .cuda()/.to()/.float() invalidate the plan
        self._engine_sig = None
        self._frozen = False
        return super()._apply(fn, *a, **kw)

    _SKIP_PQMF = True   # MultibandMelganGenerator.inference appends PQMF synthesis (multiband_melgan_generator.py:38)

    def _run(self, c, pad, skip_pqmf, tap_layers=None, profile=None, out=None, out_dtype=None, keep=None):
        eng = self._get_engine()
        c = c.detach().to(device=eng.device, dtype=torch.float32, non_blocking=True)   # async when c is pinned
        prec = _lib.PRECISIONS[self.precision]
        return eng.forward(c, int(pad), prec, skip_pqmf=skip_pqmf, tap_layers=tap_layers, profile=profile, out=out,
                           out_dtype=out_dtype, keep=keep)

    def inference_padded(self, c, pad, out=None, keep=None):
        """``inference`` with an explicit padding for this call only.  Long-form and streaming windows run with
        padding 0 (tts_b200/longform.py, streaming.py); they must not change ``inference_padding`` on a module
        that request threads share (TTS/server/server.py:64,82).  ``keep=(lo, n)`` with ``out`` stores only that
        sample range, in place (GeneratorEngine.forward)."""
        return self._run(c, int(pad), self._SKIP_PQMF, out=out, keep=keep)

    # ------------------------------------------------------------------ reference surface
    def forward(self, c):
        if self.training:
            return self.layers(c)
        return self._run(c, 0, True)

    def inference(self, c):
        return self._run(c, self.inference_padding, True)

    def inference_to_host(self, c, out=None, dtype=torch.float32):
        """inference() followed by an asynchronous device->host copy into pinned memory, all on the current
        stream, so the caller's ``.cpu().numpy()`` sync disappears (TTS/server/synthesizer.py:155-159,
        SURVEY.md §8f-2).  Returns ``(wav_host, event)``: ``wav_host`` (pinned, reusable via ``out=``) is valid
        once ``event.synchronize()`` returns.  With a pinned ``c`` the upload is asynchronous too, so calls on
        alternating streams overlap copies with compute.

        ``dtype=torch.float16`` makes the last kernel store the waveform as fp16 (|wav| <= 1 after tanh / PQMF, so
        the rounding is <= 2^-11 relative, well inside the 16-bit WAV every reference caller ends in,
        TTS/utils/audio.py:338-340) and halves the device->host bytes."""
        y = self._run(c, self.inference_padding, self._SKIP_PQMF, out_dtype=dtype)
        if out is None:
            out = torch.empty(y.shape, dtype=y.dtype, pin_memory=True)
        out.copy_(y, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(y.device))
        return out, ev

    def inference_with_taps(self, c, tap_layers, profile=None):
        """inference() that also returns {layers[i] index: output} for per-layer parity tests."""
        return self._run(c, self.inference_padding, self._SKIP_PQMF, tap_layers=list(tap_layers), profile=profile)

    def remove_weight_norm(self):
        for m in self.modules():
            if isinstance(m, (nn.Conv1d, nn.ConvTranspose1d)) and hasattr(m, "weight_g"):
                nn.utils.remove_weight_norm(m)
        self._engine_sig = None
