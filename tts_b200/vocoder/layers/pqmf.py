"""PQMF with the reference's buffers and methods (TTS/vocoder/layers/pqmf.py:10-56).

``H``, ``G`` and ``updown_filter`` are computed with the same scipy/numpy calls as the reference
so they are bit-identical (they are part of MB-MelGAN checkpoints).  ``synthesis`` / ``analysis``
on CUDA tensors run the polyphase kernels in tts_b200/csrc/pqmf.cu through the C ABI.
"""
import numpy as np
import torch
import torch.nn.functional as F
from scipy import signal as sig

from ... import _lib


class PQMF(torch.nn.Module):
    def __init__(self, N=4, taps=62, cutoff=0.15, beta=9.0):
        super().__init__()
        self.N, self.taps, self.cutoff, self.beta = N, taps, cutoff, beta
        proto = sig.firwin(taps + 1, cutoff, window=("kaiser", beta))
        n = np.arange(taps + 1)
        H = np.zeros((N, taps + 1))
        G = np.zeros((N, taps + 1))
        for k in range(N):
            # centre (taps-1)/2 kept as in pqmf.py:23-25 (the reference's own TODO); do not "fix"
            arg = (2 * k + 1) * (np.pi / (2 * N)) * (n - ((taps - 1) / 2))
            ph = (-1) ** k * np.pi / 4
            H[k] = 2 * proto * np.cos(arg + ph)
            G[k] = 2 * proto * np.cos(arg - ph)
        self.register_buffer("H", torch.from_numpy(H[:, None, :]).float())
        self.register_buffer("G", torch.from_numpy(G[None, :, :]).float())
        up = torch.zeros((N, N, N)).float()
        for k in range(N):
            up[k, k, 0] = 1.0
        self.register_buffer("updown_filter", up)
        self.pad_fn = torch.nn.ConstantPad1d(taps // 2, 0.0)

    def forward(self, x):
        return self.analysis(x)

    @staticmethod
    def _needs_autograd(x):
        return torch.is_grad_enabled() and x.requires_grad

    def _require_cuda(self, x, what):
        if not x.is_cuda or not self.G.is_cuda:
            raise RuntimeError(f"tts_b200 PQMF.{what}: input and module must be on a CUDA (sm_100a) device; "
                               "there is no CPU fallback")

    def analysis(self, x):
        if self._needs_autograd(x):  # training graph (train_vocoder.py:119-121) — out of scope path
            return F.conv1d(x, self.H, padding=self.taps // 2, stride=self.N)
        self._require_cuda(x, "analysis")
        x = x.detach().float().contiguous()
        B, one, T = x.shape
        if one != 1:
            raise RuntimeError(f"PQMF.analysis expects [B, 1, T], got {tuple(x.shape)}")
        Tout = (T + 2 * (self.taps // 2) - self.taps - 1) // self.N + 1
        y = torch.empty((B, self.N, Tout), dtype=torch.float32, device=x.device)
        st = torch.cuda.current_stream(x.device).cuda_stream
        _lib.check(_lib.load().ttsb200_pqmf_analysis(x.data_ptr(), self.H.data_ptr(), B, self.N, self.taps, T,
                                                     y.data_ptr(), st))
        return y

    def synthesis(self, x):
        if self._needs_autograd(x):
            x = F.conv_transpose1d(x, self.updown_filter * self.N, stride=self.N)
            return F.conv1d(x, self.G, padding=self.taps // 2)
        self._require_cuda(x, "synthesis")
        x = x.detach().float().contiguous()
        B, N, Ts = x.shape
        if N != self.N:
            raise RuntimeError(f"PQMF.synthesis expects [B, {self.N}, T], got {tuple(x.shape)}")
        Tout = N * Ts + 2 * (self.taps // 2) - self.taps
        y = torch.empty((B, 1, Tout), dtype=torch.float32, device=x.device)
        st = torch.cuda.current_stream(x.device).cuda_stream
        _lib.check(_lib.load().ttsb200_pqmf_synthesis(x.data_ptr(), self.G.data_ptr(), B, N, self.taps, Ts,
                                                      y.data_ptr(), st))
        return y
