"""GeneratorEngine: owns one ttsb200_plan (device-resident, kernel-ready weights) and launches
generator.inference() through the C ABI.  PyTorch is used only for device memory and streams.
"""
import collections
import ctypes
import threading

import torch

from . import _lib


def folded_tensors(module):
    """Folded fp32 host tensors in the order ttsb200_plan_create() expects (include/tts_b200.h).

    Walks the reference-shaped ``module.layers`` (melgan_generator.py:29-78).  Weight-norm is
    folded with torch's own ``_weight_norm`` so the result is bit-identical to
    ``remove_weight_norm()`` (melgan_generator.py:91-97).
    """
    def wb(conv):
        if hasattr(conv, "weight_g"):
            w = torch._weight_norm(conv.weight_v.detach(), conv.weight_g.detach(), 0)
        else:
            w = conv.weight.detach()
        return [w.float().cpu().contiguous(), conv.bias.detach().float().cpu().contiguous()]

    out = wb(module.layers[1])
    n_stages = len(module.upsample_factors)
    for s in range(n_stages):
        out += wb(module.layers[3 + 3 * s])
        stack = module.layers[4 + 3 * s]
        for blk in stack.blocks:
            out += wb(blk[2])
            out += wb(blk[4])
        for sc in stack.shortcuts:
            out += wb(sc)
    out += wb(module.layers[4 + 3 * n_stages])
    pq = getattr(module, "pqmf_layer", None)
    if pq is not None:
        out.append(pq.G.detach().float().cpu().contiguous())
    return out


class Profile:
    """Per-launch CUDA-event timings (ttsb200_profile)."""

    def __init__(self):
        self._h = ctypes.c_void_p()
        _lib.check(_lib.load().ttsb200_profile_create(ctypes.byref(self._h)))

    def reset(self):
        _lib.load().ttsb200_profile_reset(self._h)

    def read(self):
        lib = _lib.load()
        n = _lib.check(lib.ttsb200_profile_read(self._h, None, 0))
        recs = (_lib.LaunchRecord * max(n, 1))()
        n = _lib.check(lib.ttsb200_profile_read(self._h, recs, n))
        return [dict(name=recs[i].name.decode(), layer=recs[i].layer, ms=recs[i].ms, flops=recs[i].flops,
                     bytes=recs[i].bytes) for i in range(n)]

    def __del__(self):
        try:
            if self._h:
                _lib.load().ttsb200_profile_destroy(self._h)
                self._h = ctypes.c_void_p()
        except Exception:
            pass


class GeneratorEngine:
    def __init__(self, module, device):
        lib = _lib.load()
        device = torch.device(device)
        if device.type != "cuda":
            raise RuntimeError("tts_b200 runs on CUDA (sm_100a) devices only; there is no CPU fallback")
        self.device = device
        index = device.index if device.index is not None else torch.cuda.current_device()
        cfg = _lib.GeneratorConfig()
        cfg.in_channels = module.in_channels
        cfg.out_channels = module.out_channels
        cfg.proj_kernel = module.proj_kernel
        cfg.base_channels = module.base_channels
        cfg.n_stages = len(module.upsample_factors)
        if cfg.n_stages > _lib.MAX_STAGES:
            raise RuntimeError(f"at most {_lib.MAX_STAGES} upsample stages are supported")
        for i, f in enumerate(module.upsample_factors):
            cfg.upsample_factors[i] = int(f)
        cfg.res_kernel = module.res_kernel
        cfg.num_res_blocks = module.num_res_blocks
        pq = getattr(module, "pqmf_layer", None)
        cfg.pqmf_bands = pq.N if pq is not None else 0
        cfg.pqmf_taps = pq.taps if pq is not None else 0
        cfg.act_slope = 0.2
        tensors = folded_tensors(module)
        ptrs = (ctypes.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])
        self._h = ctypes.c_void_p()
        _lib.check(lib.ttsb200_plan_create(ctypes.byref(cfg), ptrs, len(tensors), index, ctypes.byref(self._h)))
        self.cfg = cfg
        self.index = index
        self._ws = collections.OrderedDict()
        self._lock = threading.Lock()

    # ---- shape helpers
    def output_length(self, frames, pad, skip_pqmf=False):
        return _lib.check(_lib.load().ttsb200_output_length(self._h, frames, pad, int(skip_pqmf)))

    def workspace_bytes(self, batch, frames, pad, precision):
        return _lib.check(_lib.load().ttsb200_workspace_bytes(self._h, batch, frames, pad, precision))

    def layer_shape(self, layer, frames, pad):
        c, l = ctypes.c_int32(), ctypes.c_int64()
        _lib.check(_lib.load().ttsb200_layer_shape(self._h, layer, frames, pad, ctypes.byref(c), ctypes.byref(l)))
        return c.value, l.value

    def launch_count(self, precision=_lib.F32, skip_pqmf=False):
        return _lib.check(_lib.load().ttsb200_launch_count(self._h, precision, int(skip_pqmf)))

    MAX_WORKSPACES = 8

    def _workspace(self, nbytes, stream_key):
        """(scratch tensor, launch lock) for one forward, keyed by the launch stream and nothing else.

        The Flask request threads that share a module (TTS/server/server.py:64,82) all launch on the default stream;
        they share ONE workspace instead of leaking one per thread.  That is safe only if the ~10 launches of one
        forward are not interleaved with another thread's on the same stream (the stream orders kernels, not launch
        sequences), so the caller holds the entry's lock while it enqueues; the enqueue takes tens of microseconds.
        At most MAX_WORKSPACES entries are kept (least recently used first out); an evicted tensor goes back to
        torch's caching allocator, which reuses the block only in stream order of the stream it was allocated on —
        the stream that may still be running kernels on it — so eviction is safe while work is in flight."""
        with self._lock:
            ent = self._ws.get(stream_key)
            if ent is None:
                ent = self._ws[stream_key] = [None, threading.Lock()]
            self._ws.move_to_end(stream_key)
            while len(self._ws) > self.MAX_WORKSPACES:
                self._ws.popitem(last=False)
        return ent

    @staticmethod
    def _grow(ent, nbytes, device):
        if ent[0] is None or ent[0].numel() < nbytes:     # under ent's lock
            ent[0] = torch.empty(nbytes, dtype=torch.uint8, device=device)
        return ent[0]

    # ---- the hot call
    def forward(self, mel, pad, precision=_lib.F32, skip_pqmf=False, tap_layers=None, profile=None, out=None,
                out_dtype=None, keep=None):
        """mel: CUDA fp32 [B, in_channels, T] on this engine's device. Returns [B, C_out, samples] (fp32, or fp16
        with ``out_dtype=torch.float16`` / an fp16 ``out``).

        ``keep=(lo, n)`` stores only samples [lo, lo + n) of every row, into ``out`` [B, C_out, n] whose batch stride
        may exceed n (a view into a larger buffer, possibly peer memory): the kept range of a long-form or
        streaming window lands in place, nothing is sliced or concatenated afterwards."""
        if mel.dim() != 3 or mel.shape[1] != self.cfg.in_channels:
            raise RuntimeError(f"expected mel of shape [B, {self.cfg.in_channels}, T], got {tuple(mel.shape)}")
        if mel.device != self.device:
            raise RuntimeError(f"mel is on {mel.device}, the plan on {self.device}")
        mel = mel.contiguous()
        B, _, T = mel.shape
        if B == 0 or T == 0:
            raise RuntimeError("empty batch or zero-length mel")
        lib = _lib.load()
        n_out = self.output_length(T, pad, skip_pqmf)
        ch = 1 if (self.cfg.pqmf_bands and not skip_pqmf) else self.cfg.out_channels
        if keep is not None:
            lo, n_keep = int(keep[0]), int(keep[1])
            if lo < 0 or n_keep <= 0 or lo + n_keep > n_out:
                raise RuntimeError(f"keep=({lo}, {n_keep}) is outside the {n_out}-sample result")
            if out is None:
                out = torch.empty((B, ch, n_keep), dtype=out_dtype or torch.float32, device=self.device)
            if tuple(out.shape) != (B, ch, n_keep) or out.stride(-1) != 1 or (ch > 1 and not out.is_contiguous()):
                raise RuntimeError(f"windowed out must be [{B}, {ch}, {n_keep}] with unit sample stride")
            out_dtype = out.dtype
            keep = (lo, n_keep, out.stride(0) if B > 1 else max(n_keep, out.stride(0)))
        elif out is not None:
            out_dtype = out.dtype
            if tuple(out.shape) != (B, ch, n_out) or not out.is_contiguous() or out.device != self.device:
                raise RuntimeError(f"out must be a contiguous {(B, ch, n_out)} tensor on {self.device}")
        out_dtype = out_dtype or torch.float32
        if out_dtype not in (torch.float32, torch.float16):
            raise RuntimeError("out dtype must be torch.float32 or torch.float16")
        # The module may live on a device that is not the caller's current one (the reference supports that):
        # allocate and look the stream up under the plan's device; the C entry point switches devices itself.
        with torch.cuda.device(self.device):
            return self._forward(lib, mel, B, T, pad, precision, skip_pqmf, tap_layers, profile, out, out_dtype, ch,
                                 n_out, keep)

    def _forward(self, lib, mel, B, T, pad, precision, skip_pqmf, tap_layers, profile, out, out_dtype, ch, n_out, keep):
        if out is None:
            out = torch.empty((B, ch, n_out), dtype=out_dtype, device=self.device)
        stream = torch.cuda.current_stream(self.device)
        ws_bytes = self.workspace_bytes(B, T, pad, precision)
        ent = self._workspace(ws_bytes, stream.cuda_stream)
        a = _lib.ForwardArgs()
        a.mel = mel.data_ptr()
        a.batch, a.frames, a.pad, a.precision = B, T, pad, precision
        a.out = out.data_ptr()
        a.stream = stream.cuda_stream
        a.skip_pqmf = int(skip_pqmf)
        a.out_dtype = _lib.OUT_F16 if out_dtype == torch.float16 else _lib.OUT_F32
        if keep is not None:
            a.out_keep_lo, a.out_keep_len, a.out_batch_stride = keep
        taps = {}
        if tap_layers:
            ids = (ctypes.c_int32 * len(tap_layers))(*tap_layers)
            bufs = []
            for layer in tap_layers:
                c, l = self.layer_shape(layer, T, pad)
                bufs.append(torch.empty((B, c, l), dtype=torch.float32, device=self.device))
                taps[layer] = bufs[-1]
            ptrs = (ctypes.c_void_p * len(bufs))(*[b.data_ptr() for b in bufs])
            a.n_taps = len(tap_layers)
            a.tap_layers = ids
            a.tap_out = ptrs
        if profile is not None:
            a.profile = profile._h
        with ent[1]:          # one launch sequence at a time per stream (see _workspace)
            ws = self._grow(ent, ws_bytes, self.device)
            a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
            _lib.check(lib.ttsb200_generator_forward(self._h, ctypes.byref(a)))
        return (out, taps) if tap_layers else out

    def forward_host(self, mel_host, pad, precision=_lib.F32):
        """Host-buffer entry (ttsb200_generator_forward_host): CPU fp32 tensor in, CPU tensor out."""
        mel_host = mel_host.detach().float().contiguous()
        B, _, T = mel_host.shape
        out = torch.empty((B, 1 if self.cfg.pqmf_bands else self.cfg.out_channels, self.output_length(T, pad)),
                          dtype=torch.float32)
        _lib.check(_lib.load().ttsb200_generator_forward_host(self._h, mel_host.data_ptr(), B, T, pad, precision,
                                                              out.data_ptr()))
        return out

    def close(self):
        if getattr(self, "_h", None):
            _lib.load().ttsb200_plan_destroy(self._h)
            self._h = ctypes.c_void_p()
            self._ws = collections.OrderedDict()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
