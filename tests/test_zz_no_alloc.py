"""ttsb200_generator_forward_static: the forward that allocates nothing (scratch = a static __device__ array inside
the library) gives what ttsb200_generator_forward gives with a caller workspace.  Through the C ABI, -m gpu.

Collected last on purpose: it must pass alone in a fresh process and after the
other tests.  The array is shared by every call on the device and keeps the previous call's contents, so the check
runs twice in a row, the second time on another shape and another input.
"""
import ctypes

import pytest
import torch

from tts_b200 import _lib
from tts_b200.vocoder.models.multiband_melgan_generator import MultibandMelganGenerator

pytestmark = pytest.mark.gpu
F32_TOL = 1e-5      # tests/test_gpu_parity.py's bar; both entry points run the same kernels, so 0 is expected


def _args(eng, mel, pad, prec, out):
    a = _lib.ForwardArgs()
    a.mel = mel.data_ptr()
    a.batch, a.frames, a.pad, a.precision = mel.shape[0], mel.shape[2], pad, prec
    a.out = out.data_ptr()
    a.stream = torch.cuda.current_stream(mel.device).cuda_stream
    return a


def _generator(seed):
    torch.manual_seed(seed)
    m = MultibandMelganGenerator()
    m.remove_weight_norm()
    return m.to("cuda:0").eval()


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_static_workspace_forward_matches_forward(precision):
    lib = _lib.load()
    m = _generator(7)
    m.precision = precision
    eng = m._get_engine()
    prec = _lib.PRECISIONS[precision]
    g = torch.Generator().manual_seed(21)
    for B, T, pad in ((2, 64, 2), (1, 150, 0)):                 # two calls in a row, different shapes
        mel = torch.randn(B, 80, T, generator=g).cuda()
        ref = m.inference_padded(mel, pad)                      # existing entry point, torch-allocated workspace
        need = eng.workspace_bytes(B, T, pad, prec)
        assert 0 < need <= _lib.STATIC_WORKSPACE_BYTES
        out = torch.full_like(ref, float("nan"))
        a = _args(eng, mel, pad, prec, out)                     # workspace / workspace_bytes left NULL / 0
        with torch.cuda.device(0):
            _lib.check(lib.ttsb200_generator_forward_static(eng._h, ctypes.byref(a)))
        torch.cuda.synchronize()
        assert torch.isfinite(out).all()
        err = float((out - ref).abs().max())
        assert err <= F32_TOL, (precision, B, T, err)


def test_static_workspace_limits_are_refused():
    lib = _lib.load()
    m = _generator(8)
    eng = m._get_engine()
    B, T = 16, 862                                              # bench shape: 82 MB of BF16 scratch
    assert eng.workspace_bytes(B, T, 2, _lib.BF16) > _lib.STATIC_WORKSPACE_BYTES
    mel = torch.zeros(B, 80, T, device="cuda:0")
    out = torch.zeros(B, 1, eng.output_length(T, 2), device="cuda:0")
    a = _args(eng, mel, 2, _lib.BF16, out)
    rc = lib.ttsb200_generator_forward_static(eng._h, ctypes.byref(a))
    assert rc == -4                                             # TTSB200_ERR_WORKSPACE
    msg = _lib.last_error()
    assert str(_lib.STATIC_WORKSPACE_BYTES) in msg and "static" in msg
    # a profile would create CUDA events: refused
    from tts_b200.engine import Profile
    prof = Profile()
    small = torch.zeros(1, 80, 32, device="cuda:0")
    out = torch.zeros(1, 1, eng.output_length(32, 0), device="cuda:0")
    a = _args(eng, small, 0, _lib.F32, out)
    a.profile = prof._h
    assert lib.ttsb200_generator_forward_static(eng._h, ctypes.byref(a)) == -1
    assert lib.ttsb200_generator_forward_static(None, None) == -1
