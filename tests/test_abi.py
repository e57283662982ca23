"""The C-ABI library loads on a CPU-only box and exports every symbol include/tts_b200.h declares.
No compute calls here (no GPU)."""
import ctypes
import os
import re

import pytest

from tts_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "tts_b200.h")).read()
    return sorted(set(re.findall(r"TTSB200_API[^;(]*?\b(ttsb200_\w+)\s*\(", text)))


def test_header_and_binding_agree():
    syms = header_symbols()
    assert len(syms) >= 18
    assert sorted(_lib.SYMBOLS) == syms


def test_library_exports_every_symbol():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in header_symbols():
        assert hasattr(lib, name), name
    assert _lib.load().ttsb200_abi_version() == _lib.ABI_VERSION


def test_struct_layout_matches_header(tmp_path):
    """Compile the header as plain C with gcc and compare sizeof/offsetof with the ctypes mirror."""
    import subprocess
    src = tmp_path / "sz.c"
    src.write_text(
        '#include <stdio.h>\n#include <stddef.h>\n#include "tts_b200.h"\n'
        'int main(void){printf("%zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(ttsb200_generator_config),'
        'sizeof(ttsb200_forward_args), sizeof(ttsb200_launch_record), offsetof(ttsb200_forward_args, out),'
        'offsetof(ttsb200_forward_args, tap_layers), offsetof(ttsb200_forward_args, profile),'
        'offsetof(ttsb200_forward_args, out_dtype), offsetof(ttsb200_forward_args, out_batch_stride));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                           str(src), "-o", str(exe)])
    got = [int(v) for v in subprocess.check_output([str(exe)]).split()]
    F = _lib.ForwardArgs
    assert got == [ctypes.sizeof(_lib.GeneratorConfig), ctypes.sizeof(F), ctypes.sizeof(_lib.LaunchRecord),
                   F.out.offset, F.tap_layers.offset, F.profile.offset, F.out_dtype.offset, F.out_batch_stride.offset]


def test_argument_validation_without_gpu():
    lib = _lib.load()
    cfg = _lib.GeneratorConfig()
    assert lib.ttsb200_expected_tensors(None) < 0
    assert "NULL" in _lib.last_error()
    cfg.n_stages, cfg.num_res_blocks, cfg.pqmf_bands = 4, 3, 4
    assert lib.ttsb200_expected_tensors(ctypes.byref(cfg)) == 2 * (2 + 4 * 10) + 1
    h = ctypes.c_void_p()
    assert lib.ttsb200_plan_create(ctypes.byref(cfg), None, 0, 0, ctypes.byref(h)) < 0
    with pytest.raises(RuntimeError):
        _lib.check(lib.ttsb200_generator_forward(None, None))
    assert lib.ttsb200_output_length(None, 1, 0, 0) < 0
    assert lib.ttsb200_generator_forward_static(None, None) < 0


def test_sass_is_sm100a_only():
    """The shipped binary carries sm_100a code and nothing else (no multi-arch fallback)."""
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_pqmf_factorisation_matches_oracle():
    """Host-side factorisation behind the generator's fast synthesis kernel (pqmf.cu): on the reference's bank
    (pqmf.py:19-32) the 8 modulated signals + 16-tap FIR reproduce PQMF.synthesis; a bank without the cosine
    structure is refused (the direct 63-tap kernel then runs)."""
    import numpy as np
    import torch
    from oracle.melgan_oracle import pqmf_filters, pqmf_synthesis

    lib = _lib.load()
    _, G, U = pqmf_filters()
    g = np.ascontiguousarray(G[0].numpy(), dtype=np.float32)
    tbl = np.zeros(96, np.float32)
    assert lib.ttsb200_pqmf_factor_4x63(g.ctypes.data, tbl.ctypes.data) == 1
    Ts = 96
    x = torch.randn(2, 4, Ts, generator=torch.Generator().manual_seed(3))
    ref = pqmf_synthesis(x, G, U)[:, 0].numpy()
    xn = x.numpy().astype(np.float64)
    C, W = tbl[:32].reshape(8, 4).astype(np.float64), tbl[32:].reshape(8, 8).astype(np.float64)
    v = np.zeros((2, 8, Ts + 16))
    v[:, :, 8:8 + Ts] = np.einsum("pk,bkm->bpm", C, xn)          # v[p][m] = sum_k C[k][p] x[k][m], window index m + 8
    y = np.zeros((2, 4 * Ts))
    for p in range(8):
        r, base = (7 - p, 2) if p >= 4 else (3 - p, 1)            # output phase fed by signal p, first window index
        for e in range(8):
            y[:, r::4] += W[p, e] * v[:, p, base + 2 * e: base + 2 * e + Ts]
    assert np.abs(y - ref).max() <= 5e-6 * np.abs(ref).max()
    # a bank that is not cosine-modulated
    rnd = np.random.default_rng(0).standard_normal((4, 63)).astype(np.float32)
    assert lib.ttsb200_pqmf_factor_4x63(rnd.ctypes.data, tbl.ctypes.data) == 0
    assert lib.ttsb200_pqmf_factor_4x63(None, tbl.ctypes.data) < 0


def test_shipped_library_is_sm100a_tcgen05_code():
    """The library that travels to the GPU box holds sm_100a code whose hot kernels carry tcgen05 MMAs (UTCHMMA), TMEM
    loads / stores (LDTM / STTM) and bulk copies (UBLKCP) — a build that lost the tensor-core path would still load and
    export every symbol.  Static check with cuobjdump (part of the CUDA toolkit; skipped where it is absent)."""
    import shutil
    import subprocess
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    elf = subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in elf, elf
    sass = subprocess.run(["cuobjdump", "-sass", _lib.LIB_PATH], capture_output=True, text=True).stdout
    per_kernel, cur = {}, None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = per_kernel.setdefault(m.group(1), set())
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m and cur is not None:
            cur.add(m.group(1))
    stack = [ops for name, ops in per_kernel.items() if "tc_stack_kernel" in name]
    convp = [ops for name, ops in per_kernel.items() if "tc_convp_kernel" in name]
    assert len(stack) >= 10 and convp
    for ops in stack + convp:
        assert {"UTCHMMA", "UBLKCP"} <= ops, ops
    assert all({"LDTM", "STTM"} <= ops for ops in stack)
    assert not any("HMMA" in ops for ops in per_kernel.values())        # no mma.sync fallback anywhere
