"""Pin the oracle to the reference: every fixture under tests/golden/ was produced by the
unmodified reference (oracle/gen_golden.py).  CPU only."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import numpy_oracle
from oracle.melgan_oracle import (GeneratorSpec, fold_weight_norm, generator_inference, pqmf_analysis,
                                  pqmf_filters, pqmf_synthesis, random_state_dict, receptive_field_frames)

TOL = 2e-6  # fp32 ATen vs fp32 ATen with possibly different thread count / fp64 numpy


def spec_from_kw(kw, cls):
    kw = dict(kw)
    if "upsample_factors" in kw:
        kw["upsample_factors"] = tuple(kw["upsample_factors"])
    return getattr(GeneratorSpec, cls)(**kw)


def weights(g, prefix):
    return {k[len(prefix):]: torch.from_numpy(g[k]) for k in g.files if k.startswith(prefix)}


def test_tiny_melgan_all_layers(golden):
    g = golden("tiny_melgan")
    spec = spec_from_kw(json.loads(str(g["kw"])), "melgan")
    sd = weights(g, "w:")
    mel = torch.from_numpy(g["mel"])
    taps = {}
    y = generator_inference(spec, sd, mel, 2, taps=taps)
    assert np.abs(y.numpy() - g["out"]).max() < TOL
    for i, v in taps.items():
        assert np.abs(v.numpy() - g[f"layer:{i}"]).max() < TOL, i
    # ATen-free restatement, float64
    taps64 = {}
    y64 = numpy_oracle.generator_inference(spec, sd, mel, 2, taps=taps64)
    assert np.abs(y64 - g["out"]).max() < TOL
    for i, v in taps64.items():
        assert np.abs(v - g[f"layer:{i}"]).max() < TOL, i


def test_tiny_mb_weight_norm_fold(golden):
    g = golden("tiny_mb_wn")
    spec = spec_from_kw(json.loads(str(g["kw"])), "multiband")
    sd_wn, sd = weights(g, "wn:"), weights(g, "w:")
    folded = fold_weight_norm(sd_wn)
    assert set(folded) == set(sd)
    for k in sd:
        assert torch.equal(folded[k], sd[k]), k  # fold is bit-identical to remove_weight_norm()
    mel = torch.from_numpy(g["mel"])
    assert np.abs(generator_inference(spec, sd_wn, mel, 2).numpy() - g["out"]).max() < TOL
    assert np.abs(numpy_oracle.generator_inference(spec, sd, mel, 2) - g["out"]).max() < TOL


SEEDED = [("melgan_cfg1", "melgan"), ("mb_small", "multiband"), ("mb_cfg2_item", "multiband"),
          ("fb_odd", "fullband"), ("mb_shipped", "multiband"), ("mb_p0_rand", "multiband")]


def seeded_model(name, cls, g):
    """Rebuild the reference's seeded default-init weights through the boundary classes."""
    from oracle.melgan_oracle import state_dict_digest
    from tts_b200.vocoder.models.fullband_melgan_generator import FullbandMelganGenerator
    from tts_b200.vocoder.models.melgan_generator import MelganGenerator
    from tts_b200.vocoder.models.multiband_melgan_generator import MultibandMelganGenerator
    C = {"melgan": MelganGenerator, "multiband": MultibandMelganGenerator, "fullband": FullbandMelganGenerator}[cls]
    kw = json.loads(str(g["kw"]))
    torch.manual_seed(54321)
    m = C(**kw)
    m.remove_weight_norm()
    assert state_dict_digest(m.state_dict()) == str(g["digest"]), "seeded init differs from the reference's"
    return m, spec_from_kw(kw, cls)


@pytest.mark.parametrize("name,cls", SEEDED)
def test_seeded_cases(golden, name, cls):
    g = golden(name)
    m, spec = seeded_model(name, cls, g)
    B, T, p, sub = int(g["B"]), int(g["T"]), int(g["p"]), int(g["sub"])
    torch.manual_seed(1234)
    mel = (torch.rand if name == "mb_p0_rand" else torch.randn)(B, 80, T)
    taps = {}
    y = generator_inference(spec, m.state_dict(), mel, p, taps=taps)
    assert tuple(y.shape) == tuple(g["out_shape"])
    assert np.abs(y.numpy()[..., ::sub] - g["out"]).max() < TOL
    for i, v in taps.items():
        key = f"lval:{i}"
        if key in g.files:
            assert tuple(v.shape) == tuple(g[f"lshape:{i}"])
            assert np.abs(v.numpy().reshape(-1)[g[f"lidx:{i}"]] - g[key]).max() < TOL, i


def test_numpy_oracle_matches_on_mb_small(golden):
    g = golden("mb_small")
    m, spec = seeded_model("mb_small", "multiband", g)
    torch.manual_seed(1234)
    mel = torch.randn(2, 80, 40)
    y = numpy_oracle.generator_inference(spec, m.state_dict(), mel, 2)
    assert np.abs(y - g["out"]).max() < TOL


def test_pqmf_buffers_and_ops(golden):
    g = golden("pqmf")
    H, G, U = pqmf_filters()
    assert np.array_equal(H.numpy(), g["H"]) and np.array_equal(G.numpy(), g["G"])
    assert np.array_equal(U.numpy(), g["updown_filter"])
    xs, xa = torch.from_numpy(g["syn_in"]), torch.from_numpy(g["ana_in"])
    assert np.abs(pqmf_synthesis(xs, G, U).numpy() - g["syn_out"]).max() < TOL
    assert np.abs(pqmf_analysis(xa, H).numpy() - g["ana_out"]).max() < TOL
    # closed forms the kernels implement
    assert np.abs(numpy_oracle.pqmf_synthesis_polyphase(g["syn_in"].astype(np.float64), g["G"][0].astype(np.float64))
                  - g["syn_out"]).max() < TOL
    assert np.abs(numpy_oracle.pqmf_analysis_direct(g["ana_in"].astype(np.float64), g["H"][:, 0].astype(np.float64))
                  - g["ana_out"]).max() < TOL


def test_reference_shape_test(golden):
    """tests/test_vocoder_melgan_generator.py:6-13 of the reference."""
    g = golden("ref_test_shapes")
    assert tuple(g["forward_shape"]) == (4, 1, 64 * 256)
    assert tuple(g["inference_shape"]) == (4, 1, (64 + 4) * 256)
    spec = GeneratorSpec.melgan()
    sd = random_state_dict(spec, 0)
    y = generator_inference(spec, sd, torch.rand(1, 80, 16), 2)
    assert tuple(y.shape) == (1, 1, (16 + 4) * 256)


def test_too_short_input_raises():
    spec = GeneratorSpec.melgan()
    sd = random_state_dict(spec, 0)
    with pytest.raises(RuntimeError):
        generator_inference(spec, sd, torch.rand(1, 80, 3), 0)   # p=0 needs T >= 4
    generator_inference(spec, sd, torch.rand(1, 80, 1), 2)       # p=2, T=1 works


def test_macs_and_receptive_field():
    assert GeneratorSpec.multiband().macs_per_item(866) == 5682464768 or \
        abs(GeneratorSpec.multiband().macs_per_item(866) - 5.6825e9) / 5.6825e9 < 1e-3   # SURVEY §8a totals
    assert abs(GeneratorSpec.melgan().macs_per_item(204) - 9.2148e9) / 9.2148e9 < 1e-3
    assert GeneratorSpec.multiband().hop == 256 and GeneratorSpec.melgan().hop == 256
    assert receptive_field_frames(GeneratorSpec.multiband()) >= 12
