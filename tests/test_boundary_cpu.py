"""Drop-in surface checks that need no GPU."""
import types

import pytest
import torch

from tts_b200.vocoder.layers.pqmf import PQMF
from tts_b200.vocoder.models.fullband_melgan_generator import FullbandMelganGenerator
from tts_b200.vocoder.models.melgan_generator import MelganGenerator
from tts_b200.vocoder.models.multiband_melgan_generator import MultibandMelganGenerator
from tts_b200.vocoder.utils.generic_utils import setup_generator, to_camel
from oracle.melgan_oracle import GeneratorSpec, conv_keys


def ref_keys(spec, weight_norm):
    keys = []
    for prefix, _, _ in conv_keys(spec):
        keys.append(prefix + ".bias")
        keys += [prefix + ".weight_g", prefix + ".weight_v"] if weight_norm else [prefix + ".weight"]
    if spec.pqmf:
        keys += ["pqmf_layer.H", "pqmf_layer.G", "pqmf_layer.updown_filter"]
    return sorted(keys)


def test_state_dict_keys_match_reference():
    m = MultibandMelganGenerator()
    assert len(m.state_dict()) == 129                      # 129 entries, 87 after removal
    assert sorted(m.state_dict()) == ref_keys(GeneratorSpec.multiband(), True)
    m.remove_weight_norm()
    assert len(m.state_dict()) == 87
    assert sorted(m.state_dict()) == ref_keys(GeneratorSpec.multiband(), False)
    assert m.layers[1].weight.device.type == "cpu"         # melgan_generator.py:84 reads this


def test_weight_norm_g_shapes():
    m = MelganGenerator()
    assert tuple(m.layers[1].weight_g.shape) == (512, 1, 1)
    assert tuple(m.layers[3].weight_g.shape) == (512, 1, 1)   # ConvTranspose1d: dim 0 = in-channels
    assert tuple(m.layers[3].weight_v.shape) == (512, 256, 16)


def test_reference_unit_test_training_mode_shapes():
    """tests/test_vocoder_melgan_generator.py:6-13 — forward() in the default (training) mode."""
    model = MelganGenerator()
    out = model(torch.rand((4, 80, 16)))
    assert out.shape == (4, 1, 16 * 256)


def test_inference_on_cpu_module_raises_loudly():
    m = MultibandMelganGenerator().eval()
    with pytest.raises(RuntimeError, match="CUDA"):
        m.inference(torch.rand(1, 80, 10))
    with pytest.raises(RuntimeError, match="CUDA"):
        m(torch.rand(1, 80, 10))
    with pytest.raises(RuntimeError, match="CUDA"):
        PQMF().synthesis(torch.rand(1, 4, 10))


def test_constructor_asserts():
    with pytest.raises(AssertionError):
        MelganGenerator(proj_kernel=6)
    with pytest.raises(AssertionError):
        MelganGenerator(res_kernel=4)


def test_factory():
    assert to_camel("multiband_melgan_generator") == "MultibandMelganGenerator"
    c = types.SimpleNamespace(generator_model="multiband_melgan_generator", audio={"num_mels": 80},
                              generator_model_params={"upsample_factors": [8, 4, 2], "num_res_blocks": 4})
    m = setup_generator(c)   # shipped config: multiband_melgan_config.json:83-87
    assert isinstance(m, MultibandMelganGenerator) and m.upsample_factors == (8, 4, 2) and m.num_res_blocks == 4
    c.generator_model = "fullband_melgan_generator"
    assert isinstance(setup_generator(c), FullbandMelganGenerator)
    c.generator_model = "melgan_generator"
    m = setup_generator(c)
    assert type(m) is MelganGenerator and m.out_channels == 1 and m.base_channels == 512
    c.generator_model = "parallel_wavegan_generator"
    with pytest.raises(NotImplementedError):
        setup_generator(c)


def test_load_state_dict_strict_roundtrip():
    a, b = MultibandMelganGenerator(), MultibandMelganGenerator()
    b.load_state_dict(a.state_dict())          # weight_g / weight_v form, strict
    a.remove_weight_norm(); b.remove_weight_norm()
    for (k, x), (_, y) in zip(a.state_dict().items(), b.state_dict().items()):
        assert torch.equal(x, y), k


def test_dropin_aliases():
    import importlib
    import sys
    from tts_b200 import dropin
    saved = {k: v for k, v in sys.modules.items() if k == "TTS" or k.startswith("TTS.")}
    try:
        dropin.install()
        mod = importlib.import_module("TTS.vocoder.models.multiband_melgan_generator")
        assert mod.MultibandMelganGenerator is MultibandMelganGenerator
        from TTS.vocoder.layers.pqmf import PQMF as P2
        assert P2 is PQMF
    finally:
        for k in [k for k in sys.modules if k == "TTS" or k.startswith("TTS.")]:
            del sys.modules[k]
        sys.modules.update(saved)


REFERENCE_ROOT = "/root/reference"


@pytest.mark.skipif(not __import__("os").path.isdir(REFERENCE_ROOT), reason="build container only: needs /root/reference")
def test_reference_setup_generator_resolves_to_this_package(monkeypatch):
    """The reference's OWN factory (TTS/vocoder/utils/generic_utils.py:40-94), unmodified, under dropin.install():
    its module imports matplotlib and TTS.tts.utils.visual at top level (:4-6), neither importable here, so those two
    names are stubbed in sys.modules; importlib + to_camel + the substring dispatch are the reference's code."""
    import importlib
    import sys
    from tts_b200 import dropin

    saved = {k: v for k, v in sys.modules.items() if k == "TTS" or k.startswith("TTS.") or k.startswith("matplotlib")}
    for k in saved:
        monkeypatch.delitem(sys.modules, k, raising=False)
    monkeypatch.syspath_prepend(REFERENCE_ROOT)
    mpl = types.ModuleType("matplotlib")
    mpl.pyplot = types.ModuleType("matplotlib.pyplot")
    monkeypatch.setitem(sys.modules, "matplotlib", mpl)
    monkeypatch.setitem(sys.modules, "matplotlib.pyplot", mpl.pyplot)
    visual = types.ModuleType("TTS.tts.utils.visual")
    visual.plot_spectrogram = lambda *a, **k: None
    try:
        importlib.import_module("TTS.tts.utils")                  # the real (empty) package
        monkeypatch.setitem(sys.modules, "TTS.tts.utils.visual", visual)
        dropin.install()
        ref = importlib.import_module("TTS.vocoder.utils.generic_utils")
        assert ref.__file__.startswith(REFERENCE_ROOT)            # the reference's file, not ours
        for name, cls, params in (
                ("multiband_melgan_generator", MultibandMelganGenerator, {"upsample_factors": [8, 4, 2], "num_res_blocks": 4}),
                ("melgan_generator", MelganGenerator, {"upsample_factors": [8, 8, 2, 2], "num_res_blocks": 3}),
                ("fullband_melgan_generator", FullbandMelganGenerator, {"upsample_factors": [2, 8, 2, 2], "num_res_blocks": 4})):
            c = types.SimpleNamespace(generator_model=name, generator_model_params=params, audio={"num_mels": 80})
            m = ref.setup_generator(c)
            assert type(m) is cls and m.upsample_factors == tuple(params["upsample_factors"])
            assert m.num_res_blocks == params["num_res_blocks"]
            ours = setup_generator(c)                             # the restated factory agrees key for key
            assert sorted(m.state_dict()) == sorted(ours.state_dict())
        from TTS.vocoder.layers.pqmf import PQMF as RefPathPQMF
        assert RefPathPQMF is PQMF
    finally:
        dropin.uninstall()
        for k in [k for k in sys.modules if k == "TTS" or k.startswith("TTS.")]:
            sys.modules.pop(k, None)
        sys.modules.update(saved)
