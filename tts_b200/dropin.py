"""Alias the tts_b200 classes into the reference's module paths.

    import tts_b200.dropin; tts_b200.dropin.install()

After this, ``importlib.import_module('TTS.vocoder.models.multiband_melgan_generator')`` — what
the reference's ``setup_generator`` does (TTS/vocoder/utils/generic_utils.py:45-49) — and
``from TTS.vocoder.layers.pqmf import PQMF`` resolve to the B200 implementations, so
TTS/server/synthesizer.py:81-91, TTS/bin/synthesize.py:131-138 and the notebooks run unchanged.
Works whether or not the TTS package is importable (a stub package tree is created if not).
"""
import importlib
import sys
import types

_ALIASES = {
    "TTS.vocoder.models.melgan_generator": "tts_b200.vocoder.models.melgan_generator",
    "TTS.vocoder.models.multiband_melgan_generator": "tts_b200.vocoder.models.multiband_melgan_generator",
    "TTS.vocoder.models.fullband_melgan_generator": "tts_b200.vocoder.models.fullband_melgan_generator",
    "TTS.vocoder.layers.melgan": "tts_b200.vocoder.layers.melgan",
    "TTS.vocoder.layers.pqmf": "tts_b200.vocoder.layers.pqmf",
}


def _ensure_package(name):
    if name in sys.modules:
        return sys.modules[name]
    try:
        return importlib.import_module(name)
    except Exception:
        pkg = types.ModuleType(name)
        pkg.__path__ = []  # mark as package
        sys.modules[name] = pkg
        if "." in name:
            parent, leaf = name.rsplit(".", 1)
            setattr(_ensure_package(parent), leaf, pkg)
        return pkg


def install():
    """Idempotent. Returns the list of aliased module names."""
    for ref_name, ours in _ALIASES.items():
        mod = importlib.import_module(ours)
        parent, leaf = ref_name.rsplit(".", 1)
        pkg = _ensure_package(parent)
        sys.modules[ref_name] = mod
        setattr(pkg, leaf, mod)
    return sorted(_ALIASES)


def uninstall():
    for ref_name in _ALIASES:
        sys.modules.pop(ref_name, None)
