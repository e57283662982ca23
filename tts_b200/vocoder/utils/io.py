"""Checkpoint loading for the vocoder generators — mirrors TTS/vocoder/utils/io.py:9-18 (``load_checkpoint``).
Writing checkpoints is the trainer's job and out of scope (the format is consumed, not rebuilt).

A reference checkpoint is a ``torch.save``d dict whose ``'model'`` entry is the generator's state dict, in the
weight-norm form the trainer writes (``layers.N.weight_g`` / ``weight_v``; ``pqmf_layer.H/G/updown_filter``
buffers for the multi-band model).  ``load_checkpoint`` loads it strictly into a tts_b200 generator — the key
set is the reference's — and leaves folding to ``remove_weight_norm()`` (TTS/server/synthesizer.py:84-85), or to
the plan builder, which folds on the fly when the norm is still attached.

Checkpoints pickled when the toolkit was called ``mozilla_voice_tts`` refer to modules of that name; the
reference retries with an unpickler that renames them (TTS/utils/io.py:6-9).  Same behaviour here, with the
name mapped onto this package when the TTS package itself is not importable.
"""
import pickle

import torch


class _RenamingUnpickler(pickle.Unpickler):
    """find_class with the toolkit's old package name mapped to an importable one."""

    def find_class(self, module, name):
        module = module.replace("mozilla_voice_tts", "TTS")
        try:
            return super().find_class(module, name)
        except ModuleNotFoundError:
            if module.startswith("TTS.vocoder."):
                return super().find_class("tts_b200." + module[len("TTS."):], name)
            raise


class _RenamingPickle:
    """The ``pickle_module`` interface torch.load expects, with the renaming unpickler plugged in."""
    __name__ = "pickle"
    Unpickler = _RenamingUnpickler
    load = staticmethod(lambda f, **kw: _RenamingUnpickler(f, **kw).load())
    loads = staticmethod(pickle.loads)
    dump = staticmethod(pickle.dump)
    dumps = staticmethod(pickle.dumps)
    Pickler = pickle.Pickler
    HIGHEST_PROTOCOL = pickle.HIGHEST_PROTOCOL


def load_checkpoint(model, checkpoint_path, use_cuda=False, allow_pickle=True):
    """(model, state) with ``state['model']`` loaded strictly; same signature as the reference's.

    A vocoder checkpoint is a plain dict of tensors and scalars, so the safe loader (``weights_only=True``) is
    tried first.  Trainer checkpoints that pickled optimizer / scheduler objects, or that were written under the
    toolkit's old package name, need the unrestricted unpickler the reference always uses
    (TTS/vocoder/utils/io.py:10-15) — which executes code from the file; pass ``allow_pickle=False`` to refuse
    that for files of unknown origin."""
    cpu = torch.device("cpu")
    try:
        state = torch.load(checkpoint_path, map_location=cpu, weights_only=True)
    except Exception:
        if not allow_pickle:
            raise
        try:
            state = torch.load(checkpoint_path, map_location=cpu, weights_only=False)
        except ModuleNotFoundError:
            state = torch.load(checkpoint_path, map_location=cpu, pickle_module=_RenamingPickle, weights_only=False)
    model.load_state_dict(state["model"])
    if use_cuda:
        model.cuda()
    return model, state
