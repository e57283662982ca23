"""Checkpoint I/O for the vocoder generators — mirrors TTS/vocoder/utils/io.py:9-18 (load) and :21-46 (save).

A reference checkpoint is a ``torch.save``d dict whose ``'model'`` entry is the generator's state dict, in the
weight-norm form the trainer writes (``layers.N.weight_g`` / ``weight_v``; ``pqmf_layer.H/G/updown_filter``
buffers for the multi-band model).  ``load_checkpoint`` loads it strictly into a tts_b200 generator — the key
set is the reference's — and leaves folding to ``remove_weight_norm()`` (TTS/server/synthesizer.py:84-85), or to
the plan builder, which folds on the fly when the norm is still attached.

Checkpoints pickled when the toolkit was called ``mozilla_voice_tts`` refer to modules of that name; the
reference retries with an unpickler that renames them (TTS/utils/io.py:6-9).  Same behaviour here, with the
name mapped onto this package when the TTS package itself is not importable.
"""
import datetime
import os
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


def load_checkpoint(model, checkpoint_path, use_cuda=False):
    """(model, state) with ``state['model']`` loaded strictly; same signature as the reference's."""
    try:
        state = torch.load(checkpoint_path, map_location=torch.device("cpu"), weights_only=False)
    except ModuleNotFoundError:
        state = torch.load(checkpoint_path, map_location=torch.device("cpu"), pickle_module=_RenamingPickle,
                           weights_only=False)
    model.load_state_dict(state["model"])
    if use_cuda:
        model.cuda()
    return model, state


def save_model(model, optimizer, scheduler, model_disc, optimizer_disc, scheduler_disc, current_step, epoch,
               output_path, **kwargs):
    def sd(m):
        return m.state_dict() if m is not None else None
    state = {
        "model": sd(model),
        "optimizer": sd(optimizer),
        "scheduler": sd(scheduler),
        "model_disc": sd(model_disc),
        "optimizer_disc": sd(optimizer_disc),
        "scheduler_disc": sd(scheduler_disc),
        "step": current_step,
        "epoch": epoch,
        "date": datetime.date.today().strftime("%B %d, %Y"),
    }
    state.update(kwargs)
    torch.save(state, output_path)


def save_checkpoint(model, optimizer, scheduler, model_disc, optimizer_disc, scheduler_disc, current_step, epoch,
                    output_folder, **kwargs):
    checkpoint_path = os.path.join(output_folder, "checkpoint_{}.pth.tar".format(current_step))
    print(" > CHECKPOINT : {}".format(checkpoint_path))
    save_model(model, optimizer, scheduler, model_disc, optimizer_disc, scheduler_disc, current_step, epoch,
               checkpoint_path, **kwargs)


def save_best_model(target_loss, best_loss, model, optimizer, scheduler, model_disc, optimizer_disc, scheduler_disc,
                    current_step, epoch, output_folder, **kwargs):
    if target_loss < best_loss:
        checkpoint_path = os.path.join(output_folder, "best_model.pth.tar")
        print(" > BEST MODEL : {}".format(checkpoint_path))
        save_model(model, optimizer, scheduler, model_disc, optimizer_disc, scheduler_disc, current_step, epoch,
                   checkpoint_path, model_loss=target_loss, **kwargs)
        best_loss = target_loss
    return best_loss
