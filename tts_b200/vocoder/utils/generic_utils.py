"""setup_generator — the reference's name-based factory for the generators on the hot path.

Restates TTS/vocoder/utils/generic_utils.py:40-78 (without its matplotlib / visualisation
imports, :4-6): module ``<generator_model>.lower()``, class = CamelCase of the string, constructor
arguments from ``c.audio['num_mels']`` and ``c.generator_model_params``.  The reference dispatches
with substring tests (``c.generator_model in 'melgan_generator'``, :50,61,70); kept.
"""
import importlib

_PACKAGE = "tts_b200.vocoder.models"


def to_camel(text):
    text = text.capitalize()
    out, up = [], False
    for ch in text:
        if ch == "_":
            up = True
        else:
            out.append(ch.upper() if up else ch)
            up = False
    return "".join(out)


def setup_generator(c):
    print(" > Generator Model: {}".format(c.generator_model))
    name = c.generator_model.lower()
    try:
        module = importlib.import_module(f"{_PACKAGE}.{name}")
    except ModuleNotFoundError as e:
        raise NotImplementedError(
            f"tts_b200 implements the MelGAN-family generators only; '{c.generator_model}' is out of scope") from e
    cls = getattr(module, to_camel(c.generator_model))
    params = c.generator_model_params
    if c.generator_model in "melgan_generator":
        return cls(in_channels=c.audio["num_mels"], out_channels=1, proj_kernel=7, base_channels=512,
                   upsample_factors=params["upsample_factors"], res_kernel=3,
                   num_res_blocks=params["num_res_blocks"])
    if c.generator_model in "multiband_melgan_generator":
        return cls(in_channels=c.audio["num_mels"], out_channels=4, proj_kernel=7, base_channels=384,
                   upsample_factors=params["upsample_factors"], res_kernel=3,
                   num_res_blocks=params["num_res_blocks"])
    if c.generator_model in "fullband_melgan_generator":
        return cls(in_channels=c.audio["num_mels"], out_channels=1, proj_kernel=7, base_channels=512,
                   upsample_factors=params["upsample_factors"], res_kernel=3,
                   num_res_blocks=params["num_res_blocks"])
    raise NotImplementedError(f"Generator model {c.generator_model} not implemented!")
