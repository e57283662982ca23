"""tts_b200 — B200-native (sm_100a) MelGAN / MB-MelGAN / FB-MelGAN vocoder inference.

Drop-in for the hot path of wjqkkky/TTS (Mozilla TTS 0.0.5): the classes under
``tts_b200.vocoder`` keep the reference's names, constructor arguments, state-dict keys and
methods (TTS/vocoder/models/*.py, TTS/vocoder/layers/{melgan,pqmf}.py) while ``inference`` and
``PQMF.synthesis`` run hand-written CUDA kernels behind the C ABI in include/tts_b200.h.
``tts_b200.dropin.install()`` aliases them into ``TTS.vocoder`` so the reference's factory,
server and notebooks pick them up unchanged.
"""
__version__ = "0.1.0"
