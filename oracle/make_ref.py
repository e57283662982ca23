"""Recipe for oracle/_ref: the UNMODIFIED reference generator modules, staged so they can be imported where
/root/reference does not exist (the GPU box).  Checker / baseline infrastructure only — the product never imports it.

    python oracle/make_ref.py            # run by __graft_entry__.build() when /root/reference is present

The reference's hot path is five pure-Python files over torch / numpy / scipy; "building" it is
staging those files, byte for byte, under oracle/_ref/TTS/vocoder/ with empty package markers.  oracle/_ref/ is
git-ignored (no reference source enters this repository's history) but travels to the GPU box with the snapshot,
where bench.py's reference arm and cpu_baseline time it (`cpu_baseline.kind = "reference"`).  A MANIFEST with the
sha256 of every staged file is written next to them so a run can show what it timed.
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("TTS_REFERENCE_ROOT", "/root/reference")
DST = os.path.join(HERE, "_ref")
FILES = [
    "TTS/vocoder/models/melgan_generator.py",
    "TTS/vocoder/models/multiband_melgan_generator.py",
    "TTS/vocoder/models/fullband_melgan_generator.py",
    "TTS/vocoder/layers/melgan.py",
    "TTS/vocoder/layers/pqmf.py",
]
PACKAGES = ["TTS", "TTS/vocoder", "TTS/vocoder/models", "TTS/vocoder/layers"]


def make(verbose=True):
    if not os.path.isdir(REF_ROOT):
        if verbose:
            print(f"oracle/make_ref: {REF_ROOT} not present; keeping whatever is in {DST}")
        return os.path.isdir(DST)
    manifest = {}
    for pkg in PACKAGES:
        os.makedirs(os.path.join(DST, pkg), exist_ok=True)
        open(os.path.join(DST, pkg, "__init__.py"), "w").close()     # the reference's package markers are empty too
    for rel in FILES:
        src, dst = os.path.join(REF_ROOT, rel), os.path.join(DST, rel)
        shutil.copyfile(src, dst)
        manifest[rel] = hashlib.sha256(open(dst, "rb").read()).hexdigest()
    json.dump({"source": REF_ROOT, "sha256": manifest}, open(os.path.join(DST, "MANIFEST.json"), "w"), indent=1)
    if verbose:
        print(f"oracle/make_ref: staged {len(FILES)} reference files under {DST}")
    return True


def load():
    """Import the staged reference classes; returns {name: class} or None when oracle/_ref is absent.

    The staged tree is put on sys.path only for the duration of the import and only if no `TTS` package is already
    loaded (tts_b200.dropin aliases that name onto the product classes — never time those as "the reference")."""
    if not os.path.isfile(os.path.join(DST, FILES[0])):
        return None
    mod = sys.modules.get("TTS")
    if mod is not None and os.path.abspath(os.path.dirname(getattr(mod, "__file__", "") or "")) != os.path.join(DST, "TTS"):
        return None
    sys.path.insert(0, DST)
    try:
        from TTS.vocoder.layers.pqmf import PQMF
        from TTS.vocoder.models.fullband_melgan_generator import FullbandMelganGenerator
        from TTS.vocoder.models.melgan_generator import MelganGenerator
        from TTS.vocoder.models.multiband_melgan_generator import MultibandMelganGenerator
    finally:
        sys.path.remove(DST)
    return {"MelganGenerator": MelganGenerator, "MultibandMelganGenerator": MultibandMelganGenerator,
            "FullbandMelganGenerator": FullbandMelganGenerator, "PQMF": PQMF}


if __name__ == "__main__":
    sys.exit(0 if make() else 1)
