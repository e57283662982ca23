"""Generate tests/golden/*.npz from the UNMODIFIED reference (build container only).

Run:  PYTHONDONTWRITEBYTECODE=1 python oracle/gen_golden.py
Needs /root/reference (read-only mount); it is put on sys.path, nothing is written there.
The GPU box has no /root/reference: tests only read the committed .npz files.

Cases (see tests/golden/MANIFEST.json after a run):
  tiny_melgan      all weights + every parametrised layer output stored (seed-independent pin)
  tiny_mb_wn       tiny multi-band model in weight_g/weight_v form (pins the weight-norm fold)
  melgan_cfg1      MelganGenerator() defaults, seed 54321, mel randn(1,80,200) seed 1234, p=2  (BASELINE config 1)
  mb_small         MultibandMelganGenerator() defaults, mel randn(2,80,40), p=2
  mb_cfg2_item     one 862-frame item of BASELINE config 2, waveform subsampled 1:16
  fb_odd           FullbandMelganGenerator(upsample_factors=(5,5,4,3)) — odd factors (BASELINE config 5 model)
  mb_shipped       MB with the shipped config [8,4,2] x 4 blocks (multiband_melgan_config.json:83-87)
  ref_test_shapes  tests/test_vocoder_melgan_generator.py:6-13 shape assertions on torch.rand(4,80,64)
  pqmf             H/G/updown_filter buffers, synthesis and analysis on seeded noise
Seeded-default-init cases store a sha256 of the folded weights instead of the weights; the
product's boundary classes construct the same torch modules in the same order, so the same
seed reproduces them (checked by digest in the tests).
"""
import json
import os
import sys
import warnings

import numpy as np
import torch

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden")
sys.dont_write_bytecode = True
sys.path.insert(0, REF)
sys.path.insert(0, ROOT)
warnings.filterwarnings("ignore")

from TTS.vocoder.models.melgan_generator import MelganGenerator  # noqa: E402
from TTS.vocoder.models.multiband_melgan_generator import MultibandMelganGenerator  # noqa: E402
from TTS.vocoder.models.fullband_melgan_generator import FullbandMelganGenerator  # noqa: E402
from TTS.vocoder.layers.pqmf import PQMF  # noqa: E402
from oracle.melgan_oracle import state_dict_digest  # noqa: E402

WEIGHT_SEED = 54321  # TTS/utils/training.py:8
MEL_SEED = 1234


def np_sd(model):
    return {k: v.detach().numpy().copy() for k, v in model.state_dict().items()}


def layer_outputs(model, c):
    """Outputs of model.layers[i] for every i (convert_melgan_torch_to_tf.py:89-101 indexing)."""
    outs = {}
    x = c
    for i, layer in enumerate(model.layers):
        x = layer(x)
        outs[i] = x.detach().numpy().copy()
    return outs


def build(cls, seed, **kw):
    torch.manual_seed(seed)
    m = cls(**kw)
    m.eval()
    return m


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(1)  # summation order is thread-count dependent (SURVEY §8b Determinism)
    manifest = {"torch": torch.__version__, "numpy": np.__version__, "threads": 1, "cases": {}}

    def save(name, **arrs):
        path = os.path.join(OUT, name + ".npz")
        np.savez_compressed(path, **arrs)
        manifest["cases"][name] = {"bytes": os.path.getsize(path), "keys": sorted(arrs)}

    with torch.no_grad():
        # ---- tiny_melgan: everything stored
        kw = dict(in_channels=8, out_channels=1, proj_kernel=7, base_channels=32,
                  upsample_factors=(2, 3), res_kernel=3, num_res_blocks=2)
        m = build(MelganGenerator, 7, **kw)
        m.remove_weight_norm()
        torch.manual_seed(8)
        mel = torch.randn(2, 8, 12)
        y = m.inference(mel)
        padded = torch.nn.functional.pad(mel, (2, 2), "replicate")
        lo = layer_outputs(m, padded)
        arrs = {"mel": mel.numpy(), "out": y.numpy(), "kw": np.array(json.dumps(kw))}
        arrs.update({"w:" + k: v for k, v in np_sd(m).items()})
        arrs.update({f"layer:{i}": v for i, v in lo.items()})
        save("tiny_melgan", **arrs)

        # ---- tiny_mb_wn: weight-norm form + fold
        kw = dict(in_channels=8, out_channels=4, proj_kernel=7, base_channels=32,
                  upsample_factors=(2, 2), res_kernel=3, num_res_blocks=2)
        m = build(MultibandMelganGenerator, 9, **kw)
        # perturb g so that g != ||v|| (default init makes the fold an identity)
        torch.manual_seed(10)
        for n, p in m.named_parameters():
            if n.endswith("weight_g"):
                p.mul_(0.5 + torch.rand_like(p))
        sd_wn = np_sd(m)
        mel = torch.randn(1, 8, 9)
        y_wn = m.inference(mel)
        m.remove_weight_norm()
        y = m.inference(mel)
        assert float((y - y_wn).abs().max()) == 0.0  # bit-identical before/after
        arrs = {"mel": mel.numpy(), "out": y.numpy(), "kw": np.array(json.dumps(kw))}
        arrs.update({"wn:" + k: v for k, v in sd_wn.items()})
        arrs.update({"w:" + k: v for k, v in np_sd(m).items()})
        save("tiny_mb_wn", **arrs)

        # ---- seeded default-init cases
        def seeded(name, cls, kw, B, T, p, sub=1, layer_stats=True, mel_fn=torch.randn):
            m = build(cls, WEIGHT_SEED, **kw)
            m.remove_weight_norm()
            m.inference_padding = p
            torch.manual_seed(MEL_SEED)
            mel = mel_fn(B, 80, T)
            y = m.inference(mel)
            arrs = {"out": y.numpy()[..., ::sub].copy(), "sub": np.array(sub), "B": np.array(B), "T": np.array(T),
                    "p": np.array(p), "kw": np.array(json.dumps(kw)),
                    "digest": np.array(state_dict_digest(m.state_dict())),
                    "out_shape": np.array(y.shape), "out_absmax": np.array(float(y.abs().max()))}
            if layer_stats:
                padded = torch.nn.functional.pad(mel, (p, p), "replicate") if p else mel
                for i, v in layer_outputs(m, padded).items():
                    flat = v.reshape(-1)
                    idx = np.linspace(0, flat.size - 1, 257).astype(np.int64)
                    arrs[f"lidx:{i}"] = idx
                    arrs[f"lval:{i}"] = flat[idx]
                    arrs[f"lshape:{i}"] = np.array(v.shape)
            save(name, **arrs)

        seeded("melgan_cfg1", MelganGenerator, {}, 1, 200, 2)
        seeded("mb_small", MultibandMelganGenerator, {}, 2, 40, 2)
        seeded("mb_cfg2_item", MultibandMelganGenerator, {}, 1, 862, 2, sub=16, layer_stats=False)
        seeded("fb_odd", FullbandMelganGenerator, {"upsample_factors": (5, 5, 4, 3)}, 1, 24, 2)
        seeded("mb_shipped", MultibandMelganGenerator, {"upsample_factors": (8, 4, 2), "num_res_blocks": 4}, 1, 32, 0)
        seeded("mb_p0_rand", MultibandMelganGenerator, {}, 1, 100, 0, layer_stats=False, mel_fn=torch.rand)

        # ---- reference's own shape test  (tests/test_vocoder_melgan_generator.py:6-13)
        m = build(MelganGenerator, WEIGHT_SEED)
        x = torch.rand((4, 80, 64))
        save("ref_test_shapes", forward_shape=np.array(m(x).shape), inference_shape=np.array(m.inference(x).shape))

        # ---- PQMF
        pq = PQMF(N=4, taps=62, cutoff=0.15, beta=9.0)
        torch.manual_seed(MEL_SEED)
        xs = torch.randn(2, 4, 100)
        xa = torch.randn(2, 1, 403)
        save("pqmf", H=pq.H.numpy(), G=pq.G.numpy(), updown_filter=pq.updown_filter.numpy(),
             syn_in=xs.numpy(), syn_out=pq.synthesis(xs).numpy(),
             ana_in=xa.numpy(), ana_out=pq.analysis(xa).numpy())

    with open(os.path.join(OUT, "MANIFEST.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)
    print(json.dumps(manifest["cases"], indent=1))


if __name__ == "__main__":
    main()
