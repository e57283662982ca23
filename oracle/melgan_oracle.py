"""Functional CPU restatement of the reference vocoder generators (oracle; test infrastructure only).

Everything here is a re-derivation, written from the reference's *behaviour*, of

* ``MelganGenerator``            /root/reference/TTS/vocoder/models/melgan_generator.py:8-97
* ``MultibandMelganGenerator``   /root/reference/TTS/vocoder/models/multiband_melgan_generator.py:7-39
* ``FullbandMelganGenerator``    /root/reference/TTS/vocoder/models/fullband_melgan_generator.py:6-30
* ``ResidualStack``              /root/reference/TTS/vocoder/layers/melgan.py:5-45
* ``PQMF``                       /root/reference/TTS/vocoder/layers/pqmf.py:10-56

The reference expresses the generator as an ``nn.Sequential`` of stock torch modules, so its
arithmetic lives in torch's ATen CPU kernels (``requirements.txt:1`` pins only ``torch>=1.5``;
this image has torch 2.11.0).  The oracle therefore works on a plain ``{key: tensor}`` state
dict with the reference's key names and calls ``torch.nn.functional`` directly, in fp32 or
fp64.  ``oracle/numpy_oracle.py`` is a second, ATen-free restatement used on small cases;
``oracle/bf16_emulation.py`` is the same graph with BF16 rounding where the CUDA path rounds; ``oracle/_ref``
(``oracle/make_ref.py``) holds the unmodified reference modules themselves for the timed CPU baseline.

Parity status: pinned by ``tests/golden/*.npz`` (made by ``oracle/gen_golden.py`` from the real
reference) — see ``tests/test_oracle_golden.py``.
"""
from __future__ import annotations

import hashlib
import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

ACT_SLOPE = 0.2  # melgan_generator.py:25 (act_slope) and melgan.py:18,25 (LeakyReLU(0.2))


@dataclass(frozen=True)
class GeneratorSpec:
    """Constructor arguments of the three generators (melgan_generator.py:9-16)."""

    in_channels: int = 80
    out_channels: int = 1
    proj_kernel: int = 7
    base_channels: int = 512
    upsample_factors: Tuple[int, ...] = (8, 8, 2, 2)
    res_kernel: int = 3
    num_res_blocks: int = 3
    pqmf: bool = False  # True for MultibandMelganGenerator (multiband_melgan_generator.py:24)

    @staticmethod
    def melgan(**kw) -> "GeneratorSpec":
        return GeneratorSpec(**kw)

    @staticmethod
    def multiband(**kw) -> "GeneratorSpec":
        # defaults: multiband_melgan_generator.py:8-16
        d = dict(out_channels=4, base_channels=384, upsample_factors=(2, 8, 2, 2), pqmf=True)
        d.update(kw)
        return GeneratorSpec(**d)

    @staticmethod
    def fullband(**kw) -> "GeneratorSpec":
        # defaults: fullband_melgan_generator.py:7-14
        d = dict(out_channels=1, base_channels=512, upsample_factors=(2, 8, 2, 2), num_res_blocks=4)
        d.update(kw)
        return GeneratorSpec(**d)

    @property
    def n_stages(self) -> int:
        return len(self.upsample_factors)

    @property
    def hop(self) -> int:
        """Output samples per mel frame."""
        h = 1
        for f in self.upsample_factors:
            h *= f
        return h * (4 if self.pqmf else 1)

    def stage_channels(self, s: int) -> Tuple[int, int]:
        """(in, out) channels of upsample stage ``s`` (melgan_generator.py:42-43)."""
        return self.base_channels // (2 ** s), self.base_channels // (2 ** (s + 1))

    # indices into the reference's ``self.layers`` nn.Sequential (melgan_generator.py:29-78)
    def idx_proj(self) -> int:
        return 1

    def idx_upsample(self, s: int) -> int:
        return 3 + 3 * s

    def idx_stack(self, s: int) -> int:
        return 4 + 3 * s

    def idx_head(self) -> int:
        return 4 + 3 * self.n_stages

    def n_layers(self) -> int:
        return 6 + 3 * self.n_stages

    def macs_per_item(self, T_padded: int) -> int:
        """Multiply-accumulates of one utterance with ``T_padded = T + 2p`` frames."""
        L = T_padded
        total = self.proj_kernel * self.in_channels * self.base_channels * L
        for s, f in enumerate(self.upsample_factors):
            cin, cout = self.stage_channels(s)
            total += 2 * f * cin * cout * L
            L *= f
            total += (self.res_kernel + 2) * self.num_res_blocks * cout * cout * L
        total += self.proj_kernel * cout * self.out_channels * L
        return total


# --------------------------------------------------------------------------------------
# weights
# --------------------------------------------------------------------------------------

def conv_keys(spec: GeneratorSpec) -> List[Tuple[str, str, Tuple[int, ...]]]:
    """[(prefix, kind, weight_shape)] of every parametrised conv, in ``state_dict()`` order.

    kind is "conv" (weight [C_out, C_in, k]) or "convT" (weight [C_in, C_out, k]).
    """
    out = [("layers.1", "conv", (spec.base_channels, spec.in_channels, spec.proj_kernel))]
    for s, f in enumerate(spec.upsample_factors):
        cin, cout = spec.stage_channels(s)
        out.append((f"layers.{spec.idx_upsample(s)}", "convT", (cin, cout, 2 * f)))
        st = f"layers.{spec.idx_stack(s)}"
        for b in range(spec.num_res_blocks):
            out.append((f"{st}.blocks.{b}.2", "conv", (cout, cout, spec.res_kernel)))
            out.append((f"{st}.blocks.{b}.4", "conv", (cout, cout, 1)))
        for b in range(spec.num_res_blocks):
            out.append((f"{st}.shortcuts.{b}", "conv", (cout, cout, 1)))
    clast = spec.stage_channels(spec.n_stages - 1)[1]
    out.append((f"layers.{spec.idx_head()}", "conv", (spec.out_channels, clast, spec.proj_kernel)))
    return out


def random_state_dict(spec: GeneratorSpec, seed: int, weight_norm: bool = False) -> Dict[str, torch.Tensor]:
    """Seeded fp32 weights with the reference's key names and torch-default-like magnitudes.

    Independent of module construction order (unlike default init under ``torch.manual_seed``);
    used for oracle-vs-CUDA parity at arbitrary sizes.  ``weight_norm=True`` emits the
    ``weight_g``/``weight_v`` form checkpoints are saved in (TTS/vocoder/utils/io.py:34-46).
    """
    g = torch.Generator().manual_seed(seed)
    sd: Dict[str, torch.Tensor] = {}
    for prefix, kind, shape in conv_keys(spec):
        fan_in = (shape[1] if kind == "conv" else shape[0]) * shape[2]
        bound = 1.0 / math.sqrt(fan_in)
        w = (torch.rand(shape, generator=g) * 2 - 1) * bound
        nb = shape[0] if kind == "conv" else shape[1]
        b = (torch.rand(nb, generator=g) * 2 - 1) * bound
        sd[f"{prefix}.bias"] = b
        if weight_norm:
            # torch weight_norm uses dim=0 for both Conv1d and ConvTranspose1d, i.e. the norm
            # is per *out*-channel for conv and per *in*-channel for convT.
            scale = 0.5 + torch.rand((shape[0], 1, 1), generator=g)
            sd[f"{prefix}.weight_g"] = w.flatten(1).norm(dim=1).view(-1, 1, 1) * scale
            sd[f"{prefix}.weight_v"] = w
        else:
            sd[f"{prefix}.weight"] = w
    if spec.pqmf:
        H, G, U = pqmf_filters()
        sd["pqmf_layer.H"], sd["pqmf_layer.G"], sd["pqmf_layer.updown_filter"] = H, G, U
    return sd


def fold_weight_norm(sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """``remove_weight_norm()`` on a state dict: w = g * v / ||v||, norm over all dims but 0.

    melgan_generator.py:91-97, melgan.py:41-45 (both defer to torch.nn.utils.remove_weight_norm,
    whose fold is ``torch._weight_norm(v, g, dim=0)``).
    """
    out: Dict[str, torch.Tensor] = {}
    for k, v in sd.items():
        if k.endswith(".weight_g"):
            continue
        if k.endswith(".weight_v"):
            g = sd[k[:-2] + "_g"]
            out[k[:-2]] = torch._weight_norm(v, g, 0)
        else:
            out[k] = v
    return out


def state_dict_digest(sd: Dict[str, torch.Tensor]) -> str:
    """sha256 over the fp32 bytes of every entry in key order — pins seeded weights in fixtures."""
    h = hashlib.sha256()
    for k in sorted(sd):
        h.update(k.encode())
        h.update(np.ascontiguousarray(sd[k].detach().cpu().float().numpy()).tobytes())
    return h.hexdigest()


# --------------------------------------------------------------------------------------
# generator forward
# --------------------------------------------------------------------------------------

def _lrelu(x: torch.Tensor) -> torch.Tensor:
    return F.leaky_relu(x, ACT_SLOPE)


def residual_stack(x: torch.Tensor, sd: Dict[str, torch.Tensor], prefix: str, spec: GeneratorSpec) -> torch.Tensor:
    """melgan.py:36-39:  x <- shortcut_b(x) + block_b(x) for each block b.

    block_b = LeakyReLU -> ReflectionPad1d(d) -> Conv1d(k, dilation d) -> LeakyReLU -> Conv1d(1)
    with d = res_kernel**b and pad = (res_kernel-1)//2 * d  (melgan.py:12-29).
    """
    half = (spec.res_kernel - 1) // 2
    for b in range(spec.num_res_blocks):
        d = spec.res_kernel ** b
        h = F.pad(_lrelu(x), (half * d, half * d), mode="reflect")
        h = F.conv1d(h, sd[f"{prefix}.blocks.{b}.2.weight"], sd[f"{prefix}.blocks.{b}.2.bias"], dilation=d)
        h = F.conv1d(_lrelu(h), sd[f"{prefix}.blocks.{b}.4.weight"], sd[f"{prefix}.blocks.{b}.4.bias"])
        x = F.conv1d(x, sd[f"{prefix}.shortcuts.{b}.weight"], sd[f"{prefix}.shortcuts.{b}.bias"]) + h
    return x


def generator_layers(spec: GeneratorSpec, sd: Dict[str, torch.Tensor], c: torch.Tensor,
                     taps: Optional[Dict[int, torch.Tensor]] = None) -> torch.Tensor:
    """``self.layers(c)`` (melgan_generator.py:80-81) on folded weights.

    ``taps``: if given, filled with {reference layer index: output of layers[index]} for the
    parametrised layers and stacks (the indexing convert_melgan_torch_to_tf.py:89-101 uses).
    """
    pp = (spec.proj_kernel - 1) // 2
    # layers[0..1]: ReflectionPad1d(3) + Conv1d(in -> C0, k7)            melgan_generator.py:30-38
    x = F.conv1d(F.pad(c, (pp, pp), mode="reflect"), sd["layers.1.weight"], sd["layers.1.bias"])
    if taps is not None:
        taps[spec.idx_proj()] = x
    for s, f in enumerate(spec.upsample_factors):
        # LeakyReLU + ConvTranspose1d(k=2f, stride f, pad f//2 + f%2, out_pad f%2)   :41-57
        iu = spec.idx_upsample(s)
        op = f % 2
        x = F.conv_transpose1d(_lrelu(x), sd[f"layers.{iu}.weight"], sd[f"layers.{iu}.bias"],
                               stride=f, padding=f // 2 + op, output_padding=op)
        if taps is not None:
            taps[iu] = x
        x = residual_stack(x, sd, f"layers.{spec.idx_stack(s)}", spec)            # :58-62
        if taps is not None:
            taps[spec.idx_stack(s)] = x
    # LeakyReLU + ReflectionPad1d(3) + Conv1d(C_last -> out, k7) + Tanh             :65-77
    ih = spec.idx_head()
    x = F.conv1d(F.pad(_lrelu(x), (pp, pp), mode="reflect"), sd[f"layers.{ih}.weight"], sd[f"layers.{ih}.bias"])
    if taps is not None:
        taps[ih] = x
    x = torch.tanh(x)
    if taps is not None:
        taps[ih + 1] = x
    return x


def generator_inference(spec: GeneratorSpec, sd: Dict[str, torch.Tensor], mel: torch.Tensor,
                        inference_padding: int = 2, dtype: torch.dtype = torch.float32,
                        taps: Optional[Dict[int, torch.Tensor]] = None) -> torch.Tensor:
    """``generator.inference(mel)`` (melgan_generator.py:83-89; MB: multiband_melgan_generator.py:32-39).

    Replicate-pads the mel by ``inference_padding`` frames per side, runs the stack, and for the
    multi-band generator applies PQMF synthesis.  ``sd`` may be in weight-norm form.
    """
    if any(k.endswith("weight_v") for k in sd):
        sd = fold_weight_norm(sd)
    sd = {k: v.detach().to("cpu", dtype) for k, v in sd.items()}
    c = mel.detach().to("cpu", dtype)
    with torch.no_grad():
        if inference_padding > 0:
            c = F.pad(c, (inference_padding, inference_padding), mode="replicate")
        y = generator_layers(spec, sd, c, taps)
        if spec.pqmf:
            H, G, U = pqmf_filters()
            y = pqmf_synthesis(y, sd.get("pqmf_layer.G", G).to(dtype), sd.get("pqmf_layer.updown_filter", U).to(dtype))
    return y


# --------------------------------------------------------------------------------------
# PQMF
# --------------------------------------------------------------------------------------

def pqmf_filters(N: int = 4, taps: int = 62, cutoff: float = 0.15, beta: float = 9.0):
    """Filter design of PQMF.__init__ (pqmf.py:19-40), float64 math then ``.float()``.

    prototype = firwin(taps+1, cutoff, kaiser(beta));  for band k and n = 0..taps:
        H[k,n] = 2 q[n] cos((2k+1) pi/(2N) (n - (taps-1)/2) + (-1)^k pi/4)
        G[k,n] = 2 q[n] cos((2k+1) pi/(2N) (n - (taps-1)/2) - (-1)^k pi/4)
    The (taps-1)/2 centre is the reference's (pqmf.py:23-25, marked TODO there) and is kept.
    Returns (H [N,1,taps+1], G [1,N,taps+1], updown_filter [N,N,N]).
    """
    from scipy.signal import firwin

    q = firwin(taps + 1, cutoff, window=("kaiser", beta))
    n = np.arange(taps + 1)
    H = np.empty((N, taps + 1))
    G = np.empty((N, taps + 1))
    for k in range(N):
        arg = (2 * k + 1) * (np.pi / (2 * N)) * (n - ((taps - 1) / 2))
        ph = (-1) ** k * np.pi / 4
        H[k] = 2 * q * np.cos(arg + ph)
        G[k] = 2 * q * np.cos(arg - ph)
    Ht = torch.from_numpy(H[:, None, :]).float()
    Gt = torch.from_numpy(G[None, :, :]).float()
    U = torch.zeros((N, N, N), dtype=torch.float32)
    for k in range(N):
        U[k, k, 0] = 1.0
    return Ht, Gt, U


def pqmf_synthesis(x: torch.Tensor, G: torch.Tensor, updown: torch.Tensor) -> torch.Tensor:
    """PQMF.synthesis (pqmf.py:51-56): zero-stuff x N with gain N, then 'same' FIR with G."""
    N = G.shape[1]
    taps = G.shape[2] - 1
    u = F.conv_transpose1d(x, updown * N, stride=N)
    return F.conv1d(u, G, padding=taps // 2)


def pqmf_analysis(x: torch.Tensor, H: torch.Tensor) -> torch.Tensor:
    """PQMF.analysis (pqmf.py:48-49): strided FIR with H."""
    N = H.shape[0]
    taps = H.shape[2] - 1
    return F.conv1d(x, H, padding=taps // 2, stride=N)


def receptive_field_frames(spec: GeneratorSpec) -> int:
    """One-sided receptive field of the generator in mel frames (ceil), by composing layer
    extents analytically.  Used to size halos for the time-split."""
    half = (spec.res_kernel - 1) // 2
    # walk from output back to input: r = one-sided extent in samples at the current rate
    r = 0.0
    if spec.pqmf:
        r = (31 + 3) / 4.0  # 63-tap synthesis FIR on the x4 zero-stuffed signal, in sub-band samples
    r += (spec.proj_kernel - 1) // 2
    for s in reversed(range(spec.n_stages)):
        f = spec.upsample_factors[s]
        r += half * sum(spec.res_kernel ** b for b in range(spec.num_res_blocks))
        r = (r + f) / f + 1  # transposed conv: each output sees 2 inputs; conservative
    r += (spec.proj_kernel - 1) // 2
    return int(math.ceil(r))
