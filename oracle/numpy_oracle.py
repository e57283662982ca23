"""ATen-free numpy restatement of the vocoder path from first principles (oracle; tests only).

Where ``oracle/melgan_oracle.py`` hands the arithmetic to ``torch.nn.functional`` (as the
reference does), this file writes every operator out from its definition in float64, using the
same closed forms the CUDA kernels use (two-tap polyphase transposed conv, 16-tap polyphase
PQMF).  Agreement of the two oracles on small cases (tests/test_oracle_golden.py) shows the
closed forms are the reference's operators and not merely close to them.

Formulas follow SURVEY.md §8a rows a2-a7; reference call sites:
  melgan_generator.py:30-38 (k7 projection), :41-57 (transposed conv), :65-77 (head),
  melgan.py:17-38 (residual stack), pqmf.py:51-56 (synthesis), pqmf.py:48-49 (analysis).
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np

from .melgan_oracle import ACT_SLOPE, GeneratorSpec


def lrelu(x: np.ndarray) -> np.ndarray:
    return np.where(x >= 0, x, ACT_SLOPE * x)


def reflect_index(i: np.ndarray, L: int) -> np.ndarray:
    """Index map of ReflectionPad1d: -1 -> 1, L -> L-2 (no edge repeat). Needs pad < L."""
    i = np.where(i < 0, -i, i)
    return np.where(i >= L, 2 * (L - 1) - i, i)


def conv1d_reflect(x: np.ndarray, W: np.ndarray, b: np.ndarray, dilation: int = 1) -> np.ndarray:
    """y[n,o,t] = b[o] + sum_{c,j} W[o,c,j] x[n,c,refl(t + (j - (k-1)/2) d)]   — 'same' length."""
    B, C, L = x.shape
    O, C2, k = W.shape
    assert C == C2 and k % 2 == 1
    half = (k - 1) // 2
    if half * dilation >= L:
        raise RuntimeError("Padding size should be less than the corresponding input dimension")
    t = np.arange(L)
    y = np.broadcast_to(b[None, :, None], (B, O, L)).astype(np.float64).copy()
    for j in range(k):
        src = reflect_index(t + (j - half) * dilation, L)
        y += np.einsum("oc,bct->bot", W[:, :, j], x[:, :, src])
    return y


def conv_transpose1d_melgan(x: np.ndarray, W: np.ndarray, b: np.ndarray, f: int) -> np.ndarray:
    """ConvTranspose1d(k=2f, stride=f, padding=f//2+f%2, output_padding=f%2), W is [C_in,C_out,2f].

    Closed form: output n takes exactly two input positions,
        tau = (n + P) // f  with tap  phi = (n + P) % f,   and   tau - 1  with tap  phi + f,
    P = f//2 + f%2, inputs outside [0, L) contributing zero.  Output length is exactly f*L.
    """
    B, C, L = x.shape
    C2, O, k = W.shape
    assert C == C2 and k == 2 * f
    P = f // 2 + f % 2
    n = np.arange(L * f)
    tau = (n + P) // f
    phi = (n + P) % f
    xp = np.concatenate([np.zeros((B, C, 1)), x, np.zeros((B, C, 1))], axis=2)  # xp[:, :, i+1] = x[:, :, i]
    hi = xp[:, :, np.clip(tau, -1, L) + 1]        # x[tau]   (tau == L  -> 0)
    lo = xp[:, :, np.clip(tau - 1, -1, L) + 1]    # x[tau-1] (tau == 0 -> 0)
    Whi = W[:, :, phi]                             # [C, O, n]
    Wlo = W[:, :, phi + f]
    y = np.einsum("con,bcn->bon", Whi, hi) + np.einsum("con,bcn->bon", Wlo, lo)
    return y + b[None, :, None]


def residual_stack(x: np.ndarray, sd: Dict[str, np.ndarray], prefix: str, spec: GeneratorSpec) -> np.ndarray:
    for blk in range(spec.num_res_blocks):
        d = spec.res_kernel ** blk
        h = conv1d_reflect(lrelu(x), sd[f"{prefix}.blocks.{blk}.2.weight"], sd[f"{prefix}.blocks.{blk}.2.bias"], d)
        h = conv1d_reflect(lrelu(h), sd[f"{prefix}.blocks.{blk}.4.weight"], sd[f"{prefix}.blocks.{blk}.4.bias"])
        s = conv1d_reflect(x, sd[f"{prefix}.shortcuts.{blk}.weight"], sd[f"{prefix}.shortcuts.{blk}.bias"])
        x = s + h
    return x


def pqmf_synthesis_polyphase(x: np.ndarray, G: np.ndarray) -> np.ndarray:
    """Closed form of PQMF.synthesis for N bands and an odd-length (taps+1) filter G [N, taps+1].

    y[N q + r] = N * sum_k sum_i G[k, N i + rho_r] x[k, q + i - D],
    rho_r = (half - r) mod N,  D = (half - r - rho_r)/N ... derived from t + j - half = N m.
    """
    B, N, Ts = x.shape
    K = G.shape[1]
    half = (K - 1) // 2
    y = np.zeros((B, N * Ts))
    q = np.arange(Ts)
    for r in range(N):
        for j in range(K):
            # t + j - half = N m  with t = N q + r  ->  m = q + (r + j - half)/N
            if (r + j - half) % N:
                continue
            m = q + (r + j - half) // N
            ok = (m >= 0) & (m < Ts)
            for k in range(N):
                y[:, N * q[ok] + r] += N * G[k, j] * x[:, k, m[ok]]
    return y[:, None, :]


def pqmf_analysis_direct(x: np.ndarray, H: np.ndarray) -> np.ndarray:
    """y[k, m] = sum_j H[k, j] x[N m + j - half], zero outside — pqmf.py:48-49."""
    B, one, T = x.shape
    N, K = H.shape
    half = (K - 1) // 2
    Tout = (T + 2 * half - K) // N + 1
    xp = np.concatenate([np.zeros((B, half)), x[:, 0], np.zeros((B, half))], axis=1)
    y = np.zeros((B, N, Tout))
    for m in range(Tout):
        y[:, :, m] = xp[:, N * m:N * m + K] @ H.T
    return y


def generator_inference(spec: GeneratorSpec, sd_t: Dict[str, "object"], mel, inference_padding: int = 2,
                        taps: Optional[Dict[int, np.ndarray]] = None) -> np.ndarray:
    """float64 numpy run of generator.inference on folded weights (torch tensors or arrays)."""
    sd = {k: np.asarray(v.detach().cpu().numpy() if hasattr(v, "detach") else v, dtype=np.float64)
          for k, v in sd_t.items()}
    c = np.asarray(mel.detach().cpu().numpy() if hasattr(mel, "detach") else mel, dtype=np.float64)
    if inference_padding > 0:  # F.pad(..., 'replicate')  melgan_generator.py:85-88
        c = np.concatenate([np.repeat(c[:, :, :1], inference_padding, 2), c,
                            np.repeat(c[:, :, -1:], inference_padding, 2)], axis=2)
    x = conv1d_reflect(c, sd["layers.1.weight"], sd["layers.1.bias"])
    if taps is not None:
        taps[1] = x
    for s, f in enumerate(spec.upsample_factors):
        iu = spec.idx_upsample(s)
        x = conv_transpose1d_melgan(lrelu(x), sd[f"layers.{iu}.weight"], sd[f"layers.{iu}.bias"], f)
        if taps is not None:
            taps[iu] = x
        x = residual_stack(x, sd, f"layers.{spec.idx_stack(s)}", spec)
        if taps is not None:
            taps[spec.idx_stack(s)] = x
    ih = spec.idx_head()
    x = np.tanh(conv1d_reflect(lrelu(x), sd[f"layers.{ih}.weight"], sd[f"layers.{ih}.bias"]))
    if spec.pqmf:
        x = pqmf_synthesis_polyphase(x, sd["pqmf_layer.G"][0])
    return x
