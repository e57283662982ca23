"""Long-form inference: one utterance split along time across ranks (BASELINE config 4).

The generator is a finite-receptive-field, shift-equivariant stack (no normalisation, no recurrence), so a time
split with a halo of at least the receptive field reproduces the full-utterance output exactly away from the
global ends.  Rank r owns mel frames ``[r*T/N, (r+1)*T/N)`` of the (globally replicate-padded) mel, runs
``generator.inference`` on those frames plus ``halo`` frames per interior side with ``inference_padding = 0``, and
keeps only the samples of its own frames.  The reference's padding rules (``F.pad(..., 'replicate')`` at
melgan_generator.py:85-88, the reflect pads inside the network) therefore act at the global ends only; what
they do at a chunk's interior ends lands in the discarded halo.

Exchange: the mel is tiny (80 x 4 B per frame; 16.5 MB for 10 minutes), so every rank holds it and the "halo
exchange" is a local slice — zero bytes on the wire.  The only collective is the final gather of the waveform
slices (``all_gather`` over NCCL; 6.6 MB per rank for 10 minutes on 8 GPUs).

Nothing here touches the arithmetic: every sample comes out of the same CUDA path as ``inference()``.
"""
import math

import torch
import torch.nn.functional as F


def receptive_halo_frames(generator):
    """Mel frames of context one output sample can depend on, per side (upper bound).

    Walks the layer list backwards (melgan_generator.py:29-78): PQMF synthesis reaches ``ceil((taps/2)/N)``
    sub-band samples, the head conv ``(k-1)/2``, each ResidualStack ``sum_b ((res_kernel-1)/2) * res_kernel**b``
    (melgan.py:14-24), each ConvTranspose1d(k=2f, stride=f) maps a reach of ``r`` output positions to
    ``ceil(r/f) + 1`` input positions, and the projection conv adds ``(k-1)/2``.
    """
    r = 0
    pq = getattr(generator, "pqmf_layer", None)
    if pq is not None:
        r = math.ceil((pq.taps // 2 + 1) / pq.N)
    r += (generator.proj_kernel - 1) // 2
    half = (generator.res_kernel - 1) // 2
    per_stack = sum(half * generator.res_kernel ** b for b in range(generator.num_res_blocks))
    for f in reversed(generator.upsample_factors):
        r += per_stack
        r = math.ceil(r / f) + 1
    return r + (generator.proj_kernel - 1) // 2


def frame_partition(n_frames, world_size):
    """Contiguous, near-equal frame ranges [(start, end)] for each rank."""
    base, rem = divmod(n_frames, world_size)
    out, s = [], 0
    for r in range(world_size):
        e = s + base + (1 if r < rem else 0)
        out.append((s, e))
        s = e
    return out


def hop_length(generator):
    hop = 1
    for f in generator.upsample_factors:
        hop *= int(f)
    pq = getattr(generator, "pqmf_layer", None)
    return hop * (pq.N if pq is not None else 1)


def _infer_unpadded(generator, mel, out=None):
    """generator.inference with padding 0 for this call only; the shared module's attribute is never written
    (request threads may be running inference() on the same module, TTS/server/server.py:64,82)."""
    return generator.inference_padded(mel, 0, out=out)


def local_slice(generator, mel_padded, start, end, halo, infer_fn=None):
    """Waveform samples of frames [start, end) of the padded mel, computed from a halo-extended window."""
    T = mel_padded.shape[-1]
    hop = hop_length(generator)
    lo, hi = max(0, start - halo), min(T, end + halo)
    infer = infer_fn or (lambda m: _infer_unpadded(generator, m))
    wav = infer(mel_padded[..., lo:hi].contiguous())
    if wav.shape[-1] != (hi - lo) * hop:
        raise RuntimeError(f"longform: expected {(hi - lo) * hop} samples for {hi - lo} frames, got {wav.shape[-1]}")
    return wav[..., (start - lo) * hop:(end - lo) * hop]


def chunked_inference(generator, mel, n_chunks, halo=None, infer_fn=None):
    """Single-process version of the split (what the ranks do, run back to back): == generator.inference(mel)."""
    p = int(generator.inference_padding)
    halo = receptive_halo_frames(generator) if halo is None else int(halo)
    mel_padded = F.pad(mel, (p, p), mode="replicate") if p > 0 else mel
    parts = [local_slice(generator, mel_padded, s, e, halo, infer_fn)
             for s, e in frame_partition(mel_padded.shape[-1], n_chunks) if e > s]
    return torch.cat(parts, dim=-1)


def longform_inference(generator, mel, group=None, halo=None, infer_fn=None, gather=True):
    """Distributed long-form inference.  Call on every rank of ``group`` with the same full ``mel`` [B, C, T].

    Returns the full waveform [B, 1, (T + 2*inference_padding) * hop] on every rank (``gather=True``) or only
    this rank's slice plus its (start_sample, end_sample) range (``gather=False``).
    """
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    p = int(generator.inference_padding)
    halo = receptive_halo_frames(generator) if halo is None else int(halo)
    mel_padded = F.pad(mel, (p, p), mode="replicate") if p > 0 else mel
    T = mel_padded.shape[-1]
    hop = hop_length(generator)
    parts = frame_partition(T, world)
    s, e = parts[rank]
    mine = local_slice(generator, mel_padded, s, e, halo, infer_fn) if e > s else None
    if not gather:
        return mine, (s * hop, e * hop)
    if world == 1:
        return mine
    # equal-size all_gather: pad every slice to the longest one, trim after
    longest = max(pe - ps for ps, pe in parts) * hop
    B = mel.shape[0]
    dev = mine.device if mine is not None else mel.device
    send = torch.zeros((B, 1, longest), dtype=torch.float32, device=dev)
    if mine is not None:
        send[..., :mine.shape[-1]] = mine
    recv = torch.empty((world * B, 1, longest), dtype=torch.float32, device=dev)   # rank-major along dim 0
    dist.all_gather_into_tensor(recv, send, group=group)
    recv = recv.view(world, B, 1, longest)
    return torch.cat([recv[r, ..., :(pe - ps) * hop] for r, (ps, pe) in enumerate(parts) if pe > ps], dim=-1)
