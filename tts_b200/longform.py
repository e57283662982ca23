"""Long-form inference: one utterance split along time across ranks (BASELINE config 4).

The generator is a finite-receptive-field, shift-equivariant stack (no normalisation, no recurrence), so a time
split with a halo of at least the receptive field reproduces the full-utterance output exactly away from the
global ends.  Rank r owns mel frames ``[r*T/N, (r+1)*T/N)`` of the (globally replicate-padded) mel, runs
``generator.inference`` on those frames plus ``halo`` frames per interior side with ``inference_padding = 0``, and
keeps only the samples of its own frames.  The reference's padding rules (``F.pad(..., 'replicate')`` at
melgan_generator.py:85-88, the reflect pads inside the network) therefore act at the global ends only; what
they do at a chunk's interior ends lands in the discarded halo.

Exchange: the mel is tiny (80 x 4 B per frame; 16.5 MB for 10 minutes), so every rank holds it and the "halo
exchange" is a local slice — zero bytes on the wire.  The waveform is not gathered either: ``LongformPlan``
registers ONE destination buffer on the consuming rank (symmetric memory / CUDA IPC, mapped over NVLink), and every
rank's last kernel (PQMF synthesis) stores the samples of its own frames — and only those; the halo samples are
computed but never written — straight into their place in that buffer.  One device-side barrier orders the stores
before the consumer reads.  No ``all_gather``, no padding buffer, no ``torch.cat``.

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

    This is the portable form (any backend, any ``infer_fn``; the gloo tests run it on the CPU oracle): slices are
    exchanged with one ``all_gather``.  On NVLink-connected GPUs use ``LongformPlan``, which places every rank's
    samples directly in the consumer's buffer instead.
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


class LongformPlan:
    """Direct-placement long-form inference on NVLink-connected GPUs (BASELINE config 4).

        plan = LongformPlan(generator, n_frames=T, batch=B)      # collective: every rank of the group
        wav = plan.run(mel)                                       # every rank; the waveform on `dst_rank`, None elsewhere

    Construction allocates the destination ``[B, 1, (T + 2 p) * hop]`` once, on every rank symmetrically, and maps
    ``dst_rank``'s copy into every other rank's address space:

      * ``symmetric``: ``torch.distributed._symmetric_memory`` (cuMem handles exchanged at rendezvous; peer stores go
        over NVLink / NVSwitch), with its signal-pad barrier ordering the stores before the consumer's reads;
      * ``ipc``: CUDA IPC handle of ``dst_rank``'s buffer broadcast to the group, NCCL one-element all_reduce as barrier;
      * ``gather``: no peer mapping available (not an NVLink box, or a CPU group) — falls back to
        ``longform_inference`` and says so in ``describe()``.

    ``run`` slices this rank's halo-extended window out of the replicated mel and calls the generator with
    ``out=`` a view of the destination at its sample offset and ``keep=`` its own sample range, so the PQMF kernel's
    stores are the only data movement of the whole gather step.  For a generator without PQMF the kept range is
    placed by one strided device-to-device copy instead (vocoder.cu finish_output).
    """

    def __init__(self, generator, n_frames, batch=1, group=None, halo=None, dst_rank=0, dtype=torch.float32,
                 mode=None):
        import torch.distributed as dist
        self.g = generator
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        if mode == "local":          # this process alone, whatever group it belongs to (a one-GPU baseline inside a job)
            self.world, self.rank, dst_rank, mode = 1, 0, 0, None
        self.dst_rank = int(dst_rank)
        self.p = int(generator.inference_padding)
        self.halo = receptive_halo_frames(generator) if halo is None else int(halo)
        self.hop = hop_length(generator)
        self.frames = int(n_frames) + 2 * self.p
        self.parts = frame_partition(self.frames, self.world)
        self.n_out = self.frames * self.hop
        self.batch = int(batch)
        self.dtype = dtype
        self.device = next(generator.parameters()).device
        self._hdl = self._flag = self._ipc_keep = None
        self.local = self.dest = None
        self.mode, self.why = self._map_destination(mode)

    # ------------------------------------------------------------------ destination mapping
    def _map_destination(self, forced):
        import torch.distributed as dist
        shape = (self.batch, 1, self.n_out)
        if self.world == 1 or self.device.type != "cuda":
            self.local = torch.empty(shape, dtype=self.dtype, device=self.device)
            self.dest = self.local
            return "local", "single rank"
        errors = []
        for mode in ([forced] if forced else ["symmetric", "ipc", "gather"]):
            try:
                if mode == "symmetric":
                    import torch.distributed._symmetric_memory as symm
                    grp = self.group if self.group is not None else dist.group.WORLD
                    buf = symm.empty(shape, dtype=self.dtype, device=self.device)
                    hdl = symm.rendezvous(buf, grp)
                    self._hdl, self.local = hdl, buf
                    self.dest = buf if self.rank == self.dst_rank else hdl.get_buffer(self.dst_rank, shape, self.dtype)
                    hdl.barrier(channel=0)
                    return mode, "torch symmetric memory, peer stores over NVLink"
                if mode == "ipc":
                    self.local = torch.empty(shape, dtype=self.dtype, device=self.device)
                    payload = [self.local.untyped_storage()._share_cuda_() if self.rank == self.dst_rank else None]
                    dist.broadcast_object_list(payload, src=self.dst_rank, group=self.group)
                    if self.rank == self.dst_rank:
                        self.dest = self.local
                    else:
                        st = torch.UntypedStorage._new_shared_cuda(*payload[0])
                        self._ipc_keep = st
                        self.dest = torch.empty(0, dtype=self.dtype, device=st.device).set_(st, 0, shape)
                        if self.dest.device != self.device:      # the mapping lives on the owner's device index
                            raise RuntimeError("peer mapping came back on another device index")
                    self._flag = torch.zeros(1, dtype=torch.float32, device=self.device)
                    return mode, "CUDA IPC handle of the consumer's buffer, NCCL flag as barrier"
                if mode == "gather":
                    return mode, "no peer mapping: " + "; ".join(errors)
            except Exception as e:   # try the next mechanism
                errors.append(f"{mode}: {type(e).__name__}: {e}")
                # every rank must take the same branch: agree on failure
            finally:
                pass
        raise RuntimeError("LongformPlan: no destination mapping available: " + "; ".join(errors))

    def describe(self):
        return {"mode": self.mode, "why": self.why, "launch": getattr(self, "graph_note", "eager"),
                "dst_rank": self.dst_rank, "halo_frames": self.halo,
                "frames_per_rank": [e - s for s, e in self.parts],
                "recomputed_halo_frames_per_rank": [min(self.frames, e + self.halo) - max(0, s - self.halo) - (e - s)
                                                    for s, e in self.parts]}

    def _barrier(self):
        import torch.distributed as dist
        if self.mode == "symmetric":
            self._hdl.barrier(channel=0)          # device-side, on the current stream
        elif self.mode == "ipc":
            dist.all_reduce(self._flag, group=self.group)   # stream-ordered after this rank's stores

    # ------------------------------------------------------------------ the hot call
    def run(self, mel):
        """mel: the full [B, C, T] mel on this rank's device (replicated).  Returns the waveform on dst_rank."""
        if mel.shape[0] != self.batch or mel.shape[-1] + 2 * self.p != self.frames:
            raise RuntimeError(f"LongformPlan was built for [{self.batch}, C, {self.frames - 2 * self.p}] mels")
        if self.mode == "gather":
            wav = longform_inference(self.g, mel, group=self.group, halo=self.halo)
            return wav if self.rank == self.dst_rank else None
        p = self.p
        s, e = self.parts[self.rank]
        if e > s:
            lo, hi = max(0, s - self.halo), min(self.frames, e + self.halo)
            # window of the replicate-padded mel without materialising the padded copy: clamp the frame indices
            if p > 0 and (lo < p or hi > self.frames - p):
                idx = torch.arange(lo - p, hi - p, device=mel.device).clamp_(0, mel.shape[-1] - 1)
                window = mel.index_select(-1, idx)
            else:
                window = mel[..., lo - p:hi - p].contiguous()
            out = self.dest[..., s * self.hop:e * self.hop]
            self.g.inference_padded(window, 0, out=out, keep=((s - lo) * self.hop, (e - s) * self.hop))
        self._barrier()
        return self.local if self.rank == self.dst_rank else None

    def graphed(self, mel):
        """``run(mel)`` captured once into a CUDA graph; returns ``replay() -> waveform (dst_rank) / None``.

        At N ranks a rank's kernels take ~1/N of the one-GPU time while the host side of ``run`` — the window slice,
        a dozen launches through the C ABI, the barrier — stays what it was: with many ranks the host, not the GPUs,
        would set the pace.  The shapes are fixed by the plan, so everything is captured for THIS ``mel`` tensor
        (callers refill it in place) and replayed with one launch.  Collective: every rank of the group calls it.
        Falls back to the eager ``run`` (and says why in ``describe()``) where the barrier cannot be captured."""
        if self.mode == "gather" or self.device.type != "cuda":
            return lambda: self.run(mel)
        dev = self.device
        cur = torch.cuda.current_stream(dev)
        try:
            side = torch.cuda.Stream(dev)
            side.wait_stream(cur)
            with torch.cuda.stream(side):      # warm-up off the capture: plan, workspace, attribute calls
                self.run(mel)
            cur.wait_stream(side)
            torch.cuda.synchronize(dev)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                self.run(mel)
        except Exception as exc:               # e.g. a barrier implementation that refuses stream capture
            self.graph_note = f"eager ({type(exc).__name__}: {str(exc)[:120]})"
            torch.cuda.synchronize(dev)
            return lambda: self.run(mel)
        self.graph_note = "cuda graph"
        self._graph = graph

        def replay():
            graph.replay()
            return self.local if self.rank == self.dst_rank else None
        return replay

    def close(self):
        self._graph = None
        self.dest = self.local = self._hdl = self._ipc_keep = None
