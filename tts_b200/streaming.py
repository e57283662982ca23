"""Chunked streaming inference (BASELINE config 5).

A stream of mel chunks goes in, waveform chunks come out; chunk k's samples are computed from its own frames plus
``context`` frames either side, where ``context`` >= the generator's receptive field, so the concatenated output
equals ``generator.inference`` of the whole utterance (same argument as tts_b200/longform.py: the network is a
finite-receptive-field, shift-equivariant stack — melgan_generator.py:29-78 has no normalisation or recurrence —
and what the padding rules do at a window's interior ends lands in the discarded context).  The reference's
padding (replicate ``inference_padding`` frames, melgan_generator.py:85-88, then the reflect pads inside) acts at
the true start and end of the stream only.

Latency: a chunk can be emitted once ``context`` frames of look-ahead have arrived; every call runs the CUDA path
on a fixed-size window (context + chunk + context frames), so plan, workspace and launch shapes are reused.
"""
import torch
import torch.nn.functional as F

from .longform import hop_length, receptive_halo_frames


class GraphedInference:
    """``generator.inference`` (padding 0) replayed from a CUDA graph per window shape.

    A streaming window is tiny (tens of frames): its ten-odd kernel launches and the allocator calls around them
    cost more host time than the kernels take on the device.  Every steady-state window has the same shape, so the
    whole forward — the C-ABI launches included, they go to the capturing stream like any other launch — is captured
    once per shape and replayed; the first chunks of a stream and the flush have their own (also cached) shapes.
    The result is copied out of the graph's static buffer, so callers may keep it.

    A captured graph holds raw pointers into the engine's plan (weight blobs) and workspace.  ``load_state_dict``,
    ``remove_weight_norm`` or ``.to()`` make the generator build a NEW engine and free the old plan, so every
    graph is tied to the engine it was captured from: each entry keeps a reference to that engine (its plan and
    workspace stay alive as long as the graph does) and the whole cache is dropped when the generator's current
    engine is a different object."""

    def __init__(self, generator, max_graphs=8):
        self.g = generator
        self.max_graphs = max_graphs
        self._graphs = {}
        self._engine = None      # the engine every cached graph was captured from

    def _capture(self, shape, device):
        static_in = torch.zeros(shape, dtype=torch.float32, device=device)
        side = torch.cuda.Stream(device)
        side.wait_stream(torch.cuda.current_stream(device))
        with torch.cuda.stream(side):          # warm-up off the capture: plan creation, attribute calls, workspace
            self.g.inference_padded(static_in, 0)
        torch.cuda.current_stream(device).wait_stream(side)
        torch.cuda.synchronize(device)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            static_out = self.g.inference_padded(static_in, 0)
        return graph, static_in, static_out

    def __call__(self, mel):
        engine = self.g._get_engine()          # rebuilt when the weights / device changed
        if engine is not self._engine:
            self._graphs.clear()               # graphs of a replaced engine point at freed weights: never replay them
            self._engine = engine
        key = (tuple(mel.shape), mel.device)
        entry = self._graphs.get(key)
        if entry is None:
            if len(self._graphs) >= self.max_graphs:
                self._graphs.pop(next(iter(self._graphs)))
            entry = self._graphs[key] = self._capture(tuple(mel.shape), mel.device)
            if self.g._get_engine() is not engine:   # the capture itself must not have swapped engines
                raise RuntimeError("GraphedInference: generator weights changed during graph capture")
        graph, static_in, static_out = entry
        static_in.copy_(mel, non_blocking=True)
        graph.replay()
        return static_out.clone()


class StreamingVocoder:
    """Feed mel frames with ``feed``; collect waveform chunks; call ``flush`` at the end of the stream.

    generator: a tts_b200 generator on a CUDA device (eval).  chunk_frames: frames per emitted chunk.
    context: frames of left/right context (default: the receptive field, ``receptive_halo_frames``).
    use_graph: replay each window shape from a CUDA graph (``GraphedInference``) instead of launching kernel by kernel.
    """

    def __init__(self, generator, chunk_frames=20, context=None, infer_fn=None, use_graph=False):
        self.g = generator
        self.chunk = int(chunk_frames)
        self.ctx = receptive_halo_frames(generator) if context is None else int(context)
        if self.chunk < 1 or self.ctx < 0:
            raise ValueError("chunk_frames must be >= 1 and context >= 0")
        self.hop = hop_length(generator)
        self.pad = int(generator.inference_padding)
        self._infer = infer_fn or (GraphedInference(generator) if use_graph else self._infer_unpadded)
        self.reset()

    def reset(self):
        self.buf = None        # frames received so far (with the left replicate padding), [B, C, n]
        self.base = 0          # padded-stream index of buf[..., 0]
        self.emitted = 0       # frames of the padded stream already turned into samples

    def _infer_unpadded(self, mel):
        return self.g.inference_padded(mel, 0)   # never touches the shared module's inference_padding

    def _window(self, start, end, total_known, final):
        """Samples of padded-stream frames [start, end)."""
        lo = max(0, start - self.ctx)
        hi = end + self.ctx
        if final:
            hi = min(hi, total_known)
        win = self.buf[..., lo - self.base:hi - self.base].contiguous()
        wav = self._infer(win)
        if wav.shape[-1] != (hi - lo) * self.hop:
            raise RuntimeError(f"streaming: expected {(hi - lo) * self.hop} samples, got {wav.shape[-1]}")
        return wav[..., (start - lo) * self.hop:(end - lo) * self.hop]

    def _drop_consumed(self):
        keep_from = max(0, self.emitted - self.ctx)   # the next window's left context starts here
        if keep_from > self.base:
            self.buf = self.buf[..., keep_from - self.base:]
            self.base = keep_from

    def feed(self, mel):
        """mel: [B, C, n] (any n >= 1).  Returns the list of waveform chunks that became computable."""
        if mel.dim() != 3 or mel.shape[-1] < 1:
            raise ValueError("feed() takes a [B, C, n>=1] mel chunk")
        if self.buf is None:
            if self.pad > 0:   # replicate padding at the true start of the stream
                mel = F.pad(mel, (self.pad, 0), mode="replicate")
            self.buf = mel
        else:
            self.buf = torch.cat([self.buf, mel.to(self.buf.device)], dim=-1)
        out = []
        known = self.base + self.buf.shape[-1]
        while self.emitted + self.chunk + self.ctx <= known:
            out.append(self._window(self.emitted, self.emitted + self.chunk, known, final=False))
            self.emitted += self.chunk
            self._drop_consumed()
        return out

    def flush(self):
        """End of stream: applies the right replicate padding and emits everything that is left."""
        if self.buf is None:
            return []
        if self.pad > 0:
            self.buf = F.pad(self.buf, (0, self.pad), mode="replicate")
        total = self.base + self.buf.shape[-1]
        out = []
        while self.emitted < total:
            end = min(self.emitted + self.chunk, total)
            out.append(self._window(self.emitted, end, total, final=True))
            self.emitted = end
            self._drop_consumed()
        self.reset()
        return out


def stream_inference(generator, mel, chunk_frames=20, feed_frames=None, context=None, infer_fn=None):
    """Whole-utterance convenience wrapper: == generator.inference(mel), computed chunk by chunk."""
    sv = StreamingVocoder(generator, chunk_frames, context, infer_fn)
    step = int(feed_frames or chunk_frames)
    parts = []
    for s in range(0, mel.shape[-1], step):
        parts += sv.feed(mel[..., s:s + step])
    parts += sv.flush()
    return torch.cat(parts, dim=-1)
