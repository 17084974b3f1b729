"""Shard-by-window scale-out of the transcription path.

The reference's own sharding precedent is `whisper_full_parallel` (sys/whisper.cpp/src/whisper.cpp:7736-7864): the
clip is cut into equal sample spans, every span is transcribed by an independent `whisper_state` with
`whisper_full_with_state` semantics, and the spans' segments are concatenated in span order with their times shifted by
`100 * start_sample / 16000` centiseconds (wc:7816-7817).  Here a span is one 30 s window (480 000 samples), spans are
dealt to ranks (one process per GPU) in contiguous blocks, every rank runs its spans as lock-step batches through
`whisper_b200_full_batch`, and the only cross-rank step is a host-side gather of the few KB of segments: there is no
collective on the math path.

No arithmetic on audio or activations happens in this file.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, replace

import numpy as np

from .api import FullParams, Segment, WhisperContext

SAMPLE_RATE = 16000
SHARD_SAMPLES = 30 * SAMPLE_RATE     # one encoder window (WHISPER_CHUNK_SIZE, whisper.h:35)
MIN_SAMPLES = SAMPLE_RATE // 10      # the engine returns no segments below 100 ms (wc:6849-6852)


@dataclass(frozen=True)
class Shard:
    index: int      # position in the clip
    start: int      # first sample
    count: int      # samples

    @property
    def t_offset(self) -> int:
        """segment-time offset in 10 ms units (wc:7816)."""
        return 100 * self.start // SAMPLE_RATE


def split_shards(n_samples: int, shard_samples: int = SHARD_SAMPLES) -> list:
    """Cut [0, n_samples) into consecutive spans of `shard_samples`; the last one may be shorter."""
    if n_samples <= 0:
        return []
    return [Shard(i, s, min(shard_samples, n_samples - s)) for i, s in enumerate(range(0, n_samples, shard_samples))]


def assign(n_shards: int, world: int, rank: int) -> range:
    """Contiguous block of shard indices owned by `rank` (blocks differ by at most one shard)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world {world}")
    base, extra = divmod(n_shards, world)
    lo = rank * base + min(rank, extra)
    return range(lo, lo + base + (1 if rank < extra else 0))


def offset_segments(segs: list, shard: Shard) -> list:
    """Shift a shard's segments to clip time (t0/t1 and the per-token t0/t1/t_dtw where set), wc:7816-7817."""
    off = shard.t_offset
    out = []
    for s in segs:
        toks = []
        for t in s.tokens:
            t = dict(t)
            for k in ("t0", "t1", "t_dtw"):
                if t.get(k, -1) >= 0:
                    t[k] += off
            toks.append(t)
        out.append(replace(s, t0=s.t0 + off, t1=s.t1 + off, tokens=toks))
    return out


def merge(per_shard: dict) -> list:
    """{shard index: segments already in clip time} -> one list in shard order (the host-side gather's reduction)."""
    out = []
    for i in sorted(per_shard):
        out.extend(per_shard[i])
    return out


def gather_segments(local: dict, group=None) -> list:
    """All ranks contribute {shard index: segments}; every rank gets the merged list.  Uses torch.distributed's object
    gather on the host (gloo or nccl process groups both work; payload is a few KB per shard)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return merge(local)
    parts = [None] * dist.get_world_size(group)
    dist.all_gather_object(parts, local, group=group)
    allsh = {}
    for p in parts:
        for k, v in p.items():
            if k in allsh:
                raise RuntimeError(f"shard {k} transcribed by two ranks")
            allsh[k] = v
    return merge(allsh)


def transcribe_shards(ctx: WhisperContext, params: FullParams, pcm: np.ndarray, shards: list, batch: int = 16,
                      states: list | None = None) -> dict:
    """Run the given shards of `pcm` on this process's GPU as lock-step batches of `batch` states through
    whisper_b200_full_batch (include/whisper_b200.h).  Returns {shard index: segments in clip time}."""
    lib = ctx.lib
    if not hasattr(lib, "whisper_b200_full_batch"):
        raise RuntimeError("transcribe_shards needs the product library (whisper_b200_full_batch)")
    from .api import whisper_full_params
    fn = lib.whisper_b200_full_batch
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.c_int, whisper_full_params, C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_int)]
    pcm = np.ascontiguousarray(pcm, dtype=np.float32)
    own = states is None
    states = states if states is not None else [ctx.create_state() for _ in range(min(batch, max(1, len(shards))))]
    out = {}
    try:
        for b0 in range(0, len(shards), len(states)):
            grp = shards[b0:b0 + len(states)]
            n = len(grp)
            st_arr = (C.c_void_p * n)(*[s.ptr for s in states[:n]])
            ptrs = (C.POINTER(C.c_float) * n)(*[C.cast(pcm[g.start:].ctypes.data, C.POINTER(C.c_float)) for g in grp])
            cnts = (C.c_int * n)(*[g.count for g in grp])
            r = fn(ctx.ptr, st_arr, n, params.fp, ptrs, cnts)
            if r != 0:
                raise RuntimeError(f"whisper_b200_full_batch failed: {r}")
            for g, st in zip(grp, states):
                out[g.index] = offset_segments(st.segments(), g)
    finally:
        if own:
            for s in states:
                s.free()
    return out


def transcribe_sharded(ctx: WhisperContext, params: FullParams, pcm: np.ndarray, rank: int = 0, world: int = 1, batch: int = 16,
                       group=None) -> list:
    """Whole-clip entry point: shard by 30 s window, run this rank's block, gather on the host."""
    shards = split_shards(len(pcm))
    mine = [shards[i] for i in assign(len(shards), world, rank)]
    return gather_segments(transcribe_shards(ctx, params, pcm, mine, batch=batch), group=group)
