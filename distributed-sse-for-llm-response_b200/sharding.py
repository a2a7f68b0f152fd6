"""Host-side multi-GPU plumbing: conversations shard by fnv1a64(conversation_id) % n_gpus (include/ssegpu.h
ssegpu_shard; Go: hash/fnv New64a), every GPU runs an independent context, nothing crosses GPUs on the data path.
The only cross-rank step is the gather of per-GPU counters that feeds the reference's Prometheus totals
(sse_messages_delivered_total, src/sse-adapter/sse_handler.go:32-35)."""
import numpy as np

from . import workload
from .binding import shard

COUNTER_KEYS = ("records", "frames", "bytes_in", "bytes_out", "err_syntax", "err_type", "filtered", "done", "fast_path",
                "batches")


def route(conversation_ids, n_gpus):
    """GPU index for each conversation id (bytes)."""
    return np.array([shard(c, n_gpus) for c in conversation_ids], dtype=np.uint32)


def rank_workload(config, n_conv_total, n_tok, rank, world):
    """The token stream of the conversations that the host routes to GPU `rank`."""
    if world == 1:
        return workload.generate(config, n_conv_total, n_tok)
    return workload.generate(config, 0, n_tok, conv_map=workload.shard_members(config, n_conv_total, rank, world))


def gather_counters(local, dist=None, device="cpu"):
    """Sum of the per-GPU counter dicts over all ranks (an all-reduce of ten integers; gloo or nccl)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return dict(local)
    import torch
    t = torch.tensor([int(local[k]) for k in COUNTER_KEYS], dtype=torch.int64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return {k: int(v) for k, v in zip(COUNTER_KEYS, t.tolist())}
