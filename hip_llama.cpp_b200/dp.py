"""Data-parallel plumbing of the batching driver across PROCESSES (one per GPU, torchrun):
the path shards by sequence with replicated weights and no data-path collective
(reference: one OpenMP thread per GPU pulling from a shared queue, src/llama.cpp:913-983).
torch.distributed is used only for rendezvous, the timing reduction and gathering results."""
from __future__ import annotations

import torch
import torch.distributed as dist


def request_shard(num_reqs: int, rank: int, world: int) -> list[int]:
    """Static round-robin deal of request ids (reproducible timing; the C++ driver's
    DriverOptions::device_offset / device_stride implement the same rule)."""
    return list(range(rank, num_reqs, world))


def max_over_ranks(value: float, device="cpu") -> float:
    """Step time of the job = the slowest rank's device time."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_outputs(local: dict[int, list[int]], num_reqs: int) -> list[list[int]] | None:
    """Every rank hands in {request id: token ids}; rank 0 gets the list ordered by request id
    (the reference writes results by request id, src/llama.cpp:1056, so the output does not depend
    on which GPU served which request)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        parts = [local]
    else:
        parts = [None] * dist.get_world_size()
        dist.all_gather_object(parts, local)
    if dist.is_initialized() and dist.get_rank() != 0:
        return None
    out: list[list[int] | None] = [None] * num_reqs
    for p in parts:
        for rid, toks in p.items():
            if out[rid] is not None:
                raise ValueError(f"request {rid} served twice")
            out[rid] = toks
    missing = [i for i, o in enumerate(out) if o is None]
    if missing:
        raise ValueError(f"requests never served: {missing[:8]}")
    return out
