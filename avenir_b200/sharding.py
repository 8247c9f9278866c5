"""Row sharding for multi-GPU runs: one process per GPU, contiguous row ranges, histograms summed over
NCCL inside the library (avt_comm_init_rank).  The only exchange step of the path is that sum -- the role
Hadoop's shuffle plays between BuilderMapper and BuilderReducer (tree/DecisionTreeBuilder.java:80-93)."""
from __future__ import annotations

from typing import Tuple


def shard_range(n_rows: int, rank: int, world: int) -> Tuple[int, int]:
    """Rows [lo, hi) of `rank`: contiguous, sizes differ by at most one, every row in exactly one shard."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside 0..{world - 1}")
    base, rem = divmod(n_rows, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def share_unique_id(dist, make_id, device=None) -> bytes:
    """Rank 0 creates the 128-byte communicator id (avt_comm_unique_id), every rank receives it through the
    already initialised torch.distributed group (any backend)."""
    import torch
    buf = torch.zeros(128, dtype=torch.uint8, device=device) if device is not None else torch.zeros(128, dtype=torch.uint8)
    if dist.get_rank() == 0:
        raw = make_id()
        buf.copy_(torch.frombuffer(bytearray(raw), dtype=torch.uint8))
    dist.broadcast(buf, 0)
    return bytes(buf.cpu().numpy().tobytes())


def connect_peers(dist, engine, device="cuda") -> None:
    """NVLink peer exchange for the level loop: gathers every rank's 64-byte IPC handle (avt_peer_export)
    through torch.distributed and maps them (avt_peer_import)."""
    import torch
    world, rank = dist.get_world_size(), dist.get_rank()
    mine = torch.frombuffer(bytearray(engine.peer_export()), dtype=torch.uint8).to(device)
    gathered = [torch.zeros(64, dtype=torch.uint8, device=device) for _ in range(world)]
    dist.all_gather(gathered, mine)
    engine.peer_import(world, rank, [bytes(g.cpu().numpy().tobytes()) for g in gathered])
