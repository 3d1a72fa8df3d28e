"""torchrun bootstrap of a multi-process job: torch.distributed only carries the NCCL id."""
from __future__ import annotations


def bootstrap_nccl(engine, rank: int, world: int) -> None:
    """Rank 0 makes the NCCL id, torch.distributed carries the 128 bytes, every
    rank joins the library's communicator."""
    import torch.distributed as dist
    from oni_ml_b200 import lda
    box = [lda.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    engine.comm_init(box[0], rank, world)
