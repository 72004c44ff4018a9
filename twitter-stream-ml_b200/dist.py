"""Row-sharded jobs: one process per GPU, torch.distributed only as the rendezvous that hands rank 0's
NCCL id to the other ranks (any backend: nccl on the GPU box, gloo in CPU tests).  The per-iteration
exchange itself is the library's own NCCL all-gather (csrc/twtml.cu), not torch's."""
import os

from . import _lib
from .api import Context


def row_range(n, rank, world_size):
    """Rows [lo, hi) of an n-row micro-batch owned by `rank`: contiguous blocks, so concatenating the
    ranks' outputs in rank order keeps predictOnValues' row order."""
    return n * rank // world_size, n * (rank + 1) // world_size


def exchange_nccl_id(make_id=None):
    """Rank 0 draws the id (twtml_nccl_unique_id), everyone receives it.  Needs an initialised
    torch.distributed process group."""
    import torch.distributed as dist
    box = [None]
    if dist.get_rank() == 0:
        box[0] = (make_id or Context.nccl_unique_id)()
    dist.broadcast_object_list(box, src=0)
    if not isinstance(box[0], (bytes, bytearray)) or len(box[0]) != _lib.NCCL_ID_BYTES:
        raise RuntimeError("bad NCCL id received from rank 0")
    return bytes(box[0])


def ranked_context(device=None):
    """Context for this process's rank of the torch.distributed job (single-process: plain context)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return Context(device if device is not None else int(os.environ.get("LOCAL_RANK", "0")))
    rank, world = dist.get_rank(), dist.get_world_size()
    dev = device if device is not None else int(os.environ.get("LOCAL_RANK", str(rank)))
    return Context(dev, rank, world, exchange_nccl_id())
