"""Multi-GPU sharding of the block stream.

Blocks are independent (ref lib/salz.c:788,821: context per call;
programs/salzcli.c:143-179: flat concatenation), so the block index space is
cut into `world` contiguous ranges, one per GPU/process.  The only exchange is
one all-gather of the per-block encoded sizes, after which every rank knows the
byte offset of each of its records in the assembled file.
"""
import numpy as np


def block_range(nblocks, rank, world):
    """Contiguous [lo, hi) range of block indices owned by `rank`."""
    base, rem = divmod(nblocks, world)
    lo = rank * base + min(rank, rem)
    hi = lo + base + (1 if rank < rem else 0)
    return lo, hi


def allgather_sizes(local_sizes, nblocks, rank, world, device=None):
    """All-gather of per-block encoded sizes (uint32) over torch.distributed.

    With the NCCL backend the tensors live on `device` and the exchange runs over
    NVLink; with gloo (CPU tests) they stay on the host.  Returns the sizes of all
    `nblocks` blocks in stream order on every rank."""
    import torch
    import torch.distributed as dist

    per = (nblocks + world - 1) // world + 1
    buf = torch.zeros(per, dtype=torch.int64)
    buf[: len(local_sizes)] = torch.from_numpy(np.asarray(local_sizes, dtype=np.int64))
    if device is not None:
        buf = buf.to(device)
    out = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(out, buf)
    sizes = np.zeros(nblocks, dtype=np.uint32)
    for r in range(world):
        lo, hi = block_range(nblocks, r, world)
        sizes[lo:hi] = out[r][: hi - lo].cpu().numpy().astype(np.uint32)
    return sizes


def record_offsets(sizes, file_header=8):
    """Byte offset of every block record (u32 length + payload) in the .salz file
    (ref programs/salzcli.c:115-121,162-169): off_b = 8 + sum_{j<b} (4 + len_j)."""
    sizes = np.asarray(sizes, dtype=np.uint64)
    offs = np.zeros(len(sizes) + 1, dtype=np.uint64)
    np.cumsum(sizes + np.uint64(4), out=offs[1:])
    return offs + np.uint64(file_header)
