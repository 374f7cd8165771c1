"""Range sharding of a check batch over the GPUs of one box.

Units are independent and the world is read-only, so rank r simply owns the contiguous unit range
``shard_range(n, r, G)`` (whole mask words) and the only exchange is one in-place all-gather of mask
words.  The functions here are pure host logic over ``torch.distributed`` (NCCL on the GPU box, gloo in
the CPU tests); the check itself is passed in as a callable so the same code drives the CUDA path.
"""
import numpy as np

ALIGN = 1024  # units; 1024 units = 32 mask words = 128 B, so every rank's slice is 128 B aligned


def shard_range(n, rank, world_size, align=ALIGN):
    """Pure-Python twin of epp_shard_range (csrc/capi.cu); the C function is the one the product uses,
    this one lets the layout be computed where the library is not loaded.  Tested equal."""
    if align <= 0 or align % 32:
        raise ValueError("align must be a positive multiple of 32")
    blocks = (n + align - 1) // align
    per = (blocks + world_size - 1) // world_size * align
    begin = min(n, per * rank)
    return begin, min(n, begin + per)


def slot_words(n, world_size, align=ALIGN):
    """Mask words of one rank's slot in the gathered buffer (equal for all ranks; the last may be partly
    or wholly padding)."""
    blocks = (n + align - 1) // align
    return (blocks + world_size - 1) // world_size * align // 32


def sharded_check(n, check_local, rank, world_size, device=None, group=None, gathered=None, align=ALIGN):
    """Run ``check_local(begin, end, out_words)`` on this rank's range, writing straight into this rank's
    row of the gathered (world_size, slot_words) int32 tensor, then all-gather in place.  Returns the
    gathered tensor; ``assemble`` turns it into the n-unit mask."""
    import torch
    import torch.distributed as dist
    sw = slot_words(n, world_size, align)
    if gathered is None:
        gathered = torch.zeros((world_size, sw), dtype=torch.int32, device=device)
    begin, end = shard_range(n, rank, world_size, align)
    mine = gathered[rank]
    words = (end - begin + 31) // 32
    if words < sw:
        mine[words:].zero_()
    if end > begin:
        check_local(begin, end, mine[:words])
    if world_size > 1:
        dist.all_gather_into_tensor(gathered.view(-1), mine, group=group)
    return gathered


def assemble(gathered, n, align=ALIGN):
    """(world_size, slot_words) gathered words -> the (n + 31) // 32 words of the whole batch."""
    g = np.ascontiguousarray(gathered.detach().cpu().numpy()).view(np.uint32)
    world_size, sw = g.shape
    out = np.zeros((n + 31) // 32, dtype=np.uint32)
    for r in range(world_size):
        b, e = shard_range(n, r, world_size, align)
        w = (e - b + 31) // 32
        out[b // 32: b // 32 + w] = g[r, :w]
    return out
