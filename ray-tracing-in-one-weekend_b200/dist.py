"""Sample-split of one frame across ranks (one process per GPU, torch.distributed plumbing).

The reference has no multi-device path; the rule here is the north-star's:
each GPU renders a slice of the samples of EVERY pixel, partial sums travel to rank 0, and the
image must not depend on the GPU count.  That is achieved without arithmetic in flight:

  * the samples of a pixel are cut into SLICES = 64 canonical slices (`slice_samples`); the path
    kernel sums each slice sequentially,
  * rank g of G (G in 1, 2, 4, 8) owns the aligned block of 64/G slices `slice_block(g, G)` and
    pre-sums exactly that sub-tree of the pairwise tree ((s0+s1)+(s2+s3))+... on its own GPU,
  * the [H*W] float4 partials are GATHERED to rank 0 in rank order (NCCL send/recv over
    NVLink: bytes only), and rank 0's finalise kernel adds the remaining upper levels of the
    same tree, fused with write_color (src/cpu/color.h:8-23).

Every fp32 addition is therefore the same, in the same order, for every G.  The functions
below hold only the host-side logic (slice ownership, gather order); they work on any
torch.distributed backend, which is how the world_size-2 `gloo` tests on CPU cover them.
"""
from __future__ import annotations

SLICES = 64


def slice_block(rank: int, world: int) -> tuple[int, int]:
    """Canonical slices [begin, end) owned by `rank` of `world` ranks."""
    if world < 1 or SLICES % world:
        raise ValueError(f"world size {world} must divide {SLICES}")
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} out of range for world size {world}")
    per = SLICES // world
    return rank * per, (rank + 1) * per


SLICE_CUM = (0, 13, 25, 35, 44, 52, 58, 61, 62)  # cumulative slice weights inside a block


def slice_samples(k: int, spp: int) -> tuple[int, int]:
    """Frame-relative sample range [begin, end) of canonical slice k — the same rule as
    slice_range() in csrc/rt_device.cuh: 8 blocks of 8 slices; block b covers
    [floor(b*spp/8), floor((b+1)*spp/8)); inside a block the slices take 13, 12, 10, 9, 8, 6, 3, 1
    sixty-secondths (exactly those counts for a 62-sample block, i.e. at 500 spp)."""
    blk, j = divmod(k, 8)
    lo, hi = (blk * spp) // 8, ((blk + 1) * spp) // 8
    n = hi - lo
    return lo + n * SLICE_CUM[j] // 62, lo + n * SLICE_CUM[j + 1] // 62


def samples_of_block(begin: int, end: int, spp: int) -> int:
    return sum(b - a for a, b in (slice_samples(k, spp) for k in range(begin, end)))


class PartialGather:
    """Moves every rank's partial buffer to rank 0, in rank order, without arithmetic.

    `mine` is this rank's partial as a tensor (device tensor aliasing the frame's partial
    buffer under NCCL; a CPU tensor under gloo).  Receive buffers are allocated once and
    re-used by every frame."""

    def __init__(self, mine, rank: int, world: int, group=None, root: int = 0):
        import torch
        self.mine, self.rank, self.world, self.group, self.root = mine, rank, world, group, root
        self.parts = None
        if world > 1 and rank == root:
            self.parts = [mine if r == rank else torch.empty_like(mine) for r in range(world)]

    def __call__(self):
        """-> list of `world` tensors in rank order on the root, None elsewhere."""
        if self.world == 1:
            return [self.mine]
        import torch.distributed as dist
        dist.gather(self.mine, self.parts, dst=self.root, group=self.group)
        return self.parts
