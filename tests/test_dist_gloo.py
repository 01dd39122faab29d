"""Host-side logic of the sample-split (ray-tracing-in-one-weekend_b200/dist.py) on CPU:
slice ownership, gather order, and that the canonical tree gives bit-identical sums for every
rank count.  world_size-2 (and 4) process groups over gloo; no GPU involved — the per-slice
buffers are synthetic float32 data and the tree is restated in numpy as the checker."""
import os
import socket

import numpy as np
import pytest


def tree(parts):
    """((s0+s1)+(s2+s3))+... pairwise in float32 — checker for csrc/rt_kernels.cu tree_sum."""
    parts = [np.asarray(p, np.float32) for p in parts]
    while len(parts) > 1:
        parts = [parts[k] + parts[k + 1] for k in range(0, len(parts), 2)]
    return parts[0]


NSLICES = 64  # == rt_b200.dist.SLICES (asserted below)


def synthetic_slices(npix, seed=5):
    rng = np.random.default_rng(seed)
    # magnitudes that make fp32 addition order matter
    return (rng.random((NSLICES, npix, 4), dtype=np.float32) * np.float32(300.0)).astype(np.float32)


def test_slice_blocks_partition_the_frame(rt):
    from rt_b200 import dist as rd
    assert rd.SLICES == NSLICES == rt.SLICES
    for G in (1, 2, 4, 8):
        blocks = [rd.slice_block(r, G) for r in range(G)]
        assert blocks[0][0] == 0 and blocks[-1][1] == rd.SLICES
        assert all(a[1] == b[0] for a, b in zip(blocks, blocks[1:]))
        for spp in (1, 7, 8, 100, 500, 4096):
            assert sum(rd.samples_of_block(b, e, spp) for b, e in blocks) == spp
    with pytest.raises(ValueError):
        rd.slice_block(0, 3)
    with pytest.raises(ValueError):
        rd.slice_block(2, 2)
    # ragged spp: slices differ by at most one sample and tile [0, spp)
    edges = [rd.slice_samples(k, 13) for k in range(rd.SLICES)]
    assert edges[0][0] == 0 and edges[-1][1] == 13
    assert all(a[1] == b[0] for a, b in zip(edges, edges[1:]))


def test_tree_is_rank_count_invariant():
    s = synthetic_slices(4096)
    full = tree(list(s))
    for G in (2, 4, 8):
        per = NSLICES // G
        partials = [tree(list(s[g * per:(g + 1) * per])) for g in range(G)]
        assert np.array_equal(tree(partials), full)
    # a plain left-to-right sum is NOT what the tree computes (so the test can fail)
    seq = s[0].copy()
    for k in range(1, NSLICES):
        seq = seq + s[k]
    assert not np.array_equal(seq, full)


def _worker(rank, world, port, npix, out_q):
    import torch
    import torch.distributed as dist
    from conftest import load_pkg
    load_pkg()
    from rt_b200 import dist as rd
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        s = synthetic_slices(npix)
        b, e = rd.slice_block(rank, world)
        mine = torch.from_numpy(tree(list(s[b:e])).copy())  # what rt_frame_reduce_local leaves
        gather = rd.PartialGather(mine, rank, world)
        for _ in range(2):  # buffers are re-used frame after frame
            parts = gather()
        if rank == 0:
            assert len(parts) == world
            out_q.put(tree([p.numpy() for p in parts]))  # what rt_finalize adds
        else:
            assert parts is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 4])
def test_gather_matches_single_rank(world):
    import torch.multiprocessing as mp
    npix = 3000
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, npix, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert np.array_equal(got, tree(list(synthetic_slices(npix))))
