"""N > 1 host logic on CPU: two gloo ranks shard independent frames, each replays its own
(the oracle stands in for the GPU here), and the union equals the single-process result;
the timing reduction is a max over ranks.  No collective touches the data path."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

import frame_cases
import oracle_lib
from hevc_b200 import shard, synth

N_FRAMES = 5


def _frame_digest(i):
    job = synth.make_job(64, 64, 10, 1, seed=shard.seed_of_frame(900, i), weighted=1)
    refs = [synth.make_picture(job.sp, 77 + s) for s in range(job.sp.dpb_slots)]
    return frame_cases.picture_digest(oracle_lib.Lib("oracle").replay(job, refs))


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = shard.frames_of_rank(N_FRAMES, rank, world)
    digests = {i: _frame_digest(i) for i in mine}
    shard.barrier()
    ms = shard.max_over_ranks([10.0 + rank, 5.0 - rank])
    gathered = [None] * world
    dist.all_gather_object(gathered, digests)          # test-side check only
    if rank == 0:
        q.put((gathered, ms))
    dist.destroy_process_group()


def test_partition_is_exact():
    for world in (1, 2, 3, 8):
        got = sorted(i for r in range(world) for i in shard.frames_of_rank(13, r, world))
        assert got == list(range(13))
    with pytest.raises(ValueError):
        shard.frames_of_rank(4, 2, 2)


def test_two_ranks_gloo():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    gathered, ms = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(gathered[0]) == [0, 2, 4] and sorted(gathered[1]) == [1, 3]
    merged = {**gathered[0], **gathered[1]}
    assert merged == {i: _frame_digest(i) for i in range(N_FRAMES)}
    assert ms == [11.0, 5.0]
    assert np.isfinite(ms).all()
