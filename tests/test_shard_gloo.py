"""N > 1 host logic on CPU: two gloo ranks shard streams / GOPs and reduce the timing."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ffmpeg_hybrid_b200 import shard


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = shard.streams_of_rank(8, world, rank)
    frames = 30 * len(mine)
    ms = 1000.0 * (1 + rank)            # rank 1 is slower
    fps = shard.aggregate_fps(frames, ms, dist)
    q.put((rank, mine, fps))
    dist.destroy_process_group()


def test_two_rank_sharding_and_reduction():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][1] == [0, 2, 4, 6] and res[1][1] == [1, 3, 5, 7]
    # 240 frames in total over the slowest rank's 2 s
    assert abs(res[0][2] - 120.0) < 1e-9 and abs(res[1][2] - 120.0) < 1e-9


def test_closed_gop_split():
    flags = [True, False, False, True, False, True, False, False, False]
    plan = shard.split_closed_gops(flags, 2)
    assert plan == [[(0, 3), (5, 9)], [(3, 5)]]
    assert shard.aggregate_fps(30, 500.0) == 60.0
