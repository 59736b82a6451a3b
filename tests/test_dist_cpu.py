"""world_size-2 gloo test of the multi-GPU host logic (frame sharding, result gather, RANSAC winner merge)."""
import os
import socket
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from instance_based_loc_b200 import dist as D
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = 7
    mine = D.shard_indices(n)
    poses = [np.arange(7, dtype=np.float64) + 10 * i for i in mine]          # stand-in for localise results
    allp = D.gather_frame_results(mine, poses, n, device="cpu")
    h0, h1 = D.hypothesis_block(1001)
    # rank 1 has the same fitness but a lower rmse -> it must win; ties on both fall to the lower hypothesis index
    fit, rmse, hyp = (0.5, 0.02, 40) if rank == 0 else (0.5, 0.01, 900)
    best = D.best_of_ranks(fit, rmse, hyp, np.eye(4) * (rank + 1), device="cpu")
    tie = D.best_of_ranks(0.25, 0.03, 77 - rank, np.eye(4), device="cpu")
    out[rank] = (mine, allp, (h0, h1), best, tie)
    dist.destroy_process_group()


def test_two_rank_sharding_and_gather():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    assert out[0][0] == [0, 2, 4, 6] and out[1][0] == [1, 3, 5]
    expect = np.stack([np.arange(7, dtype=np.float64) + 10 * i for i in range(7)])
    for r in range(world):
        assert np.array_equal(out[r][1], expect)                             # every rank holds every pose
    assert out[0][2] == (0, 501) and out[1][2] == (501, 1001)
    for r in range(world):
        fit, rmse, hyp, T, winner = out[r][3]
        assert (fit, rmse, hyp, winner) == (0.5, 0.01, 900, 1) and np.array_equal(T, np.eye(4) * 2)
        assert out[r][4][2] == 76 and out[r][4][4] == 1                      # lower hypothesis index wins the tie


def test_single_process_paths():
    sys.path.insert(0, ROOT)
    from instance_based_loc_b200 import dist as D
    assert D.world() == (0, 1)
    assert D.shard_indices(5) == [0, 1, 2, 3, 4]
    p = D.gather_frame_results([0, 1], [np.ones(7), np.zeros(7)], 2, device="cpu")
    assert p.shape == (2, 7) and p[0, 0] == 1 and p[1, 0] == 0
    assert D.best_of_ranks(0.1, 0.2, 3, np.eye(4), device="cpu")[2] == 3
