"""Host-side multi-rank logic on CPU: world_size-2 gloo ranks shard a ragged utterance list, "generate"
deterministic rows and gather them to rank 0 (the N>1 path of bench.py / infer.py without a GPU)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from zipvoice_b200 import parallel


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, lengths, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    r, w, _ = parallel.init_from_env("gloo")
    assert (r, w) == (rank, world)
    mine = parallel.shard_indices(lengths, world, rank)
    lens = torch.tensor([lengths[i] for i in mine], dtype=torch.int64)
    tmax = int(lens.max()) if len(mine) else 1
    x = torch.zeros(len(mine), tmax, 3)
    for row, i in enumerate(mine):
        x[row, : lengths[i]] = float(i) + torch.arange(lengths[i])[:, None] * 0.001
    parts = parallel.gather_ragged(x, lens, dst=0)
    if rank == 0:
        lists = [parallel.shard_indices(lengths, world, k) for k in range(world)]
        rows = parallel.scatter_back(parts, lists, len(lengths))
        ok = all(rows[i].shape == (lengths[i], 3) and float(rows[i][0, 0]) == float(i) and
                 abs(float(rows[i][-1, 2]) - (i + (lengths[i] - 1) * 0.001)) < 1e-4 for i in range(len(lengths)))
        q.put(ok)
    else:
        assert parts is None
    dist.destroy_process_group()


def test_shard_indices_partition_and_balance():
    lengths = [5, 100, 37, 64, 64, 3, 90, 12, 55]
    for world in (1, 2, 4, 8):
        parts = [parallel.shard_indices(lengths, world, r) for r in range(world)]
        flat = sorted(i for p in parts for i in p)
        assert flat == list(range(len(lengths)))
        sizes = [len(p) for p in parts]
        assert max(sizes) - min(sizes) <= 1
    two = [sum(lengths[i] for i in parallel.shard_indices(lengths, 2, r)) for r in range(2)]
    assert abs(two[0] - two[1]) <= max(lengths)


def test_gather_ragged_world2_gloo():
    lengths = [7, 19, 4, 11, 13]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, lengths, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=10) is True


def test_gather_ragged_with_an_empty_shard_gloo():
    """More ranks than utterances: one rank holds nothing and only takes part in the collectives (the empty-shard path of
    infer.generate_batch); rank 0 still gets every row back in the original order."""
    lengths = [9]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, lengths, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=10) is True
