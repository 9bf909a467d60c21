"""Data-parallel sharding of utterance lists over the GPUs of one box (one process per GPU, launched by
torchrun) and the NCCL gather of generated mels / waveforms.  The reference has no
inference-time parallelism (single cuda:0, bin/infer_zipvoice.py:568-569); utterances are independent, so
there is no collective inside the sampling loop -- only this gather.  Works with backend "gloo" on CPU
tensors too (tests/test_parallel_cpu.py)."""
import os
from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def init_from_env(backend: Optional[str] = None) -> Tuple[int, int, int]:
    """-> (rank, world, local_rank); initialises torch.distributed if WORLD_SIZE > 1."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        backend = backend or ("nccl" if torch.cuda.is_available() else "gloo")
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend)
    return rank, world, local


def shard_indices(lengths: Sequence[int], world: int, rank: int) -> List[int]:
    """Length-balanced static partition: utterances sorted by length (longest first) are dealt
    round-robin in a serpentine order, so every rank gets the same count (+-1) and a similar padded T."""
    order = sorted(range(len(lengths)), key=lambda i: (-int(lengths[i]), i))
    mine = []
    for pos, idx in enumerate(order):
        rnd, slot = divmod(pos, world)
        owner = slot if rnd % 2 == 0 else world - 1 - slot
        if owner == rank:
            mine.append(idx)
    return mine


def gather_ragged(x: torch.Tensor, lens: torch.Tensor, dst: int = 0):
    """x (B_r, T_r, F) or (B_r, N_r) padded rows with valid lengths lens (B_r,) on every rank ->
    on rank dst: list over ranks of (x_r, lens_r); None elsewhere.  Two collectives: an all_gather of the
    shapes, then one gather of the payload padded to the per-batch maximum."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    if world == 1:
        return [(x, lens)]
    rank = dist.get_rank()
    dev = x.device
    shape = torch.tensor([x.shape[0], x.shape[1]], dtype=torch.int64, device=dev)
    shapes = [torch.zeros_like(shape) for _ in range(world)]
    dist.all_gather(shapes, shape)
    bmax = int(max(s[0] for s in shapes))
    tmax = int(max(s[1] for s in shapes))
    pad_shape = (bmax, tmax) + tuple(x.shape[2:])
    xp = x.new_zeros(pad_shape)
    xp[: x.shape[0], : x.shape[1]] = x
    lp = lens.new_zeros((bmax,))
    lp[: lens.shape[0]] = lens
    outs = [torch.empty_like(xp) for _ in range(world)] if rank == dst else None
    louts = [torch.empty_like(lp) for _ in range(world)] if rank == dst else None
    dist.gather(xp, outs, dst=dst)
    dist.gather(lp, louts, dst=dst)
    if rank != dst:
        return None
    return [(o[: int(s[0]), : int(s[1])], l[: int(s[0])]) for o, l, s in zip(outs, louts, shapes)]


def scatter_back(parts, index_lists: Sequence[Sequence[int]], total: int):
    """Undo shard_indices on rank dst: parts[r] = (x_r, lens_r) in the order of index_lists[r]."""
    out = [None] * total
    for (x, lens), idxs in zip(parts, index_lists):
        for row, i in enumerate(idxs):
            out[i] = x[row, : int(lens[row])]
    return out
