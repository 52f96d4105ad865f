"""Multi-GPU plumbing: one process per GPU (torchrun), `torch.distributed` over NCCL (gloo on CPU
for tests).  Replaces the reference's `nn.DataParallel` wrap (`optimize_token.py:43-44`), which
re-broadcasts the 860 M-parameter UNet on every forward and gathers through cuda:0
(SURVEY.md §2.1 last row).

Two sharding modes:
  * extraction / evaluation / find_best_indices: images are independent -> `shard_range` /
    `shard_dataset`, no collective on the data path; results are gathered once at the end
    (`all_gather_cat`);
  * token optimisation: the image batch is sharded and the (1,T,ctx) context gradient
    (1.5-2 MB) is all-reduced once per optimiser step (`allreduce_mean_`).
Every helper degrades to a no-op when torch.distributed is not initialised (single GPU).
"""
from __future__ import annotations

import os
from typing import Tuple

import torch
import torch.distributed as dist


def is_dist() -> bool:
    return dist.is_available() and dist.is_initialized()


def world_size() -> int:
    return dist.get_world_size() if is_dist() else 1


def rank() -> int:
    return dist.get_rank() if is_dist() else 0


def init_from_env(backend: str | None = None) -> Tuple[int, int, int]:
    """Initialise from torchrun's env (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_*).
    -> (rank, local_rank, world_size).  No env -> single process, nothing initialised."""
    ws = int(os.environ.get("WORLD_SIZE", "1"))
    if ws <= 1:
        return 0, int(os.environ.get("LOCAL_RANK", "0")), 1
    r, lr = int(os.environ["RANK"]), int(os.environ.get("LOCAL_RANK", "0"))
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend == "nccl":
        torch.cuda.set_device(lr)
    if not is_dist():
        kw = {"device_id": torch.device("cuda", lr)} if backend == "nccl" else {}
        dist.init_process_group(backend=backend, rank=r, world_size=ws, **kw)
    return r, lr, ws


def shard_range(n: int, r: int | None = None, ws: int | None = None) -> range:
    """Indices of the n units this rank owns: strided, so every rank gets floor or ceil of n/ws."""
    r = rank() if r is None else r
    ws = world_size() if ws is None else ws
    return range(r, n, ws)


def shard_dataset(dataset, r: int | None = None, ws: int | None = None):
    r = rank() if r is None else r
    ws = world_size() if ws is None else ws
    if ws == 1:
        return dataset
    return torch.utils.data.Subset(dataset, list(shard_range(len(dataset), r, ws)))


def allreduce_mean_(t: torch.Tensor | None) -> None:
    """In-place mean over ranks (sum then divide: the reference averages the per-GPU losses,
    `optimize.py:428-429`, so the gradient is the mean of the per-rank gradients)."""
    if t is None or not is_dist():
        return
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    t.div_(world_size())


def broadcast_(t: torch.Tensor, src: int = 0) -> None:
    if is_dist():
        dist.broadcast(t, src=src)


def all_gather_cat(t: torch.Tensor) -> torch.Tensor:
    """Concatenate equal-shaped per-rank tensors along dim 0, rank-major."""
    if not is_dist():
        return t
    out = [torch.empty_like(t) for _ in range(world_size())]
    dist.all_gather(out, t.contiguous())
    return torch.cat(out, dim=0)


def interleave_shards(gathered: torch.Tensor, n: int) -> torch.Tensor:
    """Undo `shard_range`'s striding after `all_gather_cat` of per-rank (ceil(n/ws), ...) tensors
    (ranks with one unit fewer pad their last row): -> (n, ...) in the original unit order."""
    ws = world_size()
    per = gathered.shape[0] // ws
    return gathered.reshape(ws, per, *gathered.shape[1:]).transpose(0, 1).reshape(ws * per, *gathered.shape[1:])[:n]


def rank_generators(device, seed: int | None = None):
    """(CPU generator, generator on `device`) for this rank's random draws (augmentation thetas, DataLoader order; DDIM
    noise), seeded `base + rank`, or (None, None) = the global generators.

    The reference draws an independent affine and independent noise for every image of a batch
    (`invertable_transform.py:42-57`, `ptp_utils.py:219`).  With one process per GPU every process starts from the same
    default seeds, so without this all ranks would sample the SAME transform and noise each step and an N-GPU step would
    average N images that share them.  Single process with seed None: the global generators, i.e. a user's
    `torch.manual_seed` keeps its meaning."""
    if seed is None and world_size() == 1:
        return None, None
    base = torch.initial_seed() if seed is None else int(seed)
    s = (base + rank()) % (2 ** 63)
    g_cpu = torch.Generator().manual_seed(s)
    g_dev = torch.Generator(device=device).manual_seed(s) if torch.device(device).type == "cuda" else g_cpu
    return g_cpu, g_dev


def barrier() -> None:
    if is_dist():
        dist.barrier()
