"""Multi-GPU plumbing: shard the population, one exchange per generation.

The reference farms shards of a generation out to one worker process per GPU through a Redis task
queue (``experiment.py:162-190``; third-party ``distributed-computing``).  Individuals are independent
given the shared data set, so here each rank (one process per GPU) trains its own shard with no
data-path collective, and the only exchange is an all-gather of the per-individual fitness and the
thresholded offspring bits, which selection needs on every rank (``experiment.py:192-238``).
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np


def lpt_shards(costs: Sequence[float], world: int) -> List[List[int]]:
    """Longest-processing-time-first partition with equal shard sizes (+-1); indices sorted per shard."""
    n = len(costs)
    cap = -(-n // world)
    loads = [0.0] * world
    bins: List[List[int]] = [[] for _ in range(world)]
    for i in sorted(range(n), key=lambda i: -costs[i]):
        r = min((k for k in range(world) if len(bins[k]) < cap), key=lambda k: loads[k])
        bins[r].append(i)
        loads[r] += costs[i]
    return [sorted(b) for b in bins]


def reference_batches(n: int, workers: int, max_per_gpu: int = 112) -> np.ndarray:
    """The reference's shard rule (``experiment.py:168-175``): ``i % workers`` while a shard stays under
    ``max_serann_per_gpu``, else ``i % ceil(n / max_per_gpu)``."""
    if n / workers < max_per_gpu:
        return np.arange(n) % workers
    return (np.arange(n) % int(np.ceil(n / max_per_gpu))).astype(int)


def pack_bits(bits: np.ndarray) -> np.ndarray:
    """(..., G) 0/1 -> (..., ceil(G/32)) uint32, bit j of word j // 32 (the layout of the readout kernel)."""
    b = np.asarray(bits).astype(np.uint32)
    G = b.shape[-1]
    pad = (-G) % 32
    if pad:
        b = np.concatenate([b, np.zeros(b.shape[:-1] + (pad,), np.uint32)], axis=-1)
    b = b.reshape(b.shape[:-1] + (-1, 32))
    return (b << np.arange(32, dtype=np.uint32)).sum(axis=-1, dtype=np.uint64).astype(np.uint32)


def allgather_generation(fitness: np.ndarray, offspring_bits: np.ndarray, device=None, group=None) -> Tuple[np.ndarray, np.ndarray]:
    """All-gather per-individual fitness (fp32) and bit-packed offspring (uint32 words) across ranks.

    Shards may differ in size by one individual; they are padded to the largest and trimmed after.
    On GPUs this is one NCCL all-gather over NVLink per array (payload <= ~2 MB at paper scale); with the
    gloo backend (CPU tests) the same code runs on host tensors.  Returns rank-ordered concatenations.
    """
    import torch
    import torch.distributed as dist
    fitness = np.ascontiguousarray(np.asarray(fitness, dtype=np.float32))
    packed = pack_bits(offspring_bits)
    n_local = len(fitness)
    world = dist.get_world_size(group)
    dev = device if device is not None else torch.device('cpu')
    counts = torch.zeros(world, dtype=torch.int64, device=dev)
    mine = torch.tensor([n_local], dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(counts, mine, group=group)
    n_max = int(counts.max().item())
    row = int(np.prod(packed.shape[1:])) if packed.ndim > 1 else 1
    f = torch.zeros(n_max, dtype=torch.float32, device=dev)
    f[:n_local] = torch.from_numpy(fitness).to(dev)
    p = torch.zeros(n_max * row, dtype=torch.int32, device=dev)
    p[:n_local * row] = torch.from_numpy(packed.reshape(-1).view(np.int32)).to(dev)
    f_all = torch.empty(world * n_max, dtype=torch.float32, device=dev)
    p_all = torch.empty(world * n_max * row, dtype=torch.int32, device=dev)
    dist.all_gather_into_tensor(f_all, f, group=group)
    dist.all_gather_into_tensor(p_all, p, group=group)
    f_np = f_all.cpu().numpy().reshape(world, n_max)
    p_np = p_all.cpu().numpy().view(np.uint32).reshape((world, n_max) + packed.shape[1:])
    cnt = counts.cpu().numpy()
    return (np.concatenate([f_np[r, :cnt[r]] for r in range(world)]),
            np.concatenate([p_np[r, :cnt[r]] for r in range(world)]))
