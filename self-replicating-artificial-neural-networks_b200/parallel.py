"""Multi-GPU plumbing: shard the population, one exchange per generation.

The reference farms shards of a generation out to one worker process per GPU through a Redis task
queue (``experiment.py:162-190``; third-party ``distributed-computing``).  Individuals are independent
given the shared data set, so here each rank (one process per GPU) trains its own shard with no
data-path collective, and the only exchange is an all-gather of the per-individual fitness and the
thresholded offspring bits, which selection needs on every rank (``experiment.py:192-238``).
"""
from __future__ import annotations

import hashlib
import time
from typing import Dict, List, Optional, Sequence, Tuple

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


# ---------------------------------------------------------------------------------------------------------------------
# One generation over several ranks (one process per GPU, torchrun): every rank holds the same generation table (the
# driver's legacy np.random stream is seeded identically everywhere, run_experiment.py:55), trains its cost-balanced
# shard, and one all-gather brings every rank the whole generation's results, so that selection - which draws from that
# stream - runs redundantly and stays aligned.  No gradient or activation ever crosses NVLink.

def stable_seed(*parts) -> int:
    """64-bit seed from arbitrary parts, the same on every rank and every run (not Python's salted hash)."""
    h = hashlib.blake2b('/'.join(str(p) for p in parts).encode(), digest_size=8)
    return int.from_bytes(h.digest(), 'little')


def individual_cost(ir, batch: int = 750) -> float:
    """Relative cost of training one individual: its GEMM work plus its activation traffic weighed at the machine's
    FLOP-per-byte balance for these small-N problems (about 40: the step moves ~17 GB for ~0.7 TFLOP)."""
    act_bytes = 0
    for l in ir.layers:
        c = l.out_shape[-1]
        act_bytes += 2 * (l.out_elems // c) * (-(-c // 8) * 8)
    return float(ir.train_flops_per_sample() + 40.0 * 6.0 * act_bytes)


class ShardContext:
    """Rank / world of this process and the exchange step.  ``backend`` nccl on GPUs, gloo in the CPU tests."""

    def __init__(self, rank: int = 0, world: int = 1, device=None, group=None):
        self.rank, self.world, self.device, self.group = rank, world, device, group
        self.last_allgather_ms = 0.0

    @classmethod
    def from_env(cls, backend: Optional[str] = None):
        import os
        world = int(os.environ.get('WORLD_SIZE', 1))
        if world == 1:
            return cls()
        import torch
        import torch.distributed as dist
        local = int(os.environ.get('LOCAL_RANK', 0))
        backend = backend or ('nccl' if torch.cuda.is_available() else 'gloo')
        dev = torch.device('cuda', local) if backend == 'nccl' else torch.device('cpu')
        if not dist.is_initialized():
            if backend == 'nccl':
                torch.cuda.set_device(local)
                dist.init_process_group('nccl', device_id=dev)
            else:
                dist.init_process_group('gloo')
        return cls(dist.get_rank(), dist.get_world_size(), dev)

    def allgather_bytes(self, payload: bytes) -> List[bytes]:
        """Variable-length byte strings from every rank, in rank order: two collectives (lengths, padded data)."""
        if self.world == 1:
            return [payload]
        import torch
        import torch.distributed as dist
        t0 = time.perf_counter()
        dev = self.device
        n = torch.tensor([len(payload)], dtype=torch.int64, device=dev)
        sizes = torch.zeros(self.world, dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(sizes, n, group=self.group)
        sizes = sizes.cpu().numpy()
        cap = int(sizes.max())
        mine = torch.zeros(cap, dtype=torch.uint8, device=dev)
        if payload:
            mine[:len(payload)] = torch.frombuffer(bytearray(payload), dtype=torch.uint8).to(dev)
        everything = torch.empty(self.world * cap, dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(everything, mine, group=self.group)
        flat = everything.cpu().numpy()
        self.last_allgather_ms = 1000.0 * (time.perf_counter() - t0)
        return [flat[r * cap: r * cap + int(sizes[r])].tobytes() for r in range(self.world)]

    def barrier(self):
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier(group=self.group)


def pack_shard_results(ids: Sequence[str], result: dict, columns: Sequence[str], R: int, G: int) -> bytes:
    """One rank's results as bytes: per individual the ``models_info`` columns (float64), a has-offspring flag, and
    per offspring the bit-packed genotype (readout kernel layout), its mutation count and its NaN flag."""
    words = (G + 31) // 32
    n = len(ids)
    info = np.full((n, len(columns)), np.nan, dtype=np.float64)
    mi = result['models_info']
    for j, c in enumerate(columns):
        info[:, j] = [float(mi.at[i, c]) if i in mi.index else np.nan for i in ids]
    has = np.zeros(n, dtype=np.uint8)
    packed = np.zeros((n, R, words), dtype=np.uint32)
    mut = np.zeros((n, R), dtype=np.int32)
    nan = np.zeros((n, R), dtype=np.uint8)
    bits_by_id = result.get('offspring_bits_by_id', {})
    for k, i in enumerate(ids):
        if i in bits_by_id:
            has[k] = 1
            packed[k] = pack_bits(bits_by_id[i])
            mut[k] = result['mutations_by_id'][i]
            nan[k] = result['offspring_nan_by_id'][i]
    times = np.array([result.get('learning_time', 0.0), result.get('replication_time', 0.0)], dtype=np.float64)
    return b''.join(a.tobytes() for a in (np.array([n, R], dtype=np.int64), times, info, has, packed, mut, nan))


def unpack_shard_results(blob: bytes, n_columns: int, G: int):
    words = (G + 31) // 32
    n, R = np.frombuffer(blob, dtype=np.int64, count=2)
    n, R = int(n), int(R)
    off = 16
    def take(dtype, shape):
        nonlocal off
        cnt = int(np.prod(shape)) if shape else 1
        a = np.frombuffer(blob, dtype=dtype, count=cnt, offset=off).reshape(shape)
        off += cnt * np.dtype(dtype).itemsize
        return a
    times = take(np.float64, (2,))
    info = take(np.float64, (n, n_columns))
    has = take(np.uint8, (n,))
    packed = take(np.uint32, (n, R, words))
    mut = take(np.int32, (n, R))
    nan = take(np.uint8, (n, R))
    return times, info, has, packed, mut, nan
