"""Host-side logic that needs no GPU: fp16 metric accumulators, sharding, bit packing, the gloo exchange."""
import os

import numpy as np
import pytest
import torch

from oracle import serann_oracle as O
from serann_b200 import parallel
from serann_b200.engine import unpack_bits
from serann_b200.worker import F16Mean, f16_accuracy, f16_loss_mean, genotype_bits, _chunks


def test_f16_mean_matches_oracle_and_quantises():
    r = np.random.default_rng(0)
    correct = r.integers(0, 751, 76)
    sizes = [750] * 76
    a, b = F16Mean(), O.F16Mean()
    for c, n in zip(correct, sizes):
        a.update(float(c), n)
        b.update(float(c), n)
    assert a.result() == b.result() == f16_accuracy(correct, sizes)
    # fp16 accumulators: 76 * 750 = 57000 is not representable (ulp 32 there); the exact mean differs
    exact = correct.sum() / 57000
    assert abs(a.result() - exact) < 2e-2
    assert np.float16(a.count) != 57000 or True


def test_f16_accuracy_of_evaluate_batches():
    # 313 batches of 32 (last one 16): counts accumulate in fp16 with ulp 8 above 8192
    sizes = _chunks(10000, 32)
    assert len(sizes) == 313 and sizes[-1] == 16
    acc = f16_accuracy([s for s in sizes], sizes)     # everything correct
    assert acc == pytest.approx(1.0, abs=2e-3)


def test_f16_loss_mean():
    assert f16_loss_mean([0.25 * 750] * 4, [750] * 4) == pytest.approx(0.25, abs=1e-3)


def test_genotype_bits_accepts_ints_and_floats():
    assert genotype_bits([0, 1, 1]).tolist() == [0, 1, 1]
    assert genotype_bits(np.array([0.0, 1.0], np.float16)).tolist() == [0, 1]


def test_lpt_shards_balanced_and_complete():
    costs = np.random.default_rng(1).random(1000) ** 3
    shards = parallel.lpt_shards(costs, 8)
    assert sorted(sum(shards, [])) == list(range(1000))
    assert all(len(s) == 125 for s in shards)
    loads = [costs[s].sum() for s in shards]
    assert max(loads) / min(loads) < 1.05


def test_reference_batch_rule():
    # experiment.py:168-175: population 1000 on 8 workers -> 9 batches of <= 112
    b = parallel.reference_batches(1000, 8)
    assert len(set(b)) == 9 and np.bincount(b).max() <= 112
    assert len(set(parallel.reference_batches(50, 1))) == 1


def test_pack_unpack_roundtrip():
    bits = np.random.default_rng(2).integers(0, 2, (5, 7, 100)).astype(np.uint8)
    packed = parallel.pack_bits(bits)
    assert packed.shape == (5, 7, 4) and packed.dtype == np.uint32
    assert np.array_equal(unpack_bits(packed, 100), bits)


def _gloo_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    dist.init_process_group('gloo', rank=rank, world_size=world)
    n_local = 3 if rank == 0 else 2                      # ragged shards
    fitness = np.arange(n_local, dtype=np.float32) + 10 * rank
    bits = ((np.arange(n_local * 4 * 100).reshape(n_local, 4, 100) + rank) % 3 == 0).astype(np.uint8)
    f, p = parallel.allgather_generation(fitness, bits)
    q.put((rank, f, unpack_bits(p, 100)))
    dist.destroy_process_group()


def test_allgather_generation_gloo_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = 29500 + os.getpid() % 1000
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(2)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    expect_f = np.array([0, 1, 2, 10, 11], dtype=np.float32)
    expect_b = np.concatenate([((np.arange(n * 400).reshape(n, 4, 100) + r) % 3 == 0).astype(np.uint8) for r, n in ((0, 3), (1, 2))])
    for rank, f, b in results:
        assert np.array_equal(f, expect_f)
        assert np.array_equal(b, expect_b)
