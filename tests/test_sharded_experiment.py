"""The multi-GPU path of the evolution loop on CPU: world-size-2 gloo processes against a single process.

Every rank builds the same generation from the same seeded np.random stream, evaluates its longest-processing-time
shard, one all-gather brings all results to every rank, selection runs redundantly; rank 0 writes the DB.  The worker
here is a stand-in whose numbers depend only on (individual seed, fit seed) - exactly the contract the CUDA worker
honours - so the sharded run must reproduce the single-process run bit for bit.  (The CUDA engine has no CPU path; the
same comparison on real GPUs is tests/test_gpu_multi.py.)
"""
import os
import sqlite3

import numpy as np
import pandas as pd
import pytest

from serann_b200 import parallel
from serann_b200.experiment import Experiment, MODELS_INFO_COLUMNS
from serann_b200.experiment_db import ExperimentDB
from serann_b200.frontend import InvalidSerann, compile_cached
import parity_util as U

PARAMS = {'num_seranns': 10, 'num_generations': 2, 'max_tokens': 350, 'genotype_size': 100, 'selection_pressure': 1,
          'max_serann_parameters': 2e6, 'classification_image_dimensions': [28, 28], 'num_classification_classes': 10,
          'training_epochs': 1, 'training_batch_size': 40, 'offspring_selection_strategy': 'random',
          'error_correction_probability': 0.5, 'initial_offspring_pool_size': 6, 'offspring_pool_size_factor': 3}
SOURCES = [U.NET_A, U.NET_B, U.NET_C, 'con = oops(', U.NET_D, U.NET_E]


class StandInWorker:
    """Results are a pure function of (individual seed, fit seed, genotype): what makes a generation independent of how it is sharded."""

    def __init__(self, params, device=0):
        self._p = params
        self.calls = []

    def run(self, seranns_info, num_replications, fit_seed=None, individual_seeds=None):
        self.calls.append(list(seranns_info.index))
        info = pd.DataFrame(np.nan, index=seranns_info.index, columns=MODELS_INFO_COLUMNS, dtype=np.float64)
        out = {'offspring_by_id': {}, 'offspring_bits_by_id': {}, 'mutations_by_id': {}, 'offspring_nan_by_id': {}}
        for sid, row in seranns_info.iterrows():
            try:
                ir = compile_cached(row['source_code'])
            except InvalidSerann:
                info.loc[sid] = [np.nan, np.nan, 0.0, 0.0, np.nan, np.nan, np.nan, np.nan]
                continue
            r = np.random.default_rng([individual_seeds[sid] % (2 ** 32), fit_seed % (2 ** 32)])
            acc = r.random(3).astype(np.float16).astype(float)
            info.loc[sid] = [float(ir.parameters_count), ir.loss_balance, 1.0, 0.0, acc[0], acc[1], acc[2], float(np.float16(r.random()))]
            parent = (np.asarray(row['genotype'], dtype=np.float32) > 0.5).astype(np.uint8)
            flips = r.random((num_replications, 100)) < 0.03
            bits = np.where(flips, 1 - parent, parent).astype(np.uint8)
            out['offspring_bits_by_id'][sid] = bits
            out['offspring_by_id'][sid] = bits.astype(np.float16)
            out['mutations_by_id'][sid] = (bits != parent).sum(axis=1).astype(np.int32)
            out['offspring_nan_by_id'][sid] = np.zeros(num_replications, dtype=bool)
        info['is_valid'] = info['is_valid'].astype(bool)
        info['is_overweight'] = info['is_overweight'].astype(bool)
        out.update({'learning_time': 0.25, 'replication_time': 0.125, 'models_info': info})
        return out


class StandInDecoder:
    def decode_to_string(self, genotypes):
        return ['decoded'] * len(genotypes)


def run_experiment(db_path, ctx):
    from serann_b200.worker import InProcessPool
    np.random.seed(79375)
    state = {'k': 0}

    def source(n):
        out = [SOURCES[(state['k'] + i) % len(SOURCES)] for i in range(n)]
        state['k'] += 1
        return out
    enc = np.random.default_rng(1).integers(0, 2, (200, 100))
    db = ExperimentDB(db_path, read_only=ctx.rank != 0)
    with InProcessPool(StandInWorker, {'args': [dict(PARAMS)]}, devices=[0]) as pool:
        exp = Experiment('exp', enc, pool, db, dict(PARAMS), genetic_encoder=StandInDecoder(), population_source=source,
                         shard_context=ctx, seed_base=5)
        last = exp.execute()
        calls = pool._workers[0].calls if pool._workers else []
    return last, calls, getattr(exp, 'last_shards', None)


def _rank_main(rank, world, port, tmp, q):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import sys
    sys.stdout = open(os.devnull, 'w')
    ctx = parallel.ShardContext.from_env(backend='gloo')
    last, calls, shards = run_experiment(os.path.join(tmp, 'w2.sqlite'), ctx)
    q.put((rank, last.reset_index().to_json(), calls, shards))
    import torch.distributed as dist
    dist.destroy_process_group()


def _tables(path):
    con = sqlite3.connect(path)
    out = {}
    for t in ('generations', 'serann'):
        df = pd.read_sql(f'select * from {t}', con)
        out[t] = df.drop(columns=[c for c in df.columns if 'time' in c])      # wall-clock columns differ by nature
    con.close()
    return out


def test_world2_generation_equals_world1(tmp_path, capsys):
    import torch.multiprocessing as mp
    single, calls1, _ = run_experiment(str(tmp_path / 'w1.sqlite'), parallel.ShardContext())
    mpctx = mp.get_context('spawn')
    q = mpctx.Queue()
    port = 29600 + os.getpid() % 1000
    procs = [mpctx.Process(target=_rank_main, args=(r, 2, port, str(tmp_path), q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=180) for _ in range(2)])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # both ranks end with the same next generation as the single process: same ids, parents, genotypes, sources
    want = single.reset_index().to_json()
    assert res[0][1] == want and res[1][1] == want
    # each rank trained half of every generation, and the halves are the LPT shards
    for rank, _, calls, shards in res:
        assert len(calls) == 2 and all(len(c) == 5 for c in calls)
    assert sorted(res[0][2][0] + res[1][2][0]) == sorted(calls1[0])
    # rank 0 wrote the same tables as the single process (rank 1 wrote nothing)
    a, b = _tables(str(tmp_path / 'w1.sqlite')), _tables(str(tmp_path / 'w2.sqlite'))
    for t in a:
        pd.testing.assert_frame_equal(a[t], b[t])


def test_pack_unpack_shard_results_roundtrip():
    ids = ['a', 'b', 'c']
    info = pd.DataFrame({c: [1.5, np.nan, 3.0] for c in MODELS_INFO_COLUMNS}, index=ids)
    info['is_valid'] = [True, False, True]
    r = np.random.default_rng(0)
    bits = {i: r.integers(0, 2, (4, 100)).astype(np.uint8) for i in ('a', 'c')}
    res = {'models_info': info, 'learning_time': 1.5, 'replication_time': 0.5, 'offspring_bits_by_id': bits,
           'mutations_by_id': {i: np.arange(4, dtype=np.int32) for i in bits},
           'offspring_nan_by_id': {i: np.array([0, 1, 0, 0], bool) for i in bits}}
    blob = parallel.pack_shard_results(ids, res, MODELS_INFO_COLUMNS, 4, 100)
    t, inf, has, packed, mut, nan = parallel.unpack_shard_results(blob, len(MODELS_INFO_COLUMNS), 100)
    assert t.tolist() == [1.5, 0.5] and has.tolist() == [1, 0, 1]
    from serann_b200.engine import unpack_bits
    assert np.array_equal(unpack_bits(packed, 100)[0], bits['a']) and np.array_equal(unpack_bits(packed, 100)[2], bits['c'])
    assert mut[2].tolist() == [0, 1, 2, 3] and nan[0].tolist() == [0, 1, 0, 0]
    assert np.isnan(inf[1, 0]) and inf[0, 2] == 1.0


def test_stable_seed_is_process_independent():
    assert parallel.stable_seed(1, 'x', 2) == parallel.stable_seed(1, 'x', 2)
    assert parallel.stable_seed(1, 'x', 2) != parallel.stable_seed(1, 'x', 3)
    assert 0 <= parallel.stable_seed('a') < 2 ** 64
