"""Drop-in worker and pool for the per-generation population evaluation.

``B200PopulationWorker`` has the surface of the reference's
``EvolutionaryExperimentWorker`` (``evolutionary_experiment/logic/experiment_worker.py:22-99``):
``__init__(experiment_params)`` once, ``run(seranns_info, num_replications)`` per job, same
returned dict.  ``InProcessPool`` stands in for ``distributed_computing.api.Pool`` (third-party,
Redis; reference ``run_experiment.py:100-110``) with the four methods the drivers use:
context manager, ``get_workers_count``, ``imap_unordered``, ``update_workers``.

Nothing here computes on the CPU: every model-level number comes from the CUDA engine; the host
only emulates Keras' fp16 metric accumulators on the integer counts the kernels return.
"""
from __future__ import annotations

import os
import time
from typing import Dict, Iterable, List, Optional, Sequence

import numpy as np
import pandas as pd

from . import engine as _engine
from .frontend import InvalidSerann, SerannIR, compile_cached

VALIDATION_SPLIT = 0.05          # experiment_worker.py:118
EVALUATE_BATCH = 32              # Keras default batch size of Model.evaluate (experiment_worker.py:87)


class F16Mean:
    """Keras ``Mean`` metric whose ``total`` and ``count`` variables are fp16 (``floatx=float16``,
    experiment_worker.py:36) [TF2.6]: sums above 2048 lose integer resolution, which is part of the
    accuracies the reference stores."""

    def __init__(self):
        self.total = np.float16(0)
        self.count = np.float16(0)

    def update(self, value_sum, weight):
        with np.errstate(all='ignore'):
            self.total = np.float16(self.total + np.float16(value_sum))
            self.count = np.float16(self.count + np.float16(weight))

    def result(self) -> float:
        with np.errstate(all='ignore'):
            return float(np.float16(self.total / self.count)) if self.count != 0 else 0.0


def f16_accuracy(correct_counts: Sequence[int], batch_sizes: Sequence[int]) -> float:
    m = F16Mean()
    for c, b in zip(correct_counts, batch_sizes):
        m.update(float(c), b)
    return m.result()


def f16_loss_mean(loss_sums: Sequence[float], batch_sizes: Sequence[int]) -> float:
    """Per-output loss metric: ``Mean.update_state(batch_loss, sample_weight=batch_size)``."""
    m = F16Mean()
    with np.errstate(all='ignore'):
        for s, b in zip(loss_sums, batch_sizes):
            batch_loss = np.float16(np.float32(s) / np.float32(b))
            m.update(np.float16(batch_loss * np.float16(b)), b)
    return m.result()


def _chunks(total: int, size: int) -> List[int]:
    return [min(size, total - s) for s in range(0, total, size)]


def genotype_bits(genotype) -> np.ndarray:
    """Stored genotypes are ints for generation 0 and 0.0/1.0 floats afterwards (experiment.py:149,210)."""
    g = np.asarray(genotype, dtype=np.float32)
    return (g > 0.5).astype(np.uint8)


class B200PopulationWorker:
    """One GPU's worker.  ``dataset`` replaces the MNIST download and the encodings ``.npz`` the
    reference loads by path (experiment_worker.py:41,197-220): a dict with x_train, y_train, g_train,
    x_test, y_test, g_test (images already scaled to [0, 1])."""

    MODELS_INFO_COLUMNS = ['parameters_count', 'loss_balance', 'is_valid', 'is_overweight',
                           'classification_validation_accuracy', 'classification_training_accuracy',
                           'classification_test_accuracy', 'replication_mse']

    def __init__(self, experiment_params: dict, dataset: Optional[dict] = None, device: int = 0,
                 memory_budget_bytes: float = 120e9, shuffle_seed: Optional[int] = None, max_wave: int = 512):
        self._parameters = experiment_params
        self._device = device
        self._budget = memory_budget_bytes
        self._max_wave = int(max_wave)          # individuals that step together at most (launch tables grow with it)
        self._rng = np.random.default_rng(shuffle_seed if shuffle_seed is not None else int.from_bytes(os.urandom(8), 'little'))
        if dataset is None:
            raise ValueError('dataset is required: the MNIST download and the figshare encodings are not '
                             'available offline (pass arrays, e.g. serann_b200.synthetic.synthetic_dataset())')
        p = experiment_params
        self._batch = int(p['training_batch_size'])
        self._engine = _engine.Engine(image_dims=tuple(p['classification_image_dimensions']),
                                      genotype_size=int(p['genotype_size']),
                                      n_classes=int(p['num_classification_classes']),
                                      max_batch=self._batch, device=device)
        self._engine.set_dataset(dataset['x_train'], dataset['y_train'], dataset['g_train'],
                                 dataset['x_test'], dataset['y_test'], dataset['g_test'])
        self._n_train, self._n_test = len(dataset['x_train']), len(dataset['x_test'])
        self.last_stats: Dict[str, float] = {}
        print('Worker initialization is done.')

    def close(self):
        self._engine.close()

    # -- helpers
    def _compile(self, seranns_info: pd.DataFrame):
        p = self._parameters
        irs: Dict[str, SerannIR] = {}
        info = pd.DataFrame(index=seranns_info.index).assign(parameters_count=np.nan, loss_balance=np.nan)
        for serann_id, row in seranns_info.iterrows():
            try:
                # memoised by source string: a clonal population compiles one program, not N (row f2)
                ir = compile_cached(row['source_code'], image_dims=tuple(p['classification_image_dimensions']),
                                    genotype_size=int(p['genotype_size']), n_classes=int(p['num_classification_classes']))
                _engine.layer_descs(ir)
            except (InvalidSerann, _engine.EngineError):
                continue                                     # experiment_worker.py:238-239
            irs[serann_id] = ir
            info.at[serann_id, 'parameters_count'] = ir.parameters_count
            info.at[serann_id, 'loss_balance'] = ir.loss_balance
        return irs, info

    def _waves(self, ids: List[str], irs: Dict[str, SerannIR]) -> List[List[str]]:
        """Split the shard so that each wave's activation arena fits the memory budget (and holds at most max_wave)."""
        waves, cur, used = [], [], 0
        for i in ids:
            need = _engine.activation_bytes_estimate(irs[i], self._batch) + 16 * irs[i].parameters_count
            if cur and (used + need > self._budget or len(cur) >= self._max_wave):
                waves.append(cur)
                cur, used = [], 0
            cur.append(i)
            used += need
        if cur:
            waves.append(cur)
        return waves

    def run(self, seranns_info: pd.DataFrame, num_replications: int, fit_seed: Optional[int] = None,
            individual_seeds: Optional[Dict[str, int]] = None) -> dict:
        """Reference ``EvolutionaryExperimentWorker.run`` (experiment_worker.py:45-99).

        ``fit_seed`` fixes the row order of the epochs and ``individual_seeds`` (id -> seed) the initial weights of
        each individual, whatever shard it is in: with both given, a generation evaluated on 1, 2, 4 or 8 GPUs
        trains every individual from the same weights on the same mini-batches (Keras draws both from unseeded
        generators, experiment_worker.py:63-70,113-118).  Besides the reference's dict entries the result carries the
        readout kernel's outputs: ``offspring_bits_by_id`` {id: (R, G) uint8}, ``mutations_by_id`` {id: (R,) int32} and
        ``offspring_nan_by_id`` {id: (R,) bool} (rows holding a NaN probability)."""
        print('Processing ', len(seranns_info), 'SeRANNs')
        p = self._parameters
        t_c = time.perf_counter()
        irs, models_info = self._compile(seranns_info)
        self._compile_s = time.perf_counter() - t_c
        models_info['is_valid'] = ~models_info['loss_balance'].isna()
        models_info['is_overweight'] = False
        for col in ('classification_validation_accuracy', 'classification_training_accuracy',
                    'classification_test_accuracy', 'replication_mse'):
            models_info[col] = np.nan
        models_info['is_overweight'] = models_info['parameters_count'] > p['max_serann_parameters']
        trainable = list(models_info.index[models_info['is_valid'] & ~models_info['is_overweight']])
        if not trainable:
            return {'learning_time': 0, 'replication_time': 0, 'models_info': models_info, 'offspring_by_id': {}}

        epochs = int(p['training_epochs'])
        n_fit = int(self._n_train * (1.0 - VALIDATION_SPLIT))          # Keras: split_at = int(len * (1 - split))
        n_val = self._n_train - n_fit
        learning_time = replication_time = build_time = test_time = 0.0
        device_ms = 0.0
        offspring_by_id, bits_by_id, mut_by_id, nan_by_id = {}, {}, {}, {}
        eng = self._engine
        num_replications = int(num_replications)
        # position of each trainable individual in the reference's joined replicate batch (experiment_worker.py:139-150)
        order = {sid: k for k, sid in enumerate(trainable)}
        repeat = int(np.ceil(len(trainable) * max(num_replications, 1) / self._n_test))
        for wave in self._waves(trainable, irs):
            t_b = time.perf_counter()
            seeds = None if individual_seeds is None else [individual_seeds[i] for i in wave]
            eng.build([irs[i] for i in wave], seed=int(self._rng.integers(0, 2 ** 63)), individual_seeds=seeds)
            build_time += time.perf_counter() - t_b
            t0 = time.perf_counter()
            print('Learning started')
            shuffle = int(self._rng.integers(0, 2 ** 63)) if fit_seed is None else int(fit_seed)
            res = eng.fit(epochs, self._batch, n_val, shuffle_seed=shuffle)
            learning_time += time.perf_counter() - t0
            device_ms += res['device_ms']
            print('Learning done')
            print('Test set evaluation started')
            t_e = time.perf_counter()
            test_correct, _ = eng.evaluate(1, 0, self._n_test, EVALUATE_BATCH)
            test_time += time.perf_counter() - t_e
            step_sizes = _chunks(n_fit, self._batch)
            val_sizes = _chunks(n_val, self._batch)
            test_sizes = _chunks(self._n_test, EVALUATE_BATCH)
            for k, sid in enumerate(wave):
                # _collect_accuracies (experiment_worker.py:169-195): last epoch's history entries
                models_info.at[sid, 'classification_training_accuracy'] = f16_accuracy(res['train_correct'][-1, :, k], step_sizes)
                models_info.at[sid, 'is_valid'] = True
                if n_val:
                    models_info.at[sid, 'classification_validation_accuracy'] = f16_accuracy(res['val_correct'][-1, :, k], val_sizes)
                    models_info.at[sid, 'replication_mse'] = f16_loss_mean(res['val_mse_sum'][-1, :, k], val_sizes)
                models_info.at[sid, 'classification_test_accuracy'] = f16_accuracy(test_correct[:, k], test_sizes)
            if num_replications > 0:
                print('Num replications:', num_replications)
                t0 = time.perf_counter()
                parents = np.stack([genotype_bits(seranns_info.at[sid, 'genotype']) for sid in wave])
                rows = np.concatenate([np.arange(order[sid] * num_replications, (order[sid] + 1) * num_replications) for sid in wave])
                probs, packed, mut = eng.replicate(parents, num_replications, image_index=rows // repeat)
                bits = _engine.unpack_bits(packed, eng.G)
                nan_rows = np.isnan(probs.astype(np.float32)).any(axis=-1)
                for k, sid in enumerate(wave):
                    offspring_by_id[sid] = probs[k]
                    bits_by_id[sid], mut_by_id[sid], nan_by_id[sid] = bits[k], mut[k], nan_rows[k]
                replication_time += time.perf_counter() - t0
                print('Replication done')
        self.last_stats = {'device_ms_fit': device_ms, 'trained': len(trainable), 'steps': epochs * len(_chunks(n_fit, self._batch)),
                           'compile_s': self._compile_s, 'build_s': build_time, 'fit_s': learning_time, 'test_s': test_time,
                           'replicate_s': replication_time}
        out = {'learning_time': learning_time, 'models_info': models_info}
        if num_replications > 0:
            out.update({'replication_time': replication_time, 'offspring_by_id': offspring_by_id,
                        'offspring_bits_by_id': bits_by_id, 'mutations_by_id': mut_by_id, 'offspring_nan_by_id': nan_by_id})
        return out


class InProcessPool:
    """The four methods of ``distributed_computing.api.Pool`` the reference uses
    (``experiment.py:45,168,182``; ``serann_evaluation/logic.py:47,65,69``), without Redis: one worker
    object per visible GPU, jobs run in this process.  Extra positional arguments of the reference's
    constructor (password, address, port, timeouts; ``run_experiment.py:102-104``) are accepted and ignored.
    """

    accepts_seeds = True         # jobs may carry kwargs (fit_seed, individual_seeds): the reference's farm passes args only
    # Jobs need not be cut to the reference's 112 SeRANNs per GPU (evolutionary_experiment/config.py:8, sized for a 16 GB
    # V100): a worker splits its job into waves that fit its own memory budget, and larger waves fill a B200 better.
    max_seranns_per_job = None

    def __init__(self, worker_cls, init_args: Optional[dict] = None, *_pool_args, devices: Optional[Sequence[int]] = None,
                 **_pool_kwargs):
        self._worker_cls = worker_cls
        self._init = init_args or {}
        if devices is None:
            devices = [int(os.environ.get('LOCAL_RANK', 0))]
        self._devices = list(devices)
        self._workers = []

    def __enter__(self):
        args = list(self._init.get('args', []))
        kwargs = dict(self._init.get('kwargs', {}))
        for d in self._devices:
            self._workers.append(self._worker_cls(*args, device=d, **kwargs))
        return self

    def __exit__(self, *exc):
        for w in self._workers:
            if hasattr(w, 'close'):
                w.close()
        self._workers = []
        return False

    def get_workers_count(self) -> int:
        return len(self._devices)

    def imap_unordered(self, jobs: Iterable[dict]):
        """Jobs are dealt round-robin to the workers; each worker (one GPU) runs its jobs in its own host thread - the
        C ABI calls release the GIL, so the devices work at the same time - and results come back in completion order."""
        jobs = list(jobs)
        if len(self._workers) == 1 or len(jobs) <= 1:
            for k, job in enumerate(jobs):
                w = self._workers[k % len(self._workers)]
                yield w.run(*job.get('args', []), **job.get('kwargs', {}))
            return
        import queue
        import threading
        done: 'queue.Queue' = queue.Queue()

        def serve(widx):
            for k in range(widx, len(jobs), len(self._workers)):
                try:
                    done.put((None, self._workers[widx].run(*jobs[k].get('args', []), **jobs[k].get('kwargs', {}))))
                except BaseException as e:           # shipped to the caller like the farm ships worker exceptions
                    done.put((e, None))
        threads = [threading.Thread(target=serve, args=(i,), daemon=True) for i in range(len(self._workers))]
        for t in threads:
            t.start()
        for _ in jobs:
            err, res = done.get()
            if err is not None:
                raise err
            yield res
        for t in threads:
            t.join()

    def update_workers(self, data: dict):
        for w in self._workers:
            if hasattr(w, 'handle_update'):
                w.handle_update(*data.get('args', []), **data.get('kwargs', {}))
