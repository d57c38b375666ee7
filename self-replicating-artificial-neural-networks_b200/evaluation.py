"""SeRANN deep evaluation on the B200 engine: fertility and mutation-rate estimation of one genotype.

Mirrors ``SerannEvaluator`` / ``SerannEvaluationWorker`` of the reference
(``serann_evaluation/logic.py:90-272``): one genotype is decoded to its source, ``num_evaluations``
independently initialised copies of that architecture are trained together (here: one clonal shard of
the population engine), each copy emits ``replications_per_evaluation`` offspring from the first R
test images, and the offspring are thresholded, proof-read, decoded and checked for validity.
"""
from __future__ import annotations

import re
from collections import Counter
from typing import Dict, List, Optional

import numpy as np

from . import engine as _engine
import copy

from .frontend import InvalidSerann, compile_cached, is_valid_serann
from .worker import VALIDATION_SPLIT, _chunks, f16_accuracy, genotype_bits

_LOSS_BALANCE_RE = re.compile(r'loss_balance *= *([0-9]+\.[0-9]+?)(\s|$)')    # serann_evaluation/logic.py:165


def probabilistic_proofreading(parent_genotype, offspring_genotypes, ec_factor):
    """Revert ``int(ec_factor * #mutated loci)`` randomly chosen mutated loci to the parent's allele
    (reference ``common/logic.py:6-14``; draws from the legacy ``np.random`` stream like the reference)."""
    out = np.array(offspring_genotypes, copy=True)
    parent = np.asarray(parent_genotype)
    rows, cols = np.where(out != parent)
    picked = np.random.permutation(len(rows))[:int(ec_factor * len(rows))]
    out[rows[picked], cols[picked]] = parent[cols[picked]]
    return out


def empty_result():
    return {'classification_accuracy': np.nan, 'offspring_survival': {}, 'mutation_rate': {}}


class B200SerannEvaluator:
    def __init__(self, experiment_params: dict, dataset: dict, num_evaluations: int = 50,
                 replications_per_evaluation: int = 100, ribosomal_ae=None, device: int = 0,
                 shuffle_seed: Optional[int] = None, nan_offspring: str = 'keep'):
        if nan_offspring not in ('keep', 'zero'):
            raise ValueError("nan_offspring must be 'keep' or 'zero'")
        # 'keep': a network whose outputs are NaN (BatchNormalizationF16 overflow) yields NaN offspring, which the decode
        # refuses exactly as the reference's does (logic.py:199-207 -> ribosomal_autoencoder.py:123); 'zero': NaN loci read as 0
        self._nan_offspring = nan_offspring
        self._params = experiment_params
        self._k = int(num_evaluations)
        self._r = int(replications_per_evaluation)
        self._ae = ribosomal_ae
        self._validity_cache: Dict[str, int] = {}
        self._rng = np.random.default_rng(shuffle_seed)
        p = experiment_params
        self._batch = int(p['training_batch_size'])
        self._engine = _engine.Engine(image_dims=tuple(p['classification_image_dimensions']),
                                      genotype_size=int(p['genotype_size']), n_classes=int(p['num_classification_classes']),
                                      max_batch=self._batch, device=device)
        self._engine.set_dataset(dataset['x_train'], dataset['y_train'], dataset['g_train'],
                                 dataset['x_test'], dataset.get('y_test'), dataset['g_test'])
        self._n_train = len(dataset['x_train'])
        if self._r > len(dataset['x_test']):
            raise ValueError('replications_per_evaluation exceeds the number of test images')

    def close(self):
        self._engine.close()

    def _kwargs(self):
        p = self._params
        return dict(image_dims=tuple(p['classification_image_dimensions']), genotype_size=int(p['genotype_size']),
                    n_classes=int(p['num_classification_classes']))

    def is_valid_serann(self, source_code: str, loss_balance=None) -> bool:
        """``ast.parse`` + a model build (reference ``logic.py:139-151``)."""
        return is_valid_serann(source_code, **self._kwargs())

    def evaluate(self, genotype, source_code: Optional[str] = None, loss_balance: Optional[float] = None,
                 initial_weights=None, perms=None) -> dict:
        """Reference ``SerannEvaluator.evaluate`` (``logic.py:153-231``).

        ``initial_weights`` (one Keras-layout weight list per copy) and ``perms`` (row order per epoch) replace the
        random initialisation and Keras' unseeded shuffle when given - the parity tests feed the same values to the
        CPU oracle."""
        return self.evaluate_many([genotype], [source_code], [loss_balance], initial_weights=initial_weights, perms=perms)[0]

    def evaluate_many(self, genotypes, source_codes=None, loss_balances=None, initial_weights=None, perms=None) -> list:
        """Several genotypes in ONE engine shard: the K copies of every genotype are built, trained and read out together
        (a shard of 3 x 32 copies fills the GPU like a population of 100; one genotype's 32 copies alone do not).  Each
        genotype's result is what ``evaluate`` returns for it; copies of different genotypes never interact."""
        n = len(genotypes)
        source_codes = list(source_codes) if source_codes is not None else [None] * n
        loss_balances = list(loss_balances) if loss_balances is not None else [None] * n
        results: list = [None] * n
        items = []                                   # (position, genotype, ir, loss_balance) of the valid ones
        for pos in range(n):
            genotype = np.array(genotypes[pos])
            source_code, loss_balance = source_codes[pos], loss_balances[pos]
            if source_code is None:
                if self._ae is None:
                    raise ValueError('a ribosomal autoencoder is needed to decode the genotype')
                source_code = self._ae.decode_to_string(genotype[None])[0]
            if loss_balance is None:
                m = _LOSS_BALANCE_RE.search(source_code)
                if m is None:
                    print('Invalid/overweight SeRANN received. Returning empty result.')
                    results[pos] = empty_result()
                    continue
                loss_balance = float(m.group(1))
            try:
                ir = copy.copy(compile_cached(source_code, **self._kwargs()))     # the cached IR is shared: do not edit it
            except InvalidSerann:
                # the reference lets the Keras build exception escape here; an invalid source has no offspring either way
                results[pos] = empty_result()
                continue
            ir.loss_balance = float(loss_balance)
            items.append((pos, genotype, ir, float(loss_balance)))
        if not items:
            return results
        k, r, eng = self._k, self._r, self._engine
        eng.build([ir for _, _, ir, _ in items for _ in range(k)], seed=int(self._rng.integers(0, 2 ** 63)))
        n_fit = int(self._n_train * (1.0 - VALIDATION_SPLIT))
        n_val = self._n_train - n_fit
        if initial_weights is not None:
            for i, w in enumerate(initial_weights):
                eng.set_weights(i, w)
        res = eng.fit(int(self._params['training_epochs']), self._batch, n_val, perms=perms,
                      shuffle_seed=int(self._rng.integers(0, 2 ** 63)))
        val_sizes = _chunks(n_val, self._batch)
        # predict([x_test[:R], genotype repeated R times]) for every copy (logic.py:188-191)
        parents = np.concatenate([np.repeat(genotype_bits(g)[None], k, axis=0) for _, g, _, _ in items], axis=0)
        probs, _, _ = eng.replicate(parents, r, image_index=np.tile(np.arange(r), k * len(items)))
        probs = probs.reshape(len(items), k * r, -1)
        self.last_probs = probs[0].reshape(k, r, -1) if len(items) == 1 else probs
        if self._ae is None:
            raise ValueError('a ribosomal autoencoder is needed to decode the offspring')
        for gi, (pos, genotype, _, loss_balance) in enumerate(items):
            accuracies = [f16_accuracy(res['val_correct'][-1, :, gi * k + i], val_sizes) for i in range(k)]
            offspring = np.round(np.clip(probs[gi].astype(np.float32), 0, 1))
            if self._nan_offspring == 'zero':
                offspring = np.nan_to_num(offspring, nan=0.0)
            fixed = probabilistic_proofreading(genotype, offspring, self._params['error_correction_probability'])
            sources = self._ae.decode_to_string(fixed)
            survival, mutation = [], []
            for i in range(k):
                micro: Dict[int, int] = {}
                for j in range(r):
                    s = sources[i * r + j]
                    # the reference's `cache.get(s) or ...` re-validates every cached invalid source (logic.py:216);
                    # a proper lookup gives the same counts without the rebuilds
                    valid = self._validity_cache.get(s)
                    if valid is None:
                        valid = self._validity_cache[s] = int(self.is_valid_serann(s, loss_balance))
                    micro[valid] = micro.get(valid, 0) + 1
                survival.append(micro)
                # the reference slices [i : i + R], not [i*R : (i+1)*R] (logic.py:223) - kept for parity
                mutations = np.sum(fixed[i: i + r] != genotype, axis=1)
                mutation.append(dict(Counter(mutations / len(genotype))))
            results[pos] = {'classification_accuracy': accuracies, 'mutation_rate': mutation, 'offspring_survival': survival}
        return results


class B200EvaluationWorker(B200SerannEvaluator):
    """Pool worker surface of ``SerannEvaluationWorker`` (``logic.py:234-272``): ``run(serann_id, serann_info)``
    with a genotype-hex result cache that ``handle_update`` extends."""

    def __init__(self, *args, evaluation_cache: bool = True, **kwargs):
        super().__init__(*args, **kwargs)
        self._evaluations_cache = {} if evaluation_cache is True else None
        print('Evaluator initialization is done.')

    def run(self, serann_id, serann_info):
        key = serann_info['genotype_hex']
        if self._evaluations_cache is not None and key in self._evaluations_cache:
            result = self._evaluations_cache[key]
        elif not serann_info['is_valid'] or serann_info['is_overweight']:
            result = empty_result()
            print('Invalid/overweight SeRANN received. Returning empty result.')
        else:
            print('Evaluating SeRANN. ', end='')
            result = self.evaluate(serann_info['genotype'], source_code=serann_info.get('source_code')
                                   if self._ae is None else None)
            print('Done.')
        if self._evaluations_cache is not None:
            self._evaluations_cache[key] = result
        return serann_id, result

    def handle_update(self, update_data):
        if self._evaluations_cache is not None:
            self._evaluations_cache.update(update_data)
