"""Evolution loop around the B200 worker: one generation = evaluate -> fertility -> offspring counts ->
statistics -> store -> select -> decode.

Host-side restatement of the reference driver ``Experiment``
(``evolutionary_experiment/logic/experiment.py``) on current pandas / NumPy (the reference uses
``DataFrame.append``, ``Series.iteritems`` and ``np.NaN``, which are gone).  The random draws keep the
reference's order on the legacy ``np.random`` stream seeded by the CLI (``run_experiment.py:55``):
first-generation sampling, one multinomial per generation, then per parent one permutation for the
offspring choice and one for proof-reading.
"""
from __future__ import annotations

from datetime import datetime
from typing import Dict, Optional
import random
from uuid import UUID

import numpy as np
import pandas as pd

from .engine import levenshtein, unpack_bits
from .frontend import InvalidSerann, compile_cached, extract_layer_counts
from .parallel import (ShardContext, individual_cost, lpt_shards, pack_shard_results, reference_batches, stable_seed,
                       unpack_shard_results)

MODELS_INFO_COLUMNS = ['parameters_count', 'loss_balance', 'is_valid', 'is_overweight',
                       'classification_validation_accuracy', 'classification_training_accuracy',
                       'classification_test_accuracy', 'replication_mse']

MAX_SERANN_PER_GPU = 112        # reference evolutionary_experiment/config.py:8


def hamming(a, b, *args, **kwargs):
    return np.not_equal(a, b).sum(axis=-1) / np.shape(a)[-1]


def pairwise_genotype_distances(genotypes: np.ndarray):
    """Mean pairwise Euclidean / Hamming distance and the cdist Hamming sum ("nucleotide diversity").

    The reference calls ``pdist`` / ``cdist`` with a Python metric, ~1.5 M interpreter calls per generation at
    population 1000 (``experiment.py:246-251``).  For 0/1 genotypes both metrics are functions of the number
    of differing loci, which one Gram matrix gives for all pairs at once; non-binary or NaN genotypes fall
    back to broadcasting.
    """
    g = np.asarray(genotypes, dtype=np.float64)
    n, L = g.shape
    if n < 2:
        return float('nan'), float('nan'), 0.0
    if np.isin(g, (0.0, 1.0)).all():
        ones = g.sum(axis=1)
        diff = ones[:, None] + ones[None, :] - 2.0 * (g @ g.T)          # differing loci per pair
        iu = np.triu_indices(n, 1)
        return float(np.sqrt(diff[iu]).mean()), float((diff[iu] / L).mean()), float(diff.sum() / L)
    d2 = ((g[:, None, :] - g[None, :, :]) ** 2).sum(-1)
    ne = (g[:, None, :] != g[None, :, :]).sum(-1) / L
    iu = np.triu_indices(n, 1)
    return float(np.sqrt(d2[iu]).mean()), float(ne[iu].mean()), float(ne.sum())


class Experiment:
    def __init__(self, experiment_id, serann_dataset, workers_pool, experiment_db, parameters, genetic_encoder,
                 start_generation: int = 0, population_source=None, shard_context: Optional[ShardContext] = None,
                 seed_base: int = 0, nan_offspring: str = 'keep'):
        """``genetic_encoder``: a ``RibosomalAutoencoder`` (reference builds it itself from config paths,
        ``experiment.py:36``).  ``population_source``: optional callable ``(n) -> list of source strings``; when
        given, the next generation's *source codes* come from it instead of from the decoder - needed with a
        random-init autoencoder, whose decodes are not valid programs.  The offspring
        genotypes are still produced, selected and decoded."""
        self._experiment_db = experiment_db
        self._parameters = parameters
        self._serann_dataset = serann_dataset
        self._id = experiment_id
        self._start_generation = start_generation
        self._genetic_encoder = genetic_encoder
        self._population_source = population_source
        self._experiment_db.save_execution_info(datetime.now(), self._parameters)
        self._workers_pool = workers_pool
        # 'keep' = the reference: an individual whose BatchNorm statistics overflowed (fp16 sample size) emits NaN
        # probabilities, its offspring genotypes are NaN and decoding them raises IndexError (ribosomal_autoencoder.py:123).
        # 'zero' keeps the run alive: NaN loci become 0, as the readout kernel packs them (such offspring differ from
        # their parent in every counted locus); used by the synthetic benchmark, whose generator mix holds ~36 % such nets.
        if nan_offspring not in ('keep', 'zero'):
            raise ValueError("nan_offspring must be 'keep' or 'zero'")
        self._nan_offspring = nan_offspring
        self.generation_seconds = []
        # one process per GPU (torchrun): this rank trains its shard of every generation, one all-gather per generation
        self._ctx = shard_context or ShardContext()
        self._seed_base = seed_base
        self._generation_index = start_generation
        # individual ids: version-4 UUIDs from a generator every rank seeds alike (uuid4() would differ per process);
        # the legacy np.random stream is left to the draws the reference makes from it
        self._id_rng = random.Random(stable_seed(seed_base, 'ids', start_generation))
        self.phase_seconds = []          # per generation: evaluate / allgather / statistics / select+decode

    def _new_id(self) -> str:
        return str(UUID(int=self._id_rng.getrandbits(128), version=4))

    # ------------------------------------------------------------------ main loop (experiment.py:43-108)
    def execute(self):
        p = self._parameters
        print(f'Experiment {self._id} has started')
        print(f'Using {self._workers_pool.get_workers_count()} GPU workers')
        current = self._get_first_generation()
        offspring_pool_size = p['initial_offspring_pool_size']
        print('Initial offspring pool size:', offspring_pool_size)
        for generation in range(self._start_generation, p['num_generations']):
            t0 = datetime.now()
            print(f'### Generation {generation} execution has started ###')
            current['experiment_id'] = self._id
            current['generation'] = generation
            current['num_offspring'] = 0
            print('Starting training and replication')
            self._generation_index = generation
            ph = {}
            tp = datetime.now()
            models_info, offspring_by_id, times = self._learn_and_replicate(current, offspring_pool_size)
            ph['evaluate'] = (datetime.now() - tp).total_seconds()
            ph['allgather_ms'] = self._ctx.last_allgather_ms
            current = current.join(models_info)
            print('Calculating fecundity scores')
            current['absolute_fertility'], current['relative_fertility'] = \
                self._get_fertility(current['classification_validation_accuracy'])
            print('Sampling offspring counts')
            if current['is_valid'].sum() > 0:
                current['num_offspring'] = np.random.multinomial(p['num_seranns'], current['relative_fertility'])
            survived = current['is_valid'].astype(bool) & ~current['is_overweight'].astype(bool)
            valid = current[survived]
            print('Extracting source codes statistics')
            counts = pd.DataFrame([extract_layer_counts(s) for s in valid['source_code']], index=valid.index)
            current = current.join(counts, how='left')
            print('Saving generation information to the experiment DB')
            tp = datetime.now()
            self._report_generation_statistics(current, generation, times, t0)
            print('Saving SeRANN records to the experiment DB')
            self._experiment_db.save_seranns_info(current)
            ph['statistics_db'] = (datetime.now() - tp).total_seconds()
            if len(valid) == 0:
                print('No valid SeRANNs left! stopping...')
                break
            print('Applying offspring selection')
            tp = datetime.now()
            nxt = self._select_offspring(current, offspring_by_id)
            ph['select_decode'] = (datetime.now() - tp).total_seconds()
            self.phase_seconds.append(ph)
            print(f'Generation {generation} execution is done')
            offspring_pool_size = int(current['num_offspring'].max() * p['offspring_pool_size_factor'])
            print('Offspring pool size was updated to:', offspring_pool_size)
            current = nxt
            dt = datetime.now() - t0
            self.generation_seconds.append(dt.total_seconds())
            s = int(dt.total_seconds())
            print(f'\033[1mGeneration execution time: {s // 3600:02d}:{(s // 60) % 60:02d}:{s % 60:02d}\033[0m')
        return current

    # ------------------------------------------------------------------ fertility (experiment.py:110-117)
    def _get_fertility(self, accuracy: pd.Series):
        absolute = accuracy ** self._parameters['selection_pressure']
        with np.errstate(all='ignore'):
            relative = np.nan_to_num(absolute / np.sum(absolute))
            relative = relative / relative.sum()
        return absolute, relative

    # ------------------------------------------------------------------ generations (experiment.py:119-160)
    def _load_last_generation_from_db(self):
        last = self._experiment_db.get_serann_by_generation(self._start_generation - 1)
        print('Replicating the last generation')
        ok = (last['is_valid'] == True) & (last['is_overweight'] == False)   # noqa: E712
        if ok.sum() == 0 or last[ok]['num_offspring'].sum() == 0:
            raise Exception('No valid SeRANNs left in the last generation!')
        print('Retraining and replicating the last generation')
        pool = self._parameters['initial_offspring_pool_size']
        _, offspring_by_id, _ = self._learn_and_replicate(last, pool)
        return self._select_offspring(last, offspring_by_id)

    def _get_first_generation(self):
        if self._start_generation > 0:
            return self._load_last_generation_from_db()
        p = self._parameters
        ancestor = p.get('ancestor_genotype')
        if ancestor is None:
            sample = np.random.randint(0, len(self._serann_dataset), p['num_seranns'])
            by_id = {self._new_id(): self._serann_dataset[s] for s in sample}
        else:
            ancestor = np.array(ancestor)
            by_id = {self._new_id(): ancestor for _ in range(p['num_seranns'])}
        gen = pd.Series(by_id).rename('genotype').rename_axis('id').to_frame()
        genotypes = np.array(gen['genotype'].values.tolist())
        gen['source_code'] = self._sources_for(genotypes)
        gen['parent_id'] = None
        for c in ('genotype_euclidean_distance_from_parent', 'genotype_hamming_distance_from_parent',
                  'source_code_levenshtein_distance_from_parent'):
            gen[c] = np.nan
        return gen

    def _sources_for(self, genotypes: np.ndarray):
        decoded = self._genetic_encoder.decode_to_string(genotypes)
        if self._population_source is None:
            return decoded
        return list(self._population_source(len(genotypes)))

    # ------------------------------------------------------------------ evaluation (experiment.py:162-190)
    def _seeds_for(self, generation: pd.DataFrame):
        """Row order of this generation's epochs and initial weights per individual, independent of the sharding."""
        fit_seed = stable_seed(self._seed_base, 'fit', self._generation_index) >> 1
        return fit_seed, {i: stable_seed(self._seed_base, 'init', i) for i in generation.index}

    def _offspring_from_kernel(self, result: dict) -> dict:
        """{id: (R, G) float16 of 0.0 / 1.0} from the readout kernel's bits (what ``np.round(np.clip(p, 0, 1))`` gives,
        experiment.py:210); a row that held a NaN probability stays NaN, as it does in the reference."""
        out = {}
        for sid, bits in result.get('offspring_bits_by_id', {}).items():
            g = bits.astype(np.float16)
            nan = np.asarray(result['offspring_nan_by_id'][sid], dtype=bool)
            if nan.any() and self._nan_offspring == 'keep':
                g[nan] = np.float16(np.nan)          # row granularity, as the all-gather carries it (an F6 row is all NaN)
            out[sid] = g
        return out

    def _learn_and_replicate(self, generation: pd.DataFrame, num_offspring):
        if self._ctx.world > 1:
            return self._learn_and_replicate_sharded(generation, num_offspring)
        times: Dict[str, list] = {}
        infos, offspring_by_id = [], {}
        workers = self._workers_pool.get_workers_count()
        per_job = getattr(self._workers_pool, 'max_seranns_per_job', MAX_SERANN_PER_GPU) or 10 ** 9     # None: the worker sizes its own waves
        batch_index = reference_batches(len(generation), workers, per_job)
        batch_by_id = pd.Series(batch_index, index=generation.index).astype(int)
        total = batch_by_id.nunique()
        seeded = getattr(self._workers_pool, 'accepts_seeds', False)
        kwargs = {}
        if seeded:
            fit_seed, ind_seeds = self._seeds_for(generation)
            kwargs = {'kwargs': {'fit_seed': fit_seed, 'individual_seeds': ind_seeds}}
        jobs = [{'args': [data, num_offspring], **kwargs} for _, data in generation.groupby(batch_by_id)]
        for i, result in enumerate(self._workers_pool.imap_unordered(jobs), 1):
            print(f'{i} of {total} batches completed')
            infos.append(result['models_info'])
            if 'offspring_bits_by_id' in result:
                offspring_by_id.update(self._offspring_from_kernel(result))     # the readout kernel's thresholded bits
            else:
                offspring_by_id.update(result.get('offspring_by_id', {}))
            times.setdefault('learning_times', []).append(result['learning_time'])
            times.setdefault('replication_times', []).append(result.get('replication_time', 0))
        return pd.concat(infos), offspring_by_id, times

    def _learn_and_replicate_sharded(self, generation: pd.DataFrame, num_offspring):
        """One process per GPU: longest-processing-time-first shards of equal size, this rank evaluates its shard, one
        all-gather carries every shard's ``models_info`` columns, bit-packed offspring and mutation counts to every
        rank (reference: pickled shards through a Redis task farm, experiment.py:162-190)."""
        p = self._parameters
        ctx = self._ctx
        dims = dict(image_dims=tuple(p['classification_image_dimensions']), genotype_size=int(p['genotype_size']),
                    n_classes=int(p['num_classification_classes']))
        costs = []
        for src in generation['source_code']:
            try:
                costs.append(individual_cost(compile_cached(src, **dims), int(p['training_batch_size'])))
            except InvalidSerann:
                costs.append(0.0)
        shards = lpt_shards(costs, ctx.world)
        self.last_shards = shards
        ids_by_rank = [[generation.index[k] for k in sh] for sh in shards]
        mine = generation.loc[ids_by_rank[ctx.rank]]
        fit_seed, ind_seeds = self._seeds_for(generation)
        job = {'args': [mine, num_offspring], 'kwargs': {'fit_seed': fit_seed, 'individual_seeds': ind_seeds}}
        result = next(iter(self._workers_pool.imap_unordered([job])))
        R, G = int(num_offspring), int(p['genotype_size'])
        blobs = ctx.allgather_bytes(pack_shard_results(ids_by_rank[ctx.rank], result, MODELS_INFO_COLUMNS, R, G))
        infos, offspring_by_id, times = [], {}, {'learning_times': [], 'replication_times': []}
        for rank, blob in enumerate(blobs):
            t, info, has, packed, mut, nan = unpack_shard_results(blob, len(MODELS_INFO_COLUMNS), G)
            ids = ids_by_rank[rank]
            df = pd.DataFrame(info, index=pd.Index(ids, name=generation.index.name), columns=MODELS_INFO_COLUMNS)
            df['is_valid'] = df['is_valid'].fillna(0).astype(bool)
            df['is_overweight'] = df['is_overweight'].fillna(0).astype(bool)
            infos.append(df)
            bits = unpack_bits(packed, G)
            for k, sid in enumerate(ids):
                if has[k]:
                    g = bits[k].astype(np.float16)
                    if nan[k].any() and self._nan_offspring == 'keep':     # F6 individuals: a NaN row stays NaN in the stored genotype
                        g[nan[k].astype(bool)] = np.float16(np.nan)
                    offspring_by_id[sid] = g
            times['learning_times'].append(float(t[0]))
            times['replication_times'].append(float(t[1]))
        print(f'{ctx.world} of {ctx.world} shards gathered ({ctx.last_allgather_ms:.2f} ms)')
        # rows back in the generation's own order
        return pd.concat(infos).loc[generation.index], {i: offspring_by_id[i] for i in generation.index if i in offspring_by_id}, times

    # ------------------------------------------------------------------ selection (experiment.py:192-238, 305-321)
    def _select_offspring(self, generation: pd.DataFrame, offspring_by_id: dict):
        p = self._parameters
        strategy = {'random': self._random_offspring_selection, 'best': self._best_offspring_selection} \
            .get(p['offspring_selection_strategy'])
        if strategy is None:
            raise Exception('Unknown offspring selection strategy')
        ids, parents, genotypes = [], [], []
        for parent_id, probs in offspring_by_id.items():
            parent_genotype = np.asarray(generation.loc[parent_id, 'genotype'])
            n = int(generation.loc[parent_id, 'num_offspring'])
            bits = np.round(np.clip(probs, 0, 1))                     # experiment.py:210
            chosen = strategy(parent_genotype, n, bits)
            chosen = self._probabilistic_proofreading(parent_genotype, chosen)
            ids += [self._new_id() for _ in range(n)]
            parents += [parent_id] * n
            genotypes += list(chosen)
        offspring = np.array(genotypes)
        sources = self._sources_for(offspring) if len(offspring) else []
        nxt = pd.DataFrame({'genotype': genotypes, 'parent_id': parents, 'source_code': sources}, index=ids).rename_axis('id')
        if len(nxt) == 0:
            return nxt
        parent_genotypes = np.array(generation.loc[parents, 'genotype'].values.tolist())
        nxt['genotype_euclidean_distance_from_parent'] = np.sqrt(np.sum(np.power(offspring.astype(np.float64) - parent_genotypes, 2), axis=1))
        nxt['genotype_hamming_distance_from_parent'] = (parent_genotypes != offspring).sum(axis=1) / p['genotype_size']
        parent_sources = generation.loc[parents, 'source_code'].values
        cache: Dict[tuple, int] = {}

        def lev(a, b):
            if a == b:
                return 0
            k = (a, b)
            if k not in cache:
                cache[k] = levenshtein(a, b)
            return cache[k]
        nxt['source_code_levenshtein_distance_from_parent'] = [lev(s, ps) for s, ps in zip(nxt['source_code'].values, parent_sources)]
        return nxt

    def _random_offspring_selection(self, _, num_offspring, offspring):
        return offspring[np.random.permutation(len(offspring))[:num_offspring]]

    def _best_offspring_selection(self, parent_genotype, num_offspring, offspring):
        return offspring[np.argsort(hamming(parent_genotype, offspring))[:num_offspring]]

    def _probabilistic_proofreading(self, parent_genotype, offspring):
        ec = self._parameters['error_correction_probability']
        out = offspring.copy()
        rows, cols = np.where(out != parent_genotype)
        fixed = np.random.permutation(len(rows))[:int(ec * len(rows))]     # consumes the stream even when ec == 0
        rows, cols = rows[fixed], cols[fixed]
        out[rows, cols] = parent_genotype[cols]
        return out

    # ------------------------------------------------------------------ statistics (experiment.py:240-303)
    def get_genotype_stats(self, info: pd.DataFrame) -> dict:
        genotypes = np.array(info['genotype'].values.tolist())
        as_str = info['genotype'].map(lambda g: str(np.asarray(g).tolist()))
        dist = as_str.value_counts(normalize=True)
        mean_euclid, mean_hamming, diversity = pairwise_genotype_distances(genotypes)
        return {'mean_pairwise_euclidean_distance': mean_euclid,
                'mean_pairwise_hamming_distance': mean_hamming,
                'mean_euclidean_distance_from_parent': info['genotype_euclidean_distance_from_parent'].mean(),
                'mean_hamming_distance_from_parent': info['genotype_hamming_distance_from_parent'].mean(),
                'shannon_index': float(-np.sum(dist * np.log(dist))),
                'nucleotide_diversity': diversity,
                'species_richness': int(as_str.nunique())}

    def get_source_code_stats(self, info: pd.DataFrame) -> dict:
        dist = info['source_code'].value_counts(normalize=True)
        return {'median_levenshtein_distance_from_parent': info['source_code_levenshtein_distance_from_parent'].median(),
                'shannon_index': float(-np.sum(dist * np.log(dist))),
                'species_richness': int(info['source_code'].nunique())}

    def _report_generation_statistics(self, info: pd.DataFrame, generation, times, start_time):
        survived = info['is_valid'].astype(bool) & ~info['is_overweight'].astype(bool)
        mean = lambda c: float(pd.to_numeric(info[c], errors='coerce').mean()) if c in info else float('nan')
        row = {
            'experiment_id': self._id, 'generation': generation, 'start_time': start_time,
            'survival_rate': float(survived.mean()),
            'overweight_rate': float(info['is_overweight'].astype(float).mean()),
            'invalid_rate': float(1 - info['is_valid'].astype(float).mean()),
            'mean_parameters_count': mean('parameters_count'),
            'mean_absolute_fertility': mean('absolute_fertility'),
            'absolute_fertility_std': float(pd.to_numeric(info['absolute_fertility'], errors='coerce').std()),
            'mean_loss_balance': mean('loss_balance'),
            'mean_classification_validation_accuracy': mean('classification_validation_accuracy'),
            'mean_classification_training_accuracy': mean('classification_training_accuracy'),
            'mean_classification_test_accuracy': mean('classification_test_accuracy'),
            'max_classification_test_accuracy': float(pd.to_numeric(info['classification_test_accuracy'], errors='coerce').max()),
            'mean_replication_mse': mean('replication_mse'),
            'learning_time_seconds': float(np.mean(times['learning_times'])),
            'replication_time_seconds': float(np.mean(times['replication_times'])),
            'total_time_seconds': (datetime.now() - start_time).total_seconds(),
            'mean_classification_layers': mean('classification_layers'),
            'mean_replication_layers': mean('replication_layers'),
            'mean_merged_layers': mean('merged_layers'),
        }
        row.update({f'genotype_{k}': v for k, v in self.get_genotype_stats(info).items()})
        row.update({f'source_code_{k}': v for k, v in self.get_source_code_stats(info).items()})
        self._experiment_db.save_generation_info(row)
