"""Host-side driver logic without a GPU: DB schema round trip, statistics, selection, the reference's RNG order."""
import numpy as np
import pandas as pd
import pytest

from serann_b200.experiment import Experiment, hamming, pairwise_genotype_distances
from serann_b200.experiment_db import ExperimentDB

PARAMS = {'num_seranns': 6, 'num_generations': 1, 'max_tokens': 350, 'genotype_size': 100, 'selection_pressure': 1,
          'max_serann_parameters': 2e6, 'classification_image_dimensions': [28, 28], 'num_classification_classes': 10,
          'training_epochs': 5, 'training_batch_size': 750, 'offspring_selection_strategy': 'random',
          'error_correction_probability': 0, 'initial_offspring_pool_size': 4, 'offspring_pool_size_factor': 5,
          'ancestor_genotype': [0, 1] * 50}


class FakeEncoder:
    def decode_to_string(self, genotypes):
        return ['con=X_layer\nloss_balance=0.' + str(int(np.sum(g)) % 10) for g in np.asarray(genotypes)]


class FakePool:
    """Stands in for InProcessPool + B200PopulationWorker: deterministic fake accuracies and offspring."""

    def __init__(self, R):
        self.R = R

    def get_workers_count(self):
        return 2

    def imap_unordered(self, jobs):
        for job in jobs:
            df, R = job['args']
            info = pd.DataFrame(index=df.index)
            info['parameters_count'] = 1000.0
            info['loss_balance'] = 0.5
            info['is_valid'] = True
            info['is_overweight'] = False
            info['classification_validation_accuracy'] = np.linspace(0.2, 0.9, len(df))
            info['classification_training_accuracy'] = 0.5
            info['classification_test_accuracy'] = 0.5
            info['replication_mse'] = 0.1
            rng = np.random.default_rng(len(df))
            off = {}
            for sid, row in df.iterrows():
                g = np.asarray(row['genotype'], dtype=np.float32)
                probs = np.repeat(g[None], R, axis=0) * 0.8 + 0.1
                flip = rng.random(probs.shape) < 0.03
                probs[flip] = 1 - probs[flip]
                off[sid] = probs.astype(np.float16)
            yield {'learning_time': 1.0, 'replication_time': 0.5, 'models_info': info, 'offspring_by_id': off}


def test_pairwise_distances_match_scipy_definition():
    from scipy.spatial.distance import cdist, pdist
    g = np.random.default_rng(0).integers(0, 2, (40, 100)).astype(float)
    e, h, d = pairwise_genotype_distances(g)
    assert e == pytest.approx(np.mean(pdist(g)))
    assert h == pytest.approx(np.mean(pdist(g, metric=hamming)))
    assert d == pytest.approx(cdist(g, g, metric=hamming).sum())
    g[3, 5] = 0.5                                   # non-binary -> general path
    e, h, d = pairwise_genotype_distances(g)
    assert e == pytest.approx(np.mean(pdist(g)))
    assert h == pytest.approx(np.mean(pdist(g, metric=hamming)))


def test_one_generation_against_the_schema(tmp_path):
    np.random.seed(79375)
    db = ExperimentDB(tmp_path / 'exp.sqlite')
    exp = Experiment('exp-1', np.zeros((10, 100)), FakePool(4), db, dict(PARAMS), genetic_encoder=FakeEncoder())
    nxt = exp.execute()
    assert db.get_generations_count() == 1 and db.get_executions_count() == 1
    stored = db.get_serann_by_generation(0)
    assert len(stored) == 6
    for col in ('genotype', 'source_code', 'parent_id', 'experiment_id', 'generation', 'num_offspring', 'parameters_count',
                'loss_balance', 'is_valid', 'is_overweight', 'classification_validation_accuracy',
                'classification_training_accuracy', 'classification_test_accuracy', 'replication_mse',
                'absolute_fertility', 'relative_fertility', 'classification_layers', 'replication_layers', 'merged_layers',
                'genotype_euclidean_distance_from_parent', 'genotype_hamming_distance_from_parent',
                'source_code_levenshtein_distance_from_parent'):
        assert col in stored.columns, col
    assert stored['num_offspring'].sum() == 6                       # multinomial over num_seranns
    assert stored['genotype'].iloc[0].dtype == np.float16 and stored['genotype'].iloc[0].shape == (100,)
    import sqlite3
    gen = pd.read_sql('select * from generations', sqlite3.connect(db.db_path))
    for col in ('survival_rate', 'overweight_rate', 'invalid_rate', 'mean_absolute_fertility', 'learning_time_seconds',
                'replication_time_seconds', 'total_time_seconds', 'genotype_mean_pairwise_hamming_distance',
                'genotype_nucleotide_diversity', 'genotype_species_richness', 'source_code_shannon_index',
                'source_code_species_richness', 'source_code_median_levenshtein_distance_from_parent'):
        assert col in gen.columns, col
    # next generation: offspring of the parents, thresholded to 0/1, distances consistent
    assert len(nxt) == 6 and set(np.unique(np.stack(nxt['genotype'].values))) <= {0.0, 1.0}
    assert (nxt['genotype_hamming_distance_from_parent'] * 100).round().astype(int).tolist() == \
        [int((np.asarray(PARAMS['ancestor_genotype']) != g).sum()) for g in nxt['genotype']]
    info = db.get_last_execution_info()
    assert info['num_seranns'] == 6 and info['classification_image_dimensions'] == [28, 28]
    assert info['ancestor_genotype'] == PARAMS['ancestor_genotype']


def test_selection_consumes_the_rng_like_the_reference():
    """One permutation for the choice and one for proof-reading per parent, even with ec = 0 (experiment.py:305-321)."""
    import tempfile, pathlib
    db = ExperimentDB(pathlib.Path(tempfile.mkdtemp()) / 'x.sqlite')
    exp = Experiment('e', np.zeros((4, 100)), FakePool(3), db, dict(PARAMS), genetic_encoder=FakeEncoder())
    parent = np.array([0, 1] * 50)
    probs = np.tile(parent.astype(np.float16), (5, 1))
    probs[2, :3] = 1 - probs[2, :3]
    gen = pd.DataFrame({'genotype': [parent], 'num_offspring': [2], 'source_code': ['con=X_layer\nloss_balance=0.0']}, index=['p'])
    np.random.seed(5)
    out = exp._select_offspring(gen, {'p': probs})
    np.random.seed(5)
    perm = np.random.permutation(5)[:2]
    bits = np.round(np.clip(probs, 0, 1))
    rows, _ = np.where(bits[perm] != parent)
    np.random.permutation(len(rows))                 # consumed by proof-reading
    after = np.random.rand()
    np.random.seed(5)
    exp._select_offspring(gen, {'p': probs})
    assert np.random.rand() == after
    assert np.array_equal(np.stack(out['genotype'].values), bits[perm])
