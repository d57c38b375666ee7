"""The drop-in layer on a GPU: worker.run contract (experiment_worker.py:45-99), evaluation worker
(serann_evaluation/logic.py:153-272) and two generations of the evolution loop with the SQLite store."""
import numpy as np
import pandas as pd
import pytest

import parity_util as U

pytestmark = pytest.mark.gpu

PARAMS = {'num_seranns': 6, 'num_generations': 2, 'max_tokens': 350, 'genotype_size': 100, 'selection_pressure': 1,
          'max_serann_parameters': 300000, 'classification_image_dimensions': [28, 28], 'num_classification_classes': 10,
          'training_epochs': 2, 'training_batch_size': 50, 'offspring_selection_strategy': 'random',
          'error_correction_probability': 0.5, 'initial_offspring_pool_size': 4, 'offspring_pool_size_factor': 5}


@pytest.fixture(scope='module')
def data():
    return U.small_dataset(n_train=400, n_test=80, seed=3)


def _frame(sources, data):
    ids = [f'id-{i}' for i in range(len(sources))]
    return pd.DataFrame({'genotype': [data['g_test'][i] for i in range(len(sources))], 'source_code': sources}, index=ids)


def test_worker_run_contract(data):
    from serann_b200.worker import B200PopulationWorker
    sources = [U.NET_B, U.NET_A, 'con = Dense(units=4)(X_layer\nloss_balance = 0.5', U.NET_D, U.NET_C]
    # NET_D has ~700k parameters: overweight under this max_serann_parameters
    w = B200PopulationWorker(dict(PARAMS), dataset=data, shuffle_seed=1)
    out = w.run(_frame(sources, data), 4)
    info = out['models_info']
    assert list(info.columns) == B200PopulationWorker.MODELS_INFO_COLUMNS
    assert info['is_valid'].tolist() == [True, True, False, True, True]
    assert info['is_overweight'].tolist() == [False, False, False, True, False]
    trained = ['id-0', 'id-1', 'id-4']
    assert set(out['offspring_by_id']) == set(trained)
    for sid in trained:
        off = out['offspring_by_id'][sid]
        assert off.shape == (4, 100) and off.dtype == np.float16
        assert np.all((off >= 0) & (off <= 1))
        for col in ('classification_validation_accuracy', 'classification_training_accuracy',
                    'classification_test_accuracy', 'replication_mse'):
            assert 0 <= info.at[sid, col] <= 1, (sid, col)
    for sid in ('id-2', 'id-3'):
        assert np.isnan(info.at[sid, 'classification_validation_accuracy'])
    assert np.isnan(info.at['id-2', 'parameters_count']) and info.at['id-3', 'parameters_count'] > 300000
    assert out['learning_time'] > 0 and out['replication_time'] > 0
    # accuracies are fp16-quantised values of k / n
    assert float(np.float16(info.at['id-0', 'classification_test_accuracy'])) == info.at['id-0', 'classification_test_accuracy']
    # nothing trainable: the reference's early return (experiment_worker.py:74-76)
    out = w.run(_frame(['loss_balance = 1'], data), 4)
    assert out['learning_time'] == 0 and out['replication_time'] == 0 and out['offspring_by_id'] == {}
    assert not out['models_info']['is_valid'].any()
    # num_replications == 0: no replication keys, as in the reference
    out = w.run(_frame([U.NET_B], data), 0)
    assert 'offspring_by_id' not in out and 'replication_time' not in out
    w.close()


def test_evaluation_worker(data):
    from serann_b200.codec import RibosomalAutoencoder
    from serann_b200.evaluation import B200EvaluationWorker
    from serann_b200.synthetic import synthetic_decoder_weights
    ae = RibosomalAutoencoder(350, synthetic_decoder_weights())
    w = B200EvaluationWorker(dict(PARAMS), data, num_evaluations=3, replications_per_evaluation=5, ribosomal_ae=ae, shuffle_seed=2)
    np.random.seed(0)
    res = w.evaluate(data['g_test'][0], source_code=U.NET_B)
    assert len(res['classification_accuracy']) == 3 and all(0 <= a <= 1 for a in res['classification_accuracy'])
    assert len(res['mutation_rate']) == 3 and all(sum(c.values()) == 5 for c in res['mutation_rate'])
    assert len(res['offspring_survival']) == 3 and all(sum(c.values()) == 5 for c in res['offspring_survival'])
    # a random-init decoder yields invalid programs: every offspring counts under key 0
    assert all(set(c) <= {0, 1} for c in res['offspring_survival'])
    # no loss_balance in the source -> the reference's empty result
    empty = w.evaluate(data['g_test'][1], source_code='con = X_layer')
    assert empty['offspring_survival'] == {} and np.isnan(empty['classification_accuracy'])
    # pool surface: cache by genotype hex, invalid rows short-circuit
    sid, r = w.run('s1', {'genotype_hex': '0xab', 'is_valid': False, 'is_overweight': False, 'genotype': data['g_test'][0]})
    assert sid == 's1' and r['mutation_rate'] == {}
    w.handle_update({'0xcd': res})
    assert w.run('s2', {'genotype_hex': '0xcd', 'is_valid': True, 'is_overweight': False, 'genotype': None})[1] is res
    w.close()
    ae.close()


def test_two_generations_end_to_end(data, tmp_path):
    from serann_b200.codec import RibosomalAutoencoder
    from serann_b200.experiment import Experiment
    from serann_b200.experiment_db import ExperimentDB
    from serann_b200.synthetic import synthetic_decoder_weights
    from serann_b200.worker import B200PopulationWorker, InProcessPool
    pool_src = [U.NET_B, U.NET_C, U.NET_A]
    source = lambda n: [pool_src[i % 3] for i in range(n)]
    np.random.seed(79375)
    ae = RibosomalAutoencoder(350, synthetic_decoder_weights())
    db = ExperimentDB(tmp_path / 'e.sqlite')
    enc = np.concatenate([data['g_train'], data['g_test']])
    with InProcessPool(B200PopulationWorker, {'args': [dict(PARAMS)], 'kwargs': {'dataset': data, 'shuffle_seed': 3}}, devices=[0]) as pool:
        exp = Experiment('e', enc, pool, db, dict(PARAMS), genetic_encoder=ae, population_source=source)
        last = exp.execute()
    assert db.get_generations_count() == 2
    g0, g1 = db.get_serann_by_generation(0), db.get_serann_by_generation(1)
    assert len(g0) == 6 and len(g1) == 6
    assert set(g1['parent_id']) <= set(g0.index)
    assert g0['num_offspring'].sum() == 6
    assert len(last) == 6
    ae.close()
