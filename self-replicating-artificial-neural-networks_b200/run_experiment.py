"""CLI of the evolutionary experiment on the B200 engine.

Same flags as the reference ``evolutionary_experiment/run_experiment.py:19-29`` (``-p -r -s`` used; the workers-pool
flags ``-w -o -a`` are accepted and ignored: there is no Redis farm).  Asset paths come from ``--assets`` (an ``.npz``
with ``encodings``, ``x_train``, ``y_train``, ``x_test``, ``y_test`` and the decoder weights); without it the run is
``--synthetic``: synthetic data of MNIST's shape, a random-init autoencoder and generator-sampled source codes
(the figshare assets and the MNIST download are not available offline).

    python -m serann_b200.run_experiment -p evolutionary_experiment/parameters/example.json --synthetic --results-dir out
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 -m serann_b200.run_experiment \
        -p evolutionary_experiment/parameters/full_experiment.json --synthetic --results-dir out

Under torchrun every rank (one per GPU) builds the same generation table from the same seeded ``np.random`` stream, trains
its cost-balanced shard, and one NCCL all-gather per generation brings every rank all results; rank 0 writes the SQLite file.
"""
from __future__ import annotations

import argparse
import json
import os
import socket
from pathlib import Path
from uuid import uuid4

import numpy as np

RANDOM_SEED = 79375             # reference evolutionary_experiment/config.py:17


def get_args(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument('-p', '--parameters', required=False, help='Experiment parameters file path')
    ap.add_argument('-r', '--resume-experiment-id', required=False, help='Experiment ID to resume')
    ap.add_argument('-s', '--random-seed', required=False, help='Override random seed')
    ap.add_argument('-w', '--workers-pool-port', required=False, type=int, help='ignored (no worker farm)')
    ap.add_argument('-o', '--workers-pool-password', required=False, help='ignored (no worker farm)')
    ap.add_argument('-a', '--workers-pool-address', required=False, help='ignored (no worker farm)')
    ap.add_argument('--assets', help='.npz with encodings, MNIST arrays and decoder weights')
    ap.add_argument('--synthetic', action='store_true', help='synthetic data, random-init autoencoder, generator sources')
    ap.add_argument('--synthetic-rows', default='60000,10000', help='train,test rows of the synthetic data set')
    ap.add_argument('--results-dir', default='data/experiment_results')
    ap.add_argument('--num-generations', type=int, help='override num_generations of the parameters file')
    ap.add_argument('--num-seranns', type=int, help='override num_seranns of the parameters file')
    ap.add_argument('--devices', help='comma-separated CUDA devices for a single-process run (default: LOCAL_RANK or 0); '
                                      'under torchrun every rank takes its LOCAL_RANK device and the population is sharded over the ranks')
    ap.add_argument('--nan-offspring', choices=('keep', 'zero'), default='keep',
                    help="offspring rows a NaN network produced (BatchNormalizationF16 overflow): 'keep' stores and decodes them as the "
                         "reference does (its decode raises on them), 'zero' reads NaN loci as 0")
    return ap.parse_args(argv)


def load_assets(args, params):
    from .synthetic import synthetic_dataset, synthetic_decoder_weights
    if args.assets:
        z = np.load(args.assets)
        enc = z['encodings']
        n_train, n_test = len(z['x_train']), len(z['x_test'])
        data = {'x_train': z['x_train'].astype(np.float32), 'y_train': z['y_train'], 'g_train': enc[:n_train],
                'x_test': z['x_test'].astype(np.float32), 'y_test': z['y_test'], 'g_test': enc[n_train:n_train + n_test],
                'encodings': enc}
        weights = {k: z[k] for k in z.files if k.startswith(('conv_', 'dense_', 'bn1_', 'bn2_'))}
        return data, weights, None
    if not args.synthetic:
        raise SystemExit('pass --assets FILE.npz or --synthetic (the reference assets are not available offline)')
    from .generator import sample_population
    n_train, n_test = (int(v) for v in args.synthetic_rows.split(','))
    data = synthetic_dataset(n_train=n_train, n_test=n_test, image_dims=tuple(params['classification_image_dimensions']),
                             genotype_size=params['genotype_size'], n_classes=params['num_classification_classes'])
    weights = synthetic_decoder_weights(params['genotype_size'], params['max_tokens'])
    state = {'seed': 0}

    def source(n):          # fresh generator-sampled sources every generation
        state['seed'] += 1
        return [s for s, _ in sample_population(n, seed=state['seed'], max_tokens=params['max_tokens'],
                                                max_parameters=params['max_serann_parameters'],
                                                genotype_size=params['genotype_size'])]
    return data, weights, source


def main(argv=None):
    args = get_args(argv)
    from .codec import RibosomalAutoencoder
    from .experiment import Experiment
    from .experiment_db import ExperimentDB
    from .worker import B200PopulationWorker, InProcessPool

    from .parallel import ShardContext
    seed = int(args.random_seed or RANDOM_SEED)
    np.random.seed(seed)                      # every rank seeds the driver's stream alike: selection runs redundantly
    ctx = ShardContext.from_env()             # torchrun: one process per GPU; plain python: a single rank
    results = Path(args.results_dir)
    results.mkdir(exist_ok=True, parents=True)
    # rank 0 owns the results file; the other ranks compute the same tables and drop them
    make_db = (lambda path: ExperimentDB(path)) if ctx.rank == 0 else (lambda path: ExperimentDB(path, read_only=True))
    info = {}
    if args.resume_experiment_id is None:
        experiment_id = ctx.allgather_bytes(str(uuid4()).encode())[0].decode()      # rank 0's draw, on every rank
        with open(args.parameters) as f:
            info = {'parameters': json.load(f), 'experiment_db': make_db(results / f'{experiment_id}.sqlite')}
    else:
        experiment_id = args.resume_experiment_id
        ctx.barrier()
        db = make_db(results / f'{experiment_id}.sqlite')
        start = db.get_generations_count()
        if args.parameters is not None:
            with open(args.parameters) as f:
                params = json.load(f)
        else:
            last = db.get_last_execution_info()
            params = dict(last.drop('start_time'))
            if start == params['num_generations']:
                params['num_generations'] *= 2
        info = {'parameters': params, 'experiment_db': db, 'start_generation': start}
    params = info['parameters']
    if args.num_generations:
        params['num_generations'] = args.num_generations
    if args.num_seranns:
        params['num_seranns'] = args.num_seranns
    params['host_name'] = socket.gethostname()
    params['encodings_dataset'] = Path(args.assets).stem if args.assets else 'synthetic'
    params['tokens_vocabulary'] = 'reconstructed_40_tokens'
    params['ribosomal_autoencoder'] = Path(args.assets).stem if args.assets else 'synthetic'
    params['random_seed'] = seed

    data, decoder_weights, source = load_assets(args, params)
    device = int(os.environ.get('LOCAL_RANK', 0))
    encoder = RibosomalAutoencoder(params['max_tokens'], decoder_weights, genotype_size=params['genotype_size'], device=device)
    pool_args = (B200PopulationWorker, {'args': [params], 'kwargs': {'dataset': data}})
    devices = [device]
    if ctx.world == 1 and args.devices:        # one process driving several GPUs: one worker and one host thread per device
        devices = [int(d) for d in args.devices.split(',')]
    with InProcessPool(*pool_args, devices=devices) as pool:
        exp = Experiment(experiment_id, data['encodings'], pool, genetic_encoder=encoder, population_source=source,
                         shard_context=ctx, seed_base=seed, nan_offspring=args.nan_offspring, **info)
        exp.execute()
    ctx.barrier()
    if ctx.rank == 0:
        print('results:', info['experiment_db'].db_path)
    return exp


if __name__ == '__main__':
    main()
