#!/usr/bin/env python
"""Benchmark of the SeRANN population-evaluation hot path (see BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--config pop100|pop1000|decode100k|eval200x32]

Metric: SeRANN-train-steps/s = (individuals trained) x (training steps) / seconds.

Default line (--config pop100 = BASELINE.json configs[1]): a population of 100 SeRANNs per GPU (generator-sampled sources,
seed 0, distinct nets), training batch 750 on a synthetic 60000/10000 data set of MNIST's shape.  One bench "step" is one
training step of every individual of the shard on one mini-batch (gather, forward, losses, backward, Adam).

  value     device-timed throughput with the data set resident in HBM (CUDA events on the engine's stream), weak scaling
  e2e       the same metric through the product's own path: Experiment.execute() over the drop-in worker for whole
            generations - H2D of the data set, source compilation, build, 5 epochs x 76 steps, validation, test,
            replication + readout, all-gather, fertility, statistics, SQLite, selection, decode of the offspring
  pop1000   BASELINE's second figure: ONE population of 1000 (generator seed 0) sharded longest-processing-time-first
            over the N ranks (strong scaling) through the same product path: seconds per generation
  roofline  dominant kernel category against the roof that bounds it (HBM below the FLOP/byte ridge, else tensor)
  --impl reference   the CPU restatement of the reference's TF path (oracle/) on the host cores, same population

Other configs (one JSON line each): pop1000 alone, decode100k (configs[3]), eval200x32 (configs[4]).
"""
import argparse
import json
import os
import pathlib
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = 'SeRANN-train-steps/s'
POP_PER_GPU = 100
BATCH = 750
PARAMS = {
    'num_seranns': POP_PER_GPU, 'num_generations': 10, 'max_tokens': 350, 'genotype_size': 100,
    'selection_pressure': 1, 'max_serann_parameters': 2e6, 'classification_image_dimensions': [28, 28],
    'num_classification_classes': 10, 'training_epochs': 5, 'training_batch_size': BATCH,
    'offspring_selection_strategy': 'random', 'error_correction_probability': 0,
    'initial_offspring_pool_size': 10, 'offspring_pool_size_factor': 5,
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--config', default='pop100', choices=['pop100', 'pop1000', 'decode100k', 'eval200x32'])
    ap.add_argument('--pop', type=int, default=POP_PER_GPU, help='individuals per GPU (pop100)')
    ap.add_argument('--mix', default='diverse', choices=['diverse', 'clonal'])
    ap.add_argument('--generations', type=int, default=3, help='generations of the end-to-end leg')
    ap.add_argument('--genotypes', type=int, default=200, help='genotypes of eval200x32')
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--no-cpu', action='store_true')
    ap.add_argument('--no-pop1000', action='store_true')
    return ap.parse_args()


def population(n_total, mix, seed=0):
    from serann_b200 import generator
    return generator.sample_population(n_total, seed=seed, distinct=(mix == 'diverse'))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    Q = 'clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,' \
        'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap'

    def __init__(self, device):
        self.device, self.proc, self.lines = device, None, []

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', f'--id={self.device}', f'--query-gpu={self.Q}',
                                          '--format=csv,noheader,nounits', '-lms', '200'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True)
            self.t.start()
        except OSError:
            self.proc = None
        return self

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.25)
            self.proc.terminate()
            self.t.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], 0, set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for ln in self.lines:
            f = [x.strip() for x in ln.split(',')]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx = max(mx, float(f[1]))
            except ValueError:
                continue
            for name, v in zip(names, f[2:6]):
                if v.lower().startswith('active'):
                    reasons.add(name)
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': mx or None, 'reasons': sorted(reasons),
                'samples': len(sm)}


def peaks():
    p = ROOT / 'MEASURED_PEAKS.json'
    if p.exists():
        d = json.loads(p.read_text())
        return d.get('hbm_gbs', 6650.0), d.get('bf16_tflops_sustained', 1400.0), d.get('bf16_tflops', 1590.0), 'measured (MEASURED_PEAKS.json)'
    return 6650.0, 1400.0, 1590.0, 'fallback (B200_PROFILING.md)'


def algorithmic_work_per_step(irs, batch):
    """Algorithmic FLOPs and bytes of one training step per kernel category, summed over the individuals (DESIGN.md
    section 4).  GEMM bytes = each operand once: input activations, output activations, kernel."""
    from serann_b200 import frontend as F
    flops = {'gemm_fwd': 0, 'gemm_dgrad': 0, 'gemm_wgrad': 0}
    byts = {k: 0 for k in ('gemm_fwd', 'gemm_dgrad', 'gemm_wgrad', 'act_bwd', 'bn', 'pool', 'adam', 'loss')}
    for ir in irs:
        needs = ir.needs_input_grad()
        for l, nd in zip(ir.layers, needs):
            out_b = 2 * l.out_elems * batch
            in_b = 2 * int(np.prod(l.in_shape)) * batch if l.in_shape else 0
            if l.kind in (F.KIND_DENSE, F.KIND_CONV):
                m = 2 * l.macs_per_sample() * batch
                w_b = 2 * int(np.prod(l.param_shapes()[0][1]))
                flops['gemm_fwd'] += m
                flops['gemm_wgrad'] += m
                byts['gemm_fwd'] += in_b + out_b + w_b
                byts['gemm_wgrad'] += in_b + out_b + 2 * w_b          # fp32 gradient out
                if nd:
                    flops['gemm_dgrad'] += m
                    byts['gemm_dgrad'] += in_b + out_b + w_b
                byts['act_bwd'] += (3 if l.act else 1) * out_b           # read dY (and Y), write dY; linear: read dY for the bias sums
            elif l.kind == F.KIND_BN:
                byts['bn'] += 5 * out_b                                 # 4 B/element forward + 6 B/element backward
            elif l.kind == F.KIND_MAXPOOL:
                byts['pool'] += 2 * (in_b + out_b)
        byts['adam'] += 16 * ir.parameters_count                        # p, m, v fp16 read + write, fp32 gradient read
        byts['loss'] += 4 * (ir.n_classes + ir.genotype_size) * batch
    return flops, byts


def roofline_from_profile(prof, irs, n_steps):
    """Dominant kernel category of the profiled steps against the roof that bounds it."""
    hbm_peak, tf_sust, tf_burst, peak_kind = peaks()
    ridge = tf_sust * 1e12 / (hbm_peak * 1e9)
    flops, byts = algorithmic_work_per_step(irs, BATCH)
    step_ms = sum(m for m, _ in prof.values()) / n_steps
    shares = {k: round((m / n_steps) / step_ms, 4) for k, (m, c) in prof.items() if c}
    cats = {}
    for k, (ms, cnt) in prof.items():
        if not cnt or k not in byts:
            continue
        sec = ms / n_steps / 1000.0
        d = {'ms_per_step': ms / n_steps, 'launches_per_step': cnt // n_steps,
             'achieved_GBs': byts[k] / sec / 1e9, 'frac_hbm': byts[k] / sec / 1e9 / hbm_peak}
        if k in flops:
            d.update({'achieved_TFLOPs': flops[k] / sec / 1e12, 'frac_tensor': flops[k] / sec / 1e12 / tf_sust,
                      'flop_per_byte': flops[k] / max(byts[k], 1)})
        cats[k] = d
    top = max(cats, key=lambda k: cats[k]['ms_per_step'])
    t = cats[top]
    hbm_bound = top not in flops or t['flop_per_byte'] < ridge
    traffic, traffic_src = None, None
    tpath = ROOT / 'profiles' / 'r02_dram_traffic_per_step.json'
    if tpath.exists():
        tj = json.loads(tpath.read_text())
        if top in tj:
            traffic = tj[top] / max(t['launches_per_step'], 1)
            traffic_src = f'profiles/{tpath.name} (dram__bytes_read.sum + dram__bytes_write.sum per step from an ncu capture of this bench ' \
                          f'command, {tj.get("_commit", "this round")}), divided by the launches per step; not measured in this run'
    launches = max(t['launches_per_step'], 1)
    roof = {'bound': 'hbm' if hbm_bound else 'tensor', 'kernel': top,
            'achieved': (t['achieved_GBs'] if hbm_bound else t['achieved_TFLOPs']),
            'peak': hbm_peak if hbm_bound else tf_sust, 'unit': 'GB/s' if hbm_bound else 'TFLOP/s',
            'frac': t['frac_hbm'] if hbm_bound else t['frac_tensor'],
            'traffic': traffic, 'traffic_source': traffic_src,
            'algorithmic_bytes_per_launch': byts[top] / launches,
            'avg_launch_ms': t['ms_per_step'] / launches,
            'peak_source': f'{peak_kind}: sustained figures (the kernel is timed inside a long step)',
            'ridge_flop_per_byte': ridge,
            'both_fractions': {'hbm': t['frac_hbm'], 'tensor': t.get('frac_tensor')},
            'timing': 'CUDA events around every launch of the category on the stream it is launched on, launches serialised on one '
                      'stream while profiling; achieved = algorithmic bytes (each operand once) of all launches / their summed time',
            'share_of_step': shares, 'categories': cats}
    return roof, step_ms


def cpu_oracle_rate(pop_sample, steps, warmup, threads, budget_s=None):
    """Oracle training steps per second on the host cores; every step trains every individual of the sample once."""
    import torch
    from oracle import serann_oracle as O
    from serann_b200.synthetic import synthetic_dataset
    torch.set_num_threads(threads)
    data = synthetic_dataset(n_train=BATCH * 4, n_test=8)
    rng = np.random.default_rng(5)
    inds = [O.Individual(ir, O.init_weights(ir, rng)) for _, ir in pop_sample]
    t_total, done_steps, s = 0.0, 0, 0
    n_rows = len(data['x_train'])
    while s < steps + warmup:
        rows = (np.arange(BATCH) + s * BATCH) % n_rows
        xi, gi = O.prepare_inputs(data['x_train'][rows], data['g_train'][rows])
        labels = torch.tensor(data['y_train'][rows].astype(np.int64))
        bits = torch.tensor(data['g_train'][rows].astype(np.float32))
        t0 = time.perf_counter()
        for ind in inds:
            O.train_step(ind, xi, gi, labels, bits, s + 1)
        dt = time.perf_counter() - t0
        if s >= warmup:
            t_total += dt
            done_steps += 1
            if budget_s is not None and t_total > budget_s:
                s += 1
                break
        s += 1
    return len(inds) * done_steps / t_total, t_total, done_steps


def workload_config(args, n_ind, world):
    return {'workload': f'configs[1]: population {args.pop} SeRANNs per GPU ({args.mix} generator mix, seed 0, the same mix on every GPU), '
                        f'batch {BATCH}, 60000/10000 synthetic rows of 28x28 + 100-bit genotypes',
            'population_per_gpu': n_ind, 'batch': BATCH, 'mix': args.mix,
            'l2_policy': 'working set per step (activation arena, GBs) is far larger than the 126 MB L2',
            'parallelism': f'population sharded over {world} GPU(s), no data-path collective in training; one all-gather per generation'}


def run_reference(args):
    """--impl reference: the CPU restatement of the reference's TF path, all host threads, the bench's own population.
    The line reports what ran: `steps` are the steps actually timed, `ms_per_step` one step of the whole population."""
    rank = int(os.environ.get('RANK', 0))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    pop = population(args.pop, args.mix)
    # the whole population when it fits the time budget (~2 s per step of 100 individuals on 16 cores), else a sample
    budget = 150.0
    rate, secs, done = cpu_oracle_rate(pop, max(1, args.steps), 1 if args.warmup > 0 else 0, threads, budget_s=budget)
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': rate, 'unit': 'steps/s', 'n_gpus': args.gpus,
        'steps': done, 'warmup': 1 if args.warmup > 0 else 0, 'ms_per_step': 1000.0 * secs / done,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f16', 'data': 'synthetic',
        'config': dict(workload_config(args, len(pop), 1), workload=f'configs[1] on the host cores: population {len(pop)} SeRANNs '
                       f'({args.mix} generator mix, seed 0), batch {BATCH}; {done} of the requested {args.steps} steps timed'
                       + ('' if done == args.steps else f' (time budget {budget:.0f} s)')),
        'cpu_baseline': {'value': rate, 'unit': 'steps/s', 'cores': threads, 'kind': 'port',
                         'sample': f'all {len(pop)} individuals x {done} training steps at batch {BATCH} (torch CPU oracle; the reference '
                                   f'itself needs TensorFlow 2.6, which is not installable here)'},
        'e2e': {'value': rate, 'unit': 'steps/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    }
    emit(line)


# --------------------------------------------------------------------------------------------------------------------
class Dist:
    def __init__(self):
        import torch
        self.torch = torch
        self.world = int(os.environ.get('WORLD_SIZE', 1))
        self.rank = int(os.environ.get('RANK', 0))
        self.local = int(os.environ.get('LOCAL_RANK', 0))
        if not torch.cuda.is_available():
            raise SystemExit('bench.py needs a CUDA device: the engine has no CPU path')
        torch.cuda.set_device(self.local)
        from serann_b200.parallel import ShardContext
        self.ctx = ShardContext.from_env('nccl')

    def barrier(self):
        self.ctx.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, v, op):
        if self.world == 1:
            return v
        import torch.distributed as dist
        t = self.torch.tensor([v], dtype=self.torch.float64, device='cuda')
        dist.all_reduce(t, op={'max': dist.ReduceOp.MAX, 'min': dist.ReduceOp.MIN, 'sum': dist.ReduceOp.SUM}[op])
        return float(t.item())


def run_generations(D, params, sources_per_generation, data, decoder_weights, tmpdir, seed=79375):
    """`len(sources_per_generation)` generations through the product path (Experiment.execute over the drop-in worker, on
    this rank's device; sharded when torchrun started several ranks).  Returns timings and byte counts."""
    from serann_b200.codec import RibosomalAutoencoder
    from serann_b200.experiment import Experiment
    from serann_b200.experiment_db import ExperimentDB
    from serann_b200 import engine as E
    from serann_b200.worker import B200PopulationWorker, InProcessPool
    import contextlib
    import io
    np.random.seed(seed)
    params = dict(params, num_generations=len(sources_per_generation))
    pools = iter(list(sources_per_generation) + [sources_per_generation[-1]])     # selection after the last generation draws once more
    D.barrier()
    allocs0 = E.alloc_count()
    t0 = time.perf_counter()
    quiet = io.StringIO()
    with contextlib.redirect_stdout(quiet):
        encoder = RibosomalAutoencoder(params['max_tokens'], decoder_weights, genotype_size=params['genotype_size'], device=D.local)
        db = ExperimentDB(pathlib.Path(tmpdir) / f'bench_{os.getpid()}.sqlite', read_only=D.rank != 0)
        stats = []
        with InProcessPool(B200PopulationWorker, {'args': [params], 'kwargs': {'dataset': data, 'shuffle_seed': 4}}, devices=[D.local]) as pool:
            worker = pool._workers[0]
            orig_run = worker.run

            def run(*a, **k):
                before = E.alloc_count()
                out = orig_run(*a, **k)
                stats.append(dict(worker.last_stats, allocs=E.alloc_count() - before))
                return out
            worker.run = run
            exp = Experiment('bench', data['encodings'], pool, db, params, genetic_encoder=encoder,
                             population_source=lambda n: list(next(pools))[:n], shard_context=D.ctx, seed_base=seed,
                             nan_offspring='zero')      # the generator mix holds F6 nets whose NaN offspring the reference cannot decode
            exp.execute()
            h2d = worker._engine.h2d_bytes + encoder.h2d_bytes
            d2h = worker._engine.d2h_bytes + encoder.d2h_bytes
        encoder.close()
    D.barrier()
    total = D.reduce(time.perf_counter() - t0, 'max')
    gens = exp.generation_seconds
    return {'total_s': total, 'generation_s': gens, 'phases': exp.phase_seconds, 'worker': stats, 'h2d': h2d, 'd2h': d2h,
            'allocs_total': E.alloc_count() - allocs0, 'trained': [s['trained'] for s in stats], 'steps': [s['steps'] for s in stats]}


def pop1000_leg(D, data, decoder_weights, tmpdir, generations=1):
    """One population of 1000 (generator seed 0) LPT-sharded over the ranks, whole generations through the product."""
    from serann_b200 import generator
    pool = generator.sample_population(1000, seed=0, distinct=True)
    srcs = [s for s, _ in pool]
    params = dict(PARAMS, num_seranns=1000)
    r = run_generations(D, params, [srcs] * generations, data, decoder_weights, tmpdir)
    fit = [w['fit_s'] for w in r['worker']]
    trained = D.reduce(sum(r['trained']), 'sum')
    steps = r['steps'][0] if r['steps'] else 0
    out = {'population': 1000, 'generations': generations, 'n_gpus': D.world, 'scaling': 'strong',
           'seconds_per_generation': float(np.mean(r['generation_s'])), 'seconds_total_incl_setup': r['total_s'],
           'steps_per_s': trained * steps / max(sum(r['generation_s']), 1e-9),
           'fit_s_rank_max': D.reduce(float(np.sum(fit)), 'max') / generations, 'fit_s_rank_min': D.reduce(float(np.sum(fit)), 'min') / generations,
           'allgather_ms': [p.get('allgather_ms', 0.0) for p in r['phases']],
           'phases_rank0_s': r['phases'], 'worker_rank0_s': r['worker'],
           'what': 'sample_population(1000, seed=0) sharded longest-processing-time-first over the ranks; one generation = compile, build, '
                   '5 x 76 training steps at batch 750, validation, test, replication + readout (R = 10), all-gather, fertility, '
                   'multinomial, statistics, SQLite (rank 0), selection, decode of the 1000 offspring'}
    return out


def _claim_stdout():
    """Libraries (NCCL's version banner, torchrun children) write to fd 1: keep the real stdout for the one JSON line and
    point fd 1 at stderr for everything else."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), 'w')
    os.dup2(2, 1)


def emit(line: dict):
    _REAL_STDOUT.write(json.dumps(line) + '\n')
    _REAL_STDOUT.flush()


def main():
    args = parse()
    _claim_stdout()
    if args.impl == 'reference':
        run_reference(args)
        return
    D = Dist()
    torch = D.torch
    from serann_b200 import engine as E
    from serann_b200.synthetic import synthetic_dataset, synthetic_decoder_weights
    if args.config == 'decode100k':
        return decode_leg(D, args)
    if args.config == 'eval200x32':
        return eval_leg(D, args)
    data = synthetic_dataset()
    decoder_weights = synthetic_decoder_weights()
    tmpdir = tempfile.mkdtemp(prefix='serann_bench_')
    if args.config == 'pop1000':
        res = pop1000_leg(D, data, decoder_weights, tmpdir, generations=max(1, args.generations - 1))
        if D.rank == 0:
            steps = PARAMS['training_epochs'] * 76
            emit({'metric': METRIC, 'value': res['steps_per_s'], 'unit': 'steps/s', 'n_gpus': D.world, 'steps': steps,
                              'warmup': 0, 'ms_per_step': 1000.0 * res['seconds_per_generation'] / steps, 'higher_is_better': True,
                              'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f16', 'data': 'synthetic',
                              'config': {'workload': 'configs[2]: one population of 1000 SeRANNs sharded over the GPUs, whole generations end to end'},
                              'pop1000': res})
        return

    # weak scaling: every GPU evaluates its own args.pop individuals drawn from the same architecture mix (independent
    # initialisations), so the per-GPU work is the same at every N; the job's population is args.pop * world
    pop = population(args.pop, args.mix)
    irs = [ir for _, ir in pop]
    n_val = 3000
    warm = max(args.warmup, 3)

    # ---- device-resident throughput -------------------------------------------------------------
    eng = E.Engine(max_batch=BATCH, device=D.local)
    eng.set_dataset(data['x_train'], data['y_train'], data['g_train'], data['x_test'], data['y_test'], data['g_test'])
    eng.build(irs, seed=7)
    eng.fit(1, BATCH, n_val, shuffle_seed=1, max_steps=warm)
    D.barrier()
    launches0 = E.launch_count()
    with ClockSampler(D.local) as clocks:
        t0 = time.perf_counter()
        res = eng.fit(1, BATCH, n_val, shuffle_seed=2, max_steps=args.steps)
        D.barrier()
        wall = time.perf_counter() - t0
    launches = E.launch_count() - launches0
    dev_s = D.reduce(res['device_ms'] / 1000.0, 'max')
    total_ind = D.reduce(len(irs), 'sum')
    value = total_ind * args.steps / dev_s

    # ---- per-category device time -> roofline of the dominant kernel -----------------------------
    eng.profile(True)
    eng.fit(1, BATCH, n_val, shuffle_seed=3, max_steps=2)
    prof = eng.profile_read()
    eng.profile(False)
    roofline, serial_ms = roofline_from_profile(prof, irs, 2)
    flops, byts = algorithmic_work_per_step(irs, BATCH)
    hbm_peak, tf_sust, _, _ = peaks()
    step_s = dev_s / args.steps
    roofline['step'] = {'ms': 1000.0 * step_s, 'serialised_kernel_ms': serial_ms,
                        'algorithmic_TFLOP': sum(flops.values()) / 1e12, 'algorithmic_GB': sum(byts.values()) / 1e9,
                        'frac_of_tensor_peak': sum(flops.values()) / step_s / 1e12 / tf_sust,
                        'frac_of_hbm_peak': sum(byts.values()) / step_s / 1e9 / hbm_peak,
                        'note': 'all kernels of one step over the measured step time (launches overlap on several streams)'}
    eng.close()

    # ---- end to end through the product path ------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        G = max(1, args.generations)
        # weak scaling: the job's population is args.pop per GPU, sharded longest-processing-time-first over the ranks
        n_job = args.pop * D.world
        # (every GPU's share is the N = 1 population of that generation: the same mix repeated once per GPU, as `value` does -
        # a fresh sample of pop x N would change the work per individual with N, the seed-0 hundred being lighter than the mean)
        # and every generation is re-seeded with the benchmark population itself (generator seed 0, the workload `value` is
        # measured on): fresh samples per generation would time other, on average 1.4x heavier, mixes under the same name
        pools = [[s for s, _ in population(args.pop, args.mix, seed=0)] * D.world for g in range(G)]
        r = run_generations(D, dict(PARAMS, num_seranns=n_job), pools, data, decoder_weights, tmpdir)
        trained = D.reduce(sum(r['trained']), 'sum')
        steps_gen = r['steps'][0]
        total_steps = steps_gen * len(r['steps'])
        e2e = {'value': trained * steps_gen / r['total_s'], 'unit': 'steps/s',
               'h2d_bytes_per_step': r['h2d'] / total_steps, 'd2h_bytes_per_step': r['d2h'] / total_steps,
               'generations': len(r['generation_s']), 'seconds_total_incl_setup': r['total_s'],
               'seconds_per_generation': r['generation_s'], 'steps_per_generation': steps_gen,
               'phases_rank0_s': r['phases'], 'worker_rank0_s': r['worker'],
               'cudaMalloc_calls_per_generation': [w['allocs'] for w in r['worker']],
               'what': 'Experiment.execute() over InProcessPool(B200PopulationWorker): H2D of the data set (once), then per generation '
                       'source compilation, build, 5 x 76 steps, validation, test, replication + readout (R = 10), D2H of metrics and '
                       'offspring, all-gather (N > 1), fertility, multinomial, statistics, SQLite, selection, decode of the offspring; '
                       'every generation starts from the benchmark population (generator seed 0, pop individuals, repeated once per GPU = pop x N individuals), LPT-sharded over the N ranks (a random-init autoencoder decodes no valid programs)'}

    # ---- BASELINE's second figure: one population of 1000 over the N GPUs ---------------------------
    p1000 = None
    if not args.no_pop1000:
        p1000 = pop1000_leg(D, data, decoder_weights, tmpdir, generations=1)

    # ---- CPU baseline (rank 0, N = 1) ---------------------------------------------------------------
    cpu = None
    if D.rank == 0 and D.world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        sample = pop[::max(1, len(pop) // 25)][:25]
        rate, secs, done = cpu_oracle_rate(sample, 3, 1, threads, budget_s=25.0)
        cpu = {'value': rate, 'unit': 'steps/s', 'cores': threads, 'kind': 'port',
               'sample': f'{len(sample)} individuals evenly spaced through the workload x {done} training steps at batch {BATCH} '
                         f'({secs:.1f} s; torch CPU oracle)'}

    if D.rank == 0:
        line = {'metric': METRIC, 'value': value, 'unit': 'steps/s', 'n_gpus': D.world, 'steps': args.steps,
                'warmup': warm, 'ms_per_step': 1000.0 * dev_s / args.steps, 'higher_is_better': True,
                'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f16', 'data': 'synthetic',
                'config': workload_config(args, len(irs), D.world), 'roofline': roofline, 'cpu_baseline': cpu, 'e2e': e2e,
                'pop1000': p1000, 'gpu_launches': int(launches), 'clocks': clocks.summary(), 'wall_s_timed_region': wall}
        emit(line)
    if D.world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------------------------------------
def decode_leg(D, args):
    """configs[3]: batched autoencoder decode of 100k synthetic genotypes (each rank decodes its slice)."""
    from serann_b200.codec import RibosomalAutoencoder
    from serann_b200.synthetic import synthetic_decoder_weights
    n_total = 100_000
    per = -(-n_total // D.world)
    lo = D.rank * per
    n = max(0, min(n_total, lo + per) - lo)
    w = synthetic_decoder_weights()
    bits = np.random.Generator(np.random.PCG64(1)).integers(0, 2, (n_total, 100)).astype(np.uint8)[lo:lo + n]
    ae = RibosomalAutoencoder(350, w, device=D.local)
    ae.decode_to_sequence(bits[:4096])                       # warm-up
    D.barrier()
    with ClockSampler(D.local) as clocks:
        t0 = time.perf_counter()
        tok = ae.decode_to_sequence(bits)
        D.barrier()
        wall = D.reduce(time.perf_counter() - t0, 'max')
    dev_s = D.reduce(ae.last_device_ms / 1000.0, 'max')
    hbm_peak, tf_sust, tf_burst, peak_kind = peaks()
    flop = 2.0 * (96 * 10 * 32 + 3072 * 14000) * n_total
    stats = getattr(ae, 'last_stats', {})
    cpu = None
    if D.rank == 0 and not args.no_cpu:
        import torch
        from oracle import serann_oracle as O
        torch.set_num_threads(os.cpu_count() or 1)
        m = 2000
        t0 = time.perf_counter()
        ref, z = O.ae_decode(bits[:m], w, 350, 40)
        cs = time.perf_counter() - t0
        top2 = np.sort(z, axis=-1)[..., -2:]
        clear = (top2[..., 1] - top2[..., 0]) > 1e-4
        mism = int((tok[:m][clear] != ref[clear]).sum())
        cpu = {'value': m / cs, 'unit': 'genotypes/s', 'cores': os.cpu_count(), 'kind': 'port',
               'sample': f'{m} of the 100000 genotypes through the fp32 torch CPU oracle ({cs:.1f} s)',
               'argmax_mismatches_outside_1e-4_margin': mism, 'tokens_compared': int(clear.sum())}
    if D.rank == 0:
        emit({'metric': 'autoencoder-decode genotypes/s', 'value': n_total / dev_s, 'unit': 'genotypes/s', 'n_gpus': D.world,
                          'steps': 1, 'warmup': 1, 'ms_per_step': 1000.0 * dev_s, 'higher_is_better': True, 'scaling': 'strong',
                          'vs_baseline': None, 'dtype': 'fp16 hi/lo split table x exact one-hot operand, fp32 accumulate', 'data': 'synthetic',
                          'config': {'workload': 'configs[3]: decode 100000 synthetic genotypes (seed 1) with a random-init ribosomal autoencoder (seed 4), max_tokens 350, vocabulary 40'},
                          'roofline': {'bound': 'tensor', 'kernel': 'k_dec_gemm', 'achieved': flop / dev_s / 1e12 / D.world, 'peak': tf_sust,
                                       'unit': 'TFLOP/s', 'frac': flop / dev_s / 1e12 / D.world / tf_sust, 'traffic': None,
                                       'note': 'algorithmic fp32 FLOPs (86.1 MFLOP / genotype) over the device time of the decode kernels; the '
                                               'kernel issues 2 fp16 products per fp32 product (one-hot left operand exact in fp16, table split hi + lo)', 'kernel_stats': stats},
                          'e2e': {'value': n_total / wall, 'unit': 'genotypes/s', 'h2d_bytes_per_step': int(bits.nbytes), 'd2h_bytes_per_step': int(tok.nbytes)},
                          'cpu_baseline': cpu, 'clocks': clocks.summary()})
    ae.close()


def eval_leg(D, args):
    """configs[4]: serann_evaluation - `--genotypes` genotypes x 32 re-initialisations, R = 100 (genotypes sharded over the ranks)."""
    import contextlib
    import io
    from serann_b200.codec import RibosomalAutoencoder
    from serann_b200.evaluation import B200SerannEvaluator
    from serann_b200 import generator
    from serann_b200.synthetic import synthetic_dataset, synthetic_decoder_weights
    data = synthetic_dataset()
    K, R = 32, 100
    pool = generator.sample_population(args.genotypes, seed=0, distinct=True)
    mine = list(range(D.rank, len(pool), D.world))
    ae = RibosomalAutoencoder(350, synthetic_decoder_weights(), device=D.local)
    params = dict(PARAMS)
    quiet = io.StringIO()
    with contextlib.redirect_stdout(quiet):
        ev = B200SerannEvaluator(params, data, num_evaluations=K, replications_per_evaluation=R, ribosomal_ae=ae, device=D.local, shuffle_seed=1, nan_offspring='zero')
    np.random.seed(0)
    D.barrier()
    t0 = time.perf_counter()
    accs, rates = [], []
    group = max(1, 96 // K)              # genotypes per engine shard: 3 x 32 copies fill the GPU like a population of 100
    for g0 in range(0, len(mine), group):
        ids = mine[g0:g0 + group]
        with contextlib.redirect_stdout(quiet):
            rs = ev.evaluate_many([data['encodings'][i] for i in ids], [pool[i][0] for i in ids])
        for r in rs:
            accs += list(r['classification_accuracy'])
            rates += [sum(k * v for k, v in c.items()) / max(1, sum(c.values())) for c in r['mutation_rate']]
    D.barrier()
    secs = D.reduce(time.perf_counter() - t0, 'max')
    n = len(pool)
    if D.rank == 0:
        steps = K * n * params['training_epochs'] * 76
        emit({'metric': 'serann_evaluation genotypes/s', 'value': n / secs, 'unit': 'genotypes/s', 'n_gpus': D.world, 'steps': n,
                          'warmup': 0, 'ms_per_step': 1000.0 * secs / n, 'higher_is_better': True, 'scaling': 'strong', 'vs_baseline': None,
                          'dtype': 'f16', 'data': 'synthetic',
                          'config': {'workload': f'configs[4]: {n} genotypes x {K} re-initialisations x 380 steps at batch {BATCH}, R = {R} offspring each, decode + validity; {max(1, 96 // K)} genotypes per engine shard (evaluate_many)', 'nan_offspring': 'NaN loci of overflowing networks read as 0 (the reference decode raises on them)'},
                          'train_steps_per_s': steps / secs,
                          'rank0': {'validation_accuracy_mean': float(np.nanmean(accs)), 'validation_accuracy_std': float(np.nanstd(accs)),
                                    'mutation_rate_mean': float(np.mean(rates))}})
    ev.close()
    ae.close()


if __name__ == '__main__':
    main()
