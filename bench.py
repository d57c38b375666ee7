#!/usr/bin/env python
"""Benchmark of the SeRANN population-evaluation hot path (see BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Metric: SeRANN-train-steps/s = (individuals trained) x (training steps) / seconds, on BASELINE.json
configs[1]: a population of 100 SeRANNs per GPU (generator-sampled sources, seed 0, distinct nets),
training batch 750 on a synthetic 60000/10000 data set of MNIST's shape.  One bench "step" is one
training step of every individual of the shard on one mini-batch (gather, forward, losses, backward, Adam).

  value  device-timed throughput with the data set resident in HBM (CUDA events on the engine's stream)
  e2e    the same metric through the drop-in worker API (B200PopulationWorker.run: host data set in,
         host results out, a whole generation: build + 5 epochs + validation + test + replication)
  --impl reference   the CPU restatement of the reference's TF path (oracle/) on the host cores
"""
import argparse
import json
import os
import pathlib
import subprocess
import sys
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
    ap.add_argument('--pop', type=int, default=POP_PER_GPU, help='individuals per GPU')
    ap.add_argument('--mix', default='diverse', choices=['diverse', 'clonal'])
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--no-cpu', action='store_true')
    return ap.parse_args()


def population(n_total, mix):
    from serann_b200 import generator
    return generator.sample_population(n_total, seed=0, distinct=(mix == 'diverse'))


def shard(pop, world, rank):
    """Cost-balanced shards (longest-processing-time first on the FLOP estimate), same size per rank."""
    order = sorted(range(len(pop)), key=lambda i: -pop[i][1].train_flops_per_sample())
    loads, bins = [0.0] * world, [[] for _ in range(world)]
    cap = -(-len(pop) // world)
    for i in order:
        r = min((k for k in range(world) if len(bins[k]) < cap), key=lambda k: loads[k])
        bins[r].append(i)
        loads[r] += pop[i][1].train_flops_per_sample()
    return sorted(bins[rank])


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
        return d.get('hbm_gbs', 6650.0), d.get('bf16_tflops_sustained', 1400.0), d.get('bf16_tflops', 1590.0), 'measured'
    return 6650.0, 1400.0, 1590.0, 'fallback'


def gemm_flops_per_step(irs, batch):
    """Algorithmic FLOPs of one training step per GEMM pass."""
    fwd = dgrad = 0
    for ir in irs:
        needs = ir.needs_input_grad()
        for l, nd in zip(ir.layers, needs):
            m = 2 * l.macs_per_sample() * batch
            fwd += m
            dgrad += m if nd else 0
    return {'gemm_fwd': fwd, 'gemm_wgrad': fwd, 'gemm_dgrad': dgrad}


def hbm_bytes_per_step(irs, batch):
    """Algorithmic bytes of the HBM-bound kernels."""
    from serann_b200 import frontend as F
    adam = sum(ir.parameters_count for ir in irs) * 16          # p, m, v fp16 r+w (12) + fp32 grad read (4); re-zeroing adds 4
    bn = sum(l.out_elems * batch * (4 + 6) for ir in irs for l in ir.layers if l.kind == F.KIND_BN)
    return {'adam': adam, 'bn': bn}


def cpu_oracle_rate(pop_sample, steps, warmup, threads):
    """Oracle training steps per second on the host cores for a bounded sample of the workload."""
    import torch
    from oracle import serann_oracle as O
    from serann_b200.synthetic import synthetic_dataset
    torch.set_num_threads(threads)
    data = synthetic_dataset(n_train=BATCH * (steps + warmup), n_test=8)
    rng = np.random.default_rng(5)
    inds = [O.Individual(ir, O.init_weights(ir, rng)) for _, ir in pop_sample]
    t_total, done, s = 0.0, 0, 0
    n_rows = len(data['x_train'])
    # at least `steps` timed steps, and keep going until ~10 s of CPU work so that the rate is stable
    while s < steps + warmup or (t_total < 10.0 and s < 200):
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
            done += len(inds)
        s += 1
    return done / t_total, t_total


def run_reference(args):
    """--impl reference: the CPU restatement of the reference's TF path, all host threads, bounded sample."""
    rank = int(os.environ.get('RANK', 0))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    pop = population(args.pop, args.mix)
    sample = pop[::max(1, len(pop) // 16)][:16]      # evenly spaced through the population
    steps = max(1, min(args.steps, 3))
    warm = 1 if args.warmup > 0 else 0
    rate, secs = cpu_oracle_rate(sample, steps, warm, threads)
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': rate, 'unit': 'steps/s', 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1000.0 * secs / steps,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f16', 'data': 'synthetic',
        'config': workload_config(args, len(sample)),
        'cpu_baseline': {'value': rate, 'unit': 'steps/s', 'cores': threads, 'kind': 'port',
                         'sample': f'{len(sample)} individuals evenly spaced through the workload x >= {steps} training steps (>= 10 s) at batch {BATCH} '
                                   f'(torch CPU oracle; TensorFlow 2.6 is not installable here)'},
        'e2e': {'value': rate, 'unit': 'steps/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    }
    print(json.dumps(line))


def workload_config(args, n_ind):
    return {'workload': f'configs[1]: population {args.pop} SeRANNs per GPU ({args.mix} generator mix, seed 0, the same mix on every GPU), '
                        f'batch {BATCH}, 60000/10000 synthetic rows of 28x28 + 100-bit genotypes',
            'population_per_gpu': n_ind, 'batch': BATCH, 'mix': args.mix,
            'l2_policy': 'working set per step (activation arena, GBs) is far larger than the 126 MB L2',
            'parallelism': f'population sharded over {args.gpus} GPU(s), no data-path collective in training'}


def main():
    args = parse()
    if args.impl == 'reference':
        run_reference(args)
        return
    import torch
    import torch.distributed as dist
    world = int(os.environ.get('WORLD_SIZE', 1))
    rank = int(os.environ.get('RANK', 0))
    local = int(os.environ.get('LOCAL_RANK', 0))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a CUDA device: the engine has no CPU path')
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    from serann_b200 import engine as E
    from serann_b200.synthetic import synthetic_dataset
    from serann_b200.worker import B200PopulationWorker

    # weak scaling: every GPU evaluates its own args.pop individuals drawn from the same architecture mix (independent
    # initialisations), so the per-GPU work is the same at every N; the job's population is args.pop * world
    pop = population(args.pop, args.mix)
    mine = list(range(len(pop)))
    irs = [pop[i][1] for i in mine]
    data = synthetic_dataset()
    n_val = 3000

    # ---- device-resident throughput -------------------------------------------------------------
    eng = E.Engine(max_batch=BATCH, device=local)
    eng.set_dataset(data['x_train'], data['y_train'], data['g_train'], data['x_test'], data['y_test'], data['g_test'])
    eng.build(irs, seed=7)
    eng.fit(1, BATCH, n_val, shuffle_seed=1, max_steps=max(args.warmup, 3))
    barrier()
    launches0 = E.launch_count()
    with ClockSampler(local) as clocks:
        t0 = time.perf_counter()
        res = eng.fit(1, BATCH, n_val, shuffle_seed=2, max_steps=args.steps)
        barrier()
        wall = time.perf_counter() - t0
    launches = E.launch_count() - launches0
    dev_s = max_over_ranks(res['device_ms'] / 1000.0)
    total_ind = sum_over_ranks(len(irs))
    value = total_ind * args.steps / dev_s

    # ---- per-category device time -> roofline of the dominant kernel -----------------------------
    eng.profile(True)
    eng.fit(1, BATCH, n_val, shuffle_seed=3, max_steps=2)
    prof = eng.profile_read()
    eng.profile(False)
    hbm_peak, tf_sust, tf_burst, peak_kind = peaks()
    flops = gemm_flops_per_step(irs, BATCH)
    step_ms = sum(m for m, _ in prof.values()) / 2.0
    shares = {k: (m / 2.0) / step_ms for k, (m, c) in prof.items() if c}
    top = max((k for k in flops), key=lambda k: prof[k][0])
    top_ms, top_launches = prof[top][0] / 2.0, prof[top][1] // 2
    achieved = flops[top] / (top_ms / 1000.0) / 1e12
    traffic = None
    tpath = ROOT / 'profiles' / 'r01_dram_traffic_per_step.json'
    if tpath.exists():      # dram__bytes_read + write per step of each kernel category, from the committed ncu capture
        traffic = json.loads(tpath.read_text()).get(top)
    roofline = {'bound': 'tensor', 'kernel': top, 'achieved': achieved, 'peak': tf_sust, 'unit': 'TFLOP/s',
                'frac': achieved / tf_sust, 'traffic': traffic, 'peak_source': f'{peak_kind} sustained bf16/fp16 dense',
                'launches_per_step': top_launches, 'ms_per_step_in_kernel': top_ms,
                'timing': 'CUDA events around every launch of the category, profiled with the launches serialised on one stream',
                'share_of_step': shares,
                'note': 'population-grouped implicit GEMM on tcgen05 (fp16 operands, fp32 TMEM accumulators, cp.async-gathered '
                        'operands); achieved = algorithmic FLOPs of this pass over all individuals / its event-timed device time. '
                        'The problems are small-N and gather-bound, far from the dense-GEMM peak: see DESIGN.md section 4'}
    roofline['step'] = {'algorithmic_TFLOP_per_step': sum(flops.values()) / 1e12,
                        'achieved_TFLOPs': sum(flops.values()) / (dev_s / args.steps) / 1e12 * (total_ind / len(irs)),
                        'frac_of_tensor_peak': sum(flops.values()) / (dev_s / args.steps) / 1e12 * (total_ind / len(irs)) / (tf_sust * world),
                        'note': 'all GEMM passes of one step over the measured step time (launches overlap on several streams)'}
    hb = hbm_bytes_per_step(irs, BATCH)
    roofline['hbm_kernels'] = {k: {'achieved_GBs': hb[k] / (prof[k][0] / 2.0 / 1000.0) / 1e9 if prof[k][1] else None,
                                   'peak_GBs': hbm_peak, 'frac': (hb[k] / (prof[k][0] / 2.0 / 1000.0) / 1e9 / hbm_peak) if prof[k][1] else None}
                               for k in ('adam', 'bn')}
    eng.close()

    # ---- end to end through the drop-in worker API ------------------------------------------------
    e2e = None
    if not args.no_e2e:
        import pandas as pd
        ids = [f'serann-{rank}-{i}' for i in mine]
        info = pd.DataFrame({'genotype': [data['encodings'][i] for i in mine], 'source_code': [pop[i][0] for i in mine]}, index=ids)
        barrier()
        t0 = time.perf_counter()
        worker = B200PopulationWorker(dict(PARAMS), dataset=data, device=local, shuffle_seed=4)
        out = worker.run(info, PARAMS['initial_offspring_pool_size'])
        fitness = out['models_info']['classification_validation_accuracy'].to_numpy(dtype=np.float32)
        offspring = np.stack([out['offspring_by_id'][i] for i in ids])
        d2h = fitness.nbytes + offspring.nbytes + out['models_info'].to_numpy(dtype=np.float64).nbytes
        if world > 1:       # the one exchange step of a generation: fitness + offspring bits for selection
            from serann_b200.parallel import allgather_generation
            allgather_generation(fitness, (offspring > 0.5), device=torch.device('cuda', local))
        barrier()
        e2e_s = max_over_ranks(time.perf_counter() - t0)
        steps_e2e = worker.last_stats['steps']
        h2d = worker._engine.h2d_bytes
        worker.close()
        e2e = {'value': total_ind * steps_e2e / e2e_s, 'unit': 'steps/s',
               'h2d_bytes_per_step': h2d / steps_e2e, 'd2h_bytes_per_step': d2h / steps_e2e,
               'seconds_per_generation': e2e_s, 'steps_per_generation': steps_e2e,
               'what': 'B200PopulationWorker(params, host dataset).run(DataFrame, 10): H2D of the data set, build, '
                       '5 epochs x 76 steps, validation, test, replication, readout, D2H of metrics and offspring'}

    # ---- CPU baseline (rank 0, N = 1) ---------------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        sample = [pop[i] for i in mine[::max(1, len(mine) // 16)][:16]]
        rate, secs = cpu_oracle_rate(sample, 2, 1, threads)
        cpu = {'value': rate, 'unit': 'steps/s', 'cores': threads, 'kind': 'port',
               'sample': f'{len(sample)} individuals evenly spaced through the workload x training steps at batch {BATCH} for {secs:.1f} s (torch CPU oracle)'}

    if rank == 0:
        line = {'metric': METRIC, 'value': value, 'unit': 'steps/s', 'n_gpus': world, 'steps': args.steps,
                'warmup': max(args.warmup, 3), 'ms_per_step': 1000.0 * dev_s / args.steps, 'higher_is_better': True,
                'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f16', 'data': 'synthetic',
                'config': workload_config(args, len(irs)), 'roofline': roofline, 'cpu_baseline': cpu, 'e2e': e2e,
                'gpu_launches': int(launches), 'clocks': clocks.summary(), 'wall_s_timed_region': wall}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
