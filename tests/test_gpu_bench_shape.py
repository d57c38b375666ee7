"""GPU parity at the shape the benchmark runs (training batch 750, the default multi-stream schedule), and the
evaluator workload (K re-initialisations of one architecture) against the CPU oracle in distribution.

What only exists at B = 750 and is therefore only checked here: weight gradients split over 4096-row pieces with
fp32 atomics, multi-chunk BatchNorm with the last-CTA finalize, the fp16 sample-size overflow inside the engine
(B * H * W >= 65520 -> NaN moving variance), forward split-K under training, the floating
weight-gradient streams.
"""
from collections import Counter

import numpy as np
import pytest
import torch

from oracle import serann_oracle as O
from serann_b200 import frontend, generator
import parity_util as U

pytestmark = pytest.mark.gpu

BATCH = 750
BAND = 2e-3          # bits whose probability lies this close to 0.5 are exempt from bit-exactness (north_star)


def test_training_steps_at_batch_750_match_the_oracle():
    """Eight individuals drawn from the bench population (every 12th of generator seed 0) plus NET_A (BatchNorm on a
    13 x 13 map: 750 * 169 elements per channel overflow the fp16 sample size): three Adam steps at B = 750 through
    srn_fit with explicit permutations, then one validation pass; per-step CE / MSE, correct counts, BN moving
    statistics and weights against O.fit."""
    pop = generator.sample_population(100, seed=0, distinct=True)[::12]
    srcs = [s for s, _ in pop] + [U.NET_A]
    irs = [ir for _, ir in pop] + [frontend.compile_source(U.NET_A)]
    steps = 3
    n_val = 300
    data = U.small_dataset(n_train=steps * BATCH + n_val, n_test=64, seed=21)
    eng, inds = U.make_engine_and_oracle(irs, data, max_batch=BATCH)
    try:
        lines = ['== B = 750 gradient parity (one step, no optimiser) ==']
        # ---- the strict check: fp32 gradient accumulators of one 750-row step against the oracle's, every weight.
        # The BN moving statistics move in this step on both sides and are compared after the fit below.
        # A ReLU whose fp16 pre-activation is exactly zero on one side and one ulp above on the other (h(acc) == -bias
        # happens a few times per 10^5 elements) opens its gate on one side only; behind a BatchNorm over a nearly dead
        # unit that single element can carry a tenth of a weight gradient's largest entry.  That is not a property of
        # either implementation, so a net with such rows is compared again on the batch without them.
        rows = np.arange(BATCH)
        todo = list(range(len(inds)))
        for attempt in range(3):
            eng.debug_step(rows, apply_adam=False)
            xi, gi = O.prepare_inputs(data['x_train'][rows], data['g_train'][rows])
            labels = torch.tensor(data['y_train'][rows].astype(np.int64))
            bits_t = torch.tensor(data['g_train'][rows].astype(np.float32))
            bad, flipped_rows, again = [], set(), []
            for i in todo:
                ind = inds[i]
                tape = {}
                O.train_step(ind, xi, gi, labels, bits_t, 1, tape=tape, apply_adam=False)
                flips = set()
                for tid, t in tape.items():
                    l = irs[i].layers[tid - 2]
                    if l.kind in (frontend.KIND_DENSE, frontend.KIND_CONV) and l.act == frontend.ACT_RELU:
                        y_gpu = eng.debug_tensor(i, tid, False, len(rows)).astype(np.float32)
                        diff = ((y_gpu > 0) != (t.detach().numpy() > 0)).reshape(len(rows), -1).any(axis=1)
                        flips |= set(np.where(diff)[0].tolist())
                if flips:
                    lines.append(f'net {i}: ReLU gates differ in batch rows {sorted(flips)}: compared again without them')
                    flipped_rows |= flips
                    again.append(i)
                kscale = {}
                for li, pi, prm in ind.trainable():
                    g_ref = prm.grad.numpy()
                    g_gpu = eng.debug_grad(i, li, pi)
                    g_ref = g_ref.reshape(g_gpu.shape)
                    # a bias that feeds a BatchNorm has a zero true gradient (the batch mean is subtracted again): both
                    # sides hold rounding noise there, so a bias is judged on the scale of its layer's kernel gradient
                    scale = float(np.abs(g_ref).max())
                    if pi == 0:
                        kscale[li] = scale
                    else:
                        scale = max(scale, kscale.get(li, 0.0))
                    e = float(np.abs(g_gpu.astype(np.float64) - g_ref).max() / (scale + 1e-12))
                    prm.grad = None
                    if flips:
                        continue
                    lines.append(f'net {i} layer {li} weight {pi} grad relerr {e:.3e} (max |g| {np.abs(g_ref).max():.3e}, {len(rows)} rows)')
                    if not e < 3e-2:
                        bad.append(lines[-1])
            assert not bad, '\n'.join(bad)
            todo = again
            if not todo:
                break
            rows = np.delete(rows, sorted(flipped_rows))
        assert not todo, 'ReLU gate differences persisted over three batches'
        U.report(lines)
        # ---- three Adam steps and a validation pass from the same fresh state on both sides
        eng.close()
        eng, inds = U.make_engine_and_oracle(irs, data, max_batch=BATCH)
        perms = np.stack([np.random.default_rng(31).permutation(steps * BATCH)])
        res = eng.fit(1, BATCH, n_val, perms=perms)
        torch.set_num_threads(max(1, torch.get_num_threads()))
        ref = O.fit(inds, data['x_train'], data['y_train'], data['g_train'], list(perms), BATCH, n_val, record_steps=steps)
        lines = ['== B = 750 training parity (3 Adam steps + validation) ==']
        nan_bn_seen = False
        for i, ir in enumerate(irs):
            for s in range(steps):
                ce, mse = res['train_loss'][0, s, i]
                r = ref[i]['steps'][s]
                lines.append(f'net {i} step {s}: ce {ce:.4f}/{r["ce"]:.4f} mse {mse:.4f}/{r["mse"]:.4f} '
                             f'correct {res["train_correct"][0, s, i]}/{r["correct"]}')
                assert abs(ce - r['ce']) <= 2e-2 * max(1.0, abs(r['ce'])), (lines[-1], srcs[i])
                assert abs(mse - r['mse']) <= 2e-2 * max(0.05, abs(r['mse'])), (lines[-1], srcs[i])
                assert abs(int(res['train_correct'][0, s, i]) - r['correct']) <= 8, (lines[-1], srcs[i])
            w_gpu, w_ref = eng.get_weights(i), inds[i].export()
            for li, (l, a, b) in enumerate(zip(ir.layers, w_gpu, w_ref)):
                for which, (x, y) in enumerate(zip(a, b)):
                    frac, mean = U.weights_close(x, y)
                    lines.append(f'net {i} layer {li} kind {l.kind} weight {which}: off-fraction {frac:.4f} mean {mean:.2e}')
                    if l.kind == frontend.KIND_BN and which == 3:
                        n_elems = BATCH * int(np.prod(l.in_shape[:-1]))
                        if n_elems >= 65520:           # F6 inside the engine: NaN on both sides
                            nan_bn_seen = True
                            assert np.isnan(x.astype(np.float32)).all() and np.isnan(y.astype(np.float32)).all(), lines[-1]
                            continue
                        assert np.isfinite(x.astype(np.float32)).all(), lines[-1]
                    # Adam normalises the step: an element whose fp16 gradient is rounding noise moves by about lr per step
                    # in a direction that differs between two correct implementations, so after `steps` steps two
                    # correct weights are at most 2 * steps * lr apart (plus rounding); the gradients themselves were
                    # compared above
                    d = np.abs(np.asarray(x, np.float32) - np.asarray(y, np.float32))
                    bound = 2.2 * steps * 1e-3 + 2e-3 * float(np.abs(np.asarray(y, np.float32)).max())
                    if l.kind != frontend.KIND_BN or which < 2:
                        assert float(d.max()) <= bound, (lines[-1], float(d.max()), srcs[i])
                    else:                              # moving statistics: plain averages, no optimiser in between
                        assert frac <= 0.02 and mean <= 3e-3, (lines[-1], srcs[i])
            # validation pass in inference mode (moving statistics): NaN individuals classify nothing
            vc, vm = int(res['val_correct'][0, :, i].sum()), float(res['val_mse_sum'][0, :, i].sum())
            lines.append(f'net {i} validation correct {vc} mse_sum {vm:.3f} oracle acc {ref[i]["classification_validation_accuracy"]:.4f} '
                         f'mse {ref[i]["replication_mse"]}')
            assert abs(vc / n_val - ref[i]['classification_validation_accuracy']) <= 0.03, lines[-1]
            if np.isnan(ref[i]['replication_mse']):
                assert np.isnan(vm), lines[-1]
            else:
                assert abs(vm / n_val - ref[i]['replication_mse']) <= 2e-2 * max(0.05, ref[i]['replication_mse']), lines[-1]
        assert nan_bn_seen, 'the population must hold a BatchNorm over >= 65520 elements per channel'
        # replication after the NaN: every probability of the NaN individual is NaN -> bit 0, 100 mutations, both sides
        # (same weights on both sides first: three noisy Adam steps apart, probabilities near 0.5 would differ)
        for i in range(len(irs)):
            eng.set_weights(i, inds[i].export())
        parents = data['g_test'][:len(irs)]
        probs, packed, mut = eng.replicate(parents, 4)
        ref_probs = O.replicate(inds, parents, data['x_test'], 4)
        from serann_b200.engine import unpack_bits
        bits = unpack_bits(packed, 100)
        for i in range(len(irs)):
            rb, rm = O.readout(ref_probs[i], parents[i])
            sure = np.abs(ref_probs[i].astype(np.float32) - 0.5) > BAND
            sure |= np.isnan(ref_probs[i].astype(np.float32))
            got = bits[i].astype(np.float32)
            want = np.where(np.isnan(rb.astype(np.float32)), 0, rb.astype(np.float32))
            assert np.array_equal(got[sure], want[sure]), i
            if np.isnan(ref_probs[i].astype(np.float32)).all():
                assert (mut[i] == 100).all() and (rm == 100).all()
        U.report(lines)
    finally:
        eng.close()


EVAL_NETS = [U.NET_B, """g_layer = Conv1D(filters=8, kernel_size=5, strides=2)(g_layer)
g_layer = Dense(units=12, activation='relu')(g_layer)
X_layer = Conv2D(filters=8, kernel_size=5, strides=3, activation='relu')(X_layer)
con = concatenate([Reshape((1, -1))(X_layer), Reshape((1, -1))(g_layer)])
con = Dense(units=48, activation='sigmoid')(con)

loss_balance = 0.3"""]


@pytest.mark.parametrize('which', [0, 1])
def test_evaluator_k32_matches_the_oracle_in_distribution(which):
    """Row a14 (serann_evaluation/logic.py:153-231): K = 32 independently initialised copies of one architecture, the
    same initial weights and row order on both sides.  The K validation accuracies (fertility) agree to one sample,
    almost all to one fp16 ulp; the offspring probabilities agree to ~1e-3 on average, no bit differs outside ten
    threshold bands, and the mutation-rate statistics match in distribution (pooled histogram within 0.1 total
    variation, per-copy mean rate within one locus in a hundred)."""
    from serann_b200.codec import RibosomalAutoencoder
    from serann_b200.evaluation import B200SerannEvaluator
    from serann_b200.synthetic import synthetic_decoder_weights
    K, R = 32, 20
    params = {'genotype_size': 100, 'classification_image_dimensions': [28, 28], 'num_classification_classes': 10,
              'training_epochs': 2, 'training_batch_size': 40, 'error_correction_probability': 0}
    data = U.small_dataset(n_train=200, n_test=64, seed=40 + which)
    src = EVAL_NETS[which]
    ir = frontend.compile_source(src)
    rng = np.random.default_rng(8)
    weights = [O.init_weights(ir, rng) for _ in range(K)]
    inds = [O.Individual(ir, w) for w in weights]
    n_fit = int(200 * 0.95)
    n_val = 200 - n_fit
    perms = np.stack([np.random.default_rng(50 + e).permutation(n_fit) for e in range(2)])
    genotype = data['g_test'][3]
    dec_w = synthetic_decoder_weights()
    ae = RibosomalAutoencoder(350, dec_w)
    ev = B200SerannEvaluator(params, data, num_evaluations=K, replications_per_evaluation=R, ribosomal_ae=ae, shuffle_seed=1)
    try:
        np.random.seed(3)
        got = ev.evaluate(genotype, source_code=src, initial_weights=weights, perms=perms)
        probs_gpu = ev.last_probs.astype(np.float32)                                     # (K, R, 100)
        # ---- the oracle's evaluate
        ref = O.fit(inds, data['x_train'], data['y_train'], data['g_train'], list(perms), 40, n_val)
        acc_ref = [r['classification_validation_accuracy'] for r in ref]
        probs_ref = []
        with torch.no_grad():
            xi, gi = O.prepare_inputs(data['x_test'][:R], np.repeat(genotype[None], R, axis=0))
            for ind in inds:
                logits = O.forward(ind, xi, gi, training=False)
                probs_ref.append(O.sigmoid_q(logits[:, 10:]).numpy())
        probs_ref = np.stack(probs_ref)
        # fertility: validation accuracies
        acc_gpu = got['classification_accuracy']
        one_sample = 1.0 / n_val + 1e-3
        diffs = np.abs(np.asarray(acc_gpu) - np.asarray(acc_ref))
        assert (diffs <= one_sample).all(), (acc_gpu, acc_ref)
        assert (diffs <= 2.0 ** -11).sum() >= K - 3, diffs          # to one fp16 ulp for all but a few near-tie rows
        # offspring: the two sides are ten noisy fp16 Adam steps apart (see the weight bound in the test above), so the
        # probabilities agree closely but not to the threshold band; what must agree is the distribution
        pdiff = np.abs(probs_gpu - probs_ref)
        sure = np.abs(probs_ref - 0.5) > 10 * BAND
        bits_gpu = np.round(np.clip(probs_gpu, 0, 1))
        bits_ref = np.round(np.clip(probs_ref, 0, 1))
        bit_mismatch = float((bits_gpu[sure] != bits_ref[sure]).mean())
        flat_gpu, flat_ref = bits_gpu.reshape(K * R, -1), bits_ref.reshape(K * R, -1)
        mut_gpu = (flat_gpu != genotype).sum(axis=1).reshape(K, R)
        mut_ref = (flat_ref != genotype).sum(axis=1).reshape(K, R)
        # pooled mutation-count histogram: total-variation distance, and the per-copy mean mutation rate
        top = int(max(mut_gpu.max(), mut_ref.max())) + 1
        h_gpu = np.bincount(mut_gpu.ravel(), minlength=top) / mut_gpu.size
        h_ref = np.bincount(mut_ref.ravel(), minlength=top) / mut_ref.size
        tv = 0.5 * float(np.abs(h_gpu - h_ref).sum())
        rate_diff = np.abs(mut_gpu.mean(axis=1) - mut_ref.mean(axis=1)) / 100.0
        stats = (f'prob diff mean {pdiff.mean():.2e} max {pdiff.max():.2e}; bits differing outside 10x band {bit_mismatch:.2e}; '
                 f'mutation histogram TV {tv:.3f}; per-copy mutation-rate diff max {rate_diff.max():.4f} '
                 f'(rates around {mut_ref.mean() / 100:.3f})')
        U.report([f'== evaluator K={K} net {which}: {stats}'])
        assert pdiff.mean() < 5e-3 and bit_mismatch < 5e-3, stats
        assert tv < 0.1 and rate_diff.max() < 0.01, stats
        # the reference's own histogram rule ([i : i + R] slice of the K * R offspring, logic.py:223) on the engine's bits
        for i in range(K):
            want = dict(Counter(np.sum(flat_gpu[i: i + R] != genotype, axis=1) / 100))
            assert got['mutation_rate'][i] == want, i
        # decode + validity: every offspring accounted for, token sequences equal to the fp32 oracle's where the bits agree
        assert all(sum(c.values()) == R for c in got['offspring_survival'])
        same = np.where((flat_gpu == flat_ref).all(axis=1))[0][:40]
        tok_gpu = ae.decode_to_sequence(flat_gpu[same])
        tok_ref, z = O.ae_decode(flat_ref[same], dec_w, 350, 40)
        top2 = np.sort(z, axis=-1)[..., -2:]
        clear = (top2[..., 1] - top2[..., 0]) > 1e-4
        assert np.array_equal(tok_gpu[clear], tok_ref[clear])
        U.report([f'== evaluator K={K} net {which}: accuracy diffs max {diffs.max():.4f}, exact {int((diffs == 0).sum())}/{K}'])
    finally:
        ev.close()
        ae.close()


def test_no_kernel_writes_past_its_buffers_at_batch_750():
    """compute-sanitizer is closed on the GPU pool, so the arenas carry guard bands instead (SRN_GUARD=1, a fresh process
    because the switch is read once): build, three training steps at B = 750, validation, test evaluation and
    replication of a slice of the bench population, then every band must still hold its pattern."""
    import subprocess
    import sys
    code = '''
import sys, numpy as np
sys.path.insert(0, "tests")
import parity_util as U
from serann_b200 import generator
from serann_b200.engine import Engine
pop = generator.sample_population(100, seed=0, distinct=True)[::9]
data = U.small_dataset(n_train=3 * 750 + 300, n_test=96, seed=3)
eng = Engine(max_batch=750)
eng.set_dataset(data["x_train"], data["y_train"], data["g_train"], data["x_test"], data["y_test"], data["g_test"])
eng.build([ir for _, ir in pop], seed=2)
assert eng._lib.srn_debug_check_guards(eng._h) == 0
eng.fit(1, 750, 300, shuffle_seed=1)
eng.evaluate(1, 0, 96, 32)
eng.replicate(data["g_test"][:len(pop)], 7)
rc = eng._lib.srn_debug_check_guards(eng._h)
eng.build([ir for _, ir in pop[:5]], seed=3)          # a smaller population in the same (grow-only) buffers
eng.fit(1, 750, 300, shuffle_seed=2, max_steps=2)
rc2 = eng._lib.srn_debug_check_guards(eng._h)
print("GUARDS", rc, rc2)
'''
    import os
    import pathlib
    root = pathlib.Path(__file__).resolve().parent.parent
    r = subprocess.run([sys.executable, '-c', code], cwd=root, env=dict(os.environ, SRN_GUARD='1', PYTHONPATH=str(root)),
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert 'GUARDS 0 0' in r.stdout, r.stdout[-500:]


def test_evaluate_many_equals_separate_evaluations():
    """Two genotypes in one engine shard (evaluate_many) against the same two evaluated one after the other with the same
    initial weights and row order: the copies of different genotypes never interact, so accuracies, mutation-rate
    counters and survival counts are identical, and an invalid source in the batch gets the empty result."""
    from serann_b200.codec import RibosomalAutoencoder
    from serann_b200.evaluation import B200SerannEvaluator, empty_result
    from serann_b200.synthetic import synthetic_decoder_weights
    K, R = 4, 8
    params = {'genotype_size': 100, 'classification_image_dimensions': [28, 28], 'num_classification_classes': 10,
              'training_epochs': 1, 'training_batch_size': 40, 'error_correction_probability': 0}
    data = U.small_dataset(n_train=200, n_test=64, seed=77)
    irs = [frontend.compile_source(s) for s in EVAL_NETS]
    rng = np.random.default_rng(9)
    weights = [[O.init_weights(ir, rng) for _ in range(K)] for ir in irs]
    perms = np.stack([np.random.default_rng(60).permutation(int(200 * 0.95))])
    genotypes = [data['g_test'][1], data['g_test'][2]]
    ae = RibosomalAutoencoder(350, synthetic_decoder_weights())
    ev = B200SerannEvaluator(params, data, num_evaluations=K, replications_per_evaluation=R, ribosomal_ae=ae, shuffle_seed=1)
    try:
        single = [ev.evaluate(genotypes[i], source_code=EVAL_NETS[i], initial_weights=weights[i], perms=perms) for i in range(2)]
        many = ev.evaluate_many([genotypes[0], genotypes[0], genotypes[1]], [EVAL_NETS[0], 'con = oops(', EVAL_NETS[1]],
                                initial_weights=weights[0] + weights[1], perms=perms)
        assert many[1] == empty_result()
        for got, want in ((many[0], single[0]), (many[2], single[1])):
            assert got['classification_accuracy'] == want['classification_accuracy']
            assert got['mutation_rate'] == want['mutation_rate']
            assert got['offspring_survival'] == want['offspring_survival']
    finally:
        ev.close()
        ae.close()


def test_slab_convolution_kernel_passes_the_parity_tests():
    """The opt-in forward path for stride-1 convolutions (csrc/srn_conv_slab.cuh: one slab of input rows per tile, window
    taps as shifted shared-memory views) under the same layerwise / step / fit parity tests as the default path.  The switch
    is read once per process, hence the subprocess."""
    import os
    import subprocess
    import sys
    env = dict(os.environ, SRN_SLAB='1')
    r = subprocess.run([sys.executable, '-m', 'pytest', 'tests/test_gpu_parity.py', '-x', '-q', '-p', 'no:cacheprovider'],
                       cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))), env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert 'passed' in r.stdout
