"""GPU parity: the CUDA engine (through the C ABI) against the CPU oracle on the same seeded inputs.

Tolerances: activations and logits rel 1e-2 of the tensor's largest magnitude (north_star: "rel 1e-2");
bits, mutation counts and token ids bit-exact (bits whose probability is within 2e-3 of 0.5 excepted).
"""
import numpy as np
import pytest
import torch

from oracle import serann_oracle as O
from serann_b200 import frontend
import parity_util as U

pytestmark = pytest.mark.gpu

ACT_TOL = 1e-2
GRAD_TOL = 3e-2


@pytest.fixture(scope='module', params=['tensor_cores', 'simt_only'])
def setup(request):
    """Every engine test runs twice: aligned GEMMs on the tcgen05 kernels (the product default), and everything
    on the general CUDA-core kernels."""
    from serann_b200.engine import set_tensor_cores
    set_tensor_cores(request.param == 'tensor_cores')
    srcs, irs = U.small_population()
    data = U.small_dataset()
    eng, inds = U.make_engine_and_oracle(irs, data)
    U.report([f'#### {request.param}'])
    yield srcs, irs, data, eng, inds
    eng.close()
    set_tensor_cores(True)


def test_parameters_count(setup):
    _, irs, _, eng, _ = setup
    for i, ir in enumerate(irs):
        assert eng.parameters_count(i) == ir.parameters_count


def test_weights_roundtrip(setup):
    _, irs, _, eng, inds = setup
    for i, ind in enumerate(inds):
        got = eng.get_weights(i)
        for a, b in zip(got, ind.export()):
            for x, y in zip(a, b):
                assert np.array_equal(x, y)


def test_layerwise_forward_backward(setup):
    """One training step without the optimizer: every activation, activation gradient and weight gradient."""
    srcs, irs, data, eng, inds = setup
    rows = np.random.default_rng(3).permutation(160)[:48]
    eng.debug_step(rows, apply_adam=False)
    xi, gi = O.prepare_inputs(data['x_train'][rows], data['g_train'][rows])
    labels = torch.tensor(data['y_train'][rows].astype(np.int64))
    bits = torch.tensor(data['g_train'][rows].astype(np.float32))
    worst = 0.0
    lines = []
    for i, ind in enumerate(inds):
        tape, l1 = {}, {}
        O.train_step(ind, xi, gi, labels, bits, 1, tape=tape, apply_adam=False)
        for tid, t in tape.items():
            l = irs[i].layers[tid - 2]
            if l.kind == frontend.KIND_RESHAPE:
                continue
            e = U.relerr(eng.debug_tensor(i, tid, False, 48), t.detach().numpy())
            lines.append(f'net {i} tensor {tid} kind {l.kind} shape {l.out_shape} act relerr {e:.3e}')
            worst = max(worst, e / ACT_TOL)
            # (a concatenation's gradient buffer is also the in-place output-gradient of some of its inputs, which
            #  rewrite their slice during their own backward: it is checked through those inputs instead)
            if t.grad is not None and l.kind != frontend.KIND_CONCAT:
                # the engine rewrites a layer's output gradient in place with the activation backward
                # (TF's ReluGrad / SigmoidGrad output), so that is what the buffer holds afterwards
                ref_g = t.grad.numpy()
                y = t.detach().numpy()
                if l.kind in (frontend.KIND_DENSE, frontend.KIND_CONV) and l.act == frontend.ACT_RELU:
                    ref_g = ref_g * (y > 0)
                elif l.kind in (frontend.KIND_DENSE, frontend.KIND_CONV) and l.act == frontend.ACT_SIGMOID:
                    h = lambda a: a.astype(np.float16).astype(np.float32)
                    ref_g = h(h(ref_g * y) * h(1 - y))
                eg = U.relerr(eng.debug_tensor(i, tid, True, 48), ref_g)
                lines.append(f'net {i} tensor {tid} kind {l.kind} shape {l.out_shape} grad relerr {eg:.3e}')
                worst = max(worst, eg / GRAD_TOL)
                l1[tid] = float(np.abs(ref_g).reshape(-1, ref_g.shape[-1]).sum(axis=0).max())
        for li, pi, p in ind.trainable():
            got_g = eng.debug_grad(i, li, pi)
            ref_w = p.grad.numpy().reshape(irs[i].layers[li].param_shapes()[pi][1])
            ew = U.relerr(got_g, ref_w)
            if pi == 1 and irs[i].layers[li].kind != frontend.KIND_BN:
                # a bias in front of a BatchNormalization has a true gradient of zero: what is left is the
                # cancellation noise of sum(dOut), so measure the error against sum(|dOut|) instead
                ew = float(np.abs(got_g - ref_w).max() / max(np.abs(ref_w).max(), 2e-2 * l1[2 + li] + 1e-12))
            lines.append(f'net {i} layer {li} kind {irs[i].layers[li].kind} param {pi} wgrad relerr {ew:.3e}')
            worst = max(worst, ew / GRAD_TOL)
        # BN moving statistics after one training-mode pass
        got = eng.get_weights(i)
        for li, l in enumerate(irs[i].layers):
            if l.kind == frontend.KIND_BN:
                for which in (2, 3):
                    em = U.relerr(got[li][which], ind.params[li][which].numpy())
                    lines.append(f'net {i} layer {li} BN moving[{which}] relerr {em:.3e}')
                    worst = max(worst, em / 2e-3)
    U.report(['== layerwise =='] + lines)
    bad = [s for s in lines if float(s.split()[-1]) > (ACT_TOL if ' act ' in s else GRAD_TOL if 'grad' in s else 2e-3)]
    assert not bad, '\n'.join(bad)


def test_fit_curves_and_metrics(setup):
    """Two epochs with Adam on both sides: per-step losses, correct counts, validation metrics, final weights."""
    srcs, irs, data, _, _ = setup
    eng, inds = U.make_engine_and_oracle(irs, data)
    try:
        n_val, batch, epochs = 40, 40, 2
        perms = np.stack([np.random.default_rng(6 + e).permutation(160 - n_val) for e in range(epochs)])
        res = eng.fit(epochs, batch, n_val, perms=perms)
        ref = O.fit(inds, data['x_train'], data['y_train'], data['g_train'], list(perms), batch, n_val, record_steps=6)
        lines = []
        for i in range(len(irs)):
            for s in range(6):
                e, st = divmod(s, res['steps_per_epoch'])
                ce, mse = res['train_loss'][e, st, i]
                r = ref[i]['steps'][s]
                lines.append(f'net {i} step {s}: ce {ce:.4f}/{r["ce"]:.4f} mse {mse:.4f}/{r["mse"]:.4f} '
                             f'correct {res["train_correct"][e, st, i]}/{r["correct"]}')
                assert abs(ce - r['ce']) <= 2e-2 * max(1.0, abs(r['ce'])), lines[-1]
                assert abs(mse - r['mse']) <= 2e-2 * max(0.05, abs(r['mse'])), lines[-1]
            w_gpu, w_ref = eng.get_weights(i), inds[i].export()
            for li, (a, b) in enumerate(zip(w_gpu, w_ref)):
                for which, (x, y) in enumerate(zip(a, b)):
                    # fp16 Adam divides by eps + sqrt(v) with v often flushed to 0, which amplifies one-ulp
                    # gradient differences by a whole step: count the elements further than 1e-2 of the tensor's scale, bound the mean
                    frac, em = U.weights_close(x, y)
                    d = float(np.abs(x.astype(np.float32) - y.astype(np.float32)).max())
                    lines.append(f'net {i} layer {li} weight {which} after fit: max diff {d:.2e} off-fraction {frac:.4f} mean {em:.3e}')
                    # two correct fp16-Adam runs are at most 2 * steps * lr apart per element (each step moves an element by
                    # at most about lr, in a direction that is rounding noise where the gradient is); the mean stays tight
                    bound = 2.2 * (epochs * res['steps_per_epoch']) * 1e-3 + 2e-3 * float(np.abs(y.astype(np.float32)).max())
                    if which < 2:
                        assert d <= bound and em < 5e-3, lines[-1]
                    else:                           # BN moving statistics: averages of the batch moments, no optimiser
                        assert frac <= 0.03 and em < 5e-3, lines[-1]
        U.report(['== fit =='] + lines)
    finally:
        eng.close()


def test_inference_logits_and_eval(setup):
    srcs, irs, data, eng, inds = setup
    logits = eng.forward_logits(1, 0, 64, training=False)
    correct, mse_sum = eng.evaluate(1, 0, 64, 32)
    lines = []
    for i, ind in enumerate(inds):
        with torch.no_grad():
            xi, gi = O.prepare_inputs(data['x_test'], data['g_test'])
            ref = O.forward(ind, xi, gi, training=False).numpy()
            ce, mse, corr, _ = O.losses(torch.tensor(ref), torch.tensor(data['y_test'].astype(np.int64)),
                                        torch.tensor(data['g_test'].astype(np.float32)), 10)
        e = U.relerr(logits[i], ref)
        lines.append(f'net {i} inference logits relerr {e:.3e}')
        assert e < ACT_TOL, lines[-1]
        # metrics are tallied per chunk of 32 rows (Keras evaluate batch); allow argmax flips on near-ties
        for c in range(2):
            ref_c = int(corr[c * 32:(c + 1) * 32].sum())
            assert abs(int(correct[c, i]) - ref_c) <= 1, (i, c, correct[c, i], ref_c)
            ref_m = float(mse[c * 32:(c + 1) * 32].sum())
            assert abs(mse_sum[c, i] - ref_m) <= 1e-2 * max(1.0, ref_m)
    U.report(['== inference =='] + lines)


def test_replicate_and_readout(setup):
    srcs, irs, data, eng, inds = setup
    R = 7
    parents = np.random.default_rng(9).integers(0, 2, (len(inds), 100)).astype(np.uint8)
    probs, packed, mut = eng.replicate(parents, R)
    from serann_b200.engine import unpack_bits
    bits = unpack_bits(packed, 100)
    ref = O.replicate(inds, parents, data['x_test'], R)
    for i in range(len(inds)):
        assert U.relerr(probs[i], ref[i]) < ACT_TOL
        # the readout itself is bit-exact on the GPU's own probabilities
        rb, rm = O.readout(probs[i], parents[i])
        assert np.array_equal(bits[i], rb.astype(np.uint8))
        assert np.array_equal(mut[i], rm)
        # and agrees with the oracle's bits except where the probability sits on the threshold
        ob, _ = O.readout(ref[i], parents[i])
        differs = bits[i] != ob.astype(np.uint8)
        near = np.abs(ref[i].astype(np.float32) - 0.5) < 2e-3
        assert not (differs & ~near).any()


def test_readout_kernel_bit_exact():
    from serann_b200.engine import readout
    r = np.random.default_rng(0)
    for rows, rpp, G in [(1, 1, 100), (37, 5, 100), (1000, 10, 100), (64, 8, 33), (9, 3, 160)]:
        p = r.random((rows, G)).astype(np.float16)
        p[r.random((rows, G)) < 0.05] = np.float16(0.5)          # exact ties round to even -> 0
        p[r.random((rows, G)) < 0.02] = np.float16(1.5)          # clipped
        p[r.random((rows, G)) < 0.02] = np.float16(-0.25)
        parents = r.integers(0, 2, (-(-rows // rpp), G)).astype(np.uint8)
        bits, mut = readout(p, parents, rpp)
        rb, rm = O.readout(p, np.repeat(parents, rpp, axis=0)[:rows])
        assert np.array_equal(bits, rb.astype(np.uint8))
        assert np.array_equal(mut, rm)
    # NaN offspring (BatchNormalizationF16 overflow): bit 0, counted as mutated
    p = np.full((2, 100), np.nan, dtype=np.float16)
    bits, mut = readout(p, np.ones((2, 100), np.uint8), 1)
    assert bits.sum() == 0 and (mut == 100).all()


def test_bn_train_forward_kernel():
    from serann_b200.engine import bn_train_forward
    r = np.random.default_rng(1)
    for rows, C in [(48, 7), (3000, 32), (70000, 3), (750, 257)]:
        x = (r.standard_normal((rows, C)) * 2 + 3).astype(np.float16)
        gamma = (1 + 0.1 * r.standard_normal(C)).astype(np.float16)
        beta = (0.1 * r.standard_normal(C)).astype(np.float16)
        mm = (0.1 * r.standard_normal(C)).astype(np.float16)
        mv = np.abs(1 + 0.1 * r.standard_normal(C)).astype(np.float16)
        y, mm2, mv2 = bn_train_forward(x, gamma, beta, mm, mv)
        t = lambda a: torch.tensor(a.astype(np.float32))
        tm, tv = t(mm).clone(), t(mv).clone()
        ref = O._bn_train(t(x), t(gamma), t(beta), tm, tv).numpy()
        assert U.relerr(y, ref) < 2e-3, (rows, C)
        assert U.relerr(mm2, tm.numpy()) < 2e-3
        if rows > 65504:        # fp16 sample size overflows: moving variance becomes NaN (fact F6)
            assert np.isnan(mv2).all() and np.isnan(tv.numpy()).all()
        else:
            assert U.relerr(mv2, tv.numpy()) < 2e-3


def test_decoder_argmax_bit_exact():
    from serann_b200.codec import RibosomalAutoencoder
    from serann_b200.synthetic import synthetic_decoder_weights
    w = synthetic_decoder_weights()
    ae = RibosomalAutoencoder(350, w)
    g = np.random.default_rng(7).integers(0, 2, (300, 100))
    tok = ae.decode_to_sequence(g)
    ref, z = O.ae_decode(g, w, 350, 40)
    top2 = np.sort(z, axis=-1)[..., -2:]
    margin = top2[..., 1] - top2[..., 0]
    differs = tok != ref
    assert not (differs & (margin > 1e-4)).any(), f'{differs.sum()} token mismatches'
    assert differs.mean() < 1e-3
    strings = ae.decode_to_string(g[:3])
    assert len(strings) == 3 and all(isinstance(s, str) for s in strings)
    with pytest.raises(IndexError):
        ae.decode_to_sequence(np.full((1, 100), np.nan))
    ae.close()


@pytest.mark.parametrize('seed', [101, 202, 303])
def test_generated_population_fuzz(seed):
    """Wider net over the generator's architecture space than the fixture's population: eight generated individuals per
    seed, inference logits and the losses of the first training steps against the oracle (tensor-core product path)."""
    from serann_b200 import generator
    from serann_b200.engine import set_tensor_cores
    set_tensor_cores(True)
    pop = generator.sample_population(8, seed=seed, max_flops_per_sample=4.0e6)
    irs = [ir for _, ir in pop]
    data = U.small_dataset(seed=seed)
    eng, inds = U.make_engine_and_oracle(irs, data)
    try:
        logits = eng.forward_logits(1, 0, 64, training=False)
        lines = [f'== fuzz seed {seed} ==']
        for i, ind in enumerate(inds):
            with torch.no_grad():
                xi, gi = O.prepare_inputs(data['x_test'], data['g_test'])
                ref = O.forward(ind, xi, gi, training=False).numpy()
            e = U.relerr(logits[i], ref)
            lines.append(f'net {i} inference logits relerr {e:.3e}')
            assert e < ACT_TOL, (lines[-1], pop[i][0])
        n_val, batch = 40, 40
        perms = np.stack([np.random.default_rng(seed).permutation(160 - n_val)])
        res = eng.fit(1, batch, n_val, perms=perms)
        ref = O.fit(inds, data['x_train'], data['y_train'], data['g_train'], list(perms), batch, n_val, record_steps=3)
        for i in range(len(irs)):
            for s in range(3):
                ce, mse = res['train_loss'][0, s, i]
                r = ref[i]['steps'][s]
                lines.append(f'net {i} step {s}: ce {ce:.4f}/{r["ce"]:.4f} mse {mse:.4f}/{r["mse"]:.4f}')
                if np.isnan(r['ce']) or np.isnan(r['mse']):
                    continue                      # a NaN on the oracle side is the F6 selection pressure; bits are compared elsewhere
                assert abs(ce - r['ce']) <= 3e-2 * max(1.0, abs(r['ce'])), (lines[-1], pop[i][0])
                assert abs(mse - r['mse']) <= 3e-2 * max(0.05, abs(r['mse'])), (lines[-1], pop[i][0])
        U.report(lines)
    finally:
        eng.close()


def test_full_batch_row_independence():
    """Size-independent property at the benchmark's batch (750 rows): inference is row-wise, so the logits of one pass
    over 750 rows equal those of passes over 64-row pieces (other tile counts, split factors and sub-tile groupings)
    up to the fp32 summation order of split reductions; both agree with the oracle on the first 64 rows."""
    from serann_b200 import generator
    from serann_b200.engine import Engine, set_tensor_cores
    set_tensor_cores(True)
    srcs = [U.NET_A, U.NET_E, U.NET_G, U.NET_H] + [s for s, _ in generator.sample_population(4, seed=77, max_flops_per_sample=4.0e6)]
    irs = [frontend.compile_source(s) for s in srcs]
    data = U.small_dataset(n_train=64, n_test=750, seed=9)
    eng, inds = U.make_engine_and_oracle(irs, data, max_batch=750)
    try:
        whole = eng.forward_logits(1, 0, 750, training=False).astype(np.float32)
        pieces = np.concatenate([eng.forward_logits(1, s, min(64, 750 - s), training=False) for s in range(0, 750, 64)], axis=1).astype(np.float32)
        for i in range(len(irs)):
            e = U.relerr(whole[i], pieces[i])
            assert e < 2e-3, (i, e)
            with torch.no_grad():
                xi, gi = O.prepare_inputs(data['x_test'][:64], data['g_test'][:64])
                ref = O.forward(inds[i], xi, gi, training=False).numpy()
            assert U.relerr(whole[i][:64], ref) < ACT_TOL, i
    finally:
        eng.close()
