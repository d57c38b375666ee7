"""Autoencoder decode (row a13, BASELINE configs[3]) through the C ABI: the tcgen05 kernel with its fused epilogue against
the fp32 CPU oracle (ribosomal_autoencoder/model.py:35-48, evolutionary_experiment/logic/ribosomal_autoencoder.py:122-124)
and against the library's own exact kernel (every token recomputed from the fp32 table in double)."""
import os

import numpy as np
import pytest

from oracle import serann_oracle as O

pytestmark = pytest.mark.gpu


def _decoders(weights):
    from serann_b200.codec import RibosomalAutoencoder
    fast = RibosomalAutoencoder(350, weights)
    os.environ['SRN_DEC_EXACT'] = '1'
    try:
        exact = RibosomalAutoencoder(350, weights)
    finally:
        del os.environ['SRN_DEC_EXACT']
    return fast, exact


def test_tensor_core_decode_equals_the_exact_kernel_and_the_oracle():
    from serann_b200.synthetic import synthetic_decoder_weights
    w = synthetic_decoder_weights()
    fast, exact = _decoders(w)
    try:
        r = np.random.default_rng(11)
        g = r.integers(0, 2, (5000, 100)).astype(np.uint8)       # 19.5 row tiles: the last one is ragged
        g[0] = 0
        g[1] = 1
        tok = fast.decode_to_sequence(g)
        assert fast.last_stats['tensor_core_kernel'] and not False
        flagged = fast.last_stats['rechecked_tokens']
        ref_exact = exact.decode_to_sequence(g)
        assert not exact.last_stats['tensor_core_kernel']
        # the margin recheck makes the tensor-core result the exact arithmetic's argmax: no mismatch at all
        assert np.array_equal(tok, ref_exact), f'{(tok != ref_exact).sum()} of {tok.size} tokens differ from the exact kernel'
        assert 0 < flagged < 0.05 * tok.size
        # the CPU oracle (fp32 chain, its own rounding): equal wherever its top-two margin is clear
        m = 600
        ref, z = O.ae_decode(g[:m], w, 350, 40)
        top2 = np.sort(z, axis=-1)[..., -2:]
        margin = top2[..., 1] - top2[..., 0]
        differs = tok[:m] != ref
        assert not (differs & (margin > 1e-4)).any(), f'{differs.sum()} token mismatches, worst margin {margin[differs].max()}'
        assert differs.mean() < 1e-3
        # edge cases the reference has: one genotype, empty batch, strings with the rstrip quirk
        assert np.array_equal(fast.decode_to_sequence(g[7:8]), tok[7:8])
        assert fast.decode_to_sequence(np.zeros((0, 100), np.uint8)).shape == (0, 350)
        s = fast.decode_to_string(g[:3])
        assert len(s) == 3 and all(isinstance(x, str) for x in s)
    finally:
        fast.close()
        exact.close()


def test_decode_with_degenerate_weights_falls_back_to_exact():
    """All-equal logits (zero Dense kernel, equal biases): every margin is zero, every token is rechecked, first index wins."""
    from serann_b200.codec import RibosomalAutoencoder
    from serann_b200.synthetic import synthetic_decoder_weights
    w = synthetic_decoder_weights()
    w = dict(w, dense_kernel=np.zeros_like(w['dense_kernel']), dense_bias=np.zeros_like(w['dense_bias']))
    for k in ('gamma', 'beta', 'mean', 'var'):
        w[f'bn2_{k}'] = np.full(40, 1.0 if k in ('gamma', 'var') else 0.0, np.float32)
    ae = RibosomalAutoencoder(350, w)
    try:
        tok = ae.decode_to_sequence(np.random.default_rng(0).integers(0, 2, (300, 100)))
        assert (tok == 0).all()
        assert ae.last_stats['rechecked_tokens'] == tok.size
    finally:
        ae.close()
