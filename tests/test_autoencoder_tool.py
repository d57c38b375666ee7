"""tools/train_autoencoder.py: the PyTorch restatement of the reference's ConcreteGAE
(ribosomal_autoencoder/model.py:10-104, training.py:51-101) trains, and what it exports means the same thing to the
CPU oracle's decoder (Keras layouts, Flatten order, BatchNormalization arithmetic)."""
import importlib.util
import pathlib

import numpy as np
import pytest
import torch

from oracle import serann_oracle as O

ROOT = pathlib.Path(__file__).resolve().parent.parent
spec = importlib.util.spec_from_file_location('train_autoencoder', ROOT / 'tools' / 'train_autoencoder.py')
T = importlib.util.module_from_spec(spec)
spec.loader.exec_module(T)


@pytest.fixture(scope='module')
def small():
    torch.manual_seed(0)
    data = T.token_dataset(96, 60, seed=3)
    model = T.GeneticAutoencoder(genotype_length=100, max_tokens=60, vocab=40, embedding=12)
    hist = T.train(model, data, steps=60, batch=32, device=torch.device('cpu'), log_every=0)
    return model, data, hist


def test_training_reduces_the_reconstruction_loss(small):
    model, data, hist = small
    assert np.isfinite(hist).all()
    assert np.mean(hist[-10:]) < 0.8 * np.mean(hist[:5])


def test_exported_decoder_equals_the_oracle_decoder(small):
    model, data, _ = small
    w = T.export_weights(model)
    g = np.random.default_rng(0).integers(0, 2, (40, 100))
    tok_torch = model.decode(torch.as_tensor(g)).numpy()
    tok_oracle, z = O.ae_decode(g, w, 60, 40)
    top2 = np.sort(z, axis=-1)[..., -2:]
    clear = (top2[..., 1] - top2[..., 0]) > 1e-4
    assert np.array_equal(tok_torch[clear], tok_oracle[clear]) and clear.mean() > 0.99
    assert w['conv_kernel'].shape == (5, 2, 32) and w['dense_kernel'].shape == (96 * 32, 60 * 40)


def test_schedules_have_the_reference_shapes():
    t, lr, kw = T.schedules(1000)
    assert t[0] == pytest.approx(0.3) and t[-1] == pytest.approx(1e-3)
    assert lr[0] == pytest.approx(3e-4) and lr[-1] == pytest.approx(2e-5)
    assert kw[0] == 0 and kw[-1] == pytest.approx(0.04)


@pytest.mark.gpu
def test_cuda_decoder_runs_the_exported_weights_and_the_codec_encodes(small):
    from serann_b200.codec import RibosomalAutoencoder
    model, data, _ = small
    w = T.export_weights(model)
    ae = RibosomalAutoencoder(60, w)
    try:
        g = np.random.default_rng(1).integers(0, 2, (300, 100))
        tok = ae.decode_to_sequence(g)
        ref, z = O.ae_decode(g, w, 60, 40)
        top2 = np.sort(z, axis=-1)[..., -2:]
        clear = (top2[..., 1] - top2[..., 0]) > 1e-4
        assert np.array_equal(tok[clear], ref[clear])
        enc = ae.encode_sequence(data[:8])
        assert enc.shape == (8, 100) and set(np.unique(enc)) <= {0, 1}
        assert np.array_equal(enc, model.encode(torch.as_tensor(data[:8])).numpy())
        assert ae.encode_string(['con = X_layer\nloss_balance = 0.5']).shape == (1, 100)
    finally:
        ae.close()
