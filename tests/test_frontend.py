"""Contract layer on the CPU: source -> IR, Keras shape/parameter rules, vocabulary, tokenizer, generator."""
import numpy as np
import pytest

from serann_b200 import frontend as F
from serann_b200 import generator, vocabulary as V

# reference synthetic_serann_generator/tutorial.ipynb:3559-3575 (source) and :3606-3643 (model.summary()):
# the only known-answer data the reference holds for this path.  Old naming style: inputs X / g, g of length 350.
NOTEBOOK_SOURCE = """x_conv2d_0 = Conv2D(filters=32, kernel_size=7, strides=1)(X_layer)
x_conv2d_1 = Conv2D(filters=8, kernel_size=5, strides=1)(x_conv2d_0)
x_maxpool2d_2 = MaxPool2D(pool_size=2)(x_conv2d_1)
x_conv2d_3 = Conv2D(filters=32, kernel_size=3, strides=1)(x_maxpool2d_2)
x_maxpool2d_4 = MaxPool2D(pool_size=2)(x_conv2d_3)
g_conv1d_0 = Conv1D(filters=8, kernel_size=5, strides=1)(g_layer)
g_dense_1 = Dense(units=91, activation='relu')(g_conv1d_0)
con = concatenate([Reshape((1, -1))(x_maxpool2d_4), Reshape((1, -1))(g_dense_1)])
m_dense_0 = Dense(units=187, activation='relu')(con)
m_dense_1 = Dense(units=113, activation='relu')(m_dense_0)
m_dense_2 = Dense(units=61, activation='relu')(m_dense_1)
m_dense_3 = Dense(units=205, activation='relu')(m_dense_2)
"""


def _params(t):
    p = t.producer
    cin = t.inputs[0].shape[-1]
    if p.kind == F.KIND_DENSE:
        return cin * p.units + p.units
    if p.kind == F.KIND_CONV:
        return int(np.prod(p.kernel)) * cin * p.filters + p.filters
    return 0


def test_notebook_known_answers():
    scope, _, _ = F.trace_source(NOTEBOOK_SOURCE, (28, 28, 1), (350, 1))
    expect = {'x_conv2d_0': ((22, 22, 32), 1600), 'x_conv2d_1': ((18, 18, 8), 6408), 'x_maxpool2d_2': ((9, 9, 8), 0),
              'x_conv2d_3': ((7, 7, 32), 2336), 'g_conv1d_0': ((346, 8), 48), 'x_maxpool2d_4': ((3, 3, 32), 0),
              'g_dense_1': ((346, 91), 819), 'con': ((1, 31774), 0), 'm_dense_0': ((1, 187), 5941925)}
    for name, (shape, params) in expect.items():
        assert scope[name].shape == shape, name
        assert _params(scope[name]) == params, name
    # total of the notebook model: the listed layers + heads Dense(10) and Dense(350) on m_dense_3
    total = sum(_params(scope[k]) for k in scope if isinstance(scope[k], F._Tensor) and scope[k].producer is not None)
    total += (205 * 10 + 10) + (205 * 350 + 350)
    assert total == 6068204


def test_compile_counts_heads_and_bn():
    src = ("X_layer = BatchNormalization()(X_layer)\n"
           "con = concatenate([Reshape((1, -1))(X_layer), Reshape((1, -1))(g_layer)])\nloss_balance = 0.5")
    ir = F.compile_source(src)
    assert ir.parameters_count == 4 * 1 + 884 * 10 + 10 + 884 * 100 + 100
    assert ir.layers[-1].units == 110 and ir.layers[-1].role == 'heads'
    assert ir.loss_balance == 0.5


@pytest.mark.parametrize('src', [
    "loss_balance = 0.5",                                                         # no con
    "con = concatenate([Reshape((1, -1))(X_layer), Reshape((1, -1))(g_layer)])",   # no loss_balance
    "con = Conv2D(filters=4, kernel_size=3)(g_layer)\nloss_balance = 0.1",         # rank mismatch
    "con = Conv1D(filters=4, kernel_size=3)(X_layer)\nloss_balance = 0.1",
    "con = Conv2D(filters=4, kernel_size=29)(X_layer)\nloss_balance = 0.1",        # kernel larger than the image
    "con = MaxPool2D(pool_size=2)(g_layer)\nloss_balance = 0.1",
    "con = Dense(units=0)(X_layer)\nloss_balance = 0.1",
    "con = Dense(units=4, activation='tanh')(X_layer)\nloss_balance = 0.1",
    "con = Dense(units=4)(X_layer\nloss_balance = 0.1",                            # syntax error
    "con = concatenate([X_layer, g_layer])\nloss_balance = 0.1",                   # rank mismatch
    "con = Dense(units=4)\nloss_balance = 0.1",                                    # layer, not a tensor
    "con = X_layer\nloss_balance = 'abc'",
    "con = BatchNormalization(units=3)(X_layer)\nloss_balance = 0.1",
    "con = Reshape((3, -1))(g_layer)\nloss_balance = 0.1",                         # 100 not divisible by 3
])
def test_invalid_sources(src):
    with pytest.raises(F.InvalidSerann):
        F.compile_source(src)
    assert not F.is_valid_serann(src)


def test_valid_oddities():
    # loss_balance may exceed 1 and may be an int (reference: float(locals()['loss_balance']))
    ir = F.compile_source("con = X_layer\nloss_balance = 3")
    assert ir.loss_balance == 3.0
    # dead layers are not counted; a branch may read the other input
    ir = F.compile_source("X_layer = Dense(units=7)(X_layer)\ng_layer = Dense(units=5)(g_layer)\ncon = g_layer\nloss_balance = .2")
    assert ir.parameters_count == (1 * 5 + 5) + 500 * 110 + 110
    # tuple arguments, pooling strides, rank-2 reshape
    ir = F.compile_source("X_layer = Conv2D(filters=3, kernel_size=(3, 5), strides=(2, 1))(X_layer)\n"
                          "X_layer = MaxPool2D(pool_size=(2, 3), strides=2)(X_layer)\n"
                          "con = Reshape((-1,))(X_layer)\nloss_balance = 0.5")
    assert ir.layers[0].out_shape == (13, 24, 3) and ir.layers[1].out_shape == (6, 11, 3)


def test_vocabulary_is_reconstructed_from_generator_templates():
    toks = set()
    for src, _ in generator.sample_population(400, seed=0):
        toks.update(V.tokenize(src))
    assert sorted(toks) == V.VOCABULARY[:-1]
    assert V.VOCABULARY_SIZE == 40 and V.VOCABULARY[-1] == '<PAD>' and V.PADDING_TOKEN == 39


def test_tokenizer_matches_reference_semantics():
    s = "X_layer = Dense(units=64, activation='relu')(X_layer)\n\nloss_balance = 0.4521"
    assert V.tokenize(s) == ['X_layer', '=', 'Dense', '(', 'units', '=', '6', '4', ',', 'activation', '=', "'", 'relu', "'",
                             ')', '(', 'X_layer', ')', '\n', 'loss_balance', '=', '0', '.', '4', '5', '2', '1']
    ids = V.encode_tokens(s, 350)
    assert ids.shape == (350,) and (ids[27:] == V.PADDING_TOKEN).all()
    back = V.sequence_to_string(ids[None])[0]
    assert back == s.replace(' ', '').replace('\n\n', '\n')
    assert F.is_valid_serann("con=X_layer\n" + back) is True


def test_sequence_to_string_rstrip_quirk():
    # rstrip('<PAD>') strips a character set: a trailing MaxPool2D loses its 'D' (reference ribosomal_autoencoder.py:76)
    seq = [V.WORD2INDEX['con'], V.WORD2INDEX['='], V.WORD2INDEX['MaxPool2D']] + [V.PADDING_TOKEN] * 3
    assert V.sequence_to_string([seq]) == ['con=MaxPool2']
    # interior padding stays
    seq = [V.WORD2INDEX['con'], V.PADDING_TOKEN, V.WORD2INDEX['=']]
    assert V.sequence_to_string([seq]) == ['con<PAD>=']
    assert V.sequence_to_string(np.zeros((0, 5), dtype=int)) == []


def test_generator_population_statistics():
    pop = generator.sample_population(300, seed=0)
    assert len({s for s, _ in pop}) == 300
    pc = np.array([ir.parameters_count for _, ir in pop])
    assert pc.max() <= 2e6
    assert all(len(V.tokenize(s)) <= 350 for s, _ in pop)
    clonal = generator.sample_population(5, seed=0, distinct=False)
    assert len({s for s, _ in clonal}) == 1 and len(clonal) == 5
    gflop = np.mean([ir.train_flops_per_sample() for _, ir in pop]) * 750 / 1e9
    assert 5 < gflop < 15          # generator-mix mean ~9.6 GFLOP per individual-step
