"""Markov-chain sampler of SeRANN source codes (synthetic populations for tests and benchmarks).

Restates the sampling process of the reference's dataset generator
(``synthetic_serann_generator/generator.py:52-99`` with the tables of
``synthetic_serann_generator/layer_transitions.py:9-72``): a first-order chain
over layer types for the image branch, one for the genotype branch, the
concatenation, and a chain over merged layers; layer arguments are drawn from
clipped normals.  The figshare dataset is not available offline, so this is the
source of populations for the benchmarks.  Unlike the
reference, randomness comes from an explicit ``numpy.random.Generator``.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import numpy as np

from .frontend import InvalidSerann, SerannIR, compile_source
from .vocabulary import tokenize

# state -> [(next state, probability)]; rows of layer_transitions.py:9-23 with zero entries dropped.
# (The 'g' row sums to 1.1 in the reference; the cumulative-probability draw below
#  reproduces what that does: the last entry is simply never reached past 1.0.)
TRANSITIONS: Dict[str, List[Tuple[str, float]]] = {
    'X':            [('X_Dense', .1), ('X_MaxPool2D', .1), ('X_Conv2D', .6), ('M_Concatenate', .2)],
    'X_Dense':      [('X_Dense', .1), ('X_Conv2D', .2), ('X_BatchNorm', .2), ('M_Concatenate', .5)],
    'X_MaxPool2D':  [('X_Dense', .3), ('X_Conv2D', .4), ('M_Concatenate', .3)],
    'X_Conv2D':     [('X_Dense', .1), ('X_MaxPool2D', .4), ('X_Conv2D', .2), ('X_BatchNorm', .2),
                     ('M_Concatenate', .1)],
    'X_BatchNorm':  [('X_Dense', .2), ('X_MaxPool2D', .1), ('M_Concatenate', .7)],
    'g':            [('g_Dense', .5), ('g_Conv1D', .4), ('M_Concatenate', .2)],
    'g_Dense':      [('g_Dense', .1), ('g_Conv1D', .2), ('g_BatchNorm', .2), ('M_Concatenate', .5)],
    'g_Conv1D':     [('g_Dense', .3), ('g_Conv1D', .2), ('g_BatchNorm', .2), ('M_Concatenate', .3)],
    'g_BatchNorm':  [('g_Dense', .2), ('g_Conv1D', .15), ('M_Concatenate', .65)],
    'M_Concatenate': [('M_Dense', 1.0)],
    'M_Dense':      [('M_Dense', .2), ('M_BatchNorm', .2), ('outputs', .6)],
    'M_BatchNorm':  [('M_Dense', .1), ('outputs', .9)],
}


def _draw(rng: np.random.Generator, options: List[Tuple[object, float]]):
    """Ascending-probability cumulative draw (reference ``generator.py:24-29``)."""
    order = sorted(range(len(options)), key=lambda i: options[i][1])
    u = rng.random()
    acc = 0.0
    for i in order:
        acc += options[i][1]
        if acc > u:
            return options[i][0]
    return options[order[-1]][0]


def _clipped(rng, mean, std, lo, hi) -> int:
    return int(np.clip(int(rng.standard_normal() * std + mean), lo, hi))


def _odd_kernel(rng) -> int:
    x = _clipped(rng, 6, 1.3, 2, 10)
    return x + (x % 2) - 1


def _filters(rng) -> int:
    return 2 ** _clipped(rng, 4.5, 1.0, 0, 6)


_ACT = [('relu', .8), ('sigmoid', .2)]
_STRIDE = [(1, .8), (2, .2)]


def _emit(rng, state: str, var: str) -> str:
    kind = state.split('_', 1)[1]
    if kind == 'Dense':
        mean, std, lo, hi = {'X': (64, 8, 8, 128), 'g': (64, 15, 8, 128), 'M': (128, 30, 32, 256)}[state[0]]
        return f"{var} = Dense(units={_clipped(rng, mean, std, lo, hi)}, activation='{_draw(rng, _ACT)}')({var})"
    if kind == 'MaxPool2D':
        return f"{var} = MaxPool2D(pool_size={_clipped(rng, 2, 1.1, 2, 5)})({var})"
    if kind in ('Conv2D', 'Conv1D'):
        return (f"{var} = {kind}(filters={_filters(rng)}, kernel_size={_odd_kernel(rng)}, "
                f"strides={_draw(rng, _STRIDE)})({var})")
    if kind == 'BatchNorm':
        return f"{var} = BatchNormalization()({var})"
    raise KeyError(state)


def _walk(rng, state: str, var: str, stop: str) -> str:
    lines = []
    while True:
        state = _draw(rng, TRANSITIONS[state])
        if state == stop:
            return '\n'.join(lines)
        lines.append(_emit(rng, state, var))


def sample_source(rng: np.random.Generator) -> str:
    """One source string in the reference's format (``generator.py:80-99``), not validated."""
    x_branch = _walk(rng, 'X', 'X_layer', 'M_Concatenate')
    g_branch = _walk(rng, 'g', 'g_layer', 'M_Concatenate')
    concat = 'con = concatenate([Reshape((1, -1))(X_layer), Reshape((1, -1))(g_layer)])'
    merged = _walk(rng, 'M_Concatenate', 'con', 'outputs')
    net = '\n\n'.join([x_branch, g_branch, concat, merged])
    return net + '\n\nloss_balance = {:.4f}'.format(rng.random())


def sample_population(n: int, seed: int = 0, max_tokens: int = 350, max_parameters: float = 2e6,
                      distinct: bool = True, genotype_size: int = 100,
                      max_flops_per_sample: Optional[float] = None) -> List[Tuple[str, SerannIR]]:
    """``n`` buildable sources (<= ``max_tokens`` tokens, <= ``max_parameters`` parameters).

    ``distinct=False`` gives the clonal mix of SURVEY.md section 8d: ``n`` copies of the first net.
    """
    rng = np.random.default_rng(seed)
    out: List[Tuple[str, SerannIR]] = []
    seen = set()
    while len(out) < (n if distinct else 1):
        src = sample_source(rng)
        if src in seen or len(tokenize(src)) > max_tokens:
            continue
        try:
            ir = compile_source(src, genotype_size=genotype_size)
        except InvalidSerann:
            continue
        if ir.parameters_count > max_parameters:
            continue
        if max_flops_per_sample is not None and ir.train_flops_per_sample() > max_flops_per_sample:
            continue
        seen.add(src)
        out.append((src, ir))
    if not distinct:
        out = out * n
    return out
