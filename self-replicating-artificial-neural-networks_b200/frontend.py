"""Source code -> layer IR (the genotype -> token -> source -> model contract).

The reference turns a SeRANN source string into a Keras graph by ``exec``-ing it
with the Keras layer classes in scope (reference ``common/logic.py:17-35``; the
same body is duplicated at ``evolutionary_experiment/logic/experiment_worker.py:248-267``).
Here the same ``exec`` runs against shape-inferring recorders, so that the
result is a flat layer list the CUDA engine can consume.  Anything the Keras
build would reject raises ``InvalidSerann`` (the reference swallows every
exception and marks the individual invalid, ``experiment_worker.py:234-239``).

Keras rules restated here (Keras 2.6 defaults): 'valid' padding, MaxPool2D
stride defaults to the pool size, Dense contracts the last axis of a tensor of
any rank, Conv2D needs rank 4 and Conv1D rank 3 (batch included), concatenate
joins along the last axis, ``count_params`` includes the four BatchNormalization
vectors, and only layers on a path from the inputs to the two heads count.

Known deviations (documented, all make a source *invalid* rather than wrong):
a layer object called more than once (weight sharing) and BatchNormalization
with any constructor argument are rejected; zero-sized layers are rejected.
"""
from __future__ import annotations

import ast
import math
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

# engine-level layer kinds (must match include/serann_b200.h)
KIND_DENSE = 0      # 1x1 "convolution" on (rows, 1, C)
KIND_CONV = 1       # Conv2D, or Conv1D as a 1xk convolution
KIND_MAXPOOL = 2
KIND_BN = 3
KIND_CONCAT = 4
KIND_RESHAPE = 5

ACT_NONE, ACT_RELU, ACT_SIGMOID = 0, 1, 2
_ACTIVATIONS = {None: ACT_NONE, 'linear': ACT_NONE, 'relu': ACT_RELU, 'sigmoid': ACT_SIGMOID}

TENSOR_X = 0        # tensor id of the image input
TENSOR_G = 1        # tensor id of the genotype input


class InvalidSerann(Exception):
    """The source does not describe a buildable SeRANN."""


@dataclass
class LayerIR:
    kind: int
    inputs: List[int]                 # tensor ids (0 = X, 1 = g, 2 + i = output of layer i)
    out_shape: Tuple[int, ...]        # per-sample shape, channel last
    in_shape: Tuple[int, ...] = ()    # per-sample shape of inputs[0]
    act: int = ACT_NONE
    units: int = 0                    # Dense units / conv filters
    kernel: Tuple[int, int] = (1, 1)  # conv kernel or pool window (h, w)
    strides: Tuple[int, int] = (1, 1)
    conv_rank: int = 0                # 1 for Conv1D, 2 for Conv2D (weight layout for get/set)
    role: str = ''                    # '' | 'heads'

    @property
    def out_elems(self) -> int:
        return int(math.prod(self.out_shape))

    def param_shapes(self) -> List[Tuple[str, Tuple[int, ...]]]:
        """Keras weight list of this layer, in ``layer.get_weights()`` order."""
        cin = self.in_shape[-1] if self.in_shape else 0
        if self.kind == KIND_DENSE:
            return [('kernel', (cin, self.units)), ('bias', (self.units,))]
        if self.kind == KIND_CONV:
            kh, kw = self.kernel
            kshape = (kw, cin, self.units) if self.conv_rank == 1 else (kh, kw, cin, self.units)
            return [('kernel', kshape), ('bias', (self.units,))]
        if self.kind == KIND_BN:
            c = self.out_shape[-1]
            return [('gamma', (c,)), ('beta', (c,)), ('moving_mean', (c,)), ('moving_variance', (c,))]
        return []

    def param_count(self) -> int:
        return sum(int(math.prod(s)) for _, s in self.param_shapes())

    def macs_per_sample(self) -> int:
        """Multiply-accumulates of the forward pass for one sample."""
        if self.kind in (KIND_DENSE, KIND_CONV):
            kh, kw = self.kernel
            cin = self.in_shape[-1]
            return (self.out_elems // self.units) * kh * kw * cin * self.units
        return 0


@dataclass
class SerannIR:
    """One individual: layers in execution order, heads last.

    ``layers[-1]`` is the fused head layer: a Dense with ``n_classes +
    genotype_size`` units whose first ``n_classes`` columns are the
    classification logits (reference ``common/logic.py:29-30``) and the rest
    the replication logits (``common/logic.py:32-33``).
    """
    layers: List[LayerIR]
    loss_balance: float
    n_classes: int
    genotype_size: int
    image_shape: Tuple[int, ...]
    parameters_count: int = 0
    signature: str = ''               # architecture key: equal signature => identical shapes

    def train_flops_per_sample(self) -> int:
        """2 * sum MACs * (2 + d_l); d_l = 0 when the layer needs no input gradient."""
        needs = self.needs_input_grad()
        return sum(2 * l.macs_per_sample() * (2 + int(needs[i])) for i, l in enumerate(self.layers))

    def needs_input_grad(self) -> List[bool]:
        """Per layer: does any input of it depend on a trainable parameter?"""
        dep = {TENSOR_X: False, TENSOR_G: False}
        out = []
        for i, l in enumerate(self.layers):
            d = any(dep[t] for t in l.inputs)
            out.append(d)
            dep[2 + i] = d or l.kind in (KIND_DENSE, KIND_CONV, KIND_BN)
        return out


# --------------------------------------------------------------------------- recorders

class _Tensor:
    __slots__ = ('shape', 'producer', 'inputs')

    def __init__(self, shape, producer=None, inputs=()):
        self.shape = tuple(int(s) for s in shape)
        self.producer = producer          # _Layer or None for the model inputs
        self.inputs = tuple(inputs)       # upstream _Tensor objects


def _as_tensor(x, what):
    if not isinstance(x, _Tensor):
        raise InvalidSerann(f'{what} called on {type(x).__name__}, not a tensor')
    return x


def _int_tuple(value, n, name):
    """Keras ``conv_utils.normalize_tuple``."""
    if isinstance(value, bool):
        raise InvalidSerann(f'{name}: bool')
    if isinstance(value, int):
        return (value,) * n
    try:
        t = tuple(value)
    except TypeError:
        raise InvalidSerann(f'{name}: not an int or tuple')
    if len(t) != n:
        raise InvalidSerann(f'{name}: expected {n} integers')
    for v in t:
        if isinstance(v, bool) or not isinstance(v, int):
            raise InvalidSerann(f'{name}: non-integer entry')
    return t


def _positive_int(v, name):
    if isinstance(v, bool) or isinstance(v, str):
        raise InvalidSerann(f'{name}: bad type')
    try:
        iv = int(v)
    except (TypeError, ValueError):
        raise InvalidSerann(f'{name}: not an integer')
    if isinstance(v, float) and v != iv:
        # Keras: Dense(units=3.5) -> int(3.5) == 3 is accepted; conv filters likewise.
        pass
    if iv <= 0:
        raise InvalidSerann(f'{name} must be positive')
    return iv


def _activation(a):
    try:
        return _ACTIVATIONS[a]
    except (KeyError, TypeError):
        raise InvalidSerann(f'unsupported activation {a!r}')


class _Layer:
    kind = -1

    def __init__(self):
        self._called = False

    def __call__(self, inputs, *args, **kwargs):
        if args or kwargs:
            raise InvalidSerann('unsupported call arguments')
        if self._called:
            raise InvalidSerann('layer called twice (weight sharing is not supported)')
        self._called = True
        return self._apply(inputs)


class _Dense(_Layer):
    kind = KIND_DENSE

    def __init__(self, units, activation=None):
        super().__init__()
        self.units = _positive_int(units, 'units')
        self.act = _activation(activation)

    def _apply(self, x):
        x = _as_tensor(x, 'Dense')
        if len(x.shape) < 1:
            raise InvalidSerann('Dense needs rank >= 2')
        return _Tensor(x.shape[:-1] + (self.units,), self, (x,))


class _Conv(_Layer):
    kind = KIND_CONV
    rank = 2

    def __init__(self, filters, kernel_size, strides=1, activation=None):
        super().__init__()
        self.filters = _positive_int(filters, 'filters')
        self.kernel = _int_tuple(kernel_size, self.rank, 'kernel_size')
        self.strides = _int_tuple(strides, self.rank, 'strides')
        if any(k <= 0 for k in self.kernel) or any(s <= 0 for s in self.strides):
            raise InvalidSerann('kernel_size/strides must be positive')
        self.act = _activation(activation)

    def _apply(self, x):
        x = _as_tensor(x, 'Conv')
        if len(x.shape) != self.rank + 1:
            raise InvalidSerann(f'Conv{self.rank}D expects rank {self.rank + 2} input')
        spatial = []
        for d, k, s in zip(x.shape[:-1], self.kernel, self.strides):
            o = (d - k) // s + 1 if d >= k else 0
            if o <= 0:
                raise InvalidSerann('convolution output dimension <= 0')
            spatial.append(o)
        return _Tensor(tuple(spatial) + (self.filters,), self, (x,))


class _Conv1D(_Conv):
    rank = 1


class _MaxPool2D(_Layer):
    kind = KIND_MAXPOOL

    def __init__(self, pool_size=(2, 2), strides=None):
        super().__init__()
        self.pool = _int_tuple(pool_size, 2, 'pool_size')
        self.strides = self.pool if strides is None else _int_tuple(strides, 2, 'strides')
        if any(p <= 0 for p in self.pool) or any(s <= 0 for s in self.strides):
            raise InvalidSerann('pool_size/strides must be positive')

    def _apply(self, x):
        x = _as_tensor(x, 'MaxPool2D')
        if len(x.shape) != 3:
            raise InvalidSerann('MaxPool2D expects rank 4 input')
        spatial = []
        for d, p, s in zip(x.shape[:2], self.pool, self.strides):
            o = (d - p) // s + 1 if d >= p else 0
            if o <= 0:
                raise InvalidSerann('pooling output dimension <= 0')
            spatial.append(o)
        return _Tensor(tuple(spatial) + (x.shape[-1],), self, (x,))


class _BatchNormalization(_Layer):
    kind = KIND_BN

    def __init__(self):
        super().__init__()

    def _apply(self, x):
        x = _as_tensor(x, 'BatchNormalization')
        return _Tensor(x.shape, self, (x,))


class _Reshape(_Layer):
    kind = KIND_RESHAPE

    def __init__(self, target_shape):
        super().__init__()
        try:
            self.target = tuple(target_shape)
        except TypeError:
            raise InvalidSerann('Reshape target must be a tuple')
        for v in self.target:
            if isinstance(v, bool) or not isinstance(v, int):
                raise InvalidSerann('Reshape target must hold integers')

    def _apply(self, x):
        x = _as_tensor(x, 'Reshape')
        total = math.prod(x.shape)
        unknown = [i for i, v in enumerate(self.target) if v == -1]
        if len(unknown) > 1 or any(v < -1 or v == 0 for v in self.target):
            raise InvalidSerann('bad Reshape target')
        shape = list(self.target)
        if unknown:
            known = math.prod(v for v in self.target if v != -1)
            if known == 0 or total % known:
                raise InvalidSerann('Reshape size mismatch')
            shape[unknown[0]] = total // known
        elif math.prod(shape) != total:
            raise InvalidSerann('Reshape size mismatch')
        if not shape:
            raise InvalidSerann('Reshape to a scalar')
        return _Tensor(shape, self, (x,))


class _Concatenate(_Layer):
    kind = KIND_CONCAT

    def _apply(self, xs):
        if not isinstance(xs, (list, tuple)) or len(xs) < 1:
            raise InvalidSerann('concatenate needs a list of tensors')
        xs = [_as_tensor(x, 'concatenate') for x in xs]
        first = xs[0].shape
        for x in xs[1:]:
            if len(x.shape) != len(first) or x.shape[:-1] != first[:-1]:
                raise InvalidSerann('concatenate shape mismatch')
        return _Tensor(first[:-1] + (sum(x.shape[-1] for x in xs),), self, xs)


def _concatenate(inputs, axis=-1):
    if axis != -1:
        raise InvalidSerann('concatenate: only the last axis is supported')
    return _Concatenate()(inputs)


def _unsupported(name):
    def ctor(*a, **k):
        raise InvalidSerann(f'{name} is not reachable from the token vocabulary')
    return ctor


def _wrap(cls):
    """Constructor wrapper turning Python signature errors into InvalidSerann."""
    def ctor(*args, **kwargs):
        try:
            return cls(*args, **kwargs)
        except TypeError as e:
            raise InvalidSerann(str(e))
    return ctor


def _namespace(x_input, g_input):
    return {
        'Dense': _wrap(_Dense), 'Conv2D': _wrap(_Conv), 'Conv1D': _wrap(_Conv1D),
        'MaxPool2D': _wrap(_MaxPool2D), 'BatchNormalization': _wrap(_BatchNormalization),
        'Reshape': _wrap(_Reshape), 'concatenate': _concatenate,
        'Lambda': _unsupported('Lambda'), 'Activation': _unsupported('Activation'),
        'X_layer': x_input, 'g_layer': g_input,
    }


# --------------------------------------------------------------------------- lowering

def _lower(outputs: Sequence[_Tensor], x_input, g_input) -> List[LayerIR]:
    """Topologically order the recorded graph reachable from ``outputs``."""
    ids = {id(x_input): TENSOR_X, id(g_input): TENSOR_G}
    layers: List[LayerIR] = []

    def visit(t: _Tensor) -> int:
        if id(t) in ids:
            return ids[id(t)]
        if t.producer is None:
            raise InvalidSerann('tensor does not derive from the model inputs')
        in_ids = [visit(u) for u in t.inputs]
        p = t.producer
        ir = LayerIR(kind=p.kind, inputs=in_ids, out_shape=t.shape, in_shape=t.inputs[0].shape)
        if p.kind == KIND_DENSE:
            ir.units, ir.act = p.units, p.act
        elif p.kind == KIND_CONV:
            ir.units, ir.act, ir.conv_rank = p.filters, p.act, p.rank
            if p.rank == 1:
                ir.kernel, ir.strides = (1, p.kernel[0]), (1, p.strides[0])
            else:
                ir.kernel, ir.strides = tuple(p.kernel), tuple(p.strides)
        elif p.kind == KIND_MAXPOOL:
            ir.kernel, ir.strides = tuple(p.pool), tuple(p.strides)
        layers.append(ir)
        ids[id(t)] = 1 + len(layers)
        return ids[id(t)]

    import sys
    old = sys.getrecursionlimit()
    sys.setrecursionlimit(max(old, 10000))
    try:
        for o in outputs:
            visit(o)
    finally:
        sys.setrecursionlimit(old)
    return layers


def _signature(layers: List[LayerIR]) -> str:
    return ';'.join(f'{l.kind}:{l.inputs}:{l.in_shape}:{l.out_shape}:{l.act}:{l.units}:{l.kernel}:{l.strides}'
                    for l in layers)


def trace_source(source_code: str, x_shape, g_shape):
    """``exec`` the source against the recorders and return the resulting namespace."""
    x_input, g_input = _Tensor(x_shape), _Tensor(g_shape)
    scope = _namespace(x_input, g_input)
    try:
        exec(source_code, {'__builtins__': __builtins__}, scope)
    except InvalidSerann:
        raise
    except BaseException as e:           # the reference catches everything (experiment_worker.py:238)
        if isinstance(e, (KeyboardInterrupt, SystemExit)):
            raise
        raise InvalidSerann(f'{type(e).__name__}: {e}')
    return scope, x_input, g_input


def compile_source(source_code: str, image_dims=(28, 28), genotype_size: int = 100,
                   n_classes: int = 10) -> SerannIR:
    """Restates ``build_serann_model`` (reference ``common/logic.py:17-35``).

    Returns the layer IR with the two heads fused into one trailing Dense, the
    ``loss_balance`` float and Keras' ``count_params()``.  Raises
    ``InvalidSerann`` where the reference build would raise.
    """
    x_shape = tuple(image_dims) + (1,)
    scope, x_input, g_input = trace_source(source_code, x_shape, (genotype_size, 1))
    if 'con' not in scope:
        raise InvalidSerann("source does not define 'con'")
    con = _as_tensor(scope['con'], 'heads')
    flat = _Reshape((1, -1))(con)
    heads = _Dense(n_classes + genotype_size)(flat)
    if 'loss_balance' not in scope:
        raise InvalidSerann("source does not define 'loss_balance'")
    try:
        loss_balance = float(scope['loss_balance'])
    except (TypeError, ValueError) as e:
        raise InvalidSerann(f'loss_balance: {e}')
    layers = _lower([heads], x_input, g_input)
    layers[-1].role = 'heads'
    ir = SerannIR(layers=layers, loss_balance=loss_balance, n_classes=n_classes,
                  genotype_size=genotype_size, image_shape=x_shape)
    # two Reshape((1,-1)) + two Dense in the reference; same parameter total as the fused head
    ir.parameters_count = sum(l.param_count() for l in layers)
    ir.signature = _signature(layers)
    return ir


# Front-end cache keyed by the source string.  The reference rebuilds a Keras model per
# individual per generation (experiment_worker.py:232-246) and per distinct offspring source for validity
# (serann_evaluation/logic.py:139-151,216-219); populations are clonal at generation 0 and dominated by a few clones
# later, so the same string is compiled once.  Entries are the IR or the InvalidSerann that the build raised.
_COMPILE_CACHE: 'OrderedDict[tuple, object]' = None
_COMPILE_CACHE_MAX = 1 << 15
compile_cache_stats = {'hits': 0, 'misses': 0}


def compile_cached(source_code: str, image_dims=(28, 28), genotype_size: int = 100, n_classes: int = 10) -> SerannIR:
    """``compile_source`` memoised by (source, geometry).  The returned IR is shared: treat it as read-only."""
    global _COMPILE_CACHE
    if _COMPILE_CACHE is None:
        from collections import OrderedDict
        _COMPILE_CACHE = OrderedDict()
    key = (source_code, tuple(image_dims), int(genotype_size), int(n_classes))
    hit = _COMPILE_CACHE.get(key)
    if hit is not None:
        _COMPILE_CACHE.move_to_end(key)
        compile_cache_stats['hits'] += 1
        if isinstance(hit, InvalidSerann):
            raise hit
        return hit
    compile_cache_stats['misses'] += 1
    try:
        out = compile_source(source_code, image_dims=image_dims, genotype_size=genotype_size, n_classes=n_classes)
    except InvalidSerann as e:
        out = e
    _COMPILE_CACHE[key] = out
    if len(_COMPILE_CACHE) > _COMPILE_CACHE_MAX:
        _COMPILE_CACHE.popitem(last=False)
    if isinstance(out, InvalidSerann):
        raise out
    return out


def is_valid_serann(source_code: str, **kwargs) -> bool:
    """Reference ``serann_evaluation/logic.py:139-151``: ``ast.parse`` then a model build."""
    try:
        ast.parse(source_code)
    except (SyntaxError, ValueError):
        return False
    try:
        compile_cached(source_code, **kwargs)
        return True
    except InvalidSerann:
        return False


def extract_layer_counts(source_code: str):
    """Reference ``evolutionary_experiment/logic/experiment.py:323-329``."""
    return {'classification_layers': source_code.count('X_layer='),
            'replication_layers': source_code.count('g_layer='),
            'merged_layers': source_code.count('con=')}
