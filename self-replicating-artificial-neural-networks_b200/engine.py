"""ctypes binding of ``libserann_b200.so`` (the C ABI in ``include/serann_b200.h``).

There is no CPU fallback: importing this module without the built library, or
creating an engine without a CUDA device, raises.
"""
from __future__ import annotations

import ctypes as C
import math
import os
import pathlib
from typing import List, Optional, Sequence

import numpy as np

from .frontend import (KIND_BN, KIND_CONCAT, KIND_CONV, KIND_DENSE, KIND_MAXPOOL, KIND_RESHAPE, LayerIR,
                       SerannIR)

_PKG = pathlib.Path(__file__).resolve().parent
LIB_PATH = _PKG / 'libserann_b200.so'
MAX_CONCAT_INPUTS = 4


class EngineError(RuntimeError):
    pass


class LayerDesc(C.Structure):
    _fields_ = [('kind', C.c_int32), ('act', C.c_int32), ('n_in', C.c_int32), ('inp', C.c_int32 * MAX_CONCAT_INPUTS),
                ('units', C.c_int32), ('kh', C.c_int32), ('kw', C.c_int32), ('sh', C.c_int32), ('sw', C.c_int32),
                ('in_h', C.c_int32), ('in_w', C.c_int32), ('in_c', C.c_int32),
                ('out_h', C.c_int32), ('out_w', C.c_int32), ('out_c', C.c_int32)]


class Config(C.Structure):
    _fields_ = [('device', C.c_int32), ('image_h', C.c_int32), ('image_w', C.c_int32),
                ('genotype_size', C.c_int32), ('n_classes', C.c_int32), ('max_batch', C.c_int32)]


_lib = None


def load_library(rebuild: bool = False) -> C.CDLL:
    """Load (building first when sources are newer and nvcc is present) the CUDA library."""
    global _lib
    if _lib is not None and not rebuild:
        return _lib
    try:
        from .build import build
        build(force=rebuild)
    except RuntimeError as e:
        if not LIB_PATH.exists():
            raise EngineError(f'libserann_b200.so is not built and cannot be built here: {e}') from e
    if not LIB_PATH.exists():
        raise EngineError(f'{LIB_PATH} is missing: run `python -m serann_b200.build` (needs nvcc). '
                          'There is no CPU fallback for this path.')
    lib = C.CDLL(str(LIB_PATH))
    vp, i32, i64, u64 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64
    lib.srn_last_error.restype = C.c_char_p
    lib.srn_launch_count.restype = u64
    lib.srn_create.argtypes = [C.POINTER(Config), C.POINTER(vp)]
    lib.srn_destroy.argtypes = [vp]
    lib.srn_destroy.restype = None
    lib.srn_set_dataset.argtypes = [vp, vp, vp, vp, i32, vp, vp, vp, i32]
    lib.srn_build.argtypes = [vp, C.POINTER(LayerDesc), vp, vp, i32, u64]
    lib.srn_get_weights.argtypes = [vp, i32, i32, i32, vp, C.POINTER(i64)]
    lib.srn_set_weights.argtypes = [vp, i32, i32, i32, vp, i64]
    lib.srn_parameters_count.argtypes = [vp, i32]
    lib.srn_parameters_count.restype = i64
    lib.srn_device_bytes.argtypes = [vp]
    lib.srn_device_bytes.restype = i64
    lib.srn_fit.argtypes = [vp, i32, i32, i32, vp, u64, i32, vp, vp, vp, vp, C.POINTER(C.c_float)]
    lib.srn_evaluate.argtypes = [vp, i32, i32, i32, i32, vp, vp]
    lib.srn_replicate.argtypes = [vp, vp, i32, vp, vp, vp, vp]
    lib.srn_forward_logits.argtypes = [vp, i32, i32, i32, i32, vp]
    lib.srn_decoder_create.argtypes = [i32, i32, i32, i32, vp, vp, vp, vp, vp, vp, C.POINTER(vp)]
    lib.srn_decoder_destroy.argtypes = [vp]
    lib.srn_decoder_destroy.restype = None
    lib.srn_decoder_decode.argtypes = [vp, vp, i32, vp, C.POINTER(C.c_float)]
    lib.srn_decoder_stats.argtypes = [vp, C.POINTER(i64), C.POINTER(i32)]
    lib.srn_readout.argtypes = [i32, vp, vp, i32, i32, i32, vp, vp]
    lib.srn_bn_train_forward.argtypes = [i32, vp, i64, i32, vp, vp, vp, vp, vp]
    lib.srn_bn_train_forward_noalloc.argtypes = [i32, vp, vp, i64, i32, vp, vp, vp, vp, vp]
    lib.srn_levenshtein.argtypes = [C.c_char_p, i32, C.c_char_p, i32]
    lib.srn_levenshtein.restype = i32
    lib.srn_debug_set_tensor_cores.argtypes = [i32]
    lib.srn_alloc_count.restype = u64
    lib.srn_build_seeded.argtypes = [vp, C.POINTER(LayerDesc), vp, vp, i32, vp]
    lib.srn_profile.argtypes = [vp, i32]
    lib.srn_profile_read.argtypes = [vp, vp, vp]
    lib.srn_debug_step.argtypes = [vp, vp, i32, i32]
    lib.srn_debug_get_tensor.argtypes = [vp, i32, i32, i32, i32, vp]
    lib.srn_debug_get_grad.argtypes = [vp, i32, i32, i32, vp, i64]
    lib.srn_debug_launch_groups.argtypes = [vp]
    lib.srn_debug_group_probs.argtypes = [vp, i32, i32, vp]
    lib.srn_debug_profile_detail.argtypes = [vp, i32, i32, vp, vp, vp, vp]
    _lib = lib
    return lib


def _check(rc: int):
    if rc != 0:
        raise EngineError(f'[{rc}] {load_library().srn_last_error().decode()}')


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def set_tensor_cores(on: bool):
    """Test hook (srn_host.h, not part of the product ABI), process-wide: aligned GEMMs on tcgen05 (default) or everything on the SIMT kernels."""
    _check(load_library().srn_debug_set_tensor_cores(int(on)))


def launch_count() -> int:
    return int(load_library().srn_launch_count())


def alloc_count() -> int:
    """cudaMalloc calls made by the library so far (zero per steady-state generation)."""
    return int(load_library().srn_alloc_count())


def _hwc(shape: Sequence[int], spatial: int):
    """Per-sample shape -> (H, W, C) as the engine sees it."""
    shape = tuple(int(s) for s in shape)
    if spatial == 2:
        return shape[0], shape[1], shape[2]
    if spatial == 1:
        return 1, shape[0], shape[1]
    return int(math.prod(shape[:-1])), 1, shape[-1]


def layer_descs(ir: SerannIR) -> List[LayerDesc]:
    out = []
    for l in ir.layers:
        d = LayerDesc()
        d.kind, d.act, d.n_in, d.units = l.kind, l.act, len(l.inputs), l.units
        if len(l.inputs) > MAX_CONCAT_INPUTS:
            raise EngineError(f'concatenate of {len(l.inputs)} tensors (engine limit {MAX_CONCAT_INPUTS})')
        for j, t in enumerate(l.inputs):
            d.inp[j] = t
        d.kh, d.kw = l.kernel
        d.sh, d.sw = l.strides
        spatial = 0
        if l.kind == KIND_CONV:
            spatial = l.conv_rank
        elif l.kind == KIND_MAXPOOL:
            spatial = 2
        d.in_h, d.in_w, d.in_c = _hwc(l.in_shape, spatial)
        d.out_h, d.out_w, d.out_c = _hwc(l.out_shape, spatial)
        out.append(d)
    return out


def activation_bytes_estimate(ir: SerannIR, max_batch: int) -> int:
    """Activation + activation-gradient bytes one individual needs in the arena."""
    total = 0
    for l in ir.layers:
        if l.kind != KIND_RESHAPE:
            c = l.out_shape[-1]
            total += 2 * (l.out_elems // c) * (-(-c // 8) * 8)      # channel stride padded to 8
    return total * 2 * max_batch


class Engine:
    """One GPU shard.  Mirrors the life cycle of the reference worker (``experiment_worker.py:22-99``)."""

    def __init__(self, image_dims=(28, 28), genotype_size=100, n_classes=10, max_batch=750, device=0):
        self._lib = load_library()
        self._h = C.c_void_p()
        cfg = Config(device, image_dims[0], image_dims[1], genotype_size, n_classes, max_batch)
        _check(self._lib.srn_create(C.byref(cfg), C.byref(self._h)))
        self.image_dims, self.G, self.ncls, self.max_batch, self.device = tuple(image_dims), genotype_size, n_classes, max_batch, device
        self.n = 0
        self.irs: List[SerannIR] = []
        self.n_train = self.n_test = 0
        # bytes this handle has moved between host and device through the ABI (counted from the arrays passed)
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def close(self):
        if getattr(self, '_h', None) is not None and self._h:
            self._lib.srn_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- data
    def set_dataset(self, x_train, y_train, g_train, x_test=None, y_test=None, g_test=None):
        def prep(x, y, g):
            if x is None:
                return None, None, None, 0
            x = np.ascontiguousarray(np.asarray(x, dtype=np.float32).reshape(len(x), -1))
            if x.shape[1] != self.image_dims[0] * self.image_dims[1]:
                raise EngineError(f'image rows have {x.shape[1]} pixels')
            y = None if y is None else np.ascontiguousarray(np.asarray(y).astype(np.uint8))
            g = np.ascontiguousarray(np.asarray(g).astype(np.uint8).reshape(len(x), self.G))
            return x, y, g, len(x)
        a = prep(x_train, y_train, g_train)
        b = prep(x_test, y_test, g_test)
        _check(self._lib.srn_set_dataset(self._h, _ptr(a[0]), _ptr(a[1]), _ptr(a[2]), a[3],
                                         _ptr(b[0]), _ptr(b[1]), _ptr(b[2]), b[3]))
        self.n_train, self.n_test = a[3], b[3]
        self.h2d_bytes += sum(v.nbytes for v in (*a[:3], *b[:3]) if v is not None)

    # ---- population
    def build(self, irs: Sequence[SerannIR], seed: int = 0, individual_seeds: Optional[Sequence[int]] = None):
        """``individual_seeds``: one initialisation seed per individual (independent of shard and position)."""
        descs, offs = [], [0]
        for ir in irs:
            descs += layer_descs(ir)
            offs.append(len(descs))
        arr = (LayerDesc * len(descs))(*descs)
        offs = np.asarray(offs, dtype=np.int32)
        lb = np.asarray([ir.loss_balance for ir in irs], dtype=np.float32)
        if individual_seeds is not None:
            sd = np.asarray([int(v) & (2 ** 64 - 1) for v in individual_seeds], dtype=np.uint64)
            if len(sd) != len(irs):
                raise EngineError('one seed per individual expected')
            _check(self._lib.srn_build_seeded(self._h, arr, _ptr(offs), _ptr(lb), len(irs), _ptr(sd)))
        else:
            _check(self._lib.srn_build(self._h, arr, _ptr(offs), _ptr(lb), len(irs), C.c_uint64(seed & (2 ** 64 - 1))))
        self.n, self.irs = len(irs), list(irs)
        self.h2d_bytes += C.sizeof(arr) + offs.nbytes + lb.nbytes

    def parameters_count(self, i: int) -> int:
        return int(self._lib.srn_parameters_count(self._h, i))

    def device_bytes(self) -> int:
        return int(self._lib.srn_device_bytes(self._h))

    def get_weights(self, i: int) -> List[List[np.ndarray]]:
        """Keras-layout fp16 arrays per layer, ``layer.get_weights()`` order."""
        out = []
        for li, l in enumerate(self.irs[i].layers):
            ws = []
            for which, (_, shape) in enumerate(l.param_shapes()):
                buf = np.empty(shape, dtype=np.float16)
                cnt = C.c_int64()
                _check(self._lib.srn_get_weights(self._h, i, li, which, _ptr(buf), C.byref(cnt)))
                assert cnt.value == buf.size
                ws.append(buf)
            out.append(ws)
        return out

    def set_weights(self, i: int, weights: Sequence[Sequence[np.ndarray]]):
        for li, (l, ws) in enumerate(zip(self.irs[i].layers, weights)):
            for which, ((_, shape), w) in enumerate(zip(l.param_shapes(), ws)):
                w = np.ascontiguousarray(np.asarray(w, dtype=np.float16))
                if w.shape != tuple(shape):
                    raise EngineError(f'layer {li} weight {which}: shape {w.shape}, expected {shape}')
                _check(self._lib.srn_set_weights(self._h, i, li, which, _ptr(w), w.size))

    # ---- the hot path
    def fit(self, epochs: int, batch: int, n_val: int, perms: Optional[np.ndarray] = None, shuffle_seed: int = 0,
            max_steps: int = 0):
        n_fit = self.n_train - n_val
        S = -(-n_fit // batch)
        V = -(-n_val // batch) if n_val else 0
        tc = np.zeros((epochs, S, self.n), dtype=np.int32)
        tl = np.zeros((epochs, S, self.n, 2), dtype=np.float32)
        vc = np.zeros((epochs, V, self.n), dtype=np.int32)
        vm = np.zeros((epochs, V, self.n), dtype=np.float32)
        ms = C.c_float()
        p = None
        if perms is not None:
            p = np.ascontiguousarray(np.asarray(perms, dtype=np.int32).reshape(epochs, n_fit))
        _check(self._lib.srn_fit(self._h, epochs, batch, n_val, _ptr(p), C.c_uint64(shuffle_seed), max_steps,
                                 _ptr(tc), _ptr(tl), _ptr(vc), _ptr(vm), C.byref(ms)))
        self.h2d_bytes += 4 * epochs * n_fit                      # the epochs' row orders (drawn on the host when perms is None)
        self.d2h_bytes += tc.nbytes + tl.nbytes + vc.nbytes + vm.nbytes
        return {'train_correct': tc, 'train_loss': tl, 'val_correct': vc, 'val_mse_sum': vm, 'device_ms': ms.value,
                'steps_per_epoch': S, 'val_chunks': V}

    def evaluate(self, split: int, start: int, rows: int, chunk: int):
        slots = -(-rows // chunk)
        c = np.zeros((slots, self.n), dtype=np.int32)
        m = np.zeros((slots, self.n), dtype=np.float32)
        _check(self._lib.srn_evaluate(self._h, split, start, rows, chunk, _ptr(c), _ptr(m)))
        self.d2h_bytes += c.nbytes + m.nbytes
        return c, m

    def replicate(self, parent_bits: np.ndarray, R: int, image_index: Optional[np.ndarray] = None,
                  want_probs: bool = True):
        pb = np.ascontiguousarray(np.asarray(parent_bits).astype(np.uint8).reshape(self.n, self.G))
        words = (self.G + 31) // 32
        probs = np.empty((self.n, R, self.G), dtype=np.float16) if want_probs else None
        packed = np.empty((self.n, R, words), dtype=np.uint32)
        mut = np.empty((self.n, R), dtype=np.int32)
        idx = None if image_index is None else np.ascontiguousarray(np.asarray(image_index, dtype=np.int32))
        _check(self._lib.srn_replicate(self._h, _ptr(pb), R, _ptr(idx), _ptr(probs), _ptr(packed), _ptr(mut)))
        self.h2d_bytes += pb.nbytes + 4 * self.n * R
        self.d2h_bytes += (probs.nbytes if probs is not None else 0) + packed.nbytes + mut.nbytes
        return probs, packed, mut

    def forward_logits(self, split: int, start: int, rows: int, training: bool = False) -> np.ndarray:
        out = np.empty((self.n, rows, self.ncls + self.G), dtype=np.float16)
        _check(self._lib.srn_forward_logits(self._h, split, start, rows, int(training), _ptr(out)))
        return out

    # ---- device-time accounting
    CATEGORIES = ('gather', 'gemm_fwd', 'gemm_dgrad', 'gemm_wgrad', 'act_bwd', 'bn', 'pool', 'concat', 'loss', 'adam',
                  'readout', 'other')

    def profile(self, on: bool):
        _check(self._lib.srn_profile(self._h, int(on)))

    def profile_read(self):
        ms = np.zeros(len(self.CATEGORIES), dtype=np.float32)
        cnt = np.zeros(len(self.CATEGORIES), dtype=np.int32)
        _check(self._lib.srn_profile_read(self._h, _ptr(ms), _ptr(cnt)))
        return {k: (float(m), int(c)) for k, m, c in zip(self.CATEGORIES, ms, cnt)}

    # ---- parity hooks
    def debug_step(self, rows: np.ndarray, apply_adam: bool):
        r = np.ascontiguousarray(np.asarray(rows, dtype=np.int32))
        _check(self._lib.srn_debug_step(self._h, _ptr(r), len(r), int(apply_adam)))

    def debug_tensor(self, i: int, tensor: int, grad: bool, nb: int) -> np.ndarray:
        l = self.irs[i].layers[tensor - 2]
        buf = np.empty((nb,) + tuple(l.out_shape), dtype=np.float16)
        _check(self._lib.srn_debug_get_tensor(self._h, i, tensor, int(grad), nb, _ptr(buf)))
        return buf

    def debug_grad(self, i: int, layer: int, which: int) -> np.ndarray:
        shape = self.irs[i].layers[layer].param_shapes()[which][1]
        buf = np.empty(shape, dtype=np.float32)
        _check(self._lib.srn_debug_get_grad(self._h, i, layer, which, _ptr(buf), buf.size))
        return buf

    def launch_groups(self) -> int:
        return int(self._lib.srn_debug_launch_groups(self._h))

    def group_probs(self, group: int, cap: int = 4096):
        """Problem shapes of one launch bucket, [n, 20] (see srn_debug_group_probs in srn_engine.cu)."""
        out = np.zeros((cap, 20), np.int32)
        n = int(self._lib.srn_debug_group_probs(self._h, group, cap, _ptr(out)))
        if n < 0:
            raise RuntimeError('srn_debug_group_probs failed')
        return out[:n]

    def profile_detail(self, nb: int, cap: int = 1 << 16):
        """Per-launch records of the last profiled region: (ms[n], info[n, 8], work[n, 3]), see srn_engine.cu."""
        ms = np.zeros(cap, np.float32)
        self.profile_start_ms = start = np.zeros(cap, np.float32)
        info = np.zeros((cap, 8), np.int32)
        work = np.zeros((cap, 3), np.float64)
        n = int(self._lib.srn_debug_profile_detail(self._h, nb, cap, _ptr(ms), _ptr(start), _ptr(info), _ptr(work)))
        if n < 0:
            raise RuntimeError('srn_debug_profile_detail failed')
        return ms[:n], info[:n], work[:n]


def levenshtein(a: str, b: str) -> int:
    """Edit distance of two strings (reference ``experiment.py:16-17`` via edlib)."""
    ab, bb = a.encode(), b.encode()
    return int(load_library().srn_levenshtein(ab, len(ab), bb, len(bb)))


def unpack_bits(packed: np.ndarray, G: int) -> np.ndarray:
    """(..., ceil(G/32)) uint32 -> (..., G) uint8, bit j of word j // 32."""
    p = np.asarray(packed, dtype=np.uint32)
    shifts = np.arange(32, dtype=np.uint32)
    bits = ((p[..., :, None] >> shifts) & 1).astype(np.uint8)
    return bits.reshape(p.shape[:-1] + (-1,))[..., :G]


def readout(probs: np.ndarray, parent_bits: np.ndarray, rows_per_parent: int, device: int = 0):
    """Threshold + bit-pack + mutation count on the GPU (reference ``experiment.py:210,229``)."""
    lib = load_library()
    p = np.ascontiguousarray(np.asarray(probs, dtype=np.float16))
    G = p.shape[-1]
    p2 = p.reshape(-1, G)
    par = np.ascontiguousarray(np.asarray(parent_bits).astype(np.uint8).reshape(-1, G))
    rows = p2.shape[0]
    packed = np.empty((rows, (G + 31) // 32), dtype=np.uint32)
    mut = np.empty(rows, dtype=np.int32)
    _check(lib.srn_readout(device, _ptr(p2), _ptr(par), rows, rows_per_parent, G, _ptr(packed), _ptr(mut)))
    return unpack_bits(packed, G), mut


def bn_train_forward(x: np.ndarray, gamma, beta, moving_mean, moving_variance, device: int = 0):
    """BatchNormalizationF16 training-mode forward on a (rows, C) fp16 array; returns y and the updated moving stats."""
    lib = load_library()
    x = np.ascontiguousarray(np.asarray(x, dtype=np.float16))
    Cn = x.shape[-1]
    x2 = x.reshape(-1, Cn)
    g, b, mm, mv = (np.ascontiguousarray(np.asarray(a, dtype=np.float16)).copy() for a in (gamma, beta, moving_mean, moving_variance))
    y = np.empty_like(x2)
    _check(lib.srn_bn_train_forward(device, _ptr(x2), x2.shape[0], Cn, _ptr(g), _ptr(b), _ptr(mm), _ptr(mv), _ptr(y)))
    return y.reshape(x.shape), mm, mv
