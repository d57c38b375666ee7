"""Genotype -> token sequence -> source string, backed by the CUDA decode kernels.

Same class surface as the reference ``RibosomalAutoencoder``
(``evolutionary_experiment/logic/ribosomal_autoencoder.py:12-83``) for the decode direction; the
daemon child process and its queues are gone (the decode runs in-process on the GPU).  ``encode_*``
is outside the per-generation path (never called by the evolution loop): it runs the PyTorch restatement of the
inference net from ``tools/train_autoencoder.py`` when the weights carry an encoder.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List

import numpy as np

from . import engine as _engine
from .vocabulary import PADDING_TOKEN, VOCABULARY_SIZE, sequence_to_string


class RibosomalAutoencoder:
    def __init__(self, max_serann_tokens: int, weights: Dict[str, np.ndarray], genotype_size: int = 100,
                 vocabulary_size: int = VOCABULARY_SIZE, device: int = 0, model_name: str = 'synthetic'):
        self._lib = _engine.load_library()
        self._max_serann_tokens = int(max_serann_tokens)
        self._G = int(genotype_size)
        self._V = int(vocabulary_size)
        self._model_name = model_name
        self.PADDING_TOKEN = PADDING_TOKEN
        L = self._G - 4
        f32 = lambda a, shape: np.ascontiguousarray(np.asarray(a, dtype=np.float32).reshape(shape))
        ck = f32(weights['conv_kernel'], (5, 2, 32))
        cb = f32(weights['conv_bias'], (32,))
        bn1 = np.ascontiguousarray(np.stack([f32(weights[f'bn1_{k}'], (32,)) for k in ('gamma', 'beta', 'mean', 'var')]))
        dk = f32(weights['dense_kernel'], (L * 32, self._max_serann_tokens * self._V))
        db = f32(weights['dense_bias'], (self._max_serann_tokens * self._V,))
        bn2 = np.ascontiguousarray(np.stack([f32(weights[f'bn2_{k}'], (self._V,)) for k in ('gamma', 'beta', 'mean', 'var')]))
        self._weights = weights
        self._d = C.c_void_p()
        _engine._check(self._lib.srn_decoder_create(device, self._G, self._max_serann_tokens, self._V,
                                                    _engine._ptr(ck), _engine._ptr(cb), _engine._ptr(bn1),
                                                    _engine._ptr(dk), _engine._ptr(db), _engine._ptr(bn2),
                                                    C.byref(self._d)))
        self.last_device_ms = 0.0
        self.h2d_bytes = sum(a.nbytes for a in (ck, cb, bn1, dk, db, bn2))
        self.d2h_bytes = 0

    def close(self):
        if getattr(self, '_d', None):
            self._lib.srn_decoder_destroy(self._d)
            self._d = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def get_model_name(self) -> str:
        return self._model_name

    def remove_sequences_padding(self, sequences):
        sequences = np.asarray(sequences)
        lengths = sequences.shape[1] - np.sum(sequences == self.PADDING_TOKEN, axis=1)
        return [sequences[i, :lengths[i]] for i in range(len(sequences))]

    def decode_to_sequence(self, encoded_serann_codes) -> np.ndarray:
        g = np.asarray(encoded_serann_codes)
        if g.ndim != 2 or g.shape[1] != self._G:
            raise ValueError(f'expected (n, {self._G}) genotypes, got {g.shape}')
        gf = g.astype(np.float64)
        if np.isnan(gf).any():
            # reference: np.eye(2)[genotype.astype(int)] raises IndexError on NaN (ribosomal_autoencoder.py:123)
            raise IndexError('NaN genotype cannot be decoded')
        gi = gf.astype(np.int64)
        if ((gi < 0) | (gi > 1)).any():
            raise IndexError('genotype entries must be 0 or 1')
        bits = np.ascontiguousarray(gi.astype(np.uint8))
        tokens = np.empty((len(bits), self._max_serann_tokens), dtype=np.int32)
        ms = C.c_float()
        _engine._check(self._lib.srn_decoder_decode(self._d, _engine._ptr(bits), len(bits), _engine._ptr(tokens), C.byref(ms)))
        self.last_device_ms = ms.value
        re, tp = C.c_int64(), C.c_int32()
        _engine._check(self._lib.srn_decoder_stats(self._d, C.byref(re), C.byref(tp)))
        self.last_stats = {'rechecked_tokens': int(re.value), 'tokens': int(tokens.size), 'tensor_core_kernel': bool(tp.value)}
        self.h2d_bytes += bits.nbytes
        self.d2h_bytes += tokens.nbytes
        return tokens.astype(np.int64)

    def sequence_to_string(self, serann_sequences) -> List[str]:
        if hasattr(serann_sequences, 'values'):           # pandas Series of sequences
            serann_sequences = np.array(serann_sequences.values.tolist())
        return sequence_to_string(serann_sequences)

    def decode_to_string(self, encoded_serann_codes) -> List[str]:
        return self.sequence_to_string(self.decode_to_sequence(encoded_serann_codes))

    def encode_sequence(self, serann_token_index_sequences) -> np.ndarray:
        """Token-id sequences (n, max_tokens) -> genotypes (n, G): argmax of the inference net's logits (reference
        ``ribosomal_autoencoder.py:40-46,119-120``; ``model.py:17-33,101-104``).  Encoding is dataset preparation, never
        called by the evolution loop, and not a CUDA entry point of this library: it runs the PyTorch restatement of
        the inference net in ``tools/train_autoencoder.py`` on the ``enc_*`` arrays that tool exports."""
        enc = {k: v for k, v in self._weights.items() if k.startswith('enc_')}
        if not enc:
            raise NotImplementedError('these weights hold no encoder (enc_* arrays, written by tools/train_autoencoder.py)')
        import importlib.util
        import pathlib
        import torch
        spec = importlib.util.spec_from_file_location('train_autoencoder', pathlib.Path(__file__).resolve().parent.parent / 'tools' / 'train_autoencoder.py')
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        model = mod.GeneticAutoencoder(self._G, self._max_serann_tokens, self._V, embedding=int(enc['enc_emb'].shape[1]))
        sd = model.state_dict()
        for k in sd:
            if k.startswith(('emb', 'e_')):
                sd[k] = torch.as_tensor(enc['enc_' + k.replace('.', '_')])
        model.load_state_dict(sd)
        return model.encode(torch.as_tensor(np.asarray(serann_token_index_sequences)).long()).numpy()

    def encode_string(self, serann_codes) -> np.ndarray:
        """Reference ``ribosomal_autoencoder.py:48-57``: tokenise, truncate / pad with <PAD>, encode."""
        from .vocabulary import encode_tokens
        return self.encode_sequence(np.stack([encode_tokens(s, self._max_serann_tokens) for s in serann_codes]))
