"""Token vocabulary, tokenizer and token-sequence -> source string.

The reference reads its vocabulary from a figshare CSV that is not available
offline (reference ``README.md:23-27``).  The vocabulary is a function of the
generator templates and the tokenizer alone: tokenising everything
``synthetic_serann_generator/layer_transitions.py:61-72`` can emit with the
tokenizer of ``helpers/source_codes_to_tokens.py:21-36`` and sorting the set
(``:73-77``) gives 39 tokens; ``<PAD>`` is appended as the last index.
"""
from __future__ import annotations

from typing import Iterable, List, Sequence

import numpy as np

PAD = '<PAD>'

# sorted(unique tokens) + ['<PAD>'] -- index == position (helpers/source_codes_to_tokens.py:73-77)
VOCABULARY: List[str] = sorted([
    '\n', "'", '(', ')', ',', '-', '.', '0', '1', '2', '3', '4', '5', '6', '7', '8', '9', '=',
    'BatchNormalization', 'Conv1D', 'Conv2D', 'Dense', 'MaxPool2D', 'Reshape', 'X_layer', '[', ']',
    'activation', 'con', 'concatenate', 'filters', 'g_layer', 'kernel_size', 'loss_balance',
    'pool_size', 'relu', 'sigmoid', 'strides', 'units',
]) + [PAD]
WORD2INDEX = {t: i for i, t in enumerate(VOCABULARY)}
PADDING_TOKEN = WORD2INDEX[PAD]
VOCABULARY_SIZE = len(VOCABULARY)

SPLIT_CHARACTERS = ('\n', '=', "'", '(', ')', '[', ']', ',', '.')


def _is_int(s: str) -> bool:
    try:
        int(s)
        return True
    except ValueError:
        return False


def tokenize(source: str, split_characters: Sequence[str] = SPLIT_CHARACTERS) -> List[str]:
    """Reference ``Tokenizer.__call__`` (``evolutionary_experiment/logic/ribosomal_autoencoder.py:129-152``).

    Iterative restatement of the recursive splitter: blank lines and spaces are
    removed, the text is cut at each split character (which is kept as a
    token), and integer runs are cut into single digits.
    """
    pieces = [source]
    for c in split_characters:
        nxt: List[str] = []
        for p in pieces:
            if p in split_characters:       # an already-emitted separator token
                nxt.append(p)
                continue
            p = p.replace('\n\n', '\n').replace('\n\n', '\n').replace(' ', '')
            if p == '':
                continue
            parts = p.split(c)
            for i, part in enumerate(parts):
                nxt.append(part)
                if i + 1 < len(parts):
                    nxt.append(c)
        pieces = nxt
    out: List[str] = []
    for p in pieces:
        if p in split_characters:
            out.append(p)
            continue
        p = p.replace(' ', '')
        if _is_int(p):
            out.extend(list(p))
        else:
            out.append(p)
    return out


def encode_tokens(source: str, max_tokens: int) -> np.ndarray:
    """Token ids padded with ``<PAD>`` to ``max_tokens`` (``ribosomal_autoencoder.py:48-57``)."""
    toks = tokenize(source)[:max_tokens]
    ids = np.full(max_tokens, PADDING_TOKEN, dtype=np.int64)
    ids[:len(toks)] = [WORD2INDEX[t] for t in toks]
    return ids


def sequence_to_string(sequences: Iterable[Sequence[int]]) -> List[str]:
    """Reference ``RibosomalAutoencoder.sequence_to_string`` (``ribosomal_autoencoder.py:69-78``).

    ``str.rstrip('<PAD>')`` strips a *character set*, so trailing ``P A D < >``
    characters of a real token go too (``...MaxPool2D`` -> ``...MaxPool2``);
    interior ``<PAD>`` tokens stay.  Kept as is for parity.
    """
    lut = np.array(VOCABULARY, dtype=object)
    seqs = np.asarray(list(sequences) if not isinstance(sequences, np.ndarray) else sequences)
    if seqs.size == 0:
        return []
    words = lut[seqs.astype(np.int64)]
    return [''.join(row).rstrip(PAD) for row in words]
