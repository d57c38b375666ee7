"""The C-ABI library loads on a machine without a GPU, exports every symbol the header declares, and
the product path refuses to run without a CUDA device (no CPU fallback)."""
import ctypes
import pathlib
import re

import pytest

ROOT = pathlib.Path(__file__).resolve().parent.parent
HEADER = ROOT / 'include' / 'serann_b200.h'


def declared_functions():
    text = re.sub(r'/\*.*?\*/', '', HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r'\b(srn_[a-z0-9_]+)\s*\(', text)))


def test_header_declares_the_path():
    names = declared_functions()
    for required in ('srn_create', 'srn_destroy', 'srn_set_dataset', 'srn_build', 'srn_fit', 'srn_evaluate',
                     'srn_replicate', 'srn_get_weights', 'srn_set_weights', 'srn_decoder_create',
                     'srn_decoder_decode', 'srn_readout', 'srn_bn_train_forward', 'srn_last_error'):
        assert required in names


def test_library_exports_every_declared_symbol():
    from serann_b200.engine import load_library
    lib = load_library()
    for name in declared_functions():
        assert hasattr(lib, name), f'{name} declared in include/serann_b200.h but not exported'


def test_header_compiles_as_c(tmp_path):
    src = tmp_path / 't.c'
    src.write_text('#include "serann_b200.h"\nint main(void) { srn_config c; (void)c; return sizeof(srn_layer_desc) == 18 * 4 ? 0 : 1; }\n')
    import subprocess
    exe = tmp_path / 't'
    subprocess.run(['gcc', '-std=c99', '-Wall', '-Werror', '-I', str(ROOT / 'include'), str(src), '-o', str(exe)], check=True)
    assert subprocess.run([str(exe)]).returncode == 0


def test_layer_desc_layout_matches_ctypes():
    from serann_b200.engine import Config, LayerDesc
    assert ctypes.sizeof(LayerDesc) == 18 * 4
    assert ctypes.sizeof(Config) == 6 * 4


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    from serann_b200.engine import Engine, EngineError, readout
    import numpy as np
    with pytest.raises(EngineError, match='no CUDA device'):
        Engine()
    with pytest.raises(EngineError, match='no CUDA device'):
        readout(np.zeros((2, 100), np.float16), np.zeros((2, 100), np.uint8), 1)


def test_product_package_never_imports_the_oracle():
    pkg = ROOT / 'self-replicating-artificial-neural-networks_b200'
    for py in pkg.glob('*.py'):
        text = py.read_text()
        assert 'import oracle' not in text and 'from oracle' not in text, py.name
