import pathlib
import sys

import pytest

ROOT = pathlib.Path(__file__).resolve().parent.parent
for p in (str(ROOT), str(ROOT / 'tests'), str(ROOT / 'tests' / 'golden')):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box with -m gpu)')


def pytest_collection_modifyitems(config, items):
    # keep tests/test_zz_* last so a -x run has already exercised everything else
    items.sort(key=lambda it: (pathlib.Path(str(it.fspath)).name.startswith('test_zz_'),))
