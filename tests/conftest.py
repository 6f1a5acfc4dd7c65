import importlib.util
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REFERENCE = "/root/reference/grace_dl"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_reference(dotted):
    """Import a module of the read-only reference tree under the alias ``ref_grace_dl``.
    Its absolute ``from grace_dl.dist import Compressor`` imports resolve to this repo's
    compatibility shim, which is fine: only the codec arithmetic is under test."""
    if not os.path.isdir(REFERENCE):
        pytest.skip("reference tree not mounted")
    path = os.path.join(REFERENCE, *dotted.split(".")) + ".py"
    name = "ref_grace_dl." + dotted
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture
def reference():
    return load_reference
