import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _build_library():
    # build.py is loaded by path: importing the package dlopens the library, which may not exist yet
    import importlib.util
    spec = importlib.util.spec_from_file_location("_sfa_build", os.path.join(ROOT, "starflashattention_b200", "build.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    return b.build()


def _build_torch_ext():
    import importlib.util
    spec = importlib.util.spec_from_file_location("_sfa_build", os.path.join(ROOT, "starflashattention_b200", "build.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    return b.build_torch_ext()


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    _build_library()


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="session", autouse=True)
def built_library():
    """The C-ABI library is built in-tree (nvcc cross-compiles without a GPU); no fallback exists."""
    return _build_library()


@pytest.fixture(scope="session")
def torch_ext(built_library):
    """The reference-facing `star_flash_attn` extension module (host-only C++ over the C ABI), built in-tree."""
    _build_torch_ext()
    import star_flash_attn
    return star_flash_attn


def rand_tensor(shape, dtype, seed, scale=1.0):
    import torch
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(dtype)


@pytest.fixture
def tuning():
    """Sets one of the library's A-B switches (sfa_tuning_set) for a test and restores it afterwards:
    tuning(name, value, default)."""
    import starflashattention_b200 as sfa
    undo = []

    def set_(name, value, default):
        sfa.tuning_set(name, value)
        undo.append((name, default))
    yield set_
    for name, default in undo:
        sfa.tuning_set(name, default)
