"""Shared fixtures.  `-m "not gpu"`: oracle vs golden vectors, host logic, ABI surface (no CUDA calls).
`-m gpu`: parity tests proper -- every one goes through the C ABI of build/libbert.so."""
import importlib.util
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def _load(name, path):
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="session")
def gio():
    return _load("embedding_cpp_b200_gguf_io", os.path.join(ROOT, "embedding.cpp_b200", "gguf_io.py"))


@pytest.fixture(scope="session")
def binding():
    return _load("embedding_cpp_b200_binding", os.path.join(ROOT, "embedding.cpp_b200", "binding.py"))


@pytest.fixture(scope="session")
def orc():
    import oracle as O
    O.build(ref=True)
    return O


@pytest.fixture(scope="session")
def lib(binding):
    return binding.load_library()


@pytest.fixture(scope="session")
def model_dir(tmp_path_factory):
    return tmp_path_factory.mktemp("models")


@pytest.fixture(scope="session")
def make_model(gio, model_dir):
    """make_model(shape, ftype, seed=1, n_vocab=500, dist='bert-init') -> path (cached per session)."""
    cache = {}

    def _make(shape, ftype, seed=1, n_vocab=500, dist="bert-init"):
        key = (shape, ftype, seed, n_vocab, dist)
        if key not in cache:
            p = os.path.join(str(model_dir), f"{shape}_{ftype}_{seed}_{n_vocab}_{dist}.gguf")
            gio.make_model(p, shape, ftype, seed=seed, n_vocab=n_vocab, dist=dist)
            cache[key] = p
        return cache[key]

    return _make


def rand_tokens(rng, n_vocab, L):
    t = rng.integers(0, n_vocab, L).astype(np.int32)
    if L >= 1 and n_vocab > 102:
        t[0] = 101
        if L >= 2:
            t[-1] = 102
    return t


def cos_maxabs(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    c = float((a * b).sum() / (np.linalg.norm(a) * np.linalg.norm(b)))
    return c, float(np.abs(a - b).max())
