"""Shared fixtures.  `gpu` marks tests that need a CUDA device (run with `-m gpu` on the B200
box); everything else runs on CPU.  tests/ is one of the few places allowed to import oracle/."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200)")


def _cuda_available() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _cuda_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


def random_text(rng: np.random.Generator, n: int, sigma: int, newline_p: float = 0.05, base: int = ord("a")):
    """n bytes over `sigma` letters with '\\n' sprinkled in (the parity matrix of SURVEY 4.1)."""
    t = (rng.integers(0, sigma, size=n) + base).astype(np.uint8)
    if newline_p > 0:
        t[rng.random(n) < newline_p] = 10
    return t


# (n, sigma, L) grid shared by the oracle-vs-reference and the GPU parity tests
GRID_N = [1, 2, 3, 31, 32, 33, 34, 64, 65, 200, 1000, 3000, 20000]
GRID_SIGMA = [1, 2, 3, 8, 26]
GRID_L = [1, 4, 8, 15, 16, 17, 31, 32, 33, 64]


def safe_queries_for(text: np.ndarray, rng: np.random.Generator, nq: int, max_len: int = 40):
    """Queries without NUL/newline whose first byte is > '\\n' (so the reference never takes its
    mid == 0 wrap when rank 0 is a newline suffix): substrings of the text plus mutations."""
    n = text.size
    out = []
    for _ in range(nq):
        m = int(rng.integers(1, max_len + 1))
        p = int(rng.integers(0, max(1, n)))
        q = bytes(text[p:p + m])
        q = q.split(b"\n")[0]
        if not q:
            q = b"a"
        if rng.random() < 0.3:
            j = int(rng.integers(0, len(q)))
            q = q[:j] + bytes([int(rng.integers(ord("a"), ord("z") + 1))]) + q[j + 1:]
        out.append(q)
    return out
