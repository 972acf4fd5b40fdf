"""CPU: the C-ABI library loads, exports every symbol include/b200fhe.h declares, and refuses to compute
without a GPU (no CPU fallback)."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "b200fhe.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200fhe_[A-Za-z0-9_]+)\s*\(", src)))


def test_exports_every_declared_symbol(fhe):
    L = fhe.lib()
    names = _declared()
    assert len(names) >= 25
    for n in names:
        assert hasattr(L, n), "libb200fhe.so does not export " + n
    assert set(names) == set(fhe.EXPORTS)


def test_host_side_errors(fhe):
    with pytest.raises(fhe.B200FheError):
        fhe.Context(12, np.array([17], dtype=np.uint64), 0, host_only=True)   # not an NTT prime
    with pytest.raises(fhe.B200FheError):
        fhe.coeff_modulus_create(4096, [70])


def test_no_cpu_fallback(fhe):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    pr = fhe.coeff_modulus_create(4096, [60, 60, 60])
    with pytest.raises(fhe.B200FheError, match="no CPU fallback|CUDA"):
        fhe.Context(12, pr, 1)
    ctx = fhe.Context(12, pr, 1, host_only=True)
    a = np.zeros((3, 4096), dtype=np.uint64)
    with pytest.raises(fhe.B200FheError, match="no CPU fallback"):
        ctx.ntt_forward_host(a, 3)
