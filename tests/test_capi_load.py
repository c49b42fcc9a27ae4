"""CPU-side checks of the C ABI: the library loads, exports every symbol include/sba_b200.h declares,
and fails loudly (no CPU fallback) when there is no CUDA device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, ptr


def header_symbols():
    src = open(os.path.join(ROOT, "include", "sba_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sba_[a-z_0-9]+)\s*\(", src)))


def test_exports_every_declared_symbol(sba_lib):
    syms = header_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(sba_lib, s), f"{s} declared in include/sba_b200.h but not exported"
    from steviapp_b200 import _capi
    assert sorted(_capi.EXPORTS) == syms


def test_default_options_match_reference(sba_lib):
    from steviapp_b200 import default_options
    o = default_options()
    # setFuncTolerance / setParamsTolerance defaults (solversactions.cpp:457-525) + Ceres defaults
    assert o.function_tolerance == 1e-8 and o.parameter_tolerance == 1e-10
    assert o.initial_radius == 1e4 and o.min_lm_diagonal == 1e-6 and o.max_lm_diagonal == 1e32
    assert o.jacobi_scaling == 1 and o.pcg_max_iters == 500


def test_no_cpu_fallback(sba_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from steviapp_b200 import SBASolver, SBAError
    with pytest.raises(SBAError) as e:
        SBASolver()
    assert e.value.code == -1 and "no CPU fallback" in str(e.value)


def test_partition_points_balanced(sba_lib):
    from steviapp_b200 import partition_points
    rng = np.random.default_rng(0)
    opp = rng.integers(2, 12, 10_000).astype(np.int32)
    for n in (1, 2, 3, 8):
        b = partition_points(opp, n)
        assert b[0] == 0 and b[-1] == opp.size and np.all(np.diff(b) >= 0)
        loads = [opp[b[i]:b[i + 1]].sum() for i in range(n)]
        assert max(loads) - min(loads) <= 2 * opp.max()
    assert list(partition_points(np.zeros(0, np.int32), 2)) == [0, 0, 0]
