"""mmc_erfc_noalloc: the allocation-free variant of mmc_dbg_erfc (it has no
counterpart in the reference).  Its device scratch is a file-scope `__device__` array reached through
cudaGetSymbolAddress, so the test that matters is simply: through the C ABI, does it return what the existing entry point
returns -- on two calls in a row, because every call shares the array and it keeps the previous call's contents.

This file is named so that pytest collects it LAST (a failure here must not hide the
parity tests).  It passes when run alone in a fresh process and when run after the other tests."""
import numpy as np
import pytest

from conftest import load_system

pytestmark = pytest.mark.gpu
MAX_N = 65536   # MMC_NOALLOC_MAX_N, include/maniac_gpu.h


def test_header_limit_matches():
    import os
    import re
    from conftest import ROOT
    txt = open(os.path.join(ROOT, "include", "maniac_gpu.h")).read()
    assert int(re.search(r"#define MMC_NOALLOC_MAX_N (\d+)", txt).group(1)) == MAX_N


def test_erfc_noalloc_equals_existing_entry_point_on_two_calls_in_a_row():
    from maniac_mc_b200.lib import ManiacGPU, ManiacGpuError
    from scipy.special import erfc
    g = ManiacGPU(load_system("lj_gas"))
    rng = np.random.default_rng(7)
    # call 1: the full scratch, arguments across the table's domain and both slow paths (tiny r, alpha r > 7)
    x1 = np.concatenate([np.linspace(1e-3, 7.5, MAX_N - 4000), 1e-3 + rng.random(4000) * 27.0])
    assert len(x1) == MAX_N
    a1 = g.erfc_noalloc(x1)
    b1 = g.dbg_erfc(x1)
    # same kernel, same inputs: the tolerance of tests/test_gpu_parity.py::test_device_erfc_accuracy (2e-15 absolute) is
    # met with room to spare -- the results are bit-identical
    assert np.max(np.abs(a1 - b1)) < 2e-15 and np.array_equal(a1, b1)
    assert np.max(np.abs(a1 - erfc(x1))) < 2e-15
    # call 2, straight after: a SHORTER, different input -- stale contents of the shared array from call 1 must not leak
    x2 = 0.05 + rng.random(1237) * 6.0
    a2 = g.erfc_noalloc(x2)
    b2 = g.dbg_erfc(x2)
    assert a2.shape == (1237,) and np.array_equal(a2, b2)
    assert np.max(np.abs(a2 - erfc(x2))) < 2e-15
    assert not np.array_equal(a2, a1[:1237])
    # and once more with the first input: the array is reusable in both directions
    assert np.array_equal(g.erfc_noalloc(x1), b1)
    # sizes beyond the fixed scratch are refused with a clear error, and the handle stays usable
    with pytest.raises(ManiacGpuError, match="exceeds the fixed scratch"):
        g.erfc_noalloc(np.ones(MAX_N + 1))
    assert np.array_equal(g.erfc_noalloc(x2), b2)
    g.close()


def test_erfc_noalloc_second_handle_same_device():
    """a second handle (another alpha, another Coulomb table) on the same GPU reuses the cached symbol address"""
    from maniac_mc_b200.lib import ManiacGPU
    x = np.linspace(0.01, 7.2, 5000)
    for name in ("zif8_h2o", "mfi_co2"):
        g = ManiacGPU(load_system(name))
        assert np.array_equal(g.erfc_noalloc(x), g.dbg_erfc(x))
        g.close()
