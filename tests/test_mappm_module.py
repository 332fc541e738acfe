"""The importable ``mappm`` package (mappm/__init__.py): the three known answers of the reference's
own test of its Fortran extension, external/vcm/tests/test_mappm.py:5-44, written the way that
file writes them (float scalars, float64 inputs)."""
import inspect

import numpy as np
import pytest


def test_mappm_package_signature():
    mappm = pytest.importorskip("mappm")
    assert list(inspect.signature(mappm.mappm).parameters) == ["pe1", "q1", "pe2", "i1", "i2", "iv", "kord", "ptop"]
    assert callable(mappm.interpolate_2d)


@pytest.mark.gpu
def test_mappm_package_known_answers():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import mappm

    n_columns, iv, kord, dummy_ptop = 1.0, 1.0, 1.0, 0.0
    p_in = np.asarray([0.0, 1.0, 2.0, 3.0, 4.0, 5.0])[None, :]
    f_in = np.asarray([0.0, 1.0, 2.0, 3.0, 4.0])[None, :]
    p_out = np.asarray([0.5, 1.2, 2.4, 2.8, 3.2, 4.5])[None, :]
    result = mappm.mappm(p_in, f_in, p_out, 1, n_columns, iv, kord, dummy_ptop)
    assert result.dtype == np.float32
    np.testing.assert_almost_equal(result, np.asarray([[0.35, 1.3, 2.1, 2.5, 3.35]], dtype=np.float32), decimal=5)

    p_in = np.asarray([1.0, 2.0, 3.0, 4.0, 5.0])[None, :]
    f_in = np.asarray([1.5, 2.5, 3.5, 4.5])[None, :]
    p_out = np.asarray([0.0, 2.5, 3.5, 4.5, 50.0])[None, :]
    result = mappm.mappm(p_in, f_in, p_out, 1, n_columns, iv, kord, dummy_ptop)
    np.testing.assert_almost_equal(result, np.asarray([[1.5, 3.0, 4.0, 4.502747]], dtype=np.float32), decimal=5)

    p_in = np.asarray([1.0, 2.0, 3.0, 2.0, 5.0])[None, :]
    f_in = np.asarray([np.nan, np.nan, np.nan, np.nan])[None, :]
    result = mappm.mappm(p_in, f_in, p_out, 1, n_columns, iv, kord, dummy_ptop)
    np.testing.assert_almost_equal(result, np.asarray([[np.nan, np.nan, np.nan, np.nan]], dtype=np.float32))
