"""Cost reduction of the batched parameter-fitting path (reference src/crm_fit/predict.rs:258-285)."""
import numpy as np
import pytest

from cr_mech_coli_b200 import _capi
from cr_mech_coli_b200.crm_fit import costs_from_snapshots


def _snap(it, pos, kinds=None):
    n = len(pos)
    ids = np.zeros(n, _capi.IDENT_DTYPE)
    ids["kind"] = _capi.ID_INITIAL if kinds is None else kinds
    ids["counter"] = np.arange(n)
    return (it, pos.astype(np.float32), np.zeros_like(pos, dtype=np.float32), np.zeros(n, _capi.ROD_PARAMS_DTYPE), ids,
            ids.copy())


def test_cost_is_sum_of_squared_xy_distances():
    rng = np.random.default_rng(0)
    data = rng.normal(size=(2, 3, 8, 2)).astype(np.float32)
    p1 = rng.normal(size=(3, 8, 3)).astype(np.float32)
    p2 = rng.normal(size=(3, 8, 3)).astype(np.float32)
    # snapshots given out of order; iterations index the SORTED save iterations
    per_sim = [[_snap(20, p2), _snap(0, p1 * 0), _snap(10, p1)], [_snap(0, p1), _snap(10, p1), _snap(20, p1)]]
    c = costs_from_snapshots(per_sim, data, iterations=[1, 2])
    want0 = ((p1[:, :, :2] - data[0]) ** 2).sum() + ((p2[:, :, :2] - data[1]) ** 2).sum()
    want1 = ((p1[:, :, :2] - data[0]) ** 2).sum() + ((p1[:, :, :2] - data[1]) ** 2).sum()
    assert np.allclose(c, [want0, want1], rtol=1e-5)


def test_cost_error_rules():
    data = np.zeros((1, 2, 4, 2), np.float32)
    pos = np.zeros((2, 4, 3), np.float32)
    bad = pos.copy()
    bad[0, 0, 0] = np.nan
    c = costs_from_snapshots([[_snap(0, pos), _snap(5, bad)], [_snap(0, pos), _snap(5, pos)]], data, [1], error_cost=123.0)
    assert c[0] == 123.0 and c[1] == 0.0
    with pytest.raises(IndexError):
        costs_from_snapshots([[_snap(0, pos, kinds=[_capi.ID_INITIAL, _capi.ID_DIVISION])]], data, [0])
