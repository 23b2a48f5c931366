"""The committed fixtures (tests/golden/golden_v1.npz, made by tests/golden/make_golden.py from the CPU oracle) are
reproduced by today's oracle: integer/bit results exactly, floating-point ones to 1e-5 (libm differences between
machines only).  The reference itself cannot be built here, so these are drift guards, not reference outputs."""
import importlib.util
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def _load_maker():
    spec_ = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
    m = importlib.util.module_from_spec(spec_)
    spec_.loader.exec_module(m)
    return m


def test_oracle_reproduces_golden_vectors():
    gold = np.load(os.path.join(HERE, "golden", "golden_v1.npz"))
    now = _load_maker().build()
    assert set(now) == set(gold.files)
    for k in gold.files:
        g, n = gold[k], np.asarray(now[k])
        assert g.shape == n.shape and g.dtype == n.dtype, k
        if np.issubdtype(g.dtype, np.integer) or g.dtype == np.bool_:
            assert np.array_equal(g, n), k
        else:
            err = np.sqrt(np.mean(np.abs(g.astype(np.complex128) - n) ** 2) / (np.mean(np.abs(g) ** 2) + 1e-30))
            assert err < 1e-5, (k, err)


def test_product_packet_sizes_match_golden(product):
    gold = np.load(os.path.join(HERE, "golden", "golden_v1.npz"))
    keys = _load_maker().SIZE_KEYS
    for cfg, val in zip(gold["sizes_cfg"], gold["sizes_val"]):
        ps = product.get_packet_sizes(*[int(v) for v in cfg])
        assert ps is not None and [ps[k] for k in keys] == list(val), cfg
