"""CPU tests of the resampler restatement (oracle/resample.c) and of the product's host-side filter design.

Known answers the reference itself states: the filter / subfilter lengths baked into the names of its hand-unrolled
kernels (resampler.cpp:462-564, resampler.hpp:121-133): L10 M9 -> 223 taps / 23 per subfilter, L9 M10 -> 223 / 25 at
os = 1; 45 / 5 for both directions at os = 2."""
import ctypes as C

import numpy as np
import pytest

from oracle import resample as ors
from conftest import load_product


@pytest.mark.parametrize("L,M,os,flen,slen", [(10, 9, 1, 223, 23), (9, 10, 1, 223, 25), (10, 9, 2, 45, 5),
                                              (9, 10, 2, 45, 5)])
def test_filter_lengths_named_by_the_reference(L, M, os, flen, slen):
    r = ors.Resampler(L, M, os)
    assert (r.filter_length, r.subfilter_length) == (flen, slen)
    assert r.history_length == slen - 1 and r.n_skip == ((flen - 1) // 2) // L
    h = r.lowpass()
    assert abs(float(h.sum()) - L) < 1e-4          # unit DC gain times L (resampler.cpp:100-113)
    assert np.allclose(h[:flen], h[:flen][::-1], atol=1e-7)   # linear phase


@pytest.mark.parametrize("L,M,os", [(10, 9, 1), (9, 10, 1), (10, 9, 2), (40, 27, 1), (27, 40, 1), (9, 10, 8)])
def test_streaming_equals_closed_form(L, M, os):
    rng = np.random.default_rng(L * 100 + M)
    x = (rng.standard_normal(2000) + 1j * rng.standard_normal(2000)).astype(np.complex64)
    r = ors.Resampler(L, M, os)
    one = ors.stream(x, L, M, os)
    assert one.size == r.n_after(x.size)
    # block structure of the calls must not matter (history carries the state)
    parts = ors.stream(x, L, M, os, blocks=[700, 64, 1000, 236])
    assert np.array_equal(one, parts)
    ref = ors.closed_form(x, L, M, os)
    assert np.abs(one - ref).max() < 5e-6
    # without the flush only outputs whose newest input sample exists are produced
    nf = ors.stream(x, L, M, os, flush=False)
    fd = (r.filter_length - 1) // 2
    assert nf.size == (x.size * L - 1 - fd) // M + 1 and np.array_equal(nf, one[:nf.size])


def test_resampling_preserves_a_tone():
    """a complex exponential well inside the pass band comes out as the same tone at the new rate"""
    L, M = 10, 9
    n = np.arange(4000)
    f = 0.05
    x = np.exp(2j * np.pi * f * n).astype(np.complex64)
    y = ors.stream(x, L, M, 1)
    m = np.arange(200, 4000)
    ref = np.exp(2j * np.pi * f * M / L * m)
    # delay compensated exactly (no phase offset); the wide 14 dB filter ripples a few percent in amplitude, "the
    # expectable EVM is about 25 dB" (resampler_param.hpp:64-66)
    assert abs(np.angle(np.mean(y[m] * ref.conj()))) < 1e-6
    assert np.abs(y[m] - ref).max() < 0.06


@pytest.mark.parametrize("L,M,os", [(10, 9, 1), (9, 10, 1), (10, 9, 2), (40, 27, 1), (27, 40, 1), (10, 9, 4), (9, 10, 8)])
def test_product_design_equals_oracle(L, M, os):
    lib = load_product().lib()
    info = (C.c_uint32 * 4)()
    taps = np.zeros(8192, dtype=np.float32)
    assert lib.dnrp_resampler_design(L, M, os, taps.ctypes.data_as(C.c_void_p), taps.size, info) == 0
    r = ors.Resampler(L, M, os)
    assert (info[0], info[1], info[2]) == (r.filter_length, r.subfilter_length, (r.filter_length - 1) // 2)
    assert np.array_equal(taps[:info[3]], r.lowpass())      # same float operations in the same order: bit equal
    lib.dnrp_resampler_n_after.restype = C.c_uint32
    for n in (r.history_length, 100, 1152, 11520, 12345):
        assert lib.dnrp_resampler_n_after(L, M, os, n) == r.n_after(n)


def test_design_rejects_unknown_oversampling():
    lib = load_product().lib()
    info = (C.c_uint32 * 4)()
    assert lib.dnrp_resampler_design(10, 9, 3, None, 0, info) != 0
