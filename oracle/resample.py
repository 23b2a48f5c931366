"""ORACLE (test infrastructure, not product code) -- Python front of oracle/resample.c (reference
lib/src/phy/resample/resampler.cpp, filter/kaiser.cpp, resampler_param.hpp).  Parity unpinned (see resample.c)."""
import ctypes as C

import numpy as np

from . import lib

# resampler_param.hpp:71-86, index = os_min (identical for TX / SYNC / RX_SYNCED in the reference snapshot)
F_PASS_NORM = {1: 0.48, 2: 0.30, 4: 0.20, 8: 0.15}
F_STOP_NORM = {1: 0.499, 2: 0.499, 4: 0.499, 8: 0.499}
F_STOP_ATT_DB = {1: 14.0, 2: 20.0, 4: 20.0, 8: 20.0}
PASSBAND_RIPPLE_DONT_CARE = 100.0


class Resampler:
    """resampler_t for ONE antenna stream (the reference keeps one history per antenna and shares the counters)"""

    def __init__(self, L, M, os=1):
        self._l = lib()
        self._l.orc_resampler_new.restype = C.c_void_p
        self._l.orc_resampler_new.argtypes = [C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float]
        self._l.orc_resampler_n_after.restype = C.c_uint32
        self._l.orc_resampler_n_after.argtypes = [C.c_void_p, C.c_uint32]
        self._l.orc_resampler_resample.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]
        self._l.orc_resampler_final.argtypes = [C.c_void_p, C.c_void_p]
        for f in (self._l.orc_resampler_free, self._l.orc_resampler_reset):
            f.argtypes = [C.c_void_p]
            f.restype = None
        self._l.orc_resampler_info.argtypes = [C.c_void_p, C.c_void_p]
        self._l.orc_resampler_lowpass.argtypes = [C.c_void_p, C.c_void_p]
        self.L, self.M, self.os = L, M, os
        self._r = self._l.orc_resampler_new(L, M, F_PASS_NORM[os], F_STOP_NORM[os], PASSBAND_RIPPLE_DONT_CARE,
                                            F_STOP_ATT_DB[os])
        assert self._r
        info = np.zeros(5, dtype=np.uint32)
        self._l.orc_resampler_info(self._r, info.ctypes.data)
        self.filter_length, self.subfilter_length, self.n_skip, self.history_length, self.rows = [int(v) for v in info]

    def __del__(self):
        try:
            self._l.orc_resampler_free(self._r)
        except Exception:
            pass

    def lowpass(self):
        out = np.zeros(self.subfilter_length * self.L, dtype=np.float32)
        self._l.orc_resampler_lowpass(self._r, out.ctypes.data)
        return out

    def reset(self):
        self._l.orc_resampler_reset(self._r)

    def n_after(self, n_in):
        return int(self._l.orc_resampler_n_after(self._r, n_in))

    def resample(self, x):
        x = np.ascontiguousarray(x, dtype=np.complex64)
        out = np.zeros(x.size * self.L // self.M + 2 * self.L + 8, dtype=np.complex64)
        n = self._l.orc_resampler_resample(self._r, x.ctypes.data, out.ctypes.data, x.size)
        if n < 0:
            raise ValueError("fewer input samples than the resampler's history")
        return out[:n]

    def final(self):
        out = np.zeros((self.history_length + 2) * self.L // self.M + 2 * self.L + 8, dtype=np.complex64)
        n = self._l.orc_resampler_final(self._r, out.ctypes.data)
        return out[:n]


def stream(x, L, M, os=1, blocks=None, flush=True):
    """resample one antenna stream the way the reference's callers do: reset, resample() per block (default: one
    block), resample_final_samples().  blocks = list of block lengths summing to len(x)."""
    if L == M:
        return np.array(x, dtype=np.complex64)
    r = Resampler(L, M, os)
    x = np.ascontiguousarray(x, dtype=np.complex64)
    blocks = [x.size] if blocks is None else blocks
    assert sum(blocks) == x.size
    out, pos = [], 0
    for b in blocks:
        out.append(r.resample(x[pos:pos + b]))
        pos += b
    if flush:
        out.append(r.final())
    return np.concatenate(out)


def closed_form(x, L, M, os=1, n_out=None):
    """the closed form the CUDA kernel implements, in float64 (used by CPU tests to validate the restatement of the
    bookkeeping against the defining convolution)"""
    r = Resampler(L, M, os)
    h = r.lowpass().astype(np.float64)
    fd = (r.filter_length - 1) // 2
    n_out = r.n_after(len(x)) if n_out is None else n_out
    xx = np.concatenate([np.zeros(r.subfilter_length, dtype=np.complex128), np.asarray(x, dtype=np.complex128),
                         np.zeros(r.subfilter_length + 2, dtype=np.complex128)])
    m = np.arange(n_out)
    p = fd + m * M
    l, i0 = p % L, p // L
    y = np.zeros(n_out, dtype=np.complex128)
    for j in range(r.subfilter_length):
        y += h[l + j * L] * xx[i0 - j + r.subfilter_length]
    return y
