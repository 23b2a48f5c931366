"""The written specification of the frequency-domain sync stage (oracle/sync_fdomain.py) on packets of the oracle's own
transmitter: the beta of packets sent at one sample rate is named, an integer carrier offset of 4 s subcarriers is found
whatever the timing error of the coarse peak.  (The CUDA kernel is held against this file in test_sync_fdomain_gpu.py.)"""
import numpy as np
import pytest

from oracle import spec, sync_fdomain as ofd, tx


def _window(b, os, cfo_subc, t0, snr_db, rng, N_eff_TX_cfg=0):
    cfg = (1, b, 1, 1, N_eff_TX_cfg, 2, 2048)
    ps = spec.get_packet_sizes(*cfg)
    tb = rng.integers(0, 256, ps["N_TB_byte"], dtype=np.uint8)
    iq = tx.tx_packet(ps, np.zeros(10, dtype=np.uint8), tb, os=os)          # [N_TX][samples]
    N = 64 * b * os
    n = np.arange(iq.shape[1])
    x = iq[0] * np.exp(2j * np.pi * cfo_subc * n / N)
    win = np.zeros(t0 + len(x) + 64, dtype=np.complex128)
    win[t0:t0 + len(x)] = x
    sigma = 10 ** (-snr_db / 20) / np.sqrt(os)
    win += sigma * (rng.standard_normal(len(win)) + 1j * rng.standard_normal(len(win))) / np.sqrt(2)
    return win[None, :], ps


@pytest.mark.parametrize("b,os", [(1, 8), (2, 4), (4, 2), (8, 1)])
def test_beta_is_named(b, os):
    rng = np.random.default_rng(b)
    win, ps = _window(b, os, 0.0, 50, 15.0, rng)
    for err in (-3, 0, 5):      # the coarse peak is a few samples off
        got, s = ofd.fdomain(win, 50 + err, 0.0, 8, 1, 3 * 64 * 8 // 4, beta_threshold_dB=-10.0)
        assert got == b and s == 0, (b, err, got)


@pytest.mark.parametrize("s_true", [-2, -1, 0, 1, 2])
def test_integer_cfo_is_found(s_true):
    rng = np.random.default_rng(10 + s_true)
    b, frac = 4, 0.6
    win, ps = _window(b, 1, 4 * s_true + frac, 40, 12.0, rng)
    N = 64 * b
    # the fractional part as the pattern autocorrelation would report it (a correction increment)
    cfo_frac = -2 * np.pi * frac / N
    for err in (-2, 0, 3):
        got, s = ofd.fdomain(win, 40 + err, cfo_frac, b, 1, 3 * N // 4, cfo_range=8)
        assert got == b and s == 4 * s_true, (s_true, err, s)
