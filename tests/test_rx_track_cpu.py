"""oracle/rx_track.py (float64 specification of the residual CFO / STO stage, own specification -- the reference leaves the
stage compiled out, rx_synced_param.hpp:156-181) on synthetic pilots: a frequency-selective channel sampled on the DRS grid,
rotated by a known common phase advance and a known per-subcarrier advance per N_step symbols; the estimates return both,
the rotation of the specification undoes them."""
import numpy as np
import pytest

from oracle import rx_track


def synthetic_pilots(rng, N_RX, N_TS, n_drs, b, N_step, dphi, dpsi, snr_db=40.0):
    N_OCC, NP = 56 * b, 14 * b
    taps = rng.standard_normal((N_RX, N_TS, 3)) + 1j * rng.standard_normal((N_RX, N_TS, 3))
    delays = np.array([0.0, 0.25, 0.6])   # in samples of the 64 b point grid: the midpoint interpolation of the estimator is first order
    h = np.zeros((N_RX, N_TS, n_drs, NP), dtype=np.complex128)
    i = np.arange(NP)
    for ts in range(N_TS):
        for n in range(n_drs):
            k = rx_track.pilot_k(i, ts, n, N_OCC) - N_OCC // 2
            H = (taps[:, ts, :, None] * np.exp(-2j * np.pi * delays[None, :, None] * k[None, None, :] / (64 * b))).sum(axis=1)
            l = 1 + n * N_step + (1 if ts >= 4 else 0)
            h[:, ts, n, :] = H * np.exp(1j * (l - 1) / N_step * (dphi + k * dpsi))[None, :]
    h += 10 ** (-snr_db / 20) * (rng.standard_normal(h.shape) + 1j * rng.standard_normal(h.shape)) / np.sqrt(2)
    return h, N_OCC


@pytest.mark.parametrize("N_RX,N_TS,n_drs,b,dphi,dpsi", [
    (1, 1, 2, 1, 0.5, 0.0), (2, 2, 4, 2, -1.2, 0.004), (4, 4, 3, 4, 2.0, -0.003), (2, 8, 2, 1, -0.3, 0.01),
    (1, 1, 6, 16, 0.0, 0.0015), (1, 1, 1, 1, 0.7, 0.01)])
def test_estimates_recover_the_rotation(N_RX, N_TS, n_drs, b, dphi, dpsi):
    rng = np.random.default_rng(N_RX * 100 + N_TS * 10 + n_drs)
    N_step = 5 if N_TS <= 2 else 10
    h, N_OCC = synthetic_pilots(rng, N_RX, N_TS, n_drs, b, N_step, dphi, dpsi)
    e_phi, e_psi = rx_track.estimate(h, N_OCC)
    if n_drs < 2:
        assert e_phi == 0.0 and e_psi == 0.0
        return
    assert abs(e_psi - dpsi) < 1e-3, (e_psi, dpsi)
    assert abs(e_phi - dphi) < 0.02, (e_phi, dphi)
    seed = N_RX * 1000 + N_TS * 10 + n_drs
    rotated, _ = synthetic_pilots(np.random.default_rng(seed), N_RX, N_TS, n_drs, b, N_step, dphi, dpsi, snr_db=400.0)
    clean, _ = synthetic_pilots(np.random.default_rng(seed), N_RX, N_TS, n_drs, b, N_step, 0.0, 0.0, snr_db=400.0)
    back = rx_track.derotate_pilots(rotated, dphi, dpsi, N_OCC, N_step)
    assert np.sqrt(np.mean(np.abs(back - clean) ** 2) / np.mean(np.abs(clean) ** 2)) < 1e-9


def test_switches():
    rng = np.random.default_rng(3)
    h, N_OCC = synthetic_pilots(rng, 2, 2, 4, 2, 5, 0.4, 0.005)
    assert rx_track.estimate(h, N_OCC, cfo=False, sto=False) == (0.0, 0.0)
    p, s = rx_track.estimate(h, N_OCC, cfo=True, sto=False)
    assert s == 0.0 and abs(p - 0.4) < 0.05
    p, s = rx_track.estimate(h, N_OCC, cfo=False, sto=True)
    assert p == 0.0 and abs(s - 0.005) < 1e-3
