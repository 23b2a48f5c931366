"""Frequency-domain stage of the packet search between coarse and fine peak (k_sync_fdomain; reference
coarse_peak_f_domain_t, compiled out there -- sync_param.hpp:244, 252): beta search and integer CFO.  CUDA against the
written specification oracle/sync_fdomain.py on the same windows (integers: equal), and the function itself: packets of
every beta at one sample rate are told apart, packets several subcarriers off frequency decode once the search is on."""
import numpy as np
import pytest

from oracle import spec, sync_fdomain as ofd

pytestmark = pytest.mark.gpu
NET = 123456789


def _packets(product, cfg, os, n, N_RX, snr, seed, inc=0.0):
    ps = spec.get_packet_sizes(*cfg)
    phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=NET, oversampling=os)
    rng = np.random.default_rng(seed)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    phy.tx(plcf, tb, iq_phase_increment=inc, fetch=False)
    t0 = rng.integers(32 * cfg[1] * os + 64, 2 * ps["N_samples_STF"] * os - 64, n)   # behind the two patterns the search skips
    win = phy.channel(n, snr_dB=snr, t0=t0, mode=0, seed=seed + 1, fetch=True)
    stride = phy.rx_stride()
    phy.close()
    return win, tb, plcf, t0, stride


def test_beta_search_tells_every_beta_apart(product):
    """a receiver at beta 16 and packets of beta 1, 2, 4, 8, 12 (oversampled to the same rate) and 16: same STF length in
    samples, the detection triggers on all of them (sync_param.hpp:238-243), the beta search names the beta"""
    N_RX, n = 2, 3
    rx = product.Phy((1, 16, 1, 1, 0, 2, 2048), max_batch=n, N_RX=N_RX, network_id=NET)
    rx.sync_set_fdomain(beta_threshold_dB=-10.0)
    for b, os in ((1, 16), (2, 8), (4, 4), (8, 2), (16, 1)):
        win, _, _, t0, stride = _packets(product, (1, b, 1, 1, 0, 2, 2048), os, n, N_RX, 20.0, 40 + b)
        assert stride == rx.rx_stride()
        res = rx.sync(win)
        assert res.detected.all(), b
        bg, cfo = rx.sync_fdomain_read(n)
        assert (bg == b).all() and (cfo == 0).all(), (b, bg, cfo)
        for i in range(n):
            bo, so = ofd.fdomain(win[i], res.coarse_peak_time_local[i], res.cfo_fractional_rad[i], 16, 1, 768,
                                 beta_threshold_dB=-10.0)
            assert bo == bg[i] and so == 0, (b, i, bo)
    with pytest.raises(product.DnrpError):
        rx.sync_set_fdomain(cfo_search_range_subc=64)
    rx.close()


@pytest.mark.parametrize("cfg,N_RX,mode", [((1, 4, 1, 1, 0, 3, 2048), 1, 0), ((1, 16, 1, 2, 5, 7, 6144), 4, 0),
                                           ((1, 16, 1, 2, 5, 4, 6144), 4, 2)])
def test_integer_cfo_search(product, cfg, N_RX, mode):
    """the transmitter sits 4 s + f subcarriers off (s = -1, 0, 1, 2; f inside the +-2 the STF resolves): with the search
    on, the integer part is found (equal to the specification), fine peak and decoding work; without it only s = 0 does"""
    ps = spec.get_packet_sizes(*cfg)
    N = 64 * cfg[1]
    shifts = [(-1, 0.3), (0, -1.2), (1, 0.7), (2, -0.4), (-2, 1.1), (0, 0.0)]
    n = len(shifts)
    inc = np.array([2 * np.pi * (4 * s + f) / N for s, f in shifts])
    phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=NET)
    rng = np.random.default_rng(9)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = np.zeros((n, 10), dtype=np.uint8)
    phy.tx(plcf, tb, iq_phase_increment=inc, fetch=False)
    t0 = rng.integers(32 * cfg[1] + 64, 2 * ps["N_samples_STF"] - 64, n)   # behind the two patterns the search skips
    kw = dict(pdp_idx=1, tau_rms_ns=100.0, fD_Hz=50.0, time_64=1000) if mode == 2 else {}
    win = phy.channel(n, snr_dB=25.0, t0=t0, mode=mode, seed=3, fetch=True, **kw)
    # off: the fine peak search decorrelates under an integer CFO (sync_param.hpp:246-252), those packets are lost
    sres, pcc, ok, tbo, _ = phy.rx(win)
    for i, (s, f) in enumerate(shifts):
        assert bool(ok[i]) == (s == 0), (i, s, bool(ok[i]))
    phy.sync_set_fdomain(cfo_search_range_subc=8)
    sres, pcc, ok, tbo, _ = phy.rx(win)
    b, cfo = phy.sync_fdomain_read(n)
    assert (b == cfg[1]).all()
    for i, (s, f) in enumerate(shifts):
        assert cfo[i] == 4 * s, (i, s, cfo)
        assert abs(sres.cfo_integer_rad[i] + 2 * np.pi * 4 * s / N) < 1e-6
        assert abs(sres.cfo_fractional_rad[i] + 2 * np.pi * f / N) < (2e-4 if mode == 0 else 1e-3), \
            (i, f, sres.cfo_fractional_rad[i] * N / 2 / np.pi)
        if mode == 0:   # a dispersive channel moves the fine peak by a few samples
            assert int(sres.fine_peak_time_local[i]) == int(t0[i]), (i, sres.fine_peak_time_local[i], t0[i])
        else:
            assert abs(int(sres.fine_peak_time_local[i]) - int(t0[i])) <= 8
        bo, so = ofd.fdomain(win[i], sres.coarse_peak_time_local[i], sres.cfo_fractional_rad[i], cfg[1], 1, 12 * cfg[1] * 4,
                             cfo_range=8, N_eff_TX_max=ps["N_eff_TX"])
        assert so == cfo[i], (i, so, cfo[i])
    assert ok.all() and np.array_equal(tbo, tb)
    # switched off again: the stage is skipped
    phy.sync_set_fdomain()
    _, _, ok2, _, _ = phy.rx(win)
    assert [bool(v) for v in ok2] == [s == 0 for s, _ in shifts]
    phy.close()
