"""Residual CFO / STO tracking on the DRS (dnrp_rx_set_tracking, k_rx_track; off by default).  The reference plans the
stage and leaves it compiled out (rx_synced_param.hpp:156-181 "not functional yet", rx_synced.cpp:1165-1169 empty todo), so
the specification is the repo's own (oracle/rx_track.py) and this parity is CUDA == oracle, not CUDA ==
reference.

  * the two estimates per packet against the float64 specification evaluated on the pilots the same context produced with
    tracking off (|diff| <= 5e-4 rad: fp32 sums over up to a few thousand pilot products), the pilots after the stage
    against the specification's rotation (relative error <= 1e-4);
  * the estimates against the impairment put into the windows: a frequency step right behind the STF (which neither the
    packet synchronisation nor the STF-based estimate of the receiver can see) and a sampling clock offset (the window
    resampled by 1 + zeta with a windowed sinc);
  * packets that only decode with the stage switched on; with no impairment the stage changes no outcome."""
import numpy as np
import pytest

from oracle import rx_track, spec
from test_rx_gpu import make_windows

pytestmark = pytest.mark.gpu
NET = 123456789


def frequency_step(win, start, delta):
    """e^{-j delta (n - start)} from sample start on"""
    n = np.arange(win.shape[-1])
    ph = np.where(n >= start, -delta * (n - start), 0.0)
    return (win * np.exp(1j * ph)[None, None, :]).astype(np.complex64)


def clock_offset(win, zeta, centre, taps=32):
    """y[m] = x(centre + (m - centre) (1 + zeta)): windowed-sinc interpolation, float64"""
    n, A, S = win.shape
    m = np.arange(S)
    t = centre + (m - centre) * (1.0 + zeta)
    j0 = np.floor(t).astype(np.int64)
    out = np.zeros((n, A, S), dtype=np.complex128)
    x = win.astype(np.complex128)
    for d in range(-taps + 1, taps + 1):
        j = j0 + d
        w = np.sinc(t - j) * (0.5 + 0.5 * np.cos(np.pi * (t - j) / taps))
        ok = (j >= 0) & (j < S)
        out[:, :, ok] += x[:, :, j[ok]] * w[ok][None, None, :]
    return out.astype(np.complex64)


def dims_of(ps, os_=1):
    N = ps["N_b_DFT"] * os_
    cp = ps["N_b_CP"] * os_
    stf = ps["N_samples_STF"] * os_
    return N, cp, stf


TRACK_CASES = {
    # cfg = (u, b, PacketLengthType, PacketLength, tm_mode, mcs, Z)
    "siso_b2_2slots": dict(cfg=(1, 2, 1, 2, 0, 3, 2048), N_RX=1, snr_db=30.0, t0=6, step_turns=0.12, zeta=150e-6),
    "txdiv2_b1_2rx_3slots": dict(cfg=(1, 1, 1, 3, 1, 3, 2048), N_RX=2, snr_db=30.0, t0=3, step_turns=-0.08, zeta=-200e-6),
    "txdiv4_b2_4rx_4slots": dict(cfg=(1, 2, 1, 4, 5, 3, 2048), N_RX=4, snr_db=30.0, t0=9, step_turns=0.15, zeta=80e-6),
    "txdiv8_b1_2rx_4slots": dict(cfg=(1, 1, 1, 4, 10, 1, 2048), N_RX=2, snr_db=30.0, t0=0, step_turns=-0.1, zeta=-100e-6),
    "sm_4x4_b1_4slots": dict(cfg=(1, 1, 1, 4, 6, 1, 2048), N_RX=4, snr_db=30.0, t0=4, step_turns=0.1, zeta=100e-6),
}


@pytest.mark.parametrize("name", list(TRACK_CASES))
def test_estimates_match_the_specification_and_the_impairment(product, name):
    case = dict(TRACK_CASES[name])
    cfg, N_RX, t0 = case["cfg"], case["N_RX"], case["t0"]
    ps = spec.get_packet_sizes(*cfg)
    n = 3
    phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=NET, plcf_type=1)
    rng = np.random.default_rng(sum(map(ord, name)))
    win, tbs, plcfs = make_windows(ps, case, n, phy.rx_stride(), rng, 1)
    N, cp, stf = dims_of(ps)
    N_eff_TX = ps["N_eff_TX"]
    N_step = 5 if N_eff_TX <= 2 else 10
    N_OCC = 56 * cfg[1]
    # step_turns: phase advance per N_step symbols, in turns
    delta = -2 * np.pi * case["step_turns"] / (N_step * (N + cp))
    # what the stage sees without impairment is the error of the receiver's STF-based CFO estimate (k_rx_stf) times the
    # distance of the DRS symbols -- at beta = 1 (an STF of 112 samples) some 0.1 rad per 10 symbols at 30 dB: the
    # impairments are measured against it
    phy.rx_set_tracking(True, True)
    phy.rx_pcc(win, t0=t0)
    est0 = phy.rx_tracking_read(n)
    for kind in ("step", "clock"):
        w = frequency_step(win, t0 + stf, delta) if kind == "step" else clock_offset(win, case["zeta"], t0 + stf // 2)
        phy.rx_set_tracking(False, False)
        phy.rx_pcc(w, t0=t0)
        raw = phy.tap(product.TAP_DRS_ZF, n).reshape(n, N_RX, N_eff_TX, -1, N_OCC // 4)
        phy.rx_set_tracking(True, True)
        phy.rx_pcc(w, t0=t0)
        est = phy.rx_tracking_read(n)
        rot = phy.tap(product.TAP_DRS_ZF, n).reshape(raw.shape)
        for i in range(n):
            dphi, dpsi = rx_track.estimate(raw[i], N_OCC)
            assert abs(est[i, 0] - dphi) <= 5e-4 and abs(est[i, 1] - dpsi) <= 5e-5, (name, kind, i, est[i], dphi, dpsi)
            ref = rx_track.derotate_pilots(raw[i], float(est[i, 0]), float(est[i, 1]), N_OCC, N_step)
            err = np.sqrt(np.mean(np.abs(rot[i] - ref) ** 2) / np.mean(np.abs(ref) ** 2))
            assert err <= 1e-4, (name, kind, i, err)
            if kind == "step":
                # e^{-j delta n} behind the STF: advance per N_step symbols = -delta N_step (N + CP)
                d = est[i, 0] - est0[i, 0] - 2 * np.pi * case["step_turns"]
                assert abs(d - 2 * np.pi * np.round(d / (2 * np.pi))) <= 0.03, (name, i, est[i], est0[i], case["step_turns"])
                assert abs(est[i, 1] - est0[i, 1]) <= 1e-3, (name, i, est[i], est0[i])
            else:
                # sampling the signal at (1 + zeta) n: the symbol timing advances by zeta N_step (N + CP) samples per
                # N_step symbols = 2 pi zeta N_step (N + CP) / N rad per subcarrier
                want = 2 * np.pi * case["zeta"] * N_step * (N + cp) / N
                assert abs(est[i, 1] - est0[i, 1] - want) <= 0.25 * abs(want) + 5e-4, (name, i, est[i], est0[i], want)   # 28 pilots per symbol in the smallest case
    phy.close()


def _decode(phy, win, t0, n):
    pcc = phy.rx_pcc(win, t0=t0)
    ok, tb, snr, nit = phy.rx_pdc(n)
    return pcc, ok, tb


@pytest.mark.parametrize("name,cfg,N_RX,step_turns,lr", [
    ("txdiv4_64qam_mode_lr", (1, 2, 1, 4, 5, 6, 2048), 2, 0.2, True),
    ("siso_64qam_mode_lr", (1, 4, 1, 2, 0, 6, 2048), 1, 0.22, True),
    ("sm_2x2_16qam", (1, 2, 1, 4, 2, 4, 2048), 2, 0.2, True),
    ("txdiv2_16qam_mode_l", (1, 2, 1, 3, 1, 4, 2048), 2, 0.12, False),     # mode l holds the last DRS estimate: nothing follows the drift
])
def test_frequency_step_decodes_only_with_tracking(product, name, cfg, N_RX, step_turns, lr):
    ps = spec.get_packet_sizes(*cfg)
    n, t0 = 8, 5
    case = dict(cfg=cfg, N_RX=N_RX, snr_db=33.0, t0=t0)
    phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=NET, plcf_type=1, chestim_mode_lr=lr)
    rng = np.random.default_rng(sum(map(ord, name)))
    win, tbs, plcfs = make_windows(ps, case, n, phy.rx_stride(), rng, 1)
    N, cp, stf = dims_of(ps)
    N_step = 5 if ps["N_eff_TX"] <= 2 else 10
    w = frequency_step(win, t0 + stf, -2 * np.pi * step_turns / (N_step * (N + cp)))
    # no impairment: the stage changes no outcome
    _, ok0, tb0 = _decode(phy, win, t0, n)
    phy.rx_set_tracking(True, True)
    _, ok1, tb1 = _decode(phy, win, t0, n)
    assert ok0.all() and ok1.all() and np.array_equal(tb0, tbs) and np.array_equal(tb1, tbs), (name, ok0, ok1)
    # the frequency step: lost without the stage, received with it
    _, ok_on, tb_on = _decode(phy, w, t0, n)
    phy.rx_set_tracking(False, False)
    _, ok_off, _ = _decode(phy, w, t0, n)
    assert ok_on.all() and np.array_equal(tb_on, tbs), (name, ok_on)
    assert ok_off.sum() <= 1, (name, ok_off)
    phy.close()


def test_clock_offset_decodes_only_with_tracking(product):
    # zeta < 0: the symbols arrive later and later, the FFT windows slide into the cyclic prefix (4 of its 32 samples)
    cfg, N_RX, zeta = (1, 4, 1, 4, 0, 6, 2048), 1, -350e-6
    ps = spec.get_packet_sizes(*cfg)
    n, t0 = 6, 8
    phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=NET, plcf_type=1)
    rng = np.random.default_rng(77)
    win, tbs, plcfs = make_windows(ps, dict(cfg=cfg, N_RX=N_RX, snr_db=33.0, t0=t0), n, phy.rx_stride(), rng, 1)
    N, cp, stf = dims_of(ps)
    w = clock_offset(win, zeta, t0 + stf // 2)
    phy.rx_set_tracking(True, True)
    _, ok_on, tb_on = _decode(phy, w, t0, n)
    est = phy.rx_tracking_read(n)
    phy.rx_set_tracking(False, False)
    _, ok_off, _ = _decode(phy, w, t0, n)
    assert ok_on.all() and np.array_equal(tb_on, tbs), (ok_on, est)
    assert ok_off.sum() <= 1, ok_off
    phy.close()


def test_one_drs_occurrence_leaves_the_packet_alone(product):
    """a packet of one subslot has a single DRS symbol: nothing to compare, the stage reports (0, 0) and changes nothing"""
    cfg, N_RX = (1, 2, 0, 1, 0, 2, 2048), 1
    ps = spec.get_packet_sizes(*cfg)
    n, t0 = 4, 3
    phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=NET, plcf_type=1)
    rng = np.random.default_rng(9)
    win, tbs, plcfs = make_windows(ps, dict(cfg=cfg, N_RX=N_RX, snr_db=25.0, t0=t0), n, phy.rx_stride(), rng, 1)
    _, ok0, tb0 = _decode(phy, win, t0, n)
    llr0 = phy.tap(product.TAP_LLR_PDC, n).copy()
    phy.rx_set_tracking(True, True)
    _, ok1, tb1 = _decode(phy, win, t0, n)
    assert np.array_equal(phy.rx_tracking_read(n), np.zeros((n, 2), dtype=np.float32))
    assert ok0.all() and ok1.all() and np.array_equal(tb1, tbs) and np.array_equal(phy.tap(product.TAP_LLR_PDC, n), llr0)
    phy.close()
