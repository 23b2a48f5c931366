"""GPU parity of the synchronised receiver (dnrp_rx_pcc_batch / dnrp_rx_pdc_batch) against the oracle's
sequential restatement of rx_synced_t (oracle/rx.c), on identical IQ windows.

Tolerances (floating point path, fp32 on both sides but different summation orders / FFT algorithms):
  OFDM symbols Y, DRS/STF zero-forced pilots   relative error <= 2e-4 of the RMS
  SNR estimates                                 <= 0.05 dB (above 60 dB, where the noise term is fp32 rounding: both > 60)
  int16 LLRs                                    |diff| <= 2 LSB and <= 2 % of the LLRs differ at all
  Wiener class sequence, PLCF/TB CRC flags, decoded bytes   identical
"""
import numpy as np
import pytest

from oracle import fec, rx, spec, tx
from conftest import CONFIGS

pytestmark = pytest.mark.gpu
NET = 123456789

RX_CASES = {
    "C1_loopback_ideal": dict(cfg=CONFIGS["C1_loopback"], snr_db=None),
    "C2_b16_mcs7_30dB_cfo": dict(cfg=CONFIGS["C2_b16_siso_mcs7"], snr_db=30.0, cfo=2e-4, cfo_err=3e-5, t0=17),
    "C3_b16_txdiv4_25dB": dict(cfg=CONFIGS["C3_b16_txdiv4_mcs7"], snr_db=25.0, t0=5),
    "txdiv2_b4_2rx_12dB": dict(cfg=CONFIGS["txdiv2_b4_mcs3"], snr_db=12.0, N_RX=2, cfo=1e-3, cfo_err=-1e-4, t0=40),
    "b2_mcs4_mode_l": dict(cfg=CONFIGS["small_b2_mcs4"], snr_db=22.0, chestim_lr=False, t0=3),
    "b2_long_2rx_stride1_3dB": dict(cfg=(1, 2, 1, 4, 0, 1, 2048), snr_db=3.0, N_RX=2, stride=1, t0=9),
    "b1_long_stride3_17dB": dict(cfg=(1, 1, 1, 3, 0, 3, 2048), snr_db=17.0, stride=3),
    "txdiv8_u8_b2_20dB": dict(cfg=(8, 2, 1, 2, 10, 1, 2048), snr_db=20.0, N_RX=2),
    "b12_256qam_38dB": dict(cfg=(4, 12, 0, 3, 0, 9, 6144), snr_db=38.0, t0=11),
    "bpsk_b1_os2_6dB": dict(cfg=(1, 1, 1, 2, 0, 0, 2048), snr_db=6.0, os=2, cfo=5e-4, t0=2),
    "b8_os4_20dB": dict(cfg=(1, 8, 0, 1, 0, 5, 6144), snr_db=20.0, os=4, t0=30),
    # FFT sizes above 4096: k_rx_stf<BIG>, k_rx_fft<BIG> (four-step in place)
    "b12_os8_fft6144_18dB": dict(cfg=(1, 12, 0, 2, 0, 3, 2048), snr_db=18.0, os=8, t0=21, cfo=2e-5),
    "b16_os8_fft8192_txdiv2_2rx_22dB": dict(cfg=(1, 16, 1, 1, 1, 4, 6144), snr_db=22.0, os=8, N_RX=2, t0=100),
    "b12_os16_fft12288_15dB": dict(cfg=(1, 12, 0, 2, 0, 2, 2048), snr_db=15.0, os=16, t0=7),
    "b16_os16_fft16384_25dB": dict(cfg=(1, 16, 1, 1, 0, 5, 6144), snr_db=25.0, os=16, t0=64, cfo=1e-5, cfo_err=1e-6),
}


def make_windows(ps, case, n, stride, rng, os_):
    """oracle TX -> flat complex gains per (rx, tx) + integer delay + CFO + AWGN -> [n][N_RX][stride]"""
    N_RX = case.get("N_RX", 1)
    t0 = case.get("t0", 0)
    cfo = case.get("cfo", 0.0)
    tbs = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcfs = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    win = np.zeros((n, N_RX, stride), dtype=np.complex64)
    for i in range(n):
        iq = tx.tx_packet(ps, plcfs[i], tbs[i], network_id=NET, plcf_type=1, rv=0, os=os_)
        N_TX, L = iq.shape
        g = (rng.standard_normal((N_RX, N_TX)) + 1j * rng.standard_normal((N_RX, N_TX))) / np.sqrt(2 * N_TX)
        if case.get("snr_db") is None:
            g = np.ones((N_RX, N_TX)) / N_TX
        y = (g @ iq).astype(np.complex128)
        sig_pow = np.mean(np.abs(y[:, :ps["N_samples_packet_no_GI"] * os_]) ** 2)
        full = np.zeros((N_RX, stride), dtype=np.complex128)
        full[:, t0:t0 + L] = y
        if case.get("snr_db") is not None:
            # noise power per sample scaled by the oversampling so that the in-band SNR is snr_db
            sigma2 = sig_pow / 10 ** (case["snr_db"] / 10) * os_ * 64 / 56
            full += np.sqrt(sigma2 / 2) * (rng.standard_normal(full.shape) + 1j * rng.standard_normal(full.shape))
        full *= np.exp(-1j * cfo * (np.arange(stride) - t0))[None, :]
        win[i] = full.astype(np.complex64)
    return win, tbs, plcfs


def rel_err(a, b):
    return float(np.sqrt(np.mean(np.abs(a - b) ** 2) / (np.mean(np.abs(b) ** 2) + 1e-30)))


@pytest.mark.parametrize("name", list(RX_CASES))
def test_rx_matches_oracle(product, name):
    case = dict(RX_CASES[name])
    cfg = case["cfg"]
    os_ = case.get("os", 1)
    N_RX = case.get("N_RX", 1)
    lr, stride_t = case.get("chestim_lr", True), case.get("stride", 2)
    ps = spec.get_packet_sizes(*cfg)
    assert ps is not None
    n = 3
    phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=NET, plcf_type=1, oversampling=os_,
                      chestim_mode_lr=lr, chestim_stride=stride_t)
    rng = np.random.default_rng(sum(map(ord, name)))
    win, tbs, plcfs = make_windows(ps, case, n, phy.rx_stride(), rng, os_)
    t0 = case.get("t0", 0)
    cfo_given = case.get("cfo", 0.0) + case.get("cfo_err", 0.0)

    pcc = phy.rx_pcc(win, t0=t0, cfo_fractional_rad=cfo_given)
    Y = phy.tap(product.TAP_Y, n).reshape(n, N_RX, ps["N_DF_symb"], 56 * ps["b"] + 1)
    cls = phy.tap(product.TAP_WIENER_CLASS, n)
    llr_pcc = phy.tap(product.TAP_LLR_PCC, n)
    ok, tb_out, snr, nit = phy.rx_pdc(n)
    llr_pdc = phy.tap(product.TAP_LLR_PDC, n)

    orc = rx.RxConfig(ps, N_RX=N_RX, os=os_, chestim_lr=lr, stride=stride_t)
    gold = fec.pdc_gold(NET, 1, ps["G"])
    for i in range(n):
        r = orc.rx_packet(win[i], t0=t0, cfo_fractional_rad=cfo_given)
        rep = r["report"]
        assert rel_err(Y[i], r["Y"]) < 2e-4, (name, i, "Y", rel_err(Y[i], r["Y"]))
        assert abs(pcc[i].cfo_residual_rad - rep.cfo_left) < 1e-6 + 1e-3 * abs(rep.cfo_left)
        assert abs(pcc[i].sto_fractional - rep.sto_fractional) < 2e-3
        if rep.snr_db_pdc < 60:
            assert abs(pcc[i].snr_dB - rep.snr_db_pcc) < 0.05, (pcc[i].snr_dB, rep.snr_db_pcc)
            assert abs(snr[i] - rep.snr_db_pdc) < 0.05, (snr[i], rep.snr_db_pdc)
        else:  # noise term is fp32 rounding only
            assert pcc[i].snr_dB > 60 and snr[i] > 60
        assert np.array_equal(cls[i, :len(r["cls"])], r["cls"]), (name, i, cls[i], r["cls"])
        for got, ref, what in ((llr_pcc[i], r["llr_pcc"], "pcc"), (llr_pdc[i], r["llr_pdc"], "pdc")):
            d = np.abs(got.astype(np.int32) - ref.astype(np.int32))
            assert d.max() <= 2, (name, i, what, int(d.max()), int(np.argmax(d)))
            # noiseless QPSK sits exactly on a truncation boundary (|llr| = 100.0): skip the count there
            if case.get("snr_db") is not None:
                assert np.mean(d > 0) < 0.02, (name, i, what, float(np.mean(d > 0)))
        # decoding outcomes: the oracle decodes its own LLRs
        o_ok, o_tb, o_nit = fec.pdc_decode(r["llr_pdc"], ps, 0, gold)
        assert bool(ok[i]) == o_ok, (name, i, ok[i], o_ok)
        if o_ok:
            assert np.array_equal(tb_out[i], o_tb) and np.array_equal(tb_out[i], tbs[i])
        p_ok, p_plcf, p_cl, p_bf, _ = fec.pcc_decode(r["llr_pcc"], 1)
        assert bool(pcc[i].plcf_ok[0]) == p_ok
        if p_ok:
            assert bytes(pcc[i].plcf[0])[:5] == bytes(p_plcf) == bytes(plcfs[i][:5])
    phy.close()


def test_rx_device_loopback(product):
    """dnrp_tx_batch -> dnrp_loopback_tx_to_rx -> dnrp_rx_pcc_batch/_pdc_batch without leaving the device"""
    cfg = CONFIGS["C2_b16_siso_mcs7"]
    ps = spec.get_packet_sizes(*cfg)
    n = 8
    phy = product.Phy(cfg, max_batch=n, network_id=NET, plcf_type=1)
    rng = np.random.default_rng(5)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    phy.tx(plcf, tb, fetch=False)
    phy.loopback_tx_to_rx(n, t0=13)
    pcc = phy.rx_pcc(None, t0=13, n=n)
    ok, tb_out, snr, nit = phy.rx_pdc(n)
    assert ok.all() and np.array_equal(tb_out, tb)
    assert all(p.plcf_ok[0] and bytes(p.plcf[0])[:5] == bytes(plcf[i][:5]) for i, p in enumerate(pcc))
    assert (snr > 40).all()
    phy.close()


@pytest.mark.parametrize("cfg,N_RX", [(CONFIGS["txdiv2_b4_mcs3"], 2), (CONFIGS["C3_b16_txdiv4_mcs7"], 4),
                                      (CONFIGS["C3_b16_txdiv4_mcs7"], 2), (CONFIGS["C1_loopback"], 4),
                                      ((8, 2, 1, 2, 10, 1, 2048), 2)])
def test_mimo_report(product, cfg, N_RX):
    """mimo_report_t of the PDC report against the oracle's codebook search (estimator_mimo.cpp) run on the same
    zero-forced DRS cells (the device's DRS tap, itself covered by test_rx_matches_oracle)"""
    ps = spec.get_packet_sizes(*cfg)
    n = 16
    phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=NET)
    rng = np.random.default_rng(21)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    phy.tx(plcf, tb, fetch=False)
    phy.channel(n, snr_dB=25.0, t0=7, mode=1, seed=31)
    phy.rx_pcc(None, t0=7, n=n)
    phy.rx_pdc(n)
    rep = phy.last_pdc_report
    N_TS = ps["N_eff_TX"]
    hzf = phy.tap(product.TAP_DRS_ZF, n).reshape(n, N_RX, N_TS, -1, ps["N_b_OCC"] // 4)
    assert (rep.mimo_N_RX == N_RX).all() and (rep.mimo_N_TS_other == N_TS).all()
    seen = set()
    for i in range(n):
        idx, rec = rx.mimo_report(hzf[i, :, :, -1, :])
        assert (int(rep.tm_3_7_beamforming_idx[i]), int(rep.tm_3_7_beamforming_reciprocal_idx[i])) == (idx, rec), i
        seen.add((idx, rec))
    if N_TS in (2, 4) and N_RX in (2, 4):
        assert len(seen) > 1          # flat fading draws: the recommendation depends on the channel
    if N_TS == 1:
        assert (rep.tm_3_7_beamforming_idx == 0).all()
    phy.close()


@pytest.mark.parametrize("name", ["C3_b16_txdiv4_25dB", "b2_mcs4_mode_l", "txdiv2_b4_2rx_12dB"])
def test_chestim_tap_matches_oracle(product, name):
    """DNRP_TAP_CHESTIM (what the reference exports per packet as ch_<rx>_<ts>, rx_synced.cpp:444-480): the interpolated
    channel after the last data field symbol, against the oracle's snapshot of chestim at that symbol"""
    case = dict(RX_CASES[name])
    cfg, N_RX = case["cfg"], case.get("N_RX", 1)
    lr, stride_t = case.get("chestim_lr", True), case.get("stride", 2)
    ps = spec.get_packet_sizes(*cfg)
    n = 2
    phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=NET, plcf_type=1, chestim_mode_lr=lr, chestim_stride=stride_t)
    rng = np.random.default_rng(4)
    win, tbs, plcfs = make_windows(ps, case, n, phy.rx_stride(), rng, 1)
    t0 = case.get("t0", 0)
    cfo = case.get("cfo", 0.0) + case.get("cfo_err", 0.0)
    phy.rx_pcc(win, t0=t0, cfo_fractional_rad=cfo)
    phy.rx_pdc(n)
    Nk = 56 * ps["b"] + 1
    got = phy.tap(product.TAP_CHESTIM, n).reshape(n, N_RX, ps["N_eff_TX"], Nk)
    orc = rx.RxConfig(ps, N_RX=N_RX, chestim_lr=lr, stride=stride_t)
    for i in range(n):
        r = orc.rx_packet(win[i], t0=t0, cfo_fractional_rad=cfo, want_H=True)
        ref = r["H"][:, :, ps["N_DF_symb"] - 1, :]
        assert rel_err(got[i], ref) < 2e-4, (name, i, rel_err(got[i], ref))
    phy.close()
