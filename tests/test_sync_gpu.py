"""GPU parity of the packet synchronisation (dnrp_sync_batch) against oracle/sync.c, which restates the
reference's streaming algorithm (moving sums, per-sample peak search, cross-correlation).
Integer results (detection step, antenna, coarse peak, fine peak, N_eff_TX) must be equal; float results agree to
fp32 summation-order tolerance (the device evaluates every position from prefix sums instead of pop/push sums):
metrics 1e-3 relative, CFO 2e-6 rad/sample absolute, rms 1e-4 relative."""
import numpy as np
import pytest

from oracle import spec, sync as osync
from conftest import CONFIGS

pytestmark = pytest.mark.gpu
NET = 123456789


def _compare(r, o, N_RX, where):
    assert int(r.detected) == int(o.detected), where
    if not o.detected:
        return
    assert int(r.detection_time_local) == o.detection_time_local, (where, r.detection_time_local, o.detection_time_local)
    assert int(r.detection_ant_idx) == o.detection_ant_idx, where
    assert abs(float(r.detection_metric) - o.detection_metric) <= 1e-3 * o.detection_metric, where
    assert abs(float(r.detection_rms) - o.detection_rms) <= 1e-4 * o.detection_rms, where
    assert int(r.coarse_peak_time_local) == o.coarse_peak_time_local, (where, r.coarse_peak_time_local,
                                                                        o.coarse_peak_time_local)
    for a in range(N_RX):
        assert abs(float(r.coarse_peak_array[a]) - o.coarse_peak_array[a]) <= 1e-3 * max(o.coarse_peak_array[a], 1e-3), where
        assert abs(float(r.rms_array[a]) - o.rms_array[a]) <= 1e-4 * max(o.rms_array[a], 1e-6), where
    assert abs(float(r.cfo_fractional_rad) - o.cfo_fractional_rad) <= 2e-6, (where, r.cfo_fractional_rad,
                                                                            o.cfo_fractional_rad)
    assert int(r.N_eff_TX) == o.N_eff_TX, (where, r.N_eff_TX, o.N_eff_TX)
    assert int(r.fine_peak_time_local) == o.fine_peak_time_local, (where, r.fine_peak_time_local,
                                                                    o.fine_peak_time_local)
    assert int(r.n_detections) == o.n_detections, where


EXTRA = {"u8_b2_txdiv8": (8, 2, 1, 2, 10, 1, 2048), "u2_b8_mcs4": (2, 8, 1, 2, 0, 4, 2048)}
CASES = [
    # name, N_RX, oversampling, channel mode, snr_dB, cfo (rad/sample at the oversampled rate)
    ("C1_loopback", 1, 1, 0, 15.0, 0.0),
    ("C1_loopback", 2, 2, 1, 12.0, 4e-3),
    ("C2_b16_siso_mcs7", 1, 1, 0, 25.0, 1e-4),
    ("txdiv2_b4_mcs3", 2, 1, 1, 15.0, -5e-4),
    ("C3_b16_txdiv4_mcs7", 2, 1, 1, 20.0, 2e-4),
    ("u8_b2_txdiv8", 4, 1, 1, 20.0, 1e-3),
    ("u2_b8_mcs4", 1, 2, 0, 6.0, 0.0),
]


@pytest.mark.parametrize("name,N_RX,os,mode,snr,cfo", CASES)
def test_sync_matches_oracle(product, name, N_RX, os, mode, snr, cfo):
    cfg = {**CONFIGS, **EXTRA}[name]
    ps = spec.get_packet_sizes(*cfg)
    n = 12
    phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=NET, oversampling=os)
    rng = np.random.default_rng(11)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    phy.tx(plcf, tb, iq_phase_increment=-cfo, fetch=False)
    stf_len = ps["N_samples_STF"] * os
    pat = 16 * cfg[1] * os
    t0 = rng.integers(2 * pat, 2 * stf_len - 4, n)
    t0[0] = 2 * stf_len - 4
    t0[1] = pat  # early packet: detection is held back by the ignore-before gate
    win = phy.channel(n, snr_dB=snr, t0=t0, mode=mode, seed=77, fetch=True)
    res = phy.sync(n=n)
    n_hit = 0
    for i in range(n):
        o = osync.search(win[i], cfg[0], cfg[1], os=os, N_eff_TX_max=ps["N_eff_TX"])
        _compare(res[i], o, N_RX, (name, i, int(t0[i])))
        n_hit += int(o.detected and o.fine_peak_time_local == t0[i] and o.N_eff_TX == ps["N_eff_TX"])
    # the algorithm itself has to find most packets exactly at these SNRs (flat fading draws may lose some)
    assert n_hit >= (n * 2) // 3, (name, n_hit)
    # host windows through the same entry point give the same records
    res2 = phy.sync(iq=win)
    assert res2.tobytes() == res.tobytes()
    phy.close()


def test_sync_noise_only_and_handover(product):
    """noise only windows are not detected; detected packets run through PCC/PDC from the device-side hand-over"""
    cfg = CONFIGS["C1_loopback"]
    ps = spec.get_packet_sizes(*cfg)
    n = 8
    phy = product.Phy(cfg, max_batch=n, N_RX=1, network_id=NET)
    rng = np.random.default_rng(5)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = np.zeros((n, 10), dtype=np.uint8)
    plcf[:, 0] = 0x20  # PLCF type 1 header format 1 (any content decodes; the CRC decides)
    phy.tx(plcf, tb, DAC_scale=0.0, fetch=False)
    win = phy.channel(n, snr_dB=20.0, t0=100, mode=0, seed=3, fetch=True)
    res = phy.sync(n=n)
    assert not res.detected.any()
    for i in range(2):
        assert not osync.search(win[i], cfg[0], cfg[1]).detected
    phy.tx(plcf, tb, fetch=False)
    t0 = rng.integers(40, 2 * ps["N_samples_STF"] - 4, n)
    phy.channel(n, snr_dB=25.0, t0=t0, mode=0, seed=4)
    res = phy.sync(n=n)
    assert res.detected.all() and (res.fine_peak_time_local == t0).all(), (res.fine_peak_time_local, t0)
    pcc = phy.rx_pcc(None, n=n, sync="device")
    assert pcc.plcf_ok[:, 0].all()
    ok, tb_out, snr_db, _ = phy.rx_pdc(n)
    assert ok.all() and (tb_out == tb).all()
    phy.close()


@pytest.mark.parametrize("resample,N_RX", [((1, 1), 1), ((10, 9), 2)])
def test_rx_batch_pipelined_equals_three_calls(product, resample, N_RX):
    """dnrp_rx_batch (chunked, copies overlapped with kernels) returns exactly what sync -> rx_pcc -> rx_pdc return"""
    cfg = CONFIGS["C1_loopback"]
    ps = spec.get_packet_sizes(*cfg)
    n = 300                                   # four chunks, the last one shorter
    phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=NET, resample=resample)
    rng = np.random.default_rng(9)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    phy.tx(plcf, tb, fetch=False)
    L, M = resample
    stf = ps["N_samples_STF"] * L // M
    t0 = rng.integers(stf // 2, 2 * stf - 8, n)
    win = phy.channel(n, snr_dB=14.0, t0=t0, mode=0, seed=5, fetch=True)
    s1 = phy.sync(iq=win)
    p1 = phy.rx_pcc(None, n=n, sync="device")
    ok1, tb1, snr1, _ = phy.rx_pdc(n)
    rep1 = phy.last_pdc_report.copy()
    # scramble the device state in between so that stale results cannot pass
    phy.channel(n, snr_dB=-20.0, t0=0, mode=0, seed=6)
    phy.sync(n=n)
    s2, p2, ok2, tb2, rep2 = phy.rx(win)
    assert s2.tobytes() == s1.tobytes()
    assert p2.tobytes() == p1.tobytes()
    assert np.array_equal(ok1, ok2) and np.array_equal(tb1, tb2) and rep2.tobytes() == rep1.tobytes()
    assert ok2.mean() > 0.9 and (tb2[ok2] == tb[ok2]).all()
    phy.close()


def test_rx_batch_sc16_equals_cf32_on_dequantised_windows(product):
    """dnrp_rx_batch_sc16 (int16 I/Q in, converted on the device) returns bit for bit what dnrp_rx_batch returns for
    the float windows int16 * scale -- the conversion is exact, everything after it is the same code path"""
    cfg = CONFIGS["C1_loopback"]
    ps = spec.get_packet_sizes(*cfg)
    n = 300
    phy = product.Phy(cfg, max_batch=n, N_RX=2, network_id=NET)
    rng = np.random.default_rng(19)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    phy.tx(plcf, tb, fetch=False)
    stf = ps["N_samples_STF"]
    t0 = rng.integers(stf // 2, 2 * stf - 8, n)
    win = phy.channel(n, snr_dB=14.0, t0=t0, mode=0, seed=5, fetch=True)
    gain = np.float32(0.1 / np.sqrt(np.mean(np.abs(win) ** 2)) * 32768.0)
    w16 = np.clip(np.rint(win.view(np.float32).reshape(n, 2, -1, 2) * gain), -32768, 32767).astype(np.int16)
    scale = np.float32(1.0) / gain
    deq = np.ascontiguousarray((w16.astype(np.float32) * scale).reshape(n, 2, -1)).view(np.complex64)
    s1, p1, ok1, tb1, rep1 = phy.rx(deq)
    rep1 = rep1.copy()
    phy.channel(n, snr_dB=-20.0, t0=0, mode=0, seed=6)   # scramble the device state in between
    phy.sync(n=n)
    s2, p2, ok2, tb2, rep2 = phy.rx(w16, sc16_scale=float(scale))
    assert s2.tobytes() == s1.tobytes() and p2.tobytes() == p1.tobytes()
    assert np.array_equal(ok1, ok2) and np.array_equal(tb1, tb2) and rep2.tobytes() == rep1.tobytes()
    assert ok2.mean() > 0.9 and (tb2[ok2] == tb[ok2]).all()
    phy.close()


def test_rx_batch_sc8_equals_cf32_on_dequantised_windows(product):
    """dnrp_rx_batch_sc8 (UHD's 8-bit wire format, an option beyond the reference's sc16) returns bit for bit what
    dnrp_rx_batch returns for the float windows int8 * scale; at 14 dB the 8-bit quantisation (12 dB back-off) still leaves
    the packets decodable"""
    cfg = CONFIGS["C1_loopback"]
    ps = spec.get_packet_sizes(*cfg)
    n = 300
    phy = product.Phy(cfg, max_batch=n, N_RX=2, network_id=NET)
    if phy.rx_stride() % 8:
        pytest.skip("window stride not a multiple of 8")
    rng = np.random.default_rng(23)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    phy.tx(plcf, tb, fetch=False)
    stf = ps["N_samples_STF"]
    t0 = rng.integers(stf // 2, 2 * stf - 8, n)
    win = phy.channel(n, snr_dB=14.0, t0=t0, mode=0, seed=5, fetch=True)
    gain = np.float32(0.25 / np.sqrt(np.mean(np.abs(win) ** 2)) * 128.0)       # rms at -12 dBFS of the int8 range
    w8 = np.clip(np.rint(win.view(np.float32).reshape(n, 2, -1, 2) * gain), -128, 127).astype(np.int8)
    scale = np.float32(1.0) / gain
    deq = np.ascontiguousarray((w8.astype(np.float32) * scale).reshape(n, 2, -1)).view(np.complex64)
    s1, p1, ok1, tb1, rep1 = phy.rx(deq)
    rep1 = rep1.copy()
    phy.channel(n, snr_dB=-20.0, t0=0, mode=0, seed=6)   # scramble the device state in between
    phy.sync(n=n)
    s2, p2, ok2, tb2, rep2 = phy.rx(w8, sc16_scale=float(scale))
    for a, b in ((s1, s2), (p1, p2), (rep1, rep2)):
        for f in a.dtype.names:
            assert np.array_equal(a[f], b[f]), f
    assert np.array_equal(ok1, ok2) and np.array_equal(tb1, tb2)
    assert ok2.mean() > 0.9 and (tb2[ok2] == tb[ok2]).all()
    phy.close()


def test_sync_refused_with_a_reason_at_very_high_oversampling(product):
    """FFT sizes above 4096 are supported by the transmitter and the synchronised receiver; the packet search keeps two
    STFs of prefix sums in shared memory and says so instead of failing at launch"""
    phy = product.Phy((1, 16, 1, 1, 0, 4, 6144), max_batch=1, oversampling=8)
    win = np.zeros((1, 1, phy.rx_stride()), dtype=np.complex64)
    with pytest.raises(product.DnrpError, match="packet synchronisation is not available"):
        phy.sync(win)
    phy.close()
