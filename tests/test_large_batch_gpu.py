"""The batch sizes and chunk schedules bench.py times (10 000 packets through dnrp_rx_batch / dnrp_rx_batch_sc16, 16 resp. 7
pipelined chunks) against the oracle, at a waterfall SNR where about half of the packets fail: for 128 sampled packets
(failing and passing) the oracle re-runs sync + receiver on the window (LLRs within 2 LSB) and re-decodes the DEVICE's
LLRs: CRC flag, per-code-block iteration counts and bytes must be identical."""
import numpy as np
import pytest

from oracle import fec, rx, spec, sync as osync
from conftest import CONFIGS

pytestmark = pytest.mark.gpu
NET = 123456789


@pytest.mark.parametrize("fmt", ["cf32", "sc16"])
def test_10k_packets_at_the_waterfall(product, fmt):
    cfg = CONFIGS["C2_b16_siso_mcs7"]
    ps = spec.get_packet_sizes(*cfg)
    n = 10000
    phy = product.Phy(cfg, max_batch=n, network_id=NET, plcf_type=1)
    rng = np.random.default_rng(77)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    stf = ps["N_samples_STF"]
    t0 = rng.integers(2 * 16 * cfg[1], 2 * stf - 4, n).astype(np.uint32)
    snr = rng.uniform(15.0, 21.0, n).astype(np.float32)   # the waterfall of MCS 7 lies inside this range
    phy.tx(plcf, tb, fetch=False)
    win = phy.channel(n, snr_dB=snr, t0=t0, mode=0, seed=4242, fetch=True)
    if fmt == "sc16":
        gain = np.float32(0.1 / np.sqrt(np.mean(np.abs(win[:64]) ** 2)) * 32768.0)
        w16 = np.empty(win.shape + (2,), dtype=np.int16)
        for i0 in range(0, n, 500):
            v = win[i0:i0 + 500].view(np.float32).reshape(-1, 1, win.shape[2], 2) * gain
            w16[i0:i0 + 500] = np.clip(np.rint(v), -32768, 32767).astype(np.int16)
        sres, pcc, ok, tb_out, pdc = phy.rx(w16, sc16_scale=float(1.0 / gain))
        win_eff = lambda i: (w16[i].astype(np.float32) * np.float32(1.0 / gain)).view(np.complex64).reshape(1, -1)
    else:
        sres, pcc, ok, tb_out, pdc = phy.rx(win)
        win_eff = lambda i: win[i]
    llr = phy.tap(product.TAP_LLR_PDC, n)
    per = 1.0 - ok.mean()
    assert 0.15 < per < 0.85, per                      # the batch straddles the waterfall
    assert np.array_equal(tb_out[ok], tb[ok])
    fails, passes = np.nonzero(~ok)[0], np.nonzero(ok)[0]
    sample = np.concatenate([rng.choice(fails, 64, replace=False), rng.choice(passes, 64, replace=False)])
    orc = rx.RxConfig(ps)
    gold = fec.pdc_gold(NET, 1, ps["G"])
    for i in sample:
        w = win_eff(i)
        s = osync.search(w, cfg[0], cfg[1], N_eff_TX_max=1)
        assert s.detected and sres[i].detected and int(s.fine_peak_time_local) == int(sres[i].fine_peak_time_local) == int(t0[i]), i
        r = orc.rx_packet(w, t0=s.fine_peak_time_local, cfo_fractional_rad=s.cfo_fractional_rad,
                          cfo_integer_rad=s.cfo_integer_rad, rms=np.array(s.rms_array[:], dtype=np.float32), want_Y=False)
        d = np.abs(llr[i].astype(np.int32) - r["llr_pdc"].astype(np.int32))
        assert d.max() <= 2 and (d > 0).mean() < 0.02, (i, d.max())
        o_ok, o_tb, o_nit = fec.pdc_decode(llr[i], ps, 0, gold)      # the device's own LLRs: outcome must be identical
        assert bool(ok[i]) == o_ok, (i, ok[i], o_ok)
        if o_ok:
            assert np.array_equal(tb_out[i], o_tb)
    phy.close()


def test_chunk_schedules_give_identical_results(product, monkeypatch):
    """the pipelined chunk schedules of dnrp_rx_batch_sc16 change when a stage runs, not what it computes: 2100 windows of
    the north-star geometry (four antennas, 0.9 GB: the ten-chunk schedule with the long taper) against the same call in
    one chunk and in six -- sync results, PCC reports, CRC flags, transport blocks and LLRs bit for bit"""
    cfg, N_RX = CONFIGS["C3_b16_txdiv4_mcs7"], 4
    ps = spec.get_packet_sizes(*cfg)
    n = 2100
    phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=NET, plcf_type=1)
    assert n * N_RX * phy.rx_stride() * 4 >= (768 << 20)
    rng = np.random.default_rng(5)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    stf = ps["N_samples_STF"]
    t0 = rng.integers(2 * 16 * cfg[1], 2 * stf - 4, n).astype(np.uint32)
    snr = rng.uniform(14.0, 24.0, n).astype(np.float32)
    phy.tx(plcf, tb, fetch=False)
    win = phy.channel(n, snr_dB=snr, t0=t0, mode=1, seed=99, fetch=True)   # flat Rayleigh per antenna pair
    gain = np.float32(0.1 / np.sqrt(np.mean(np.abs(win[:16]) ** 2)) * 32768.0)
    w16 = np.empty(win.shape + (2,), dtype=np.int16)
    for i0 in range(0, n, 100):
        v = win[i0:i0 + 100].view(np.float32).reshape(-1, N_RX, win.shape[2], 2) * gain
        w16[i0:i0 + 100] = np.clip(np.rint(v), -32768, 32767).astype(np.int16)
    del win
    out = {}
    for name, env in (("taper10", None), ("one", "1"), ("six", "6")):
        if env is None:
            monkeypatch.delenv("DNRP_RX_CHUNKS", raising=False)
        else:
            monkeypatch.setenv("DNRP_RX_CHUNKS", env)
        sres, pcc, ok, tb_out, pdc = phy.rx(w16, sc16_scale=float(1.0 / gain))
        out[name] = (sres.copy(), pcc.copy(), ok.copy(), tb_out.copy(), phy.tap(product.TAP_LLR_PDC, n).copy())
    ref = out["one"]
    assert 0.05 < 1.0 - ref[2].mean() < 0.95 and np.array_equal(ref[3][ref[2]], tb[ref[2]])
    for name in ("taper10", "six"):
        o = out[name]
        for k in (0, 1):     # field by field: the records carry padding bytes
            for f in ref[k].dtype.names:
                assert np.array_equal(o[k][f], ref[k][f]), (name, k, f)
        assert np.array_equal(o[2], ref[2]) and np.array_equal(o[3][o[2]], ref[3][ref[2]]) and np.array_equal(o[4], ref[4]), name
    phy.close()
