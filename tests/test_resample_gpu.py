"""GPU parity of the rational resampler (k_resample) inside the TX chain, the synchronisation and the synchronised
receiver against the oracle's streaming restatement of resampler_t (oracle/resample.c).
Tolerances: resampled TX IQ EVM <= -80 dB (fp32 dot products in a different order); sync integers equal; OFDM symbols
after resampling 2e-4 relative; decoded transport blocks equal."""
import numpy as np
import pytest

from oracle import fec, resample as ors, rx as orx, spec, sync as osync, tx as otx
from conftest import CONFIGS

pytestmark = pytest.mark.gpu
NET = 123456789

CASES = [
    # name, psdef, N_RX, os, (L, M)
    ("C1_loopback_1.92MSps", CONFIGS["C1_loopback"], 1, 1, (10, 9)),
    ("b4_txdiv2_os2", CONFIGS["txdiv2_b4_mcs3"], 2, 2, (10, 9)),
    ("b2_generic_40_27", CONFIGS["small_b2_mcs4"], 1, 1, (40, 27)),
]


def evm_db(a, b):
    return 10 * np.log10(np.sum(np.abs(a - b) ** 2) / np.sum(np.abs(b) ** 2) + 1e-30)


@pytest.mark.parametrize("name,cfg,N_RX,os,LM", CASES)
def test_tx_resampled_matches_oracle(product, name, cfg, N_RX, os, LM):
    ps = spec.get_packet_sizes(*cfg)
    n = 3
    phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=NET, oversampling=os, resample=LM)
    rng = np.random.default_rng(2)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    inc = 1e-3
    iq, ns = phy.tx(plcf, tb, iq_phase_rad=0.3, iq_phase_increment=inc, GI_percentage=50)
    for i in range(n):
        ref = otx.tx_packet(ps, plcf[i], tb[i], network_id=NET, os=os, iq_phase_rad=0.3, iq_phase_increment=inc,
                            GI_percentage=50, resample=LM)
        assert ns[i] == ref.shape[1], (ns[i], ref.shape)
        got = iq[i, :, :ns[i]]
        assert evm_db(got, ref) <= -80.0, (name, i, evm_db(got, ref))
        assert not iq[i, :, ns[i]:].any()
    phy.close()


@pytest.mark.parametrize("name,cfg,N_RX,os,LM", CASES)
def test_sync_and_rx_resampled(product, name, cfg, N_RX, os, LM):
    L, M = LM
    ps = spec.get_packet_sizes(*cfg)
    n = 6
    phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=NET, oversampling=os, resample=LM)
    rng = np.random.default_rng(8)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    phy.tx(plcf, tb, fetch=False)
    stf_rs = ps["N_samples_STF"] * os * L // M
    t0 = rng.integers(stf_rs // 2, 2 * stf_rs - 8, n)
    win = phy.channel(n, snr_dB=30.0, t0=t0, mode=0, seed=5, fetch=True)
    res = phy.sync(n=n)
    for i in range(n):
        o = osync.search(win[i], cfg[0], cfg[1], os=os, N_eff_TX_max=ps["N_eff_TX"], resample=LM)
        assert int(res[i].detected) == int(o.detected) == 1, (name, i)
        assert int(res[i].detection_time_local) == o.detection_time_local
        assert int(res[i].coarse_peak_time_local) == o.coarse_peak_time_local
        assert int(res[i].coarse_peak_time_hw) == o.coarse_peak_time_hw
        assert int(res[i].fine_peak_time_local) == o.fine_peak_time_local, (name, i, res[i].fine_peak_time_local,
                                                                            o.fine_peak_time_local, t0[i])
        assert int(res[i].N_eff_TX) == o.N_eff_TX == ps["N_eff_TX"]
        assert abs(float(res[i].cfo_fractional_rad) - o.cfo_fractional_rad) <= 2e-6
        assert abs(int(res[i].fine_peak_time_local) - int(t0[i])) <= 1, (name, i, res[i].fine_peak_time_local, t0[i])
    # synchronised receiver from the device-side hand-over: resampled from the fine peak, OFDM symbols vs oracle
    pcc = phy.rx_pcc(None, n=n, sync="device")
    Y = phy.tap(product.TAP_Y, n)
    ocfg = orx.RxConfig(ps, N_RX=N_RX, os=os)
    n_no_gi = ps["N_samples_packet_no_GI"] * os
    for i in range(n):
        f0 = int(res[i].fine_peak_time_local)
        x = np.stack([ors.stream(win[i, a, f0:], M, L, os, flush=False)[:n_no_gi] for a in range(N_RX)])
        r = ocfg.rx_packet(x, t0=0, cfo_fractional_rad=float(res[i].cfo_fractional_rad),
                           cfo_integer_rad=float(res[i].cfo_integer_rad), rms=np.array(res[i].rms_array, np.float32))
        ref = r["Y"].reshape(-1)
        got = Y[i].reshape(-1)
        assert np.abs(got - ref).max() <= 2e-4 * np.abs(ref).max(), (name, i, float(np.abs(got - ref).max()))
    assert pcc.plcf_ok[:, 0].all() if phy.plcf_type == 1 else True
    ok, tb_out, snr_db, _ = phy.rx_pdc(n)
    assert ok.all() and (tb_out == tb).all(), (name, ok)
    phy.close()


def test_c4_geometry_with_resampling_and_harq(product):
    """BASELINE.json configs[3]: mu=8 beta=16, eight transmit streams (TM 10), 256-QAM R=5/6 (MCS 9, C = 59 code blocks of
    6080 bits), polyphase resampling 10/9 to the device rate, HARQ retransmission combining.  First transmission (rv 0)
    at an SNR where 256-QAM R=5/6 does not decode, retransmission (rv 2) combined in the same soft buffer slots decodes;
    LLRs and outcomes of both rounds equal the oracle's on the same IQ."""
    cfg, N_RX, LM = (8, 16, 1, 1, 10, 9, 6144), 2, (10, 9)
    L, M = LM
    ps = spec.get_packet_sizes(*cfg)
    assert ps["C"] == 59 and ps["N_eff_TX"] == 8 and ps["G"] == 429296
    n = 2
    phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=NET, resample=LM)
    rng = np.random.default_rng(21)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    gold = fec.pdc_gold(NET, 1, ps["G"])
    ocfg = orx.RxConfig(ps, N_RX=N_RX)
    n_no_gi = ps["N_samples_packet_no_GI"]
    harq = [fec.HarqRx(ps) for _ in range(n)]
    slots = np.arange(n, dtype=np.uint32)
    outcomes = []
    for rnd, (rv, snr) in enumerate(((0, 19.0), (2, 19.0))):
        phy.tx(plcf, tb, rv=rv, fetch=False)
        stf_rs = ps["N_samples_STF"] * L // M
        t0 = rng.integers(stf_rs // 2, 2 * stf_rs - 8, n)
        win = phy.channel(n, snr_dB=snr, t0=t0, mode=0, seed=77 + rnd, fetch=True)
        res = phy.sync(n=n)
        assert (res.detected == 1).all() and (res.N_eff_TX == 8).all()
        assert (np.abs(res.fine_peak_time_local.astype(np.int64) - t0) <= 1).all()
        pcc = phy.rx_pcc(None, n=n, sync="device")
        assert pcc.plcf_ok[:, 0].all()
        ok, tb_out, _, nit = phy.rx_pdc(n, harq_slot=slots, rv=np.full(n, rv, np.uint8),
                                        harq_reset=np.full(n, 1 if rnd == 0 else 0, np.uint8))
        llr = phy.tap(product.TAP_LLR_PDC, n)
        for i in range(n):
            f0 = int(res[i].fine_peak_time_local)
            x = np.stack([ors.stream(win[i, a, f0:], M, L, 1, flush=False)[:n_no_gi] for a in range(N_RX)])
            r = ocfg.rx_packet(x, t0=0, cfo_fractional_rad=float(res[i].cfo_fractional_rad),
                               cfo_integer_rad=float(res[i].cfo_integer_rad), rms=np.array(res[i].rms_array, np.float32),
                               want_Y=False)
            d = np.abs(llr[i].astype(np.int32) - r["llr_pdc"].astype(np.int32))
            assert d.max() <= 2 and np.count_nonzero(d) <= 0.02 * d.size, (rnd, i, int(d.max()), np.count_nonzero(d))
            # decoder parity on the product's own LLRs (the <= 2 LSB differences must not decide the comparison)
            o_ok, o_tb, o_nit = fec.pdc_decode(llr[i], ps, rv, gold, harq=harq[i])
            assert bool(ok[i]) == o_ok and np.array_equal(nit[i], o_nit.astype(np.uint8)), (rnd, i, ok[i], o_ok)
            if o_ok:
                assert np.array_equal(tb_out[i], o_tb) and np.array_equal(tb_out[i], tb[i])
        outcomes.append(ok.copy())
    assert not outcomes[0].any() and outcomes[1].all(), outcomes
    phy.close()
