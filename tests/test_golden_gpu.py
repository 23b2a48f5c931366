"""CUDA product against the committed golden vectors (tests/golden/golden_v1.npz), through the C ABI, without the oracle
in the loop: coded bits, decoder outcomes and iteration counts, synchronisation integers exactly; IQ at 1e-4 relative
(-80 dB EVM); LLRs within 2 LSB (< 2 % differing)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
NET = 123456789
GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_v1.npz"))


def _bits(b, n):
    return np.unpackbits(np.asarray(b, dtype=np.uint8))[:n]


@pytest.mark.parametrize("name,N_RX", [("c1", 1), ("txdiv2_b4", 2)])
def test_chain_against_golden(product, name, N_RX):
    g = {k[len(name) + 1:]: GOLD[k] for k in GOLD.files if k.startswith(name + "_")}
    cfg = tuple(int(v) for v in g["cfg"])
    phy = product.Phy(cfg, max_batch=4, N_RX=N_RX, network_id=NET, plcf_type=1)
    ps = phy.ps
    tb, plcf = g["tb"][None], g["plcf"][None]
    # FEC
    for rv in range(4):
        pcc, pdc = phy.fec_encode(plcf, tb, rv=np.array([rv], dtype=np.uint8))
        assert np.array_equal(_bits(pdc[0], ps["G"]), _bits(g["pdc_bits_rv%d" % rv], ps["G"])), rv
        assert np.array_equal(_bits(pcc[0], 196), _bits(g["pcc_bits"], 196))
    # TX
    iq, ns = phy.tx(plcf, tb)
    ref = g["tx_iq"]
    assert int(ns[0]) == ref.shape[1]
    err = np.sqrt(np.mean(np.abs(iq[0, :, :ref.shape[1]] - ref) ** 2) / np.mean(np.abs(ref) ** 2))
    assert err < 1e-4, err
    # synchronisation of the golden window
    win = g["win"][None]
    assert win.shape[-1] == phy.rx_stride()
    s = phy.sync(iq=win)[0]
    si, sf = g["sync_int"], g["sync_flt"]
    got = [s.detected, s.detection_ant_idx, s.detection_time_local, s.coarse_peak_time_local, s.N_eff_TX,
           s.fine_peak_time_local, s.n_detections]
    assert [int(v) for v in got] == [int(v) for v in si], (got, si)
    assert int(s.fine_peak_time_local) == int(g["t0"])
    assert abs(float(s.detection_rms) - sf[0]) <= 1e-4 * sf[0] and abs(float(s.detection_metric) - sf[1]) <= 1e-3 * sf[1]
    assert abs(float(s.cfo_fractional_rad) - sf[2]) <= 2e-6
    # receiver from the golden synchronisation values
    pcc_rep = phy.rx_pcc(win, t0=int(si[5]), cfo_fractional_rad=float(sf[2]), rms=sf[3 + N_RX:3 + 2 * N_RX][None])
    ok, tb_out, snr, nit = phy.rx_pdc(1)
    for got, ref, what in ((phy.tap(product.TAP_LLR_PCC, 1)[0], g["llr_pcc"], "pcc"),
                           (phy.tap(product.TAP_LLR_PDC, 1)[0], g["llr_pdc"], "pdc")):
        d = np.abs(got[:ref.size].astype(np.int32) - ref.astype(np.int32))
        assert d.max() <= 2 and np.mean(d > 0) < 0.02, (what, int(d.max()), float(np.mean(d > 0)))
    assert abs(float(snr[0]) - float(g["snr_db"])) < 0.05
    assert bool(pcc_rep[0].plcf_ok[0]) and bytes(pcc_rep[0].plcf[0])[:5] == bytes(g["plcf"][:5])
    assert bool(ok[0]) == bool(g["dec_ok"])
    if ok[0]:
        assert np.array_equal(tb_out[0], g["dec_tb"]) and np.array_equal(tb_out[0], g["tb"])
    phy.close()


def test_decoder_against_golden(product):
    cfg = tuple(int(v) for v in GOLD["dec_cfg"])
    llr = GOLD["dec_llr"]
    phy = product.Phy(cfg, max_batch=llr.shape[0], network_id=NET, plcf_type=1)
    ok, out, nit = phy.fec_decode_tb(llr)
    assert np.array_equal(ok.astype(bool), GOLD["dec_ok"])
    assert np.array_equal(nit, GOLD["dec_nit"])
    assert np.array_equal(out, GOLD["dec_tb"])
    assert GOLD["dec_ok"].any() and np.array_equal(out[GOLD["dec_ok"]], GOLD["dec_tb_in"][GOLD["dec_ok"]])
    phy.close()
