"""Generates tests/golden/golden_v1.npz: input/output vectors of the hot path, produced by the CPU oracle (oracle/).

The reference cannot be built or imported in this environment (C++ with srsRAN/VOLK/FFTW/Eigen dependencies, DESIGN.md
section 1) and its tests hold no vectors for this path, so these fixtures are NOT reference outputs: they freeze the
oracle's results (whose KAT anchors are in tests/test_oracle_cpu.py) so that neither the oracle nor the CUDA product can
drift silently.  Run from the repository root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import channel, fec, resample, rx, spec, sync, tx  # noqa: E402

NET = 123456789
SIZES = [(1, 1, 1, 1, 0, 2, 2048), (1, 16, 1, 1, 0, 7, 6144), (1, 16, 1, 2, 5, 7, 6144), (8, 16, 1, 1, 10, 9, 6144),
         (1, 2, 1, 2, 0, 4, 2048), (1, 4, 1, 2, 1, 3, 2048), (2, 8, 1, 2, 0, 4, 2048), (4, 12, 0, 3, 0, 9, 6144),
         (8, 2, 1, 2, 10, 1, 2048), (1, 1, 1, 4, 0, 0, 2048), (2, 4, 0, 6, 3, 5, 6144), (4, 8, 1, 3, 6, 6, 6144)]
SIZE_KEYS = ["N_PACKET_symb", "N_PDC_subc", "G", "N_TB_bits", "C", "N_samples_STF", "N_samples_packet",
             "N_samples_packet_no_GI", "N_TS", "N_TX", "N_eff_TX", "N_bps"]
CHAIN = {"c1": ((1, 1, 1, 1, 0, 2, 2048), 1), "txdiv2_b4": ((1, 4, 1, 2, 1, 3, 2048), 2)}


def build():
    out = {}
    rng = np.random.default_rng(20261104)
    # ---- packet sizes
    out["sizes_cfg"] = np.array(SIZES, dtype=np.int32)
    out["sizes_val"] = np.array([[spec.get_packet_sizes(*c)[k] for k in SIZE_KEYS] for c in SIZES], dtype=np.int64)
    # ---- resampler: design and a streamed block
    x = (rng.standard_normal(700) + 1j * rng.standard_normal(700)).astype(np.complex64)
    out["rs_x"] = x
    out["rs_y_10_9"] = resample.stream(x, 10, 9, 1)
    out["rs_y_9_10"] = resample.stream(x, 9, 10, 1)
    # ---- whole chain per geometry: FEC bits, TX IQ, channel window, sync report, receiver LLRs, decoded TB
    for name, (cfg, N_RX) in CHAIN.items():
        ps = spec.get_packet_sizes(*cfg)
        tb = rng.integers(0, 256, ps["N_TB_byte"], dtype=np.uint8)
        plcf = rng.integers(0, 256, 10, dtype=np.uint8)
        gold = fec.pdc_gold(NET, 1, ps["G"])
        out[name + "_cfg"] = np.array(cfg, dtype=np.int32)
        out[name + "_tb"], out[name + "_plcf"] = tb, plcf
        for rv in range(4):
            out[name + "_pdc_bits_rv%d" % rv] = fec.pdc_encode(tb, ps, rv, gold)
        out[name + "_pcc_bits"] = fec.pcc_encode(plcf[:5], 1, False, False)
        iq = tx.tx_packet(ps, plcf, tb, network_id=NET, plcf_type=1, rv=0)
        out[name + "_tx_iq"] = iq
        stride = ps["N_samples_packet"] + 2 * ps["N_samples_STF"]
        stride += (-stride) % 4
        t0 = 37 + 16 * cfg[1]
        win = channel.channel(iq, stride, N_RX, t0=t0, mode=0, snr_dB=18.0, net_bandwidth_norm=1.0, seed=99, pkt=0)
        out[name + "_t0"] = np.int32(t0)
        out[name + "_win"] = win
        s = sync.search(win, cfg[0], cfg[1], os=1, N_eff_TX_max=ps["N_eff_TX"])
        out[name + "_sync_int"] = np.array([s.detected, s.detection_ant_idx, s.detection_time_local,
                                            s.coarse_peak_time_local, s.N_eff_TX, s.fine_peak_time_local,
                                            s.n_detections], dtype=np.int32)
        out[name + "_sync_flt"] = np.array([s.detection_rms, s.detection_metric, s.cfo_fractional_rad]
                                           + list(s.coarse_peak_array[:N_RX]) + list(s.rms_array[:N_RX]), dtype=np.float32)
        r = rx.RxConfig(ps, N_RX=N_RX).rx_packet(win, t0=s.fine_peak_time_local,
                                                 cfo_fractional_rad=s.cfo_fractional_rad,
                                                 rms=np.array(s.rms_array[:N_RX], dtype=np.float32))
        out[name + "_llr_pcc"], out[name + "_llr_pdc"] = r["llr_pcc"], r["llr_pdc"]
        out[name + "_snr_db"] = np.float32(r["report"].snr_db_pdc)
        ok, tb_dec, nit = fec.pdc_decode(r["llr_pdc"], ps, 0, gold)
        out[name + "_dec_ok"], out[name + "_dec_tb"], out[name + "_dec_nit"] = np.bool_(ok), tb_dec, np.array(nit)
    # ---- turbo decoder near the waterfall (iteration counts are part of the contract)
    cfg = (1, 2, 1, 2, 0, 4, 2048)
    ps = spec.get_packet_sizes(*cfg)
    gold = fec.pdc_gold(NET, 1, ps["G"])
    tbs = rng.integers(0, 256, (6, ps["N_TB_byte"]), dtype=np.uint8)
    llr = np.zeros((6, ps["G"]), dtype=np.int16)
    oks, nits, decs = [], [], []
    for i in range(6):
        bits = fec.unpack_bits(fec.pdc_encode(tbs[i], ps, 0, gold), ps["G"]).astype(np.float64)
        sigma = 10 ** (-(2.5 + 0.5 * i) / 20)
        y = (2.0 * bits - 1.0) + sigma * rng.standard_normal(bits.shape)
        llr[i] = np.clip(np.round(y * 2 / sigma ** 2 * 16), -1000, 1000).astype(np.int16)
        ok, dec, nit = fec.pdc_decode(llr[i], ps, 0, gold)
        oks.append(ok), nits.append(nit), decs.append(dec)
    out["dec_cfg"], out["dec_llr"], out["dec_tb_in"] = np.array(cfg, dtype=np.int32), llr, tbs
    out["dec_ok"], out["dec_nit"], out["dec_tb"] = np.array(oks), np.array(nits), np.array(decs)
    return out


if __name__ == "__main__":
    d = build()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_v1.npz")
    np.savez_compressed(path, **d)
    print(path, os.path.getsize(path), "bytes;", len(d), "arrays")
