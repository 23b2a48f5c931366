"""The oracle (and the product's host-only functions) against oracle/_ref: the reference's OWN sources -- sections_part3,
tx_t, resampler_t, sync_chunk_t, rx_synced_t, fec_t -- compiled unmodified from /root/reference (oracle/ref_recipe).
This is what pins the oracle: every in-tree algorithm of the path is checked against the reference's object code; only
the arithmetic of the absent third-party libraries (srsRAN FEC / demapper, FFTW, VOLK, Eigen) is restated in the shim."""
import numpy as np
import pytest

from oracle import fec, ref, resample, rx, spec, sync, tx
from conftest import CONFIGS

pytestmark = pytest.mark.skipif(not (ref.available() or ref.can_build()),
                                reason="oracle/_ref is neither prebuilt nor buildable (no /root/reference)")
NET = 123456789


def _awgn(x, snr_db, rng):
    s = np.float32(10 ** (-snr_db / 20) / np.sqrt(2))
    return (x + s * (rng.standard_normal(x.shape) + 1j * rng.standard_normal(x.shape))).astype(np.complex64)


def test_packet_sizes_full_reference_sweep(product):
    """sp3::get_packet_sizes over the sweep of the reference's own tb2pdc test (tb2pdc.cpp:78-99, every class of
    radio_device_class.cpp) with step 1 instead of 2: reference == oracle == product, including WHICH configurations
    are impossible (std::nullopt)"""
    n_valid = n_invalid = 0
    for Z in (2048, 6144):
        for u in (1, 2, 4, 8):
            for b in (1, 2, 4, 8, 12, 16):
                for t in range(12):
                    for p_type in (0, 1):
                        for p in (1, 2, 3, 5, 8, 13, 16):
                            for mcs in range(0, 12):
                                cfg = (u, b, p_type, p, t, mcs, Z)
                                r = ref.get_packet_sizes(*cfg) if mcs <= 9 else None
                                o = spec.get_packet_sizes(*cfg) if mcs <= 9 else None
                                if mcs > 9:
                                    continue   # MCS 10/11 (1024-QAM) assert in the reference's mcs.cpp; widened in the product
                                d = product.get_packet_sizes(*cfg)
                                assert (r is None) == (o is None) == (d is None), cfg
                                if r is None:
                                    n_invalid += 1
                                    continue
                                n_valid += 1
                                for k, v in r.items():
                                    if k in o:
                                        assert o[k] == v, (cfg, k, o[k], v)
                                    if k in d:
                                        assert d[k] == v, (cfg, k, d[k], v)
    assert n_valid > 20000 and n_invalid > 1000


@pytest.mark.parametrize("b", [1, 2, 4, 8, 12, 16])
@pytest.mark.parametrize("N_TS", [1, 2, 4, 8])
def test_pcc_pdc_cell_lists(b, N_TS):
    """pcc_t::get_k_i_linear / pdc_t::get_k_i_linear (pcc.cpp:110-246, pdc.cpp:221-310) == oracle"""
    assert np.array_equal(ref.k_i_linear("pcc", b, N_TS), np.asarray(spec.pcc_k_i_linear(b, N_TS)))
    r = ref.k_i_linear("pdc", b, N_TS)
    # the reference lists the cells of its longest virtual frame; a packet's list is the prefix inside its symbols
    for u, n_symb in ((1, 10), (1, 20), (2, 15), (8, 40)):
        if N_TS == 4 and n_symb < 15 or N_TS == 8 and n_symb < 20:
            continue
        o = np.asarray(spec.pdc_k_i_linear(b, N_TS, u, n_symb))
        if o.size <= r.size:
            assert np.array_equal(r[:o.size], o), (b, N_TS, u, n_symb)


@pytest.mark.parametrize("b", [1, 2, 4, 8, 12, 16])
def test_stf_sequences(b):
    for n_eff in (1, 2, 4, 8):
        r = ref.stf(b, n_eff)
        o = spec.stf_filled(b, n_eff)
        assert r.shape == o.shape and np.abs(r - o).max() < 1e-6, (b, n_eff)


def test_codebooks():
    for N_TS, N_TX in sorted(tx.W_TABLES):   # every (N_TS, N_TX) of Table 7.2-1
        _, n = ref.W(N_TS, N_TX, 0)
        assert n == len(tx.W_TABLES[(N_TS, N_TX)])
        for idx in range(n):
            r, _ = ref.W(N_TS, N_TX, idx)
            o, scale = tx.codebook(N_TS, N_TX, idx)   # get_W holds {0, +-1, +-j}; the 1/sqrt(#nonzero) scale is checked by test_tx_iq
            assert np.abs(r - o).max() < 1e-6, (N_TS, N_TX, idx)


@pytest.mark.parametrize("name", list(CONFIGS))
@pytest.mark.parametrize("plcf_type", [1, 2])
def test_fec_bits_all_redundancy_versions(name, plcf_type):
    """fec_t::encode_plcf / encode_tb inside tx_t (pcc_enc.cpp:145-213, pdc_enc.cpp:127-229): coded + scrambled bits of
    the reference's HARQ buffers == oracle, rv 0 -> 2 -> 3 -> 1 on one HARQ process"""
    cfg = CONFIGS[name]
    ps = spec.get_packet_sizes(*cfg)
    rng = np.random.default_rng(7)
    t = ref.Tx(cfg)
    gold = fec.pdc_gold(NET, plcf_type, ps["G"])
    for trial in range(2):
        plcf = rng.integers(0, 256, 10, dtype=np.uint8)
        tb = rng.integers(0, 256, ps["N_TB_byte"], dtype=np.uint8)
        for i, rv in enumerate((0, 2, 3, 1)):
            _, pcc_d, pdc_d = t.packet(cfg, plcf, tb, network_id=NET, plcf_type=plcf_type, rv=rv, new_tb=(i == 0))
            o = fec.pdc_encode(tb, ps, rv, gold)
            assert np.array_equal(fec.unpack_bits(pdc_d, ps["G"]), fec.unpack_bits(o, ps["G"])), (name, rv)
            nb = 5 if plcf_type == 1 else 10
            op = fec.pcc_encode(plcf[:nb], plcf_type, bool(ps["cl"]), False)
            assert np.array_equal(fec.unpack_bits(pcc_d, 196), fec.unpack_bits(op, 196)), (name, rv)
    t.close()


TX_CASES = {
    "siso_b1": dict(cfg=(1, 1, 1, 1, 0, 2, 2048)),
    "siso_b16_mcs7": dict(cfg=(1, 16, 1, 1, 0, 7, 6144)),
    "txdiv2_b4": dict(cfg=(2, 4, 1, 1, 1, 3, 2048)),
    "txdiv4_b16": dict(cfg=(1, 16, 1, 2, 5, 7, 6144)),
    "txdiv8_b2": dict(cfg=(4, 2, 1, 1, 10, 4, 2048)),
    "beamforming_1x2": dict(cfg=(1, 2, 1, 1, 3, 2, 2048), codebook_index=3),
    "beamforming_1x4": dict(cfg=(1, 4, 1, 1, 7, 1, 2048), codebook_index=17),
    "mimo_2x2_open": dict(cfg=(1, 4, 1, 2, 2, 3, 2048)),
    "b12": dict(cfg=(2, 12, 0, 4, 0, 5, 6144)),
    "os2_mixer": dict(cfg=(1, 2, 1, 1, 0, 1, 2048), os=2, iq_phase_rad=0.7, iq_phase_increment=0.013, DAC_scale=0.5),
    "resample_10_9": dict(cfg=(1, 1, 1, 1, 0, 2, 2048), resample=(10, 9), iq_phase_increment=-0.004),
    "resample_10_9_os2": dict(cfg=(1, 2, 1, 1, 1, 3, 2048), os=2, resample=(10, 9)),
}


@pytest.mark.parametrize("case", list(TX_CASES))
def test_tx_iq(case):
    """tx_t::generate_tx_packet (tx.cpp:165-314) == oracle/tx.py: every sample of every antenna, incl. the sample count"""
    c = dict(TX_CASES[case])
    cfg = c.pop("cfg")
    os_, rs = c.pop("os", 1), c.pop("resample", (1, 1))
    ps = spec.get_packet_sizes(*cfg)
    rng = np.random.default_rng(11)
    plcf = rng.integers(0, 256, 10, dtype=np.uint8)
    tb = rng.integers(0, 256, ps["N_TB_byte"], dtype=np.uint8)
    t = ref.Tx(cfg, os=os_, resample=rs)
    iq, _, _ = t.packet(cfg, plcf, tb, network_id=NET, plcf_type=2, **c)
    o = tx.tx_packet(ps, plcf, tb, network_id=NET, plcf_type=2, os=os_, resample=rs, **c)
    assert iq.shape == o.shape, (iq.shape, o.shape)
    rms = np.sqrt((np.abs(o) ** 2).mean())
    assert np.abs(iq - o).max() < 2e-5 * max(rms, 1e-3) + 2e-6, (case, np.abs(iq - o).max(), rms)
    t.close()


@pytest.mark.parametrize("L,M,os_,user", [(10, 9, 1, 0), (9, 10, 1, 2), (9, 10, 1, 1), (10, 9, 2, 0), (9, 10, 2, 2), (1, 2, 1, 0)])
def test_resampler(product, L, M, os_, user):
    """resampler_t (resampler.cpp:56-562, the unrolled 223/23, 223/25 and 45/5 kernels included) == oracle's streaming
    restatement; sample counts equal; the product's filter design (dnrp_resampler_design) is the oracle's"""
    rng = np.random.default_rng(5)
    n = 4000
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
    for unit, flush in ((0, True), (800, True), (250, False)):   # every call must bring at least the filter history
        if unit and unit < 64:
            continue
        r = ref.resample(x, L, M, os=os_, user=user, unit=unit, flush=flush)
        o = resample.stream(x, L, M, os_, blocks=None if unit == 0 else [unit] * (n // unit) + ([n % unit] if n % unit else []), flush=flush)
        assert r.shape == o.shape, (r.shape, o.shape)
        assert np.abs(r - o).max() < 2e-5, np.abs(r - o).max()
    assert ref.resample_n_after(L, M, n, os=os_, user=user) == resample.Resampler(L, M, os_).n_after(n)


RX_CASES = {
    "siso_b1_mcs2": dict(cfg=(1, 1, 1, 1, 0, 2, 2048), snr=10.0),
    "siso_b2_mcs4": dict(cfg=(1, 2, 0, 3, 0, 4, 2048), snr=18.0),
    "siso_b16_mcs7": dict(cfg=(1, 16, 1, 1, 0, 7, 6144), snr=24.0),
    "txdiv2_b4_mcs3": dict(cfg=(2, 4, 1, 1, 1, 3, 2048), snr=14.0),
    "txdiv4_b4_mcs5": dict(cfg=(1, 4, 1, 2, 5, 5, 2048), snr=20.0),
    "siso_b4_mcs9_cfo": dict(cfg=(1, 4, 1, 1, 0, 9, 6144), snr=32.0, cfo=0.002),
}


@pytest.mark.parametrize("case", list(RX_CASES))
def test_rx_synced(case):
    """rx_synced_t::demoddecod_rx_pcc / demoddecod_rx_pdc (rx_synced.cpp:186-436) on noisy windows == oracle/rx.c:
    PCC / PDC LLRs within 1 LSB (fp32 rounding of the combiner), PLCF bytes, SNR / STO reports"""
    c = RX_CASES[case]
    cfg = c["cfg"]
    ps = spec.get_packet_sizes(*cfg)
    N_RX = ps["N_TX"]
    rng = np.random.default_rng(21)
    plcf = np.zeros(10, dtype=np.uint8)
    plcf[:5] = rng.integers(0, 256, 5, dtype=np.uint8)
    tb = rng.integers(0, 256, ps["N_TB_byte"], dtype=np.uint8)
    cfo = c.get("cfo", 0.0)
    iq = tx.tx_packet(ps, plcf, tb, network_id=NET, plcf_type=1, iq_phase_increment=cfo)
    t0 = 29
    nwin = iq.shape[1] + 300
    win = np.zeros((N_RX, nwin), dtype=np.complex64)
    h = (rng.standard_normal((N_RX, ps["N_TX"])) + 1j * rng.standard_normal((N_RX, ps["N_TX"]))).astype(np.complex64) / np.sqrt(2)
    if N_RX == 1:
        h[:] = 1
    win[:, t0:t0 + iq.shape[1]] = h @ iq
    win = _awgn(win, c["snr"], rng)
    stf_len = ps["N_samples_STF"]
    rms = np.sqrt((np.abs(win[:, t0:t0 + stf_len]) ** 2).mean(axis=1)).astype(np.float32)
    r = ref.Rx(cfg, nwin)
    p = r.pcc(win, t0, cfg[0], cfg[1], ps["N_eff_TX"], cfo_fractional_rad=-cfo, rms=rms)
    d = r.pdc(cfg, network_id=NET, plcf_type=1)
    o = rx.RxConfig(ps, N_RX=N_RX).rx_packet(win, t0=t0, cfo_fractional_rad=-cfo, rms=rms)
    dl = np.abs(p["llr_pcc"].astype(np.int32) - o["llr_pcc"].astype(np.int32))
    assert dl.max() <= 1, (case, dl.max())
    gold_bits = np.asarray(fec.pdc_gold(NET, 1, ps["G"]))[:ps["G"]].astype(np.int32)   # one bit per entry
    o_desc = np.where(gold_bits == 1, -o["llr_pdc"].astype(np.int32), o["llr_pdc"].astype(np.int32))
    r_desc = d["llr_pdc_descrambled"].astype(np.int32)
    dd = np.abs(r_desc - o_desc)
    # the reference descrambles in place and, for code block C - gamma, N_bps LLRs fewer than the transmitter scrambled
    # (pdc_enc.cpp:325-328 '>' against :155 '<='; restated in oracle/fec_chain.c): those stay scrambled in its buffer
    odd = np.nonzero(dd > 1)[0]
    assert odd.size <= ps["N_bps"] and np.abs(np.abs(r_desc[odd]) - np.abs(o_desc[odd])).max(initial=0) <= 1, (case, odd)
    dd[odd] = 0
    assert dd.max() <= 1 and (dd > 0).mean() < 0.02, (case, dd.max(), (dd > 0).mean())
    assert p["plcf_ok"][0] and np.array_equal(p["plcf"][0, :5], plcf[:5])
    rep = o["report"]
    assert abs(p["snr_dB"] - rep.snr_db_pcc) < 0.05 and abs(d["snr_dB"] - rep.snr_db_pdc) < 0.05, (p["snr_dB"], rep.snr_db_pcc)
    assert abs(p["sto_fractional"] - rep.sto_fractional) < 1e-3
    assert np.abs(p["rms"][:N_RX] - np.array(rep.rms[:N_RX])).max() < 1e-4
    assert d["crc_ok"] and np.array_equal(d["tb"], tb), case
    r.close()


@pytest.mark.parametrize("cfg,snr_db", [((1, 1, 1, 1, 0, 2, 2048), 8.0), ((1, 2, 1, 1, 0, 4, 2048), 15.0), ((1, 16, 1, 1, 0, 7, 6144), 20.0)])
def test_sync_chunk(cfg, snr_db):
    """sync_chunk_t::search (sync_chunk.cpp:146-283: autocorrelator_detection_t, autocorrelator_peak_t,
    crosscorrelator_t) over a stream holding three packets at random offsets == oracle/sync.c's sequential search: the
    same packets at the same fine-peak samples with the same N_eff_TX, fractional CFO within 2e-5"""
    ps = spec.get_packet_sizes(*cfg)
    rng = np.random.default_rng(31)
    pk = ps["N_samples_packet"]
    stream = np.zeros((1, 3 * pk + 3000), dtype=np.complex64)
    starts = []
    pos = 700
    for i in range(3):
        plcf = rng.integers(0, 256, 10, dtype=np.uint8)
        tb = rng.integers(0, 256, ps["N_TB_byte"], dtype=np.uint8)
        iq = tx.tx_packet(ps, plcf, tb, network_id=NET, iq_phase_increment=0.001 * (i - 1))
        pos += int(rng.integers(0, 300))
        stream[0, pos:pos + iq.shape[1]] += iq[0]
        starts.append(pos)
        pos += pk
    stream = _awgn(stream, snr_db, rng)
    found = ref.sync_stream(stream, cfg)
    fine = [f["fine_peak_time_64"] for f in found]
    assert len(found) >= 3, fine
    for s in starts:
        assert min(abs(f - s) for f in fine) <= 1, (s, fine)
    o = sync.stream_search(stream, cfg[0], cfg[1], N_eff_TX_max=1)
    o_fine = [int(q[0]) for q in o]
    assert o_fine == fine, (o_fine, fine)
    for (o_f, o_c, rep), f in zip(o, found):
        assert int(o_c) == f["coarse_peak_time_64"]
        assert int(rep.N_eff_TX) == f["N_eff_TX"]
        assert abs(float(rep.cfo_fractional_rad) - f["cfo_fractional_rad"]) < 2e-5
