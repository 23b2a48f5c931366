"""GPU parity tests of the FEC kernels against the CPU oracle, through the C ABI.  Bit-exact: coded
bits, decoded transport blocks, CRC outcomes AND iteration counts must be identical (the decoder is an
integer algorithm with a fixed contract)."""
import numpy as np
import pytest

from oracle import fec, spec
from conftest import CONFIGS

pytestmark = pytest.mark.gpu
NET = 123456789


def _noisy_llr(bits, snr_db, rng, scale=16, clip=1000):
    sigma = 10 ** (-snr_db / 20)
    y = (2.0 * bits - 1.0) + sigma * rng.standard_normal(bits.shape)
    return np.clip(np.round(y * 2 / sigma ** 2 * scale), -clip, clip).astype(np.int16)


@pytest.mark.parametrize("name", list(CONFIGS))
@pytest.mark.parametrize("plcf_type", [1, 2])
def test_encode_bit_exact(product, name, plcf_type):
    cfg = CONFIGS[name]
    ps = spec.get_packet_sizes(*cfg)
    n = 6
    phy = product.Phy(cfg, max_batch=n, network_id=NET, plcf_type=plcf_type)
    rng = np.random.default_rng(10)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    masks = np.array([0, 0x5555, 0xAAAA, 0xFFFF, 0, 0x5555], dtype=np.uint32)
    rv = np.array([0, 1, 2, 3, 0, 2], dtype=np.uint8)
    pcc, pdc = phy.fec_encode(plcf, tb, cl_bf_mask=masks, rv=rv)
    gold = fec.pdc_gold(NET, plcf_type, ps["G"])
    nb = 5 if plcf_type == 1 else 10
    for i in range(n):
        ref = fec.pdc_encode(tb[i], ps, int(rv[i]), gold)
        assert (fec.unpack_bits(pdc[i], ps["G"]) == fec.unpack_bits(ref, ps["G"])).all(), (name, i)
        m = int(masks[i])
        refp = fec.pcc_encode(plcf[i, :nb], plcf_type, m in (0x5555, 0xFFFF), m in (0xAAAA, 0xFFFF))
        assert (fec.unpack_bits(pcc[i], 196) == fec.unpack_bits(refp, 196)).all(), (name, i)
    phy.close()


@pytest.mark.parametrize("name,snrs", [("C1_loopback", (2.0, 4.0, 8.0)), ("C2_b16_siso_mcs7", (5.5, 6.0, 7.0)),
                                       ("C3_b16_txdiv4_mcs7", (5.5, 6.5)), ("small_b2_mcs4", (3.0, 5.0)),
                                       ("txdiv2_b4_mcs3", (0.0, 2.0))])
def test_decode_tb_bit_exact(product, name, snrs):
    cfg = CONFIGS[name]
    ps = spec.get_packet_sizes(*cfg)
    gold = fec.pdc_gold(NET, 1, ps["G"])
    rng = np.random.default_rng(11)
    per = 4
    n = per * len(snrs)
    phy = product.Phy(cfg, max_batch=n, network_id=NET, plcf_type=1)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    llr = np.zeros((n, ps["G"]), dtype=np.int16)
    for i in range(n):
        bits = fec.unpack_bits(fec.pdc_encode(tb[i], ps, 0, gold), ps["G"]).astype(np.float64)
        llr[i] = _noisy_llr(bits, snrs[i // per], rng)
    ok, out, nit = phy.fec_decode_tb(llr)
    n_ok = 0
    for i in range(n):
        rok, rout, rnit = fec.pdc_decode(llr[i], ps, 0, gold)
        assert bool(ok[i]) == rok, (name, i)
        assert (nit[i] == rnit).all(), (name, i, nit[i], rnit)
        assert (out[i] == rout).all(), (name, i)
        n_ok += rok
    assert 0 < n_ok  # the SNR grid straddles the waterfall: at least the high-SNR packets decode
    phy.close()


@pytest.mark.parametrize("name", ["C2_b16_siso_mcs7", "small_b2_mcs4"])
def test_decode_tb_saturated_llrs_bit_exact(product, name):
    """worst case for the path-metric range (the decoder normalises only every fourth step in its training and
    backward passes): saturated LLRs with random / constant / alternating signs -- hard decisions, CRC outcome and
    iteration counts still equal the oracle's, whose 32-bit arithmetic range-checks every value"""
    cfg = CONFIGS[name]
    ps = spec.get_packet_sizes(*cfg)
    gold = fec.pdc_gold(NET, 1, ps["G"])
    rng = np.random.default_rng(13)
    G = ps["G"]
    pats = [rng.choice([-1023, 1023], G), np.full(G, 1023), np.full(G, -1023), np.where(np.arange(G) % 2, 1023, -1023),
            rng.integers(-1023, 1024, G), rng.choice([-32767, 32767], G)]
    llr = np.stack(pats).astype(np.int16)
    phy = product.Phy(cfg, max_batch=len(pats), network_id=NET, plcf_type=1)
    ok, out, nit = phy.fec_decode_tb(llr)
    for i in range(len(pats)):
        rok, rout, rnit = fec.pdc_decode(llr[i], ps, 0, gold)
        assert bool(ok[i]) == rok and (nit[i] == rnit).all() and (out[i] == rout).all(), (name, i)
    phy.close()


def test_decode_tb_harq_combining_bit_exact(product):
    cfg = CONFIGS["C2_b16_siso_mcs7"]
    ps = spec.get_packet_sizes(*cfg)
    gold = fec.pdc_gold(NET, 1, ps["G"])
    rng = np.random.default_rng(12)
    n = 4
    phy = product.Phy(cfg, max_batch=8, network_id=NET, plcf_type=1)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    harq = [fec.HarqRx(ps) for _ in range(n)]
    slots = np.array([5, 2, 7, 0], dtype=np.uint32)
    done = np.zeros(n, dtype=bool)
    for t, rv in enumerate((0, 2, 3, 1)):
        llr = np.zeros((n, ps["G"]), dtype=np.int16)
        for i in range(n):
            bits = fec.unpack_bits(fec.pdc_encode(tb[i], ps, rv, gold), ps["G"]).astype(np.float64)
            llr[i] = _noisy_llr(bits, 3.0 + i, rng)
        ok, out, nit = phy.fec_decode_tb(llr, harq_slot=slots, rv=np.full(n, rv, np.uint8),
                                         harq_reset=np.full(n, 1 if t == 0 else 0, np.uint8))
        for i in range(n):
            if done[i]:
                continue
            rok, rout, rnit = fec.pdc_decode(llr[i], ps, rv, gold, harq=harq[i])
            assert bool(ok[i]) == rok and (nit[i] == rnit).all() and (out[i] == rout).all(), (t, i)
            done[i] = rok
    assert done.any()
    phy.close()


def test_decode_plcf_bit_exact(product):
    cfg = CONFIGS["C1_loopback"]
    n = 32
    phy = product.Phy(cfg, max_batch=n, network_id=NET, plcf_type=1)
    rng = np.random.default_rng(13)
    llr = np.zeros((n, 196), dtype=np.int16)
    for i in range(n):
        ty = 1 + (i % 2)
        plcf = rng.integers(0, 256, 5 if ty == 1 else 10, dtype=np.uint8)
        bits = fec.unpack_bits(fec.pcc_encode(plcf, ty, i % 4 == 1, i % 8 >= 4), 196).astype(np.float64)
        llr[i] = _noisy_llr(bits, -4.0 + 0.5 * (i // 2), rng, scale=8, clip=127)
    ok, plcf_out, mask, nit = phy.fec_decode_plcf(llr)
    n_ok = 0
    for i in range(n):
        for ty in (1, 2):
            rok, rout, cl, bf, rnit = fec.pcc_decode(llr[i], ty)
            assert bool(ok[i, ty - 1]) == rok, (i, ty)
            assert nit[i, ty - 1] == rnit, (i, ty)
            if rok:
                nb = 5 if ty == 1 else 10
                assert (plcf_out[i, ty - 1, :nb] == rout).all()
                assert mask[i, ty - 1] == {(0, 0): 0, (1, 0): 0x5555, (0, 1): 0xAAAA, (1, 1): 0xFFFF}[(int(cl), int(bf))]
                n_ok += 1
    assert n_ok >= n // 2
    phy.close()


def test_duplicate_or_out_of_range_harq_slots_are_refused(product):
    """two packets of one batch on the same HARQ slot would combine into one soft buffer concurrently: the C entry points
    refuse that (the reference hands out each HARQ process once, process_pool.cpp) and stay usable afterwards"""
    cfg = CONFIGS["small_b2_mcs4"]
    ps = spec.get_packet_sizes(*cfg)
    phy = product.Phy(cfg, max_batch=4, network_id=123456789)
    llr = np.zeros((3, ps["G"]), dtype=np.int16)
    with pytest.raises(product.DnrpError, match="appears twice"):
        phy.fec_decode_tb(llr, harq_slot=[1, 2, 1])
    with pytest.raises(product.DnrpError, match="out of range"):
        phy.fec_decode_tb(llr, harq_slot=[0, 1, 4])
    win = np.zeros((3, 1, phy.rx_stride()), dtype=np.complex64)
    with pytest.raises(product.DnrpError, match="appears twice"):
        phy.rx(win, harq_slot=[3, 3, 0])
    phy.rx_pcc(win, t0=0)
    with pytest.raises(product.DnrpError, match="appears twice"):
        phy.rx_pdc(3, harq_slot=[2, 0, 2])
    ok, tb, nit = phy.fec_decode_tb(llr, harq_slot=[2, 0, 1])   # still usable
    assert ok.shape == (3,) and not tb.any()
    phy.close()


@pytest.mark.parametrize("cfg,snrs", [((1, 16, 1, 1, 1, 3, 6144), (0.5, 1.5, 3.0)), ((1, 12, 1, 1, 1, 5, 6144), (2.0, 3.0, 4.5))])
def test_decode_tb_bit_exact_largest_code_blocks(product, cfg, snrs):
    """K = 6144 (two code blocks) and K = 6144 / 6080 mixed (three): the sizes whose decoder keeps its first two backward
    checkpoints in the boundary rows to stay at four blocks per SM (k_turbo_decode<.., CKALIAS>) --
    outcomes, iteration counts and bytes equal the oracle's across the waterfall"""
    ps = spec.get_packet_sizes(*cfg)
    assert ps["seg"]["K1"] == 6144
    gold = fec.pdc_gold(NET, 1, ps["G"])
    rng = np.random.default_rng(12)
    per = 4
    n = per * len(snrs)
    phy = product.Phy(cfg, max_batch=n, network_id=NET, plcf_type=1)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    llr = np.zeros((n, ps["G"]), dtype=np.int16)
    for i in range(n):
        bits = fec.unpack_bits(fec.pdc_encode(tb[i], ps, 0, gold), ps["G"]).astype(np.float64)
        llr[i] = _noisy_llr(bits, snrs[i // per], rng)
    ok, out, nit = phy.fec_decode_tb(llr)
    n_ok = 0
    for i in range(n):
        rok, rout, rnit = fec.pdc_decode(llr[i], ps, 0, gold)
        assert bool(ok[i]) == rok, (cfg, i)
        assert (nit[i] == rnit).all(), (cfg, i, nit[i], rnit)
        assert (out[i] == rout).all(), (cfg, i)
        n_ok += rok
    assert 0 < n_ok < n, n_ok   # the SNR grid straddles the waterfall
    phy.close()
