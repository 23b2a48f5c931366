"""CPU tests (no GPU): oracle pinned against published known answers, product host tables against the
oracle, and the C ABI export list against include/dnrp_c_api.h."""
import os
import re
import subprocess

import numpy as np
import pytest

from oracle import fec, spec
from conftest import CONFIGS, ROOT


def test_crc_known_answers():
    # check values of the CRC catalogue for the ASCII string "123456789":
    # CRC-24/LTE-A 0xCDE703, CRC-24/LTE-B 0x23EF52, CRC-16/XMODEM 0x31C3 (TS 36.212 5.1.1 polynomials)
    msg = np.unpackbits(np.frombuffer(b"123456789", dtype=np.uint8))
    assert fec.crc(msg, 0x1864CFB, 24) == 0xCDE703
    assert fec.crc(msg, 0x1800063, 24) == 0x23EF52
    assert fec.crc(msg, 0x11021, 16) == 0x31C3


def test_qpp_table_is_permutation_and_contention_free():
    for K in spec.TC_CB_SIZES:
        pi = spec.qpp(K)
        assert len(set(pi.tolist())) == K
        # pi(j + K/2) = pi(j) + K/2 (mod K): the property the CUDA decoder's int16x2 packing relies on
        assert ((pi[K // 2:] - pi[:K // 2]) % K == K // 2).all()


def test_gold_sequence_definition():
    # TS 36.211 7.2: x1(n+31) = x1(n+3)+x1(n), Nc = 1600; with c_init = 0 the output is x1 alone
    c = spec.gold_sequence(0, 64)
    x1 = np.zeros(1600 + 64 + 31, dtype=np.uint8)
    x1[0] = 1
    for i in range(len(x1) - 31):
        x1[i + 31] = x1[i + 3] ^ x1[i]
    assert (c == x1[1600:1664]).all()
    # linearity in c_init
    a, b = spec.gold_sequence(0x1234, 256), spec.gold_sequence(0x4321, 256)
    assert ((a ^ b) == (spec.gold_sequence(0x1234 ^ 0x4321, 256) ^ c.repeat(4)[:256] * 0 ^ spec.gold_sequence(0, 256))).all()


def test_turbo_encoder_termination_and_systematic():
    rng = np.random.default_rng(0)
    for K in (40, 56, 96, 512, 6144):
        c = rng.integers(0, 2, K, dtype=np.uint8)
        d = fec.turbo_encode(c)
        assert (d[:K] == c).all()
        # all-zero input -> all-zero codeword (linear code, zero state needs no termination bits)
        assert not fec.turbo_encode(np.zeros(K, dtype=np.uint8)).any()
        # linearity
        c2 = rng.integers(0, 2, K, dtype=np.uint8)
        assert (fec.turbo_encode(c ^ c2) == (d ^ fec.turbo_encode(c2))).all()


def test_rate_matching_roundtrip_all_rv():
    rng = np.random.default_rng(1)
    for K in (56, 96, 480, 5952):
        d = rng.integers(0, 2, 3 * (K + 4), dtype=np.uint8)
        N = 3 * (K + 4)
        for rv in range(4):
            e = fec.rm_tx(d, K, N, rv)       # exactly one lap of the circular buffer: a permutation
            assert sorted(e.tolist()) == sorted(d.tolist())
            soft = fec.rm_rx((e.astype(np.int16) * 2 - 1), K, rv)
            assert ((soft > 0).astype(np.uint8) == d).all()
        e2 = fec.rm_tx(d, K, 2 * N + 5, 0)   # repetition wraps around
        assert (e2[:N] == e2[N:2 * N]).all()


@pytest.mark.parametrize("name", list(CONFIGS))
def test_pdc_chain_roundtrip(name):
    ps = spec.get_packet_sizes(*CONFIGS[name])
    rng = np.random.default_rng(3)
    gold = fec.pdc_gold(123456789, 1, ps["G"])
    tb = rng.integers(0, 256, ps["N_TB_byte"], dtype=np.uint8)
    d = fec.unpack_bits(fec.pdc_encode(tb, ps, 0, gold), ps["G"])
    llr = ((d.astype(np.int16) * 2 - 1) * 10).astype(np.int16)   # reference tb2pdc.cpp:173-176
    ok, out, nit = fec.pdc_decode(llr, ps, 0, gold)
    assert ok and (out == tb).all() and (nit == 2).all()
    # HARQ (reference tb2pdc_awgn.cpp:214-228, rv order 0,2,3,1): a first transmission with 45 % of
    # the LLRs erased fails, the soft-combined retransmissions recover the block
    harq = fec.HarqRx(ps)
    results = []
    for i, rv in enumerate((0, 2, 3, 1)):
        d = fec.unpack_bits(fec.pdc_encode(tb, ps, rv, gold), ps["G"])
        llr = ((d.astype(np.int16) * 2 - 1) * 10).astype(np.int16)
        llr[rng.random(ps["G"]) < (0.45 if i == 0 else 0.2)] = 0
        ok, out, nit = fec.pdc_decode(llr, ps, rv, gold, harq=harq)
        results.append(ok)
        if ok:   # the reference finalizes (resets) the HARQ process after a success
            assert (out == tb).all()
            break
    assert results[-1] and not results[0]


def test_plcf_chain_roundtrip_all_masks():
    rng = np.random.default_rng(4)
    for plcf_type in (1, 2):
        for cl in (False, True):
            for bf in (False, True):
                plcf = rng.integers(0, 256, 5 if plcf_type == 1 else 10, dtype=np.uint8)
                d = fec.unpack_bits(fec.pcc_encode(plcf, plcf_type, cl, bf), 196)
                llr = ((d.astype(np.int16) * 2 - 1) * 127).astype(np.int16)  # reference plcf2pcc.cpp
                ok, out, cl2, bf2, _ = fec.pcc_decode(llr, plcf_type)
                assert ok and (out == plcf).all() and cl2 == cl and bf2 == bf


def test_decoder_int16_range_never_violated():
    """the decoder's int16 metric contract, including the product's normalise-every-4th-step schedule in the
    training / backward passes (oracle/fec.c norm_train, norm_back): no value leaves 16 bits, even for saturated
    and adversarial inputs, and for window lengths with L % 4 != 0"""
    L = fec._L()
    rng = np.random.default_rng(5)
    t = L.orc_tdec_alloc()
    for K in (40, 72, 1056, 6144):
        f1, f2 = spec.QPP_F1F2[spec.TC_CB_SIZES.index(K)]
        bits = np.zeros(K, dtype=np.uint8)
        n = 3 * (K + 4)
        pats = [rng.integers(-a, a + 1, n) for a in (50, 1000, 32767)]
        pats += [rng.choice([-32767, 32767], n), np.full(n, 32767), np.full(n, -32767),
                 np.where(np.arange(n) % 2, 32767, -32767), np.where(np.arange(n) % 3 == 0, 32767, -32767)]
        for soft in pats:
            L.orc_tdec_new_cb(t, np.ascontiguousarray(soft, dtype=np.int16), K, f1, f2)
            for _ in range(4):
                L.orc_tdec_iteration(t, bits)
            assert L.orc_tdec_range_violations(t) == 0, K
    L.orc_tdec_free(t)


def test_packet_sizes_match_survey_table():
    # SURVEY.md section 8 "Reference sizes" (derived from the reference's formulas)
    ps = spec.get_packet_sizes(1, 16, 1, 1, 0, 7, 6144)
    assert (ps["N_PDC_subc"], ps["G"], ps["N_TB_bits"], ps["C"]) == (6622, 39732, 32920, 6)
    ps = spec.get_packet_sizes(1, 16, 1, 2, 5, 7, 6144)
    assert (ps["N_PDC_subc"], ps["G"], ps["N_TB_bits"], ps["C"]) == (14238, 85428, 70856, 12)
    ps = spec.get_packet_sizes(8, 16, 1, 1, 10, 9, 6144)
    assert (ps["N_PDC_subc"], ps["G"], ps["N_TB_bits"], ps["C"]) == (53662, 429296, 356256, 59)
    assert spec.get_packet_sizes(1, 16, 1, 1, 5, 7, 6144) is None  # N_eff_TX = 4 needs >= 15 symbols


def test_re_allocation_counts():
    # the counts the reference asserts at the end of every TX packet (tx.cpp:299-313)
    for (u, b, plt, pl, tm, mcs, Z) in [(1, 1, 1, 1, 0, 2, 2048), (1, 16, 1, 2, 5, 7, 6144), (2, 4, 1, 1, 1, 3, 2048),
                                        (8, 2, 1, 2, 10, 1, 2048), (1, 12, 0, 5, 1, 3, 6144)]:
        ps = spec.get_packet_sizes(u, b, plt, pl, tm, mcs, Z)
        assert ps is not None
        pcc = spec.pcc_k_i_linear(b, ps["N_TS"])
        pdc = spec.pdc_k_i_linear(b, ps["N_TS"], u, ps["N_PACKET_symb"])
        assert len(pcc) == 98 and len(set(pcc)) == 98
        assert len(pdc) == ps["N_PDC_subc"] and not set(pdc) & set(pcc)


def test_product_packet_sizes_match_oracle(product):
    # the sweep of reference lib/src/phy/fec/test/tb2pdc.cpp:78-99 (u x b x tm x type x len(odd) x mcs(even))
    n = 0
    for u in (1, 2, 4, 8):
        for b in (1, 2, 4, 8, 12, 16):
            for tm in range(12):
                for plt in (0, 1):
                    for pl in (1, 3, 5, 7, 16):
                        for mcs in (0, 2, 4, 6, 8, 9):
                            for Z in (2048, 6144):
                                a = spec.get_packet_sizes(u, b, plt, pl, tm, mcs, Z)
                                p = product.get_packet_sizes(u, b, plt, pl, tm, mcs, Z)
                                assert (a is None) == (p is None), (u, b, plt, pl, tm, mcs, Z)
                                if a is None:
                                    continue
                                n += 1
                                for k in ("N_PACKET_symb", "N_DF_symb", "N_PDC_subc", "N_DRS_subc", "G", "N_TB_bits",
                                          "N_TB_byte", "C", "N_samples_packet", "N_samples_STF", "N_samples_GI"):
                                    assert a[k] == p[k], (k, u, b, plt, pl, tm, mcs, Z)
                                assert a["seg"]["K1"] == p["K_plus"] and a["seg"]["C2"] == p["C_minus"]
    assert n > 5000


def test_c_abi_exports_every_declared_symbol(product):
    hdr = open(product.HEADER_PATH).read()
    declared = set(re.findall(r"\b(dnrp_[a-z0-9_]+)\s*\(", hdr))
    out = subprocess.check_output(["nm", "-D", "--defined-only", product.LIB_PATH]).decode()
    exported = set(l.split()[-1] for l in out.splitlines() if " T " in l)
    missing = declared - exported
    assert not missing, missing


def test_tdl_profiles_rms_delay_spread():
    """link.hpp:84-88 quotes the measured rms delay spreads of the three generic profiles"""
    import ctypes as C
    import oracle
    from oracle import channel as ochan
    L = C.CDLL(oracle.build())
    L.orc_tdl_tau_rms_ns.restype = C.c_double
    for idx, quoted in enumerate((43.13, 356.65, 990.93)):
        assert abs(L.orc_tdl_tau_rms_ns(idx) - quoted) < 0.011
    # with the profile's own spread the delays are the generic ones floored to samples, powers sum to one
    d, p = ochan.tdl_pdp(1.0e9, 1, L.orc_tdl_tau_rms_ns(1))
    assert len(d) == 9 and all(abs(int(a) - b) <= 1 for a, b in zip(d, [0, 30, 150, 310, 370, 710, 1090, 1730, 2510]))
    assert abs(p.sum() - 1.0) < 1e-12
    # equal sample delays are merged (link.cpp:103-121)
    d2, p2 = ochan.tdl_pdp(1.728e6, 0, 43.13)
    assert len(d2) < 7 and len(set(d2.tolist())) == len(d2) and abs(p2.sum() - 1.0) < 1e-12


def test_binding_structs_have_the_layout_of_the_header(product, tmp_path):
    """the ctypes / numpy mirrors of the C ABI's structs (dect-nr-plus-sdr_b200/__init__.py) against sizeof / offsetof of
    include/dnrp_c_api.h compiled by gcc: a field added to the header and not to the binding shows up here, without a GPU"""
    import ctypes as C
    import subprocess
    src = tmp_path / "layout.c"
    src.write_text('''#include <stdio.h>
#include <stddef.h>
#include "dnrp_c_api.h"
int main(void) {
    printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(dnrp_psdef_t), sizeof(dnrp_packet_sizes_t), sizeof(dnrp_config_t),
           offsetof(dnrp_config_t, N_soft), offsetof(dnrp_config_t, M_DL_HARQ), sizeof(dnrp_tx_desc_t), sizeof(dnrp_sync_report_t),
           sizeof(dnrp_pcc_report_t), sizeof(dnrp_pdc_report_t), sizeof(dnrp_sync_result_t), sizeof(dnrp_stream_packet_t));
    return 0;
}
''')
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-std=c11", "-I" + os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(v) for v in subprocess.check_output([str(exe)], text=True).split()]
    p = product
    want = [C.sizeof(p.Psdef), C.sizeof(p.PacketSizes), C.sizeof(p.Config), p.Config.N_soft.offset, p.Config.M_DL_HARQ.offset,
            C.sizeof(p.TxDesc), C.sizeof(p.SyncReport), C.sizeof(p.PccReport), C.sizeof(p.PdcReport), p.SYNC_RESULT_DTYPE.itemsize,
            p.STREAM_PACKET_DTYPE.itemsize]
    assert got == want, (got, want)
