"""Limited soft buffer in the rate matching (dnrp_config_t::N_soft / M_DL_HARQ; the reference lists "channel coding with a
limited number of soft bits" as not implemented, README.md:180, ToDo at pdc_enc.cpp:122-123 -> ETSI TS 103 636-3 6.1.5.3,
which follows TS 36.212 5.1.4.1.2).  The oracle's circular-buffer walk with N_cb < K_w against a textbook restatement of
TS 36.212 5.1.4.1.1 / 5.1.4.1.2 written here from the clause (sub-block interleaver, bit collection, k0(N_cb), NULL skipping)."""
import ctypes as C

import numpy as np
import pytest

from oracle import fec, spec

P = [0, 16, 8, 24, 4, 20, 12, 28, 2, 18, 10, 26, 6, 22, 14, 30, 1, 17, 9, 25, 5, 21, 13, 29, 3, 19, 11, 27, 7, 23, 15, 31]


def textbook_selection(K, N_cb, rv, E):
    """indices into d = [d0 | d1 | d2] (each K + 4 long) of the E selected bits"""
    D = K + 4
    R = -(-D // 32)
    Kp = 32 * R
    ND = Kp - D

    def interleave(stream, last):
        y = [None] * ND + [stream * D + i for i in range(D)]
        if not last:      # matrix written row by row, columns permuted, read column by column
            return [y[r * 32 + P[c]] for c in range(32) for r in range(R)]
        return [y[(P[k // R] + 32 * (k % R) + 1) % Kp] for k in range(Kp)]

    v0, v1, v2 = interleave(0, False), interleave(1, False), interleave(2, True)
    w = v0 + [x for pair in zip(v1, v2) for x in pair]
    N_cb = min(N_cb, 3 * Kp)
    k0 = R * (2 * -(-N_cb // (8 * R)) * rv + 2)
    out, j = [], 0
    while len(out) < E:
        a = w[(k0 + j) % N_cb]
        if a is not None:
            out.append(a)
        j += 1
    return out


def _set_ncb(v):
    C.c_uint32.in_dll(fec._L(), "orc_N_cb").value = v


@pytest.mark.parametrize("K,frac,E", [(40, 1.0, 100), (40, 0.5, 200), (512, 0.7, 900), (1024, 0.4, 3000), (6144, 0.6, 9000),
                                      (6144, 0.34, 12000), (2048, 2.0, 5000)])
def test_circular_buffer_walk_matches_the_clause(K, frac, E):
    Kw = 3 * 32 * -(-(K + 4) // 32)
    N_cb = int(frac * Kw)
    for rv in range(4):
        want = textbook_selection(K, N_cb, rv, E)
        e = (np.arange(E) % 251 + 1).astype(np.int16)
        _set_ncb(N_cb)
        try:
            soft = fec.rm_rx(e, K, rv)
            d = np.random.default_rng(K + rv).integers(0, 2, 3 * (K + 4)).astype(np.uint8)
            sel = fec.rm_tx(d, K, E, rv)
        finally:
            _set_ncb(0)
        ref = np.zeros(3 * (K + 4), dtype=np.int64)
        np.add.at(ref, want, e.astype(np.int64))
        assert np.array_equal(soft.astype(np.int64), ref), (K, N_cb, rv)
        assert np.array_equal(sel, d[want]), (K, N_cb, rv)
        if frac < 1.0:   # nothing outside the limited buffer is ever touched
            assert len(set(want)) <= N_cb


def test_transport_block_round_trip_with_a_limited_buffer():
    cfg = (1, 4, 1, 2, 0, 3, 2048)     # 16-QAM 1/2: n_e about 2 K per code block
    ps = spec.get_packet_sizes(*cfg)
    gold = fec.pdc_gold(1234, 1, ps["G"])
    rng = np.random.default_rng(5)
    tb = rng.integers(0, 256, ps["N_TB_byte"], dtype=np.uint8)
    ref_bits = fec.unpack_bits(fec.pdc_encode(tb, ps, 0, gold), ps["G"])
    try:
        fec.set_soft_limit(10 ** 9)                       # no effect
        assert np.array_equal(fec.unpack_bits(fec.pdc_encode(tb, ps, 0, gold), ps["G"]), ref_bits)
        fec.set_soft_limit(int(1.4 * ps["N_TB_bits"]), 1)    # rate-5/7 buffers: the selection wraps around
        bits = fec.unpack_bits(fec.pdc_encode(tb, ps, 0, gold), ps["G"])
        assert not np.array_equal(bits, ref_bits)
        llr = (200 * (2 * bits.astype(np.int32) - 1)).astype(np.int16)
        ok, out, nit = fec.pdc_decode(llr, ps, 0, gold)
        assert ok and np.array_equal(out[:ps["N_TB_byte"]], tb)
        # HARQ: rv 2 on top of rv 0 inside the same limited buffer
        h = fec.HarqRx(ps)
        noisy = (llr.astype(np.int32) // 40 + rng.integers(-9, 10, llr.shape)).astype(np.int16)
        fec.pdc_decode(noisy, ps, 0, gold, harq=h)
        bits2 = fec.unpack_bits(fec.pdc_encode(tb, ps, 2, gold), ps["G"])
        ok2, out2, _ = fec.pdc_decode((60 * (2 * bits2.astype(np.int32) - 1)).astype(np.int16), ps, 2, gold, harq=h)
        assert ok2 and np.array_equal(out2[:ps["N_TB_byte"]], tb)
    finally:
        fec.set_soft_limit(0)
