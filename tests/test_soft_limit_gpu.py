"""Limited soft buffer (dnrp_config_t::N_soft / M_DL_HARQ) on the GPU against the oracle, through the C ABI: coded bits,
decoded transport blocks, CRC outcomes and iteration counts identical for redundancy versions 0..3 and over a HARQ sequence.
The reference does not implement the limitation (README.md:180, pdc_enc.cpp:122-123); the clause it points to follows TS 36.212
5.1.4.1.2, against which tests/test_soft_limit_cpu.py checks the oracle."""
import numpy as np
import pytest

from oracle import fec, spec
from test_fec_gpu import _noisy_llr

pytestmark = pytest.mark.gpu
NET = 123456789

CASES = {
    # cfg, N_soft as a multiple of the transport block size, M_DL_HARQ, SNRs
    "b4_16qam_r12_2cb_wraps": ((1, 4, 1, 2, 0, 3, 2048), 1.4, 1, (3.0, 6.0)),        # n_e > N_cb: selection wraps (repetition loops)
    "b16_64qam_r56_6144": ((1, 16, 1, 2, 0, 7, 6144), 8 * 1.25, 8, (6.0, 8.5)),      # eight HARQ processes share N_soft
    "b8_64qam_r34_txdiv": ((1, 8, 1, 2, 1, 6, 2048), 2.0, 1, (7.0, 10.0)),
    "b2_qpsk_r12_single_cb": ((1, 2, 1, 1, 0, 1, 2048), 1.7, 1, (1.0, 4.0)),
}


@pytest.mark.parametrize("name", list(CASES))
def test_limited_soft_buffer_bit_exact(product, name):
    cfg, mult, M, snrs = CASES[name]
    ps = spec.get_packet_sizes(*cfg)
    N_soft = int(mult * ps["N_TB_bits"])
    gold = fec.pdc_gold(NET, 1, ps["G"])
    rng = np.random.default_rng(sum(map(ord, name)))
    per = 3
    n = per * len(snrs)
    phy = product.Phy(cfg, max_batch=n, network_id=NET, plcf_type=1, N_soft=N_soft, M_DL_HARQ=M)
    free = product.Phy(cfg, max_batch=n, network_id=NET, plcf_type=1)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    try:
        fec.set_soft_limit(N_soft, M)
        # transmitter: all redundancy versions; the limited selection differs from the unlimited one
        for rv in range(4):
            rvs = np.full(n, rv, dtype=np.uint8)
            _, pdc = phy.fec_encode(plcf, tb, rv=rvs)
            _, pdc_free = free.fec_encode(plcf, tb, rv=rvs)
            for i in range(n):
                ref = fec.pdc_encode(tb[i], ps, rv, gold)
                assert (fec.unpack_bits(pdc[i], ps["G"]) == fec.unpack_bits(ref, ps["G"])).all(), (name, rv, i)
            if rv > 0:    # k0 follows N_cb; rv 0 only differs where the selection wraps around
                assert not np.array_equal(pdc, pdc_free), (name, rv)
        # receiver: HARQ sequence rv 0 -> 2 -> 3 -> 1 at SNRs around the waterfall of the first transmission
        harq = [fec.HarqRx(ps) for _ in range(n)]
        done = np.zeros(n, dtype=bool)
        slots = np.arange(n, dtype=np.uint32)[::-1].copy()
        n_first = 0
        for t, rv in enumerate((0, 2, 3, 1)):
            llr = np.zeros((n, ps["G"]), dtype=np.int16)
            for i in range(n):
                bits = fec.unpack_bits(fec.pdc_encode(tb[i], ps, rv, gold), ps["G"]).astype(np.float64)
                llr[i] = _noisy_llr(bits, snrs[i // per] - 3.0, rng)
            ok, out, nit = phy.fec_decode_tb(llr, harq_slot=slots, rv=np.full(n, rv, np.uint8),
                                             harq_reset=np.full(n, 1 if t == 0 else 0, np.uint8))
            for i in range(n):
                if done[i]:
                    continue
                rok, rout, rnit = fec.pdc_decode(llr[i], ps, rv, gold, harq=harq[i])
                assert bool(ok[i]) == rok and (nit[i] == rnit).all() and (out[i] == rout).all(), (name, t, i)
                done[i] = rok
                if rok:
                    assert np.array_equal(out[i][:ps["N_TB_byte"]], tb[i])
            if t == 0:
                n_first = int(done.sum())
        assert done.all() and n_first < n, (name, done, n_first)
    finally:
        fec.set_soft_limit(0)
        phy.close()
        free.close()


def test_buffer_smaller_than_the_systematic_part_is_refused(product):
    cfg = (1, 4, 1, 2, 0, 3, 2048)
    ps = spec.get_packet_sizes(*cfg)
    with pytest.raises(Exception, match="systematic"):
        product.Phy(cfg, max_batch=1, network_id=NET, plcf_type=1, N_soft=ps["N_TB_bits"] // 2, M_DL_HARQ=1)
