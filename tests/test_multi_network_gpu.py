"""Per-packet network id and PLCF type in one batch (dnrp_add_network_id, dnrp_tx_batch_ex, dnrp_rx_pdc_batch_ex), as
the reference takes them from each packet's HARQ process (rx_synced.cpp:325-350, tx.cpp:176-186), and dnrp_harq_reset."""
import numpy as np
import pytest

from oracle import fec, spec, tx
from conftest import CONFIGS

pytestmark = pytest.mark.gpu
NETS = (123456789, 0x11223344, 100)


@pytest.mark.parametrize("name", ["small_b2_mcs4", "txdiv2_b4_mcs3"])
def test_mixed_network_ids_and_plcf_types_in_one_batch(product, name):
    cfg = CONFIGS[name]
    ps = spec.get_packet_sizes(*cfg)
    n = 12
    N_RX = ps["N_TX"]
    phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=NETS[0], plcf_type=1)
    rng = np.random.default_rng(3)
    ids = np.array([NETS[i % 3] for i in range(n)], dtype=np.uint32)
    tys = np.array([1 + (i // 3) % 2 for i in range(n)], dtype=np.uint8)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    with pytest.raises(product.DnrpError, match="not registered"):
        phy.tx_ex(plcf, tb, network_id=ids, plcf_type=tys)
    for net in NETS[1:]:
        phy.add_network_id(net)
    iq, ns = phy.tx_ex(plcf, tb, network_id=ids, plcf_type=tys)
    # transmitter: every packet equals the oracle's packet for ITS network id and PLCF type
    for i in range(n):
        o = tx.tx_packet(ps, plcf[i], tb[i], network_id=int(ids[i]), plcf_type=int(tys[i]))
        err = np.abs(iq[i, :, :ns[i]] - o).max()
        assert ns[i] == o.shape[1] and err < 1e-4, (name, i, err)
    # receiver: loop the packets back, decode with the per-packet sequence
    phy.loopback_tx_to_rx(n, t0=7)
    pcc = phy.rx_pcc(None, t0=7, n=n)
    ok, tb_out, _, _ = phy.rx_pdc_ex(n, network_id=ids, plcf_type=tys)
    assert ok.all() and np.array_equal(tb_out, tb)
    for i in range(n):
        t = int(tys[i]) - 1
        assert pcc[i].plcf_ok[t] and bytes(pcc[i].plcf[t])[:5 * (t + 1)] == bytes(plcf[i][:5 * (t + 1)])
    # the wrong sequence must not decode (ids with distinct c_init for both types)
    wrong = np.roll(ids, 1)
    ok_w, _, _, _ = phy.rx_pdc_ex(n, network_id=wrong, plcf_type=tys)
    assert not ok_w.any()
    # and the plain call still means "the context's network id and PLCF type"
    ok_d, _, _, _ = phy.rx_pdc(n)
    expect = (ids == NETS[0]) & (tys == 1)
    assert np.array_equal(ok_d, expect)
    phy.close()


def test_harq_reset_clears_one_slot(product):
    """rv 0 at a hopeless SNR leaves soft values in two slots; dnrp_harq_reset on one of them: a retransmission (rv 2,
    no reset flag) then decodes like a first transmission there and like a combined one in the other"""
    cfg = CONFIGS["small_b2_mcs4"]
    ps = spec.get_packet_sizes(*cfg)
    phy = product.Phy(cfg, max_batch=2, network_id=NETS[0], plcf_type=1)
    rng = np.random.default_rng(9)
    gold = fec.pdc_gold(NETS[0], 1, ps["G"])
    tb = rng.integers(0, 256, ps["N_TB_byte"], dtype=np.uint8)

    def llr(rv, sigma):
        bits = fec.unpack_bits(fec.pdc_encode(tb, ps, rv, gold), ps["G"]).astype(np.float64)
        y = (2 * bits - 1) + sigma * rng.standard_normal(ps["G"])
        return np.clip(np.round(y * 40), -500, 500).astype(np.int16)
    l0, l2 = llr(0, 1.3), llr(2, 1.3)
    first = np.stack([l0, l0])
    ok, _, _ = phy.fec_decode_tb(first, harq_slot=[0, 1], rv=[0, 0], harq_reset=[1, 1])
    phy.harq_reset(1)
    second = np.stack([l2, l2])
    ok2, out2, nit2 = phy.fec_decode_tb(second, harq_slot=[0, 1], rv=[2, 2], harq_reset=[0, 0])
    h = fec.HarqRx(ps)
    fec.pdc_decode(l0, ps, 0, gold, harq=h)
    o_comb = fec.pdc_decode(l2, ps, 2, gold, harq=h)
    o_fresh = fec.pdc_decode(l2, ps, 2, gold, harq=fec.HarqRx(ps))
    assert bool(ok2[0]) == o_comb[0] and (nit2[0] == o_comb[2]).all()
    assert bool(ok2[1]) == o_fresh[0] and (nit2[1] == o_fresh[2]).all()
    phy.close()
