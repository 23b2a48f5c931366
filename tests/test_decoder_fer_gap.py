"""The product's turbo decoder is windowed (P = K / L windows decoded in parallel with training passes);
the reference delegates to srsRAN's full-trellis decoder.  Bit-exact parity between CUDA and oracle/fec.c says nothing
about a systematic FER loss of the windowed formulation, so this test bounds it: at the waterfall (about 40 % of the
packets decode) the windowed decoder may lose at most 5 % of the packets against the same arithmetic run over ONE
undivided window (the oracle's test hook orc_window_override: no window borders, no training, exact start and
termination states = a full-trellis max-log-MAP), for a large (K = 5504) and a small (K = 480) code block.
Measured when written (K = 5504, 300 packets each): 31 / 31 at 3.9 dB, 102 / 109 at 4.0 dB, 192 / 197 at 4.1 dB, 273 / 271 at
4.2 dB (windowed / full): the waterfall moves by about 0.01 dB."""
import numpy as np
import pytest

from oracle import fec, spec

NET = 123456789
CASES = [((1, 4, 1, 2, 0, 2, 6144), 4.0), ((1, 1, 1, 1, 0, 2, 2048), 4.0)]   # single code block, K = 5504 / 480; QPSK R = 3/4


def _llrs(ps, snr_db, n, seed):
    gold = fec.pdc_gold(NET, 1, ps["G"])
    rng = np.random.default_rng(seed)
    tbs = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    llr = np.zeros((n, ps["G"]), dtype=np.int16)
    sigma = 10 ** (-snr_db / 20)
    for i in range(n):
        bits = fec.unpack_bits(fec.pdc_encode(tbs[i], ps, 0, gold), ps["G"]).astype(np.float64)
        y = (2 * bits - 1) + sigma * rng.standard_normal(bits.shape)
        llr[i] = np.clip(np.round(y * 2 / sigma ** 2 * 8), -1000, 1000).astype(np.int16)
    return tbs, llr, gold


def _count(ps, tbs, llr, gold, window):
    fec.set_window_override(window)
    try:
        return sum(int(ok and np.array_equal(out, tb))
                   for tb, (ok, out, _) in ((tbs[i], fec.pdc_decode(llr[i], ps, 0, gold)) for i in range(len(tbs))))
    finally:
        fec.set_window_override(0)


@pytest.mark.parametrize("cfg,snr_db", CASES)
def test_windowed_decoder_fer_close_to_full_trellis_cpu(cfg, snr_db):
    ps = spec.get_packet_sizes(*cfg)
    assert ps["C"] == 1
    n = 300
    tbs, llr, gold = _llrs(ps, snr_db, n, seed=7)
    n_win = _count(ps, tbs, llr, gold, 0)
    n_full = _count(ps, tbs, llr, gold, 1 << 20)
    assert 0.1 * n < n_full < 0.9 * n, n_full          # the operating point is the waterfall
    assert n_win >= n_full - 0.05 * n, (n_win, n_full)


@pytest.mark.gpu
@pytest.mark.parametrize("cfg,snr_db", CASES)
def test_windowed_decoder_fer_close_to_full_trellis_gpu(product, cfg, snr_db):
    ps = spec.get_packet_sizes(*cfg)
    n = 300
    tbs, llr, gold = _llrs(ps, snr_db, n, seed=7)
    phy = product.Phy(cfg, max_batch=n, network_id=NET, plcf_type=1)
    ok, out, nit = phy.fec_decode_tb(llr)
    n_gpu = int(sum(bool(ok[i]) and np.array_equal(out[i], tbs[i]) for i in range(n)))
    n_full = _count(ps, tbs, llr, gold, 1 << 20)
    assert n_gpu == _count(ps, tbs, llr, gold, 0)      # the contract: CUDA == windowed oracle, packet by packet
    assert n_gpu >= n_full - 0.05 * n, (n_gpu, n_full)
    phy.close()
