"""GPU parity of the transmitter (FEC + OFDM) against the oracle's symbol-by-symbol restatement of
tx_t::generate_tx_packet.  IQ tolerance: EVM <= -80 dB w.r.t. the oracle (fp32 FFT vs float64 FFT)."""
import numpy as np
import pytest

from oracle import spec, tx
from conftest import CONFIGS

pytestmark = pytest.mark.gpu
NET = 123456789

TX_CASES = {
    "C1_loopback": dict(cfg=CONFIGS["C1_loopback"]),
    "C2_b16_siso_mcs7": dict(cfg=CONFIGS["C2_b16_siso_mcs7"]),
    "C3_b16_txdiv4_mcs7": dict(cfg=CONFIGS["C3_b16_txdiv4_mcs7"]),
    "txdiv2_b4_u2": dict(cfg=CONFIGS["txdiv2_b4_mcs3"]),
    "txdiv8_u8_b2": dict(cfg=(8, 2, 1, 2, 10, 1, 2048)),
    "beamform_1x2_b1": dict(cfg=(1, 1, 1, 1, 3, 1, 2048), codebook_index=4),
    "beamform_1x4_b2": dict(cfg=(1, 2, 1, 1, 7, 3, 2048), codebook_index=19),
    "mimo2x2_b4_cb1": dict(cfg=(1, 4, 1, 1, 2, 4, 2048), codebook_index=1),
    "b12_siso_256qam": dict(cfg=(4, 12, 0, 3, 0, 9, 6144)),
    "bpsk_b1_os2": dict(cfg=(1, 1, 1, 2, 0, 0, 2048), os=2),
    "b8_os4_mix": dict(cfg=(1, 8, 0, 1, 0, 5, 6144), os=4, iq_phase_rad=0.3, iq_phase_increment=0.0123, DAC_scale=0.5),
    # 1024-QAM (MCS 10 / 11): outside the reference (tx.cpp:53-60), own extension of the modulation recursion
    "b2_siso_1024qam_mcs10": dict(cfg=(1, 2, 1, 2, 0, 10, 2048)),
    "txdiv2_b1_1024qam_mcs11": dict(cfg=(1, 1, 1, 3, 1, 11, 2048)),
    "mimo4x4_b2_cb3": dict(cfg=(1, 2, 1, 2, 9, 3, 2048), codebook_index=3),
    "mimo8x8_b1": dict(cfg=(1, 1, 1, 6, 11, 2, 2048)),
    # FFT sizes above 4096 (tx_rx.hpp:66-69: 6144, 8192, 12288, 16384): four-step in-place transforms
    "b12_os8_fft6144": dict(cfg=(1, 12, 0, 2, 0, 2, 2048), os=8),
    "b16_os8_fft8192_txdiv2": dict(cfg=(1, 16, 1, 1, 1, 3, 2048), os=8),
    "b12_os16_fft12288": dict(cfg=(1, 12, 0, 2, 0, 1, 2048), os=16),
    "b16_os16_fft16384": dict(cfg=(1, 16, 1, 1, 0, 4, 6144), os=16, iq_phase_rad=0.1, iq_phase_increment=0.0007),
}


def evm_db(a, ref):
    return 10 * np.log10((np.abs(a - ref) ** 2).sum() / (np.abs(ref) ** 2).sum() + 1e-30)


@pytest.mark.parametrize("name", list(TX_CASES))
def test_tx_iq_matches_oracle(product, name):
    case = dict(TX_CASES[name])
    cfg = case.pop("cfg")
    os_ = case.pop("os", 1)
    ps = spec.get_packet_sizes(*cfg)
    assert ps is not None
    n = 3
    phy = product.Phy(cfg, max_batch=n, network_id=NET, plcf_type=2 if "mimo" in name else 1, oversampling=os_)
    rng = np.random.default_rng(21)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    rv = np.array([0, 2, 3])
    iq, ns = phy.tx(plcf, tb, rv=rv, GI_percentage=np.array([5, 50, 100]), **case)
    for i in range(n):
        ref = tx.tx_packet(ps, plcf[i], tb[i], network_id=NET, plcf_type=phy.plcf_type, rv=int(rv[i]), os=os_,
                           GI_percentage=[5, 50, 100][i], **case)
        assert ns[i] == ref.shape[1]
        got = iq[i, :, :ns[i]]
        assert evm_db(got, ref) < -80.0, (name, i, evm_db(got, ref))
        assert not iq[i, :, ns[i]:].any()
    phy.close()


def test_tx_batch_sc16_is_the_rounded_float_output(product):
    """dnrp_tx_batch_sc16: the same packets as dnrp_tx_batch, every sample round-to-nearest(sample x scale) saturated to
    int16 (what UHD's fc32 -> sc16 conversion does on the host in the reference, hw_usrp.cpp:422-424)"""
    cfg = (1, 4, 1, 2, 5, 3, 2048)       # four transmit antennas
    ps = spec.get_packet_sizes(*cfg)
    n = 5
    phy = product.Phy(cfg, max_batch=n, network_id=123456789, plcf_type=1)
    rng = np.random.default_rng(3)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    iq, ns = phy.tx(plcf, tb)
    for scale in (32767.0, 2.0e5):     # the second one saturates the peaks
        iq16, ns16 = phy.tx(plcf, tb, sc16_scale=scale)
        assert np.array_equal(ns, ns16)
        want = np.clip(np.rint(iq.view(np.float32).reshape(iq16.shape).astype(np.float32) * np.float32(scale)), -32768, 32767)
        d = np.abs(iq16.astype(np.int32) - want.astype(np.int32))
        assert d.max() <= 1 and (d > 0).mean() < 1e-3      # ties and the product's rounding in float32
        assert scale < 1e5 or (np.abs(iq16) == 32767).any() or (iq16 == -32768).any()
    phy.close()
