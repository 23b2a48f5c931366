"""GPU parity of the simulation channel (dnrp_channel_batch) against oracle/channel.c, plus the statistics
the reference's channel guarantees (noise power = 10^(n0_dBFS/10), unit-power flat coefficients).
Tolerance: same Philox stream on both sides, so samples agree to fp32 rounding of logf/sincos:
max |diff| <= 2e-5 * (sigma + max|x|)."""
import numpy as np
import pytest

from oracle import channel as ochan, spec
from conftest import CONFIGS

pytestmark = pytest.mark.gpu
NET = 123456789


@pytest.mark.parametrize("name,N_RX,mode,snr", [("C1_loopback", 1, 0, 20.0), ("txdiv2_b4_mcs3", 2, 1, 7.0),
                                                 ("C3_b16_txdiv4_mcs7", 4, 1, 25.0)])
def test_channel_matches_oracle(product, name, N_RX, mode, snr):
    cfg = CONFIGS[name]
    ps = spec.get_packet_sizes(*cfg)
    n = 3
    phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=NET)
    rng = np.random.default_rng(3)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    iq_tx, ns = phy.tx(plcf, tb)
    t0 = np.array([0, 7, 33])
    seed = 0x1234567811223344
    rx = phy.channel(n, snr_dB=snr, t0=t0, mode=mode, seed=seed, fetch=True)
    sigma = 10 ** (-snr / 20)
    for i in range(n):
        ref = ochan.channel(iq_tx[i], phy.rx_stride(), N_RX, t0=int(t0[i]), mode=mode, snr_dB=snr, seed=seed, pkt=i,
                            n_samples=ps["N_samples_packet"])
        tol = 2e-5 * (sigma + np.abs(iq_tx[i]).max() * ps["N_TX"])
        assert np.abs(rx[i] - ref).max() <= tol, (name, i, float(np.abs(rx[i] - ref).max()), tol)
    phy.close()


def test_channel_statistics(product):
    """noise only windows: E|n|^2 = 10^(-snr/10) / net_bandwidth_norm; flat coefficients: E|g|^2 = 1"""
    cfg = CONFIGS["C1_loopback"]
    ps = spec.get_packet_sizes(*cfg)
    n = 512
    phy = product.Phy(cfg, max_batch=n, N_RX=2, network_id=NET, oversampling=2)
    tb = np.zeros((n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = np.zeros((n, 10), dtype=np.uint8)
    iq_tx, ns = phy.tx(plcf, tb, fetch=True)
    rx = phy.channel(n, snr_dB=10.0, t0=0, mode=1, seed=99, fetch=True)
    tail = rx[:, :, ps["N_samples_packet"] * 2 + 8:]          # behind the packet: noise only
    p_noise = float(np.mean(np.abs(tail) ** 2))
    assert abs(p_noise / (10 ** -1.0 * 2) - 1.0) < 0.01, p_noise
    assert abs(float(np.mean(tail.real * tail.imag))) < 2e-3 * p_noise * 50
    # link coefficients from the noise free part: y = g x + n  ->  g ~ <y, x> / <x, x>
    L = ps["N_samples_packet_no_GI"] * 2
    x = iq_tx[:, 0, :L]
    g = np.einsum("pal,pl->pa", rx[:, :, :L], np.conj(x)) / np.einsum("pl,pl->p", x, np.conj(x))[:, None]
    assert abs(float(np.mean(np.abs(g) ** 2)) - 1.0) < 0.12
    assert abs(np.mean(g)) < 0.12
    phy.close()


@pytest.mark.parametrize("name,N_RX,pdp,tau,fD,os", [("C1_loopback", 1, 0, 43.13, 0.0, 1),
                                                      ("txdiv2_b4_mcs3", 2, 1, 300.0, 200.0, 1),
                                                      ("C3_b16_txdiv4_mcs7", 2, 2, 500.0, 2000.0, 1),
                                                      ("C1_loopback", 2, 1, 2000.0, 50.0, 2)])
def test_channel_doubly_matches_oracle(product, name, N_RX, pdp, tau, fD, os):
    """doubly selective channel (tapped delay line x 40 sinusoids): same Philox draws on both sides; the device
    advances phasors by fixed rotations inside a block, the oracle evaluates every sample's phase in closed form:
    tolerance 1e-4 of the signal peak"""
    cfg = CONFIGS[name]
    ps = spec.get_packet_sizes(*cfg)
    n = 2
    phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=NET, oversampling=os)
    rng = np.random.default_rng(4)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    iq_tx, ns = phy.tx(plcf, tb)
    t0 = np.array([5, 77])
    seed, time_64 = 0xABCDEF0123456789, 987654321012
    snr = 30.0
    rx = phy.channel(n, snr_dB=snr, t0=t0, mode=2, seed=seed, fetch=True, pdp_idx=pdp, tau_rms_ns=tau, fD_Hz=fD,
                     time_64=time_64)
    fs = cfg[0] * cfg[1] * os * 1.728e6
    for i in range(n):
        ref = ochan.channel_doubly(iq_tx[i], phy.rx_stride(), N_RX, fs, t0=int(t0[i]), snr_dB=snr,
                                   net_bandwidth_norm=1.0 / os, seed=seed, pkt=i, n_samples=ps["N_samples_packet"] * os,
                                   pdp_idx=pdp, tau_rms_ns=tau, fD_Hz=fD, time_64=time_64)
        tol = 1e-4 * np.abs(ref).max()
        assert np.abs(rx[i] - ref).max() <= tol, (name, i, float(np.abs(rx[i] - ref).max()), tol)
    # mixed batch: packet 0 awgn, packet 1 doubly selective
    mode = np.array([0, 2])
    rx2 = phy.channel(n, snr_dB=snr, t0=t0, mode=mode, seed=seed, fetch=True, pdp_idx=pdp, tau_rms_ns=tau, fD_Hz=fD,
                      time_64=time_64)
    ref0 = ochan.channel(iq_tx[0], phy.rx_stride(), N_RX, t0=int(t0[0]), mode=0, snr_dB=snr, seed=seed, pkt=0,
                         net_bandwidth_norm=1.0 / os, n_samples=ps["N_samples_packet"] * os)
    assert np.abs(rx2[0] - ref0).max() <= 2e-5 * (np.abs(ref0).max() + 1)
    assert np.array_equal(rx2[1], rx[1])
    phy.close()


def test_channel_doubly_decodes(product):
    """a mildly dispersive, slowly fading link: sync + receiver still decode most packets"""
    cfg = CONFIGS["C1_loopback"]
    ps = spec.get_packet_sizes(*cfg)
    n = 64
    phy = product.Phy(cfg, max_batch=n, N_RX=2, network_id=NET)
    rng = np.random.default_rng(6)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = np.zeros((n, 10), dtype=np.uint8)
    phy.tx(plcf, tb, fetch=False)
    t0 = rng.integers(40, 200, n)
    phy.channel(n, snr_dB=25.0, t0=t0, mode=2, seed=11, pdp_idx=0, tau_rms_ns=100.0, fD_Hz=20.0, time_64=5000)
    res = phy.sync(n=n)
    phy.rx_pcc(None, n=n, sync="device")
    ok, tb_out, _, _ = phy.rx_pdc(n)
    assert res.detected.sum() >= n * 3 // 4
    assert ok.sum() >= n // 2 and (tb_out[ok] == tb[ok]).all()
    phy.close()


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_large_scale_clip_and_quantise_match_oracle(product, mode):
    """row a38 of SURVEY.md section 8: large-scale gain from the path loss (channel.cpp:53-73, pathloss.cpp:28-38) and the
    simulated radio's converters (hw_simulator.cpp:560-566, 685-691, 710-731: symmetric clipping at 1.0, DAC / ADC
    quantisation).  The ADC step is 2 / 2^12: a sample may fall on the other side of a rounding boundary because of fp32
    rounding in front of it -> at most one step, on fewer than 0.1 % of the samples"""
    cfg = CONFIGS["txdiv2_b4_mcs3"]
    ps = spec.get_packet_sizes(*cfg)
    n, N_RX = 2, 2
    phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=NET)
    rng = np.random.default_rng(6)
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    iq_tx, ns = phy.tx(plcf, tb, DAC_scale=2.5)        # hot enough that the DAC clips
    assert (np.abs(iq_tx.real) > 1.0).mean() > 0.01
    ls = product.large_scale_via_pathloss(10.0, -30.0, 25.0, 1.9e9)
    assert abs(ls - ochan.large_scale_via_pathloss(10.0, -30.0, 25.0, 1.9e9)) < 1e-6 * ls
    assert abs(20 * np.log10(ls) - (10.0 + 30.0 - (20 * np.log10(25.0) + 20 * np.log10(1.9e9) - 147.55))) < 1e-3
    leak = product.large_scale_via_pathloss(10.0, -30.0, 0.0, 0.0, same_node=True, tx_into_rx_leakage_dB=60.0)
    assert abs(leak - ochan.large_scale_via_pathloss(10.0, -30.0, 0.0, 0.0, True, 60.0)) < 1e-6 * leak
    t0, seed, snr = 11, 0x55AA55AA12345678, 35.0
    kw = dict(pdp_idx=1, tau_rms_ns=200.0, fD_Hz=100.0, time_64=123456) if mode == 2 else {}
    rx = phy.channel(n, snr_dB=snr, t0=t0, mode=mode, seed=seed, fetch=True, large_scale=ls, clip_limit=1.0, dac_bits=14,
                     adc_bits=12, **kw)
    step = 2.0 / 2 ** 12
    fs = cfg[0] * cfg[1] * 1.728e6
    for i in range(n):
        if mode == 2:
            ref = ochan.with_hardware(ochan.channel_doubly, iq_tx[i], phy.rx_stride(), N_RX, fs, t0=t0, snr_dB=snr, seed=seed,
                                      pkt=i, n_samples=ps["N_samples_packet"], large_scale=ls, clip_limit=1.0, dac_bits=14,
                                      adc_bits=12, **kw)
        else:
            ref = ochan.with_hardware(ochan.channel, iq_tx[i], phy.rx_stride(), N_RX, t0=t0, mode=mode, snr_dB=snr, seed=seed,
                                      pkt=i, n_samples=ps["N_samples_packet"], large_scale=ls, clip_limit=1.0, dac_bits=14,
                                      adc_bits=12)
        # every sample sits on the ADC grid and inside the clipping range
        q = rx[i].view(np.float32) / np.float32(step)
        assert np.array_equal(q, np.round(q)) and np.abs(rx[i].view(np.float32)).max() <= 1.0
        d = np.abs(rx[i].view(np.float32) - ref.view(np.float32))
        assert d.max() <= step * 1.001 and (d > 0).mean() < 1e-3, (mode, i, float(d.max()), float((d > 0).mean()))
    # the fields left at 0 change nothing
    a = phy.channel(n, snr_dB=snr, t0=t0, mode=mode, seed=seed, fetch=True, **kw)
    b = phy.channel(n, snr_dB=snr, t0=t0, mode=mode, seed=seed, fetch=True, large_scale=1.0, **kw)
    assert np.array_equal(a, b)
    phy.close()


_TDL_SCRIPT = r"""
import sys, numpy as np
sys.path.insert(0, {root!r}); sys.path.insert(0, {root!r} + "/tests")
from conftest import load_product, CONFIGS
from oracle import spec
product = load_product()
cfg = CONFIGS["C3_b16_txdiv4_mcs7"]
ps = spec.get_packet_sizes(*cfg)
n = 2
phy = product.Phy(cfg, max_batch=n, N_RX=4, network_id=123456789)
rng = np.random.default_rng(21)
tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
phy.tx(plcf, tb, fetch=False)
out = []
for fD in (50.0, 400.0, 2000.0):
    out.append(phy.channel(n, snr_dB=60.0, t0=np.array([5, 300]), mode=2, seed=77, fetch=True, pdp_idx=1, tau_rms_ns=100.0,
                           fD_Hz=fD, time_64=987654321))
np.save(sys.argv[1], np.stack(out))
"""


def test_tdl_interpolated_coefficients_equal_the_exact_evaluation(tmp_path):
    """k_channel_tdl evaluates the tap coefficients at knots and interpolates where the Doppler is slow (knot distance
    from dnrp_channel_batch: 128 samples at 50 Hz / 27.648 MS/s, 16 at 400 Hz, 0 = every sample at 2000 Hz); DNRP_TDL_EXACT=1
    forces the exact path (read once per process: two subprocesses).  Bound of the design: 2e-6 of the coefficient's rms"""
    import os
    import subprocess
    import sys
    from conftest import ROOT
    script = tmp_path / "tdl.py"
    script.write_text(_TDL_SCRIPT.format(root=ROOT))
    res = {}
    for tag, env in (("interp", {}), ("exact", {"DNRP_TDL_EXACT": "1"})):
        e = dict(os.environ)
        e.pop("DNRP_TDL_EXACT", None)
        e.update(env)
        f = str(tmp_path / (tag + ".npy"))
        subprocess.run([sys.executable, str(script), f], check=True, env=e, timeout=600)
        res[tag] = np.load(f)
    for i, fD in enumerate((50.0, 400.0, 2000.0)):
        a, b = res["interp"][i], res["exact"][i]
        err = np.abs(a - b).max() / np.abs(b).max()
        assert err <= 2e-5, (fD, float(err))
    assert np.array_equal(res["interp"][2], res["exact"][2])   # 2000 Hz: both runs take the exact path
    assert not np.array_equal(res["interp"][0], res["exact"][0])   # 50 Hz: the interpolated path did run
