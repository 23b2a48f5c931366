"""ORACLE (test infrastructure, not product code) -- Python front of oracle/channel.c (simulation channel,
reference lib/src/simulation/wireless/channel_{awgn,flat}.cpp + hardware/noise.cpp)."""
import ctypes as C

import numpy as np

from . import lib


def channel(tx_iq, rx_stride, N_RX, t0=0, mode=0, snr_dB=100.0, net_bandwidth_norm=1.0, seed=1, pkt=0,
            n_samples=None):
    """tx_iq [N_TX][tx_stride] complex64 -> rx [N_RX][rx_stride]"""
    L = lib()
    tx_iq = np.ascontiguousarray(tx_iq, dtype=np.complex64)
    N_TX, tx_stride = tx_iq.shape
    n_samples = tx_stride if n_samples is None else n_samples
    rx = np.zeros((N_RX, rx_stride), dtype=np.complex64)
    L.orc_channel.argtypes = [C.c_void_p, C.c_long, C.c_int, C.c_int, C.c_void_p, C.c_long, C.c_int, C.c_int, C.c_int,
                              C.c_float, C.c_float, C.c_uint64, C.c_uint32]
    L.orc_channel.restype = None
    L.orc_channel(tx_iq.ctypes.data, tx_stride, N_TX, n_samples, rx.ctypes.data, rx_stride, N_RX, t0, mode,
                  snr_dB, net_bandwidth_norm, seed, pkt)
    return rx


def tdl_pdp(samp_rate, pdp_idx, tau_rms_ns):
    """link_t::set_pdp (link.cpp:72-146): merged tap delays in samples and normalised linear powers"""
    L = lib()
    d = np.zeros(9, dtype=np.uint32)
    p = np.zeros(9, dtype=np.float64)
    L.orc_tdl_pdp.argtypes = [C.c_double, C.c_int, C.c_float, C.c_void_p, C.c_void_p]
    n = L.orc_tdl_pdp(float(samp_rate), pdp_idx, tau_rms_ns, d.ctypes.data, p.ctypes.data)
    return d[:n].copy(), p[:n].copy()


def channel_doubly(tx_iq, rx_stride, N_RX, samp_rate, t0=0, snr_dB=100.0, net_bandwidth_norm=1.0, seed=1, pkt=0,
                   n_samples=None, pdp_idx=0, tau_rms_ns=100.0, fD_Hz=50.0, time_64=0):
    """channel_doubly_t (tapped delay line, 40 sinusoids per tap): tx_iq [N_TX][tx_stride] -> rx [N_RX][rx_stride]"""
    L = lib()
    tx_iq = np.ascontiguousarray(tx_iq, dtype=np.complex64)
    N_TX, tx_stride = tx_iq.shape
    n_samples = tx_stride if n_samples is None else n_samples
    rx = np.zeros((N_RX, rx_stride), dtype=np.complex64)
    L.orc_channel_doubly.argtypes = [C.c_void_p, C.c_long, C.c_int, C.c_int, C.c_void_p, C.c_long, C.c_int, C.c_int,
                                     C.c_float, C.c_float, C.c_uint64, C.c_uint32, C.c_double, C.c_int, C.c_float,
                                     C.c_float, C.c_int64]
    L.orc_channel_doubly.restype = None
    L.orc_channel_doubly(tx_iq.ctypes.data, tx_stride, N_TX, n_samples, rx.ctypes.data, rx_stride, N_RX, t0, snr_dB,
                         net_bandwidth_norm, seed, pkt, float(samp_rate), pdp_idx, tau_rms_ns, fD_Hz, time_64)
    return rx


def hw_clip_quantize(x, clip_limit, n_bits):
    """hw_simulator_t::clip_and_quantize on a copy of x (complex64, any shape)"""
    L = lib()
    y = np.array(x, dtype=np.complex64, copy=True, order="C")
    L.orc_hw_clip_quantize.argtypes = [C.c_void_p, C.c_long, C.c_float, C.c_uint32]
    L.orc_hw_clip_quantize.restype = None
    L.orc_hw_clip_quantize(y.ctypes.data, y.size, float(clip_limit), int(n_bits))
    return y


def large_scale_via_pathloss(tx_power_ant_0dBFS, rx_power_ant_0dBFS, distance_m, freq_Hz, same_node=False,
                             tx_into_rx_leakage_dB=0.0):
    L = lib()
    L.orc_large_scale_via_pathloss.restype = C.c_float
    L.orc_large_scale_via_pathloss.argtypes = [C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, C.c_float]
    return float(L.orc_large_scale_via_pathloss(tx_power_ant_0dBFS, rx_power_ant_0dBFS, distance_m, freq_Hz, int(same_node),
                                                tx_into_rx_leakage_dB))


def with_hardware(fn, tx_iq, *args, large_scale=1.0, clip_limit=0.0, dac_bits=0, adc_bits=0, **kw):
    """the simulated radio around a channel function (channel / channel_doubly): DAC clip + quantise on the TX samples
    (hw_simulator.cpp:560-566), large-scale gain (channel.cpp:53-73), channel + noise, ADC clip + quantise on the RX
    samples (hw_simulator.cpp:685-691)"""
    x = hw_clip_quantize(tx_iq, clip_limit, dac_bits)
    x = (x * np.float32(large_scale)).astype(np.complex64)
    return hw_clip_quantize(fn(x, *args, **kw), clip_limit, adc_bits)
