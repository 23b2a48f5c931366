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
