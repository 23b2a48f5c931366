"""ORACLE (test infrastructure, not product code) -- Python front of oracle/rx.c: builds the receiver
configuration of one packet geometry from oracle/spec.py and runs the sequential restatement of
rx_synced_t::demoddecod_rx_pcc / demoddecod_rx_pdc (reference lib/src/phy/rx/rx_synced/rx_synced.cpp)
for single packets.  Parity unpinned (see rx.c header)."""
import ctypes as C

import numpy as np

from . import lib, spec

# rx_synced_param.hpp:216-232
NU_MAX_HZ = (100.0, 100.0, 500.0)
TAU_RMS_SEC = (0.1e-6, 0.1e-6, 1.0e-6)
SNR_DB = (-5.0, 15.0, 35.0)
N_INTERP_LR = (14, 8, 3)
N_INTERP_L = (7, 4, 2)


class _Cfg(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("b", "u", "N_eff_TX", "N_RX", "N_os", "cp_os", "stf_cp_os", "n_pat",
                                         "N_DF", "N_bps", "chestim_lr", "stride", "n_pdc", "literal_neff8", "N_SS")] + \
               [(n, C.c_void_p) for n in ("stf_sign", "cell_type", "cell_ts", "cell_sign", "pcc_l", "pcc_k",
                                          "pdc_l", "pdc_k")] + \
               [("lut_l", C.c_void_p * 3), ("lut_lr", C.c_void_p * 3)]


class Report(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("cfo_left", "sto_inc", "sto_fractional", "snr_db_pcc", "snr_db_pdc")] + \
               [("rms", C.c_float * 8), ("n_cls", C.c_int32), ("cls", C.c_uint8 * 512)]


_LUT_CACHE = {}


def _setup():
    L = lib()
    L.orc_wiener_build.restype = C.c_void_p
    L.orc_wiener_build.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int]
    L.orc_wiener_free.argtypes = [C.c_void_p]
    L.orc_wiener_dims.argtypes = [C.c_void_p] + [C.POINTER(C.c_int)] * 4
    for f in ("orc_wiener_idx_pilot", "orc_wiener_idx_weight", "orc_wiener_weights"):
        getattr(L, f).restype = C.c_void_p
        getattr(L, f).argtypes = [C.c_void_p]
    L.orc_rx_packet.restype = C.c_int
    L.orc_rx_packet.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_int, C.c_float, C.c_float, C.c_void_p,
                                C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    return L


def wiener_lut(b, u, N_step_virtual, cls, lr):
    """handle of the LUT of SNR class `cls` (0..2); cached for the process lifetime"""
    key = (b, u, N_step_virtual, cls, lr)
    if key not in _LUT_CACHE:
        L = _setup()
        n = (N_INTERP_LR if lr else N_INTERP_L)[cls]
        h = L.orc_wiener_build(b, u, N_step_virtual, NU_MAX_HZ[cls], TAU_RMS_SEC[cls], SNR_DB[cls], n)
        if not h:
            raise RuntimeError("Wiener LUT construction failed")
        _LUT_CACHE[key] = h
    return _LUT_CACHE[key]


def wiener_arrays(handle):
    """(idx_pilot [n_t,4,Nk], idx_weight [n_t,4,Nk], weights [n_vec,n_interp]) as numpy copies"""
    L = _setup()
    d = [C.c_int() for _ in range(4)]
    L.orc_wiener_dims(handle, *[C.byref(x) for x in d])
    n_interp, n_t, Nk, n_vec = [x.value for x in d]
    ip = np.ctypeslib.as_array(C.cast(L.orc_wiener_idx_pilot(handle), C.POINTER(C.c_uint32)), (n_t, 4, Nk)).copy()
    iw = np.ctypeslib.as_array(C.cast(L.orc_wiener_idx_weight(handle), C.POINTER(C.c_uint32)), (n_t, 4, Nk)).copy()
    w = np.ctypeslib.as_array(C.cast(L.orc_wiener_weights(handle), C.POINTER(C.c_float)), (n_vec, n_interp)).copy()
    return ip, iw, w


class RxConfig:
    """tables of one packet geometry in the layout oracle/rx.c expects"""

    def __init__(self, ps, N_RX=1, os=1, chestim_lr=True, stride=2, literal_neff8=False):
        self.ps = ps
        b, u = ps["b"], ps["u"]
        N_TS = ps["N_TS"]
        N_eff = ps["N_eff_TX"]
        N_DF = ps["N_DF_symb"]
        Nk = 56 * b + 1
        self.Nk, self.N_DF, self.N_RX, self.N_eff_TX = Nk, N_DF, N_RX, N_eff
        stf = spec.stf_filled(b, N_eff)
        self.stf_sign = np.sign(stf.real).astype(np.int8)
        self.cell_type = np.zeros((N_DF + 1, Nk), dtype=np.uint8)
        self.cell_ts = np.zeros((N_DF + 1, Nk), dtype=np.int8)
        self.cell_sign = np.zeros((N_DF + 1, Nk), dtype=np.int8)
        N_step = spec.get_N_step(N_TS)
        n_drs = spec.nof_drs_symbols_per_TS(ps["N_PACKET_symb"], N_TS)
        for t in range(N_TS):
            y = spec.drs_y_single(b, t)
            for n in range(n_drs):
                l = 1 + t // 4 + n * N_step
                k = spec.drs_k_i_single(b, t, n)
                self.cell_type[l, k] = 1
                self.cell_ts[l, k] = t
                self.cell_sign[l, k] = y.astype(np.int8)
        pl, pk = spec.lin2lk(spec.pcc_k_i_linear(b, N_TS), b)
        dl, dk = spec.lin2lk(spec.pdc_k_i_linear(b, N_TS, u, ps["N_PACKET_symb"]), b)
        self.pcc_l, self.pcc_k = pl.astype(np.int32), pk.astype(np.int32)
        self.pdc_l, self.pdc_k = dl.astype(np.int32), dk.astype(np.int32)
        self.cell_type[self.pcc_l, self.pcc_k] = 2
        self.cell_type[self.pdc_l, self.pdc_k] = 3
        self.n_drs = n_drs
        c = _Cfg()
        c.b, c.u, c.N_eff_TX, c.N_RX = b, u, N_eff, N_RX
        c.N_os, c.cp_os, c.stf_cp_os = 64 * b * os, 8 * b * os, ps["N_samples_STF_CP_only"] * os
        c.n_pat = 7 if u == 1 else 9
        c.N_DF, c.N_bps, c.chestim_lr, c.stride, c.n_pdc = N_DF, ps["N_bps"], int(chestim_lr), stride, len(dl)
        c.literal_neff8 = int(literal_neff8)  # see rx.c run_pdc_ps_in_chestim_mode_lr_t
        c.N_SS = ps["N_SS"]                   # > 1: spatial multiplexing, rx.c detect_sm (own specification)
        for name in ("stf_sign", "cell_type", "cell_ts", "cell_sign", "pcc_l", "pcc_k", "pdc_l", "pdc_k"):
            setattr(c, name, getattr(self, name).ctypes.data)
        for i in range(3):
            c.lut_l[i] = wiener_lut(b, u, 0, i, False)
            c.lut_lr[i] = wiener_lut(b, u, N_step, i, True)
        self.c = c
        self.N_os = c.N_os

    def rx_packet(self, iq, t0=0, cfo_fractional_rad=0.0, cfo_integer_rad=0.0, rms=None, want_Y=True, want_H=False):
        """iq [N_RX][n] complex64 -> dict(llr_pcc, llr_pdc, Y, H, report)"""
        L = _setup()
        iq = np.ascontiguousarray(iq, dtype=np.complex64).reshape(self.N_RX, -1)
        need = t0 + self.ps["N_samples_packet_no_GI"] * (self.N_os // (64 * self.ps["b"]))
        assert iq.shape[1] >= need, (iq.shape, need)
        llr_pcc = np.zeros(196, dtype=np.int16)
        llr_pdc = np.zeros(self.ps["G"], dtype=np.int16)
        Y = np.zeros((self.N_RX, self.N_DF, self.Nk), dtype=np.complex64) if want_Y else None
        H = np.zeros((self.N_RX, self.N_eff_TX, self.N_DF, self.Nk), dtype=np.complex64) if want_H else None
        rep = Report()
        rms_a = None if rms is None else np.ascontiguousarray(rms, dtype=np.float32)
        rc = L.orc_rx_packet(C.addressof(self.c), iq.ctypes.data, iq.shape[1], int(t0), float(cfo_fractional_rad),
                             float(cfo_integer_rad), None if rms_a is None else rms_a.ctypes.data,
                             llr_pcc.ctypes.data, llr_pdc.ctypes.data, None if Y is None else Y.ctypes.data,
                             None if H is None else H.ctypes.data, C.addressof(rep))
        if rc != 0:
            raise RuntimeError("oracle receiver: cell counters inconsistent")
        return dict(llr_pcc=llr_pcc, llr_pdc=llr_pdc, Y=Y, H=H, report=rep,
                    cls=np.array(rep.cls[:rep.n_cls], dtype=np.uint8))


def demap(z, N_bps):
    L = _setup()
    z = np.ascontiguousarray(z, dtype=np.complex64)
    out = np.zeros(z.size * N_bps, dtype=np.int16)
    L.orc_demap.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    L.orc_demap(z.ctypes.data, z.size, N_bps, out.ctypes.data)
    return out


def fft(x):
    L = _setup()
    x = np.ascontiguousarray(x, dtype=np.complex64)
    out = np.zeros_like(x)
    L.orc_fft.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    L.orc_fft(x.ctypes.data, out.ctypes.data, x.size)
    return out


def mimo_report(hzf_last, literal=True):
    """estimator_mimo_t::process_drs (estimator_mimo.cpp:80-272) on the zero-forced DRS cells of the packet's last DRS
    occurrence, hzf_last [N_RX][N_TS][NP] complex64.  Returns (tm_3_7_beamforming_idx, reciprocal_idx).
    Constants as shipped (rx_synced_param.hpp:271-286): 4 wideband cells, only matrices without zero entries
    (first index 2 for N = 2, 12 for N = 4), metric = highest minimum RX magnitude times the matrix scale."""
    from . import tx as otx
    h = np.asarray(hzf_last, dtype=np.complex64)
    N_RX, N_TS, NP = h.shape
    step = NP // 4
    cells = h[:, :, step // 2 + step * np.arange(4)]              # [rx][ts][4]

    def search(stage, n_tx, n_rx):                                 # stage [rx_virt][tx_virt][4]
        if n_tx not in (2, 4):
            return 0
        first = 2 if n_tx == 2 else 12
        n = len(otx.W_TABLES[(1, n_tx)])
        best, ret = np.float32(-1.0e6), -1
        for wm in range(first, n):
            W, scale = otx.codebook(1, n_tx, wm)
            w = W.reshape(n_tx)
            inner = np.float32(1.0e6)
            for rx in range(n_rx):
                s = np.complex64(0)
                for tx in range(n_tx):
                    part = np.complex64(0)
                    for c in range(4):
                        part = np.complex64(part + stage[rx, tx, c] * w[tx])
                    s = np.complex64(s + part)
                inner = min(inner, np.float32(abs(s)))
            inner = np.float32(inner * scale)
            if best < inner:
                best, ret = inner, wm
        return ret

    idx = 0 if N_TS == 1 else search(cells, N_TS, N_RX)
    rec = 0 if N_RX == 1 else search(np.transpose(cells, (1, 0, 2)), N_RX, N_TS)
    return idx, rec
