"""ORACLE (test infrastructure): numpy-facing wrappers of oracle/fec.c + fec_chain.c."""
import ctypes as C
import numpy as np
from . import lib, spec

u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
i16p = np.ctypeslib.ndpointer(np.int16, flags="C_CONTIGUOUS")
u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")


class Seg(C.Structure):
    _fields_ = [(n, C.c_uint32) for n in
                ("tbs", "C", "K1", "K2", "C1", "C2", "f1_K1", "f2_K1", "f1_K2", "f2_K2")]


def make_seg(seg):
    f1a, f2a = spec.QPP_F1F2[seg["K1_idx"]]
    f1b, f2b = spec.QPP_F1F2[seg["K2_idx"]] if seg["C"] > 1 else (0, 0)
    return Seg(seg["tbs"], seg["C"], seg["K1"], seg["K2"], seg["C1"], seg["C2"], f1a, f2a, f1b, f2b)


_proto_done = False


def _L():
    global _proto_done
    l = lib()
    if not _proto_done:
        l.orc_crc.restype = C.c_uint32
        l.orc_crc.argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_uint32]
        l.orc_turbo_encode.argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_uint32, u8p]
        l.orc_rm_tx.argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_uint32, u8p]
        l.orc_rm_rx.argtypes = [i16p, C.c_uint32, C.c_uint32, C.c_uint32, i16p]
        l.orc_pdc_encode.argtypes = [u8p, C.POINTER(Seg), C.c_uint32, C.c_uint32, C.c_uint32, u8p, u8p]
        l.orc_pdc_decode.restype = C.c_int
        l.orc_pdc_decode.argtypes = [i16p, C.POINTER(Seg), C.c_uint32, C.c_uint32, C.c_uint32, u8p, i16p,
                                     u8p, u8p, u8p, u32p, C.c_uint32, C.c_uint32]
        l.orc_pcc_encode.argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, u8p, u8p]
        l.orc_pcc_decode.restype = C.c_int
        l.orc_pcc_decode.argtypes = [i16p, C.c_uint32, C.c_uint32, C.c_uint32, u8p, u8p,
                                     C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32),
                                     C.c_uint32]
        l.orc_tdec_alloc.restype = C.c_void_p
        l.orc_tdec_free.argtypes = [C.c_void_p]
        l.orc_tdec_new_cb.argtypes = [C.c_void_p, i16p, C.c_uint32, C.c_uint32, C.c_uint32]
        l.orc_tdec_iteration.argtypes = [C.c_void_p, u8p]
        l.orc_tdec_range_violations.restype = C.c_int32
        l.orc_tdec_range_violations.argtypes = [C.c_void_p]
        l.orc_tdec_window.restype = C.c_uint32
        l.orc_tdec_window.argtypes = [C.c_uint32]
        _proto_done = True
    return l


def crc(bits, poly, order):
    bits = np.ascontiguousarray(bits, dtype=np.uint8)
    return _L().orc_crc(bits, len(bits), poly, order)


def turbo_encode(c_bits):
    K = len(c_bits)
    f1, f2 = spec.QPP_F1F2[spec.TC_CB_SIZES.index(K)]
    d = np.zeros(3 * (K + 4), dtype=np.uint8)
    _L().orc_turbo_encode(np.ascontiguousarray(c_bits, dtype=np.uint8), K, f1, f2, d)
    return d


def rm_tx(d, K, E, rv):
    e = np.zeros(E, dtype=np.uint8)
    _L().orc_rm_tx(np.ascontiguousarray(d, dtype=np.uint8), K, E, rv, e)
    return e


def rm_rx(e, K, rv, soft=None):
    if soft is None:
        soft = np.zeros(3 * (K + 4), dtype=np.int16)
    _L().orc_rm_rx(np.ascontiguousarray(e, dtype=np.int16), K, len(e), rv, soft)
    return soft


def pdc_gold(network_id, plcf_type, G):
    """scrambling_pdc.cpp:43-56: c_init = id & 0xFF (type 1) or id >> 8 (type 2)."""
    c_init = (network_id & 0xFF) if plcf_type == 1 else (network_id >> 8)
    return spec.gold_sequence(c_init, G)


def pdc_encode(tb_bytes, ps, rv, gold):
    seg = make_seg(ps["seg"])
    out = np.zeros((ps["G"] + 7) // 8, dtype=np.uint8)
    _L().orc_pdc_encode(np.ascontiguousarray(tb_bytes, dtype=np.uint8), C.byref(seg), ps["N_bps"], rv, ps["G"],
                        np.ascontiguousarray(gold, dtype=np.uint8), out)
    return out


class HarqRx:
    """harq::buffer_rx_t equivalent (reference phy/harq/buffer_rx.cpp:28-66): soft buffers + CB flags."""

    def __init__(self, ps):
        seg = ps["seg"]
        self.soft = np.zeros(seg["C"] * 3 * (seg["K1"] + 4), dtype=np.int16)
        self.cb_crc = np.zeros(seg["C"], dtype=np.uint8)
        self.saved = np.zeros(seg["C"] * seg["K1"] // 8, dtype=np.uint8)


def pdc_decode(llr, ps, rv, gold, harq=None, min_it=2, max_it=10):
    """returns (crc_ok, tb_bytes(N_TB_byte), n_iter per CB).  llr is not modified."""
    seg = make_seg(ps["seg"])
    harq = harq or HarqRx(ps)
    e = np.array(llr, dtype=np.int16, copy=True)
    data = np.zeros((ps["N_TB_bits"] + 24) // 8 + ps["seg"]["K1"] // 8 + 8, dtype=np.uint8)
    n_iter = np.zeros(ps["C"], dtype=np.uint32)
    ok = _L().orc_pdc_decode(e, C.byref(seg), ps["N_bps"], rv, ps["G"], np.ascontiguousarray(gold, dtype=np.uint8),
                             harq.soft, harq.cb_crc, harq.saved, data, n_iter, min_it, max_it)
    return bool(ok), data[:ps["N_TB_byte"]].copy(), n_iter


PCC_GOLD = None


def pcc_gold():
    global PCC_GOLD
    if PCC_GOLD is None:
        PCC_GOLD = spec.gold_sequence(0x44454354, 200)
    return PCC_GOLD


def pcc_encode(plcf_bytes, plcf_type, cl, bf):
    K = (40 if plcf_type == 1 else 80) + 16
    f1, f2 = spec.QPP_F1F2[spec.TC_CB_SIZES.index(K)]
    out = np.zeros(25, dtype=np.uint8)
    _L().orc_pcc_encode(np.ascontiguousarray(plcf_bytes, dtype=np.uint8), plcf_type, int(cl), int(bf), f1, f2,
                        pcc_gold(), out)
    return out


def pcc_decode(llr196, plcf_type, max_it=5):
    K = (40 if plcf_type == 1 else 80) + 16
    f1, f2 = spec.QPP_F1F2[spec.TC_CB_SIZES.index(K)]
    out = np.zeros(10, dtype=np.uint8)
    cl, bf, nit = C.c_uint32(), C.c_uint32(), C.c_uint32()
    ok = _L().orc_pcc_decode(np.ascontiguousarray(llr196, dtype=np.int16), plcf_type, f1, f2, pcc_gold(), out,
                             C.byref(cl), C.byref(bf), C.byref(nit), max_it)
    return bool(ok), out[:5 if plcf_type == 1 else 10].copy(), bool(cl.value), bool(bf.value), nit.value


def unpack_bits(b, n):
    return np.unpackbits(np.asarray(b, dtype=np.uint8))[:n]


def set_soft_limit(N_soft, M_DL_HARQ=1):
    """limited soft buffer of the PDC chain (N_IR = N_soft / M_DL_HARQ per transport block; 0 = none, the reference's behaviour)"""
    C.c_uint32.in_dll(_L(), "orc_N_IR").value = (N_soft // max(M_DL_HARQ, 1)) if N_soft else 0


def set_window_override(L):
    C.c_uint32.in_dll(_L(), "orc_window_override").value = L
