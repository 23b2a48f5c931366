"""dect-nr-plus-sdr_b200 -- B200 (sm_100a) implementation of the DECT NR+ PCC/PDC PHY hot path.

This Python module is only a thin ctypes binding over the C ABI in include/dnrp_c_api.h
(libdnrp_cuda.so); the product is the CUDA library.  There is no CPU fallback: if the library or a GPU
is missing, calls raise.  The directory name is not a valid Python identifier, load it with

    import importlib.util, sys
    spec = importlib.util.spec_from_file_location("dectnrp_b200", ".../dect-nr-plus-sdr_b200/__init__.py")

or use `tests/conftest.py::load_product()`.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdnrp_cuda.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "dnrp_c_api.h")


class DnrpError(RuntimeError):
    pass


class Psdef(C.Structure):
    _fields_ = [(n, C.c_uint32) for n in
                ("u", "b", "PacketLengthType", "PacketLength", "tm_mode_index", "mcs_index", "Z")]


class PacketSizes(C.Structure):
    _fields_ = [("psdef", Psdef)] + [(n, C.c_uint32) for n in (
        "N_eff_TX", "N_SS", "N_TS", "N_TX", "closed_loop", "N_bps", "N_b_DFT", "N_b_CP", "N_b_OCC",
        "N_PACKET_symb", "N_DF_symb", "N_PDC_subc", "N_DRS_subc", "G", "N_PDC_bits", "N_TB_bits",
        "N_TB_byte", "C", "K_plus", "K_minus", "C_plus", "C_minus", "N_samples_STF",
        "N_samples_STF_CP_only", "N_samples_DF", "N_samples_GI", "N_samples_packet_no_GI",
        "N_samples_packet")]

    def as_dict(self):
        d = {n: getattr(self, n) for n, _ in self._fields_ if n != "psdef"}
        d.update({n: getattr(self.psdef, n) for n, _ in Psdef._fields_})
        return d


class Config(C.Structure):
    _fields_ = [("psdef", Psdef)] + [(n, C.c_uint32) for n in (
        "N_RX", "oversampling", "resample_L", "resample_M", "max_batch", "network_id", "plcf_type",
        "chestim_mode_lr", "chestim_mode_lr_t_stride", "tdec_min_iter", "tdec_max_iter", "N_soft", "M_DL_HARQ")]


class TxDesc(C.Structure):
    _fields_ = [("codebook_index", C.c_uint32), ("optimal_scaling_DAC", C.c_uint32), ("DAC_scale", C.c_float),
                ("iq_phase_rad", C.c_float), ("iq_phase_increment_s2s_post_resampling_rad", C.c_float),
                ("GI_percentage", C.c_uint32), ("rv", C.c_uint32)]


class SyncReport(C.Structure):
    _fields_ = [("fine_peak_time_local", C.c_uint32), ("cfo_fractional_rad", C.c_float),
                ("cfo_integer_rad", C.c_float), ("rms", C.c_float * 8)]


class PccReport(C.Structure):
    _fields_ = [("plcf_ok", C.c_uint8 * 2), ("plcf", (C.c_uint8 * 10) * 2), ("crc_mask", C.c_uint32 * 2),
                ("snr_dB", C.c_float), ("cfo_residual_rad", C.c_float), ("sto_fractional", C.c_float),
                ("rms", C.c_float * 8)]


class PdcReport(C.Structure):
    _fields_ = [("crc_ok", C.c_uint32), ("snr_dB", C.c_float), ("mimo_N_RX", C.c_uint32),
                ("mimo_N_TS_other", C.c_uint32), ("tm_3_7_beamforming_idx", C.c_uint32),
                ("tm_3_7_beamforming_reciprocal_idx", C.c_uint32)]


# numpy views of the same C structs (batched calls fill / read them without a Python loop)
SYNC_DTYPE = np.dtype([("fine_peak_time_local", "u4"), ("cfo_fractional_rad", "f4"), ("cfo_integer_rad", "f4"),
                       ("rms", "f4", (8,))], align=True)
PCC_DTYPE = np.dtype([("plcf_ok", "u1", (2,)), ("plcf", "u1", (2, 10)), ("crc_mask", "u4", (2,)), ("snr_dB", "f4"),
                      ("cfo_residual_rad", "f4"), ("sto_fractional", "f4"), ("rms", "f4", (8,))], align=True)
PDC_DTYPE = np.dtype([("crc_ok", "u4"), ("snr_dB", "f4"), ("mimo_N_RX", "u4"), ("mimo_N_TS_other", "u4"),
                      ("tm_3_7_beamforming_idx", "u4"), ("tm_3_7_beamforming_reciprocal_idx", "u4")], align=True)
assert SYNC_DTYPE.itemsize == C.sizeof(SyncReport) and PCC_DTYPE.itemsize == C.sizeof(PccReport)
assert PDC_DTYPE.itemsize == C.sizeof(PdcReport)


class ChannelDesc(C.Structure):
    _fields_ = [("snr_dB", C.c_float), ("net_bandwidth_norm", C.c_float), ("t0", C.c_uint32), ("mode", C.c_uint32),
                ("seed", C.c_uint64), ("pdp_idx", C.c_uint32), ("tau_rms_ns", C.c_float), ("fD_Hz", C.c_float),
                ("time_64", C.c_int64), ("large_scale", C.c_float), ("clip_limit", C.c_float), ("dac_bits", C.c_uint32),
                ("adc_bits", C.c_uint32)]


CHANNEL_DTYPE = np.dtype([("snr_dB", "f4"), ("net_bandwidth_norm", "f4"), ("t0", "u4"), ("mode", "u4"),
                          ("seed", "u8"), ("pdp_idx", "u4"), ("tau_rms_ns", "f4"), ("fD_Hz", "f4"), ("time_64", "i8"),
                          ("large_scale", "f4"), ("clip_limit", "f4"), ("dac_bits", "u4"), ("adc_bits", "u4")], align=True)
assert CHANNEL_DTYPE.itemsize == C.sizeof(ChannelDesc)
TAP_Y, TAP_DRS_ZF, TAP_STF_ZF, TAP_LLR_PCC, TAP_LLR_PDC, TAP_WIENER_CLASS, TAP_TB_OK, TAP_CHESTIM = 1, 2, 3, 4, 5, 6, 7, 8
KERNEL_TURBO_DECODE_PDC, KERNEL_RX_FFT, KERNEL_RX_DEMAP_PDC, KERNEL_RX_STF, KERNEL_DERM_PDC, KERNEL_SYNC, \
    KERNEL_RESAMPLE = range(7)
# dnrp_sync_result_t
SYNC_RESULT_DTYPE = np.dtype([("detected", "u4"), ("detection_time_local", "u4"), ("detection_ant_idx", "u4"),
                              ("detection_metric", "f4"), ("detection_rms", "f4"), ("coarse_peak_time_local", "u4"),
                              ("coarse_peak_array", "f4", (8,)), ("rms_array", "f4", (8,)),
                              ("cfo_fractional_rad", "f4"), ("cfo_integer_rad", "f4"), ("N_eff_TX", "u4"),
                              ("fine_peak_time_local", "u4"), ("n_detections", "u4"), ("strongest_ant", "u4"),
                              ("coarse_peak_time_hw", "u4")], align=True)
assert SYNC_RESULT_DTYPE.itemsize == 116
# dnrp_stream_packet_t
STREAM_PACKET_DTYPE = np.dtype({"names": ["fine_peak_time_64", "coarse_peak_time_64", "sync", "chunk", "round"],
                                "formats": ["i8", "i8", SYNC_RESULT_DTYPE, "u4", "u4"],
                                "offsets": [0, 8, 16, 132, 136], "itemsize": 144})


def build(verbose=False):
    """compile libdnrp_cuda.so in-tree for sm_100a (nvcc cross-compiles without a GPU)"""
    subprocess.check_call(["make", "-C", _HERE, "-j8", "libdnrp_cuda.so"],
                          stdout=None if verbose else subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise DnrpError("libdnrp_cuda.so is not built (run __graft_entry__.build()); no CPU fallback exists")
        _lib = C.CDLL(LIB_PATH)
        _lib.dnrp_last_error_string.restype = C.c_char_p
    return _lib


def _check(rc):
    if rc != 0:
        raise DnrpError("dnrp error %d: %s" % (rc, lib().dnrp_last_error_string().decode()))


def get_packet_sizes(u, b, PacketLengthType, PacketLength, tm_mode_index, mcs_index, Z):
    """sp3::get_packet_sizes (host only); returns dict or None if the configuration is impossible"""
    ps = PacketSizes()
    rc = lib().dnrp_get_packet_sizes(C.byref(Psdef(u, b, PacketLengthType, PacketLength, tm_mode_index, mcs_index, Z)),
                                     C.byref(ps))
    return ps.as_dict() if rc == 0 else None


def _ptr(a, dtype):
    if a is None:
        return None
    assert a.dtype == dtype and a.flags["C_CONTIGUOUS"], (a.dtype, dtype)
    return a.ctypes.data_as(C.c_void_p)


def large_scale_via_pathloss(tx_power_ant_0dBFS, rx_power_ant_0dBFS, distance_m, freq_Hz, same_node=False,
                             tx_into_rx_leakage_dB=0.0):
    """dnrp_large_scale_via_pathloss: channel_t::get_large_scale_via_pathloss (channel.cpp:53-73)"""
    f = lib().dnrp_large_scale_via_pathloss
    f.restype = C.c_float
    f.argtypes = [C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, C.c_float]
    return float(f(tx_power_ant_0dBFS, rx_power_ant_0dBFS, distance_m, freq_Hz, int(same_node), tx_into_rx_leakage_dB))


def _nccl_of_torch_first():
    # the library binds NCCL with dlopen("libnccl.so.2"); the loader keeps ONE library per soname, so the copy PyTorch
    # ships must be mapped before ours asks -- otherwise a later `import torch` would find the system's (older) NCCL
    # under that name and fail to resolve its symbols
    try:
        import torch  # noqa: F401
    except ImportError:
        pass


def comm_unique_id():
    """dnrp_comm_unique_id: the 128 bytes rank 0 hands to every other rank before comm_init_rank"""
    _nccl_of_torch_first()
    buf = C.create_string_buffer(128)
    _check(lib().dnrp_comm_unique_id(buf))
    return buf.raw


def comm_init_all(phys):
    """dnrp_comm_init_all: one communicator over several contexts (one per GPU) of this process"""
    _nccl_of_torch_first()
    arr = (C.c_void_p * len(phys))(*[p._ctx for p in phys])
    _check(lib().dnrp_comm_init_all(arr, len(phys)))


def counters_allreduce_all(phys, counters, op="sum"):
    """dnrp_counters_allreduce_all: counters [len(phys)][n], row i belongs to phys[i]; every row holds the result"""
    a = np.ascontiguousarray(counters, dtype=np.int64).copy()
    assert a.ndim == 2 and a.shape[0] == len(phys)
    arr = (C.c_void_p * len(phys))(*[p._ctx for p in phys])
    _check(lib().dnrp_counters_allreduce_all(arr, len(phys), _ptr(a, np.int64), a.shape[1], {"sum": 0, "min": 1, "max": 2}[op]))
    return a


class Phy:
    """One context = one GPU + one packet geometry + a maximum batch size."""

    def __init__(self, psdef, max_batch, N_RX=1, network_id=123456789, plcf_type=1, device=0, oversampling=1,
                 resample=(1, 1), chestim_mode_lr=True, chestim_stride=2, tdec_iters=(2, 10), N_soft=0, M_DL_HARQ=1):
        self._ctx = C.c_void_p()
        cfg = Config(Psdef(*psdef), N_RX, oversampling, resample[0], resample[1], max_batch, network_id, plcf_type,
                     int(chestim_mode_lr), chestim_stride, tdec_iters[0], tdec_iters[1], N_soft, M_DL_HARQ)
        _check(lib().dnrp_ctx_create(C.byref(cfg), device, C.byref(self._ctx)))
        ps = PacketSizes()
        _check(lib().dnrp_ctx_get_packet_sizes(self._ctx, C.byref(ps)))
        self.ps = ps.as_dict()
        self.max_batch = max_batch
        self.plcf_type = plcf_type
        self.N_RX = N_RX
        self.oversampling = oversampling
        self.resample = tuple(resample)
        self.dbits_stride = ((self.ps["G"] + 7) // 8 + 3) & ~3

    # ---- multi-GPU counter reduction (api_comm.cu; one NCCL all-reduce over the counter vector)
    def comm_init_rank(self, n_ranks, rank, unique_id):
        """dnrp_comm_init_rank: collective over all ranks; unique_id = comm_unique_id() of rank 0 (128 bytes)"""
        assert len(unique_id) == 128
        _nccl_of_torch_first()
        _check(lib().dnrp_comm_init_rank(self._ctx, int(n_ranks), int(rank), C.c_char_p(bytes(unique_id))))

    def counters_allreduce(self, counters, op="sum"):
        """dnrp_counters_allreduce: int64 vector reduced in place over all ranks (op: sum, min, max)"""
        a = np.ascontiguousarray(counters, dtype=np.int64).reshape(-1).copy()
        _nccl_of_torch_first()
        _check(lib().dnrp_counters_allreduce(self._ctx, _ptr(a, np.int64), a.size, {"sum": 0, "min": 1, "max": 2}[op]))
        return a.reshape(np.shape(counters))

    def close(self):
        if self._ctx:
            lib().dnrp_ctx_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- fec_t (reference lib/include/dectnrp/phy/fec/fec.hpp:84-166)
    def fec_encode(self, plcf, tb, cl_bf_mask=None, rv=None):
        n = tb.shape[0]
        plcf = np.ascontiguousarray(plcf, dtype=np.uint8).reshape(n, 10)
        tb = np.ascontiguousarray(tb, dtype=np.uint8).reshape(n, self.ps["N_TB_byte"])
        pcc = np.zeros((n, 28), dtype=np.uint8)
        pdc = np.zeros((n, self.dbits_stride), dtype=np.uint8)
        mask = None if cl_bf_mask is None else np.ascontiguousarray(cl_bf_mask, dtype=np.uint32)
        rvv = None if rv is None else np.ascontiguousarray(rv, dtype=np.uint8)
        _check(lib().dnrp_fec_encode(self._ctx, n, _ptr(plcf, np.uint8), _ptr(mask, np.uint32), _ptr(tb, np.uint8),
                                     _ptr(rvv, np.uint8), _ptr(pcc, np.uint8), _ptr(pdc, np.uint8)))
        return pcc, pdc

    def fec_decode_tb(self, llr, harq_slot=None, rv=None, harq_reset=None):
        n = llr.shape[0]
        llr = np.ascontiguousarray(llr, dtype=np.int16).reshape(n, self.ps["G"])
        tb = np.zeros((n, self.ps["N_TB_byte"]), dtype=np.uint8)
        ok = np.zeros(n, dtype=np.uint8)
        nit = np.zeros((n, self.ps["C"]), dtype=np.uint8)
        slot = None if harq_slot is None else np.ascontiguousarray(harq_slot, dtype=np.uint32)
        rvv = None if rv is None else np.ascontiguousarray(rv, dtype=np.uint8)
        rst = None if harq_reset is None else np.ascontiguousarray(harq_reset, dtype=np.uint8)
        _check(lib().dnrp_fec_decode_tb(self._ctx, n, _ptr(llr, np.int16), _ptr(slot, np.uint32), _ptr(rvv, np.uint8),
                                        _ptr(rst, np.uint8), _ptr(tb, np.uint8), _ptr(ok, np.uint8),
                                        _ptr(nit, np.uint8)))
        return ok.astype(bool), tb, nit

    def fec_decode_plcf(self, llr196):
        n = llr196.shape[0]
        llr196 = np.ascontiguousarray(llr196, dtype=np.int16).reshape(n, 196)
        ok = np.zeros((n, 2), dtype=np.uint8)
        plcf = np.zeros((n, 2, 10), dtype=np.uint8)
        mask = np.zeros((n, 2), dtype=np.uint32)
        nit = np.zeros((n, 2), dtype=np.uint8)
        _check(lib().dnrp_fec_decode_plcf(self._ctx, n, _ptr(llr196, np.int16), _ptr(ok, np.uint8),
                                          _ptr(plcf, np.uint8), _ptr(mask, np.uint32), _ptr(nit, np.uint8)))
        return ok.astype(bool), plcf, mask, nit

    def fec_decode_plcf_noalloc(self, n, llr196_dev, ok_dev, plcf_dev, mask_dev, n_iter_dev, stream=None):
        """dnrp_fec_decode_plcf_noalloc: all arguments are DEVICE addresses (ints) of caller-owned buffers
        (llr196 [n][196] int16, ok [n][2] u8, plcf [n][2][16] u8, mask [n][2] u32, n_iter [n][2] u8); the call only
        enqueues kernels (scratch = the library's static device array) and does not synchronise"""
        _check(lib().dnrp_fec_decode_plcf_noalloc(self._ctx, C.c_uint32(n), C.c_void_p(llr196_dev), C.c_void_p(ok_dev),
                                                  C.c_void_p(plcf_dev), C.c_void_p(mask_dev), C.c_void_p(n_iter_dev),
                                                  C.c_void_p(stream)))

    # ---- tx_t::generate_tx_packet (reference lib/src/phy/tx/tx.cpp:165-314)
    def tx(self, plcf, tb, codebook_index=0, DAC_scale=1.0, iq_phase_rad=0.0, iq_phase_increment=0.0,
           GI_percentage=5, rv=0, optimal_scaling_DAC=False, fetch=True, sc16_scale=None):
        """dnrp_tx_batch; sc16_scale: dnrp_tx_batch_sc16, the IQ comes back as int16 [n][N_TX][stride][2] = round(sample x scale)"""
        n = tb.shape[0]
        plcf = np.ascontiguousarray(plcf, dtype=np.uint8).reshape(n, 10)
        tb = np.ascontiguousarray(tb, dtype=np.uint8).reshape(n, self.ps["N_TB_byte"])

        def per(v):
            return np.broadcast_to(np.asarray(v), (n,))

        descs = (TxDesc * n)()
        cb, ds, ph, inc, gi, rvv = per(codebook_index), per(DAC_scale), per(iq_phase_rad), per(iq_phase_increment), \
            per(GI_percentage), per(rv)
        for i in range(n):
            descs[i] = TxDesc(int(cb[i]), int(optimal_scaling_DAC), float(ds[i]), float(ph[i]), float(inc[i]),
                              int(gi[i]), int(rvv[i]))
        lib().dnrp_tx_stride.restype = C.c_uint32
        stride = lib().dnrp_tx_stride(self._ctx)
        iq = np.zeros((n, self.ps["N_TX"], stride), dtype=np.complex64) if fetch else None
        ns = np.zeros(n, dtype=np.uint32)
        if sc16_scale is not None:
            iq16 = np.zeros((n, self.ps["N_TX"], stride, 2), dtype=np.int16)
            f = lib().dnrp_tx_batch_sc16
            f.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p]
            _check(f(self._ctx, n, C.cast(descs, C.c_void_p), plcf.ctypes.data_as(C.c_void_p), tb.ctypes.data_as(C.c_void_p),
                     iq16.ctypes.data_as(C.c_void_p), float(sc16_scale), ns.ctypes.data_as(C.c_void_p)))
            return iq16, ns
        _check(lib().dnrp_tx_batch(self._ctx, n, descs, _ptr(plcf, np.uint8), _ptr(tb, np.uint8),
                                   _ptr(iq, np.complex64), _ptr(ns, np.uint32)))
        return iq, ns

    # ---- several networks / PLCF types per batch (worker_pool_t::add_network_id, per-packet HARQ process fields)
    def add_network_id(self, network_id):
        _check(lib().dnrp_add_network_id(self._ctx, C.c_uint32(network_id)))

    def harq_reset(self, slot):
        _check(lib().dnrp_harq_reset(self._ctx, C.c_uint32(slot)))

    def tx_ex(self, plcf, tb, network_id=None, plcf_type=None, codebook_index=0, rv=0, GI_percentage=5, fetch=True):
        """dnrp_tx_batch_ex: network_id [n] / plcf_type [n] per packet (None = the context's)"""
        n = tb.shape[0]
        plcf = np.ascontiguousarray(plcf, dtype=np.uint8).reshape(n, 10)
        tb = np.ascontiguousarray(tb, dtype=np.uint8).reshape(n, self.ps["N_TB_byte"])
        descs = (TxDesc * n)()
        rvv, cb = np.broadcast_to(np.asarray(rv), (n,)), np.broadcast_to(np.asarray(codebook_index), (n,))
        for i in range(n):
            descs[i] = TxDesc(int(cb[i]), 0, 1.0, 0.0, 0.0, int(GI_percentage), int(rvv[i]))
        ids = None if network_id is None else np.ascontiguousarray(network_id, dtype=np.uint32)
        tys = None if plcf_type is None else np.ascontiguousarray(plcf_type, dtype=np.uint8)
        lib().dnrp_tx_stride.restype = C.c_uint32
        stride = lib().dnrp_tx_stride(self._ctx)
        iq = np.zeros((n, self.ps["N_TX"], stride), dtype=np.complex64) if fetch else None
        ns = np.zeros(n, dtype=np.uint32)
        _check(lib().dnrp_tx_batch_ex(self._ctx, n, descs, _ptr(ids, np.uint32), _ptr(tys, np.uint8), _ptr(plcf, np.uint8),
                                      _ptr(tb, np.uint8), _ptr(iq, np.complex64), _ptr(ns, np.uint32)))
        return iq, ns

    def rx_pdc_ex(self, n, network_id=None, plcf_type=None, harq_slot=None, rv=None, harq_reset=None):
        """dnrp_rx_pdc_batch_ex: network_id [n] / plcf_type [n] per packet (None = the context's)"""
        tb = np.zeros((n, self.ps["N_TB_byte"]), dtype=np.uint8)
        nit = np.zeros((n, self.ps["C"]), dtype=np.uint8)
        slot = None if harq_slot is None else np.ascontiguousarray(harq_slot, dtype=np.uint32)
        rvv = None if rv is None else np.ascontiguousarray(rv, dtype=np.uint8)
        rst = None if harq_reset is None else np.ascontiguousarray(harq_reset, dtype=np.uint8)
        ids = None if network_id is None else np.ascontiguousarray(network_id, dtype=np.uint32)
        tys = None if plcf_type is None else np.ascontiguousarray(plcf_type, dtype=np.uint8)
        out = np.zeros(n, dtype=PDC_DTYPE)
        _check(lib().dnrp_rx_pdc_batch_ex(self._ctx, n, _ptr(slot, np.uint32), _ptr(rvv, np.uint8), _ptr(rst, np.uint8),
                                          _ptr(ids, np.uint32), _ptr(tys, np.uint8), _ptr(tb, np.uint8),
                                          out.ctypes.data_as(C.c_void_p), _ptr(nit, np.uint8)))
        self.last_pdc_report = out.view(np.recarray)
        return out["crc_ok"].astype(bool), tb, out["snr_dB"].copy(), nit

    # ---- rx_synced_t::demoddecod_rx_pcc / demoddecod_rx_pdc (reference rx_synced.cpp:186-436)
    def rx_stride(self):
        lib().dnrp_rx_stride.restype = C.c_uint32
        return lib().dnrp_rx_stride(self._ctx)

    def make_sync(self, n, t0=0, cfo_fractional_rad=0.0, cfo_integer_rad=0.0, rms=None):
        """[n] dnrp_sync_report_t as a numpy structured array (same memory layout as SyncReport)"""
        sync = np.zeros(n, dtype=SYNC_DTYPE)
        sync["fine_peak_time_local"] = t0
        sync["cfo_fractional_rad"] = cfo_fractional_rad
        sync["cfo_integer_rad"] = cfo_integer_rad
        if rms is not None:
            r = np.asarray(rms, dtype=np.float32).reshape(n, -1)
            sync["rms"][:, :r.shape[1]] = r
        return sync

    def rx_pcc(self, iq, t0=0, cfo_fractional_rad=0.0, cfo_integer_rad=0.0, rms=None, n=None, sync=None):
        """iq [n][N_RX][rx_stride] complex64 host windows, or None to use the device-resident windows
        (after channel / loopback_tx_to_rx); returns [n] PccReport records (numpy structured array)"""
        if iq is not None:
            n = iq.shape[0]
            iq = np.ascontiguousarray(iq, dtype=np.complex64)
            assert iq.shape[-1] == self.rx_stride(), (iq.shape, self.rx_stride())
        if sync is None:
            sync = self.make_sync(n, t0, cfo_fractional_rad, cfo_integer_rad, rms)
        # sync="device": continue on the results the preceding sync() left on the device
        sync_p = None if isinstance(sync, str) else sync.ctypes.data_as(C.c_void_p)
        out = np.zeros(n, dtype=PCC_DTYPE).view(np.recarray)
        _check(lib().dnrp_rx_pcc_batch(self._ctx, n, _ptr(iq, np.complex64), sync_p, out.ctypes.data_as(C.c_void_p)))
        return out

    # ---- sync_chunk_t::search (reference lib/src/phy/rx/sync/sync_chunk.cpp:120-215), one search per window
    def sync(self, iq=None, n=None, search_length=0, fetch=True):
        """iq [n][N_RX][rx_stride] complex64 host windows or None (device-resident windows); returns [n]
        dnrp_sync_result_t records"""
        if iq is not None:
            n = iq.shape[0]
            iq = np.ascontiguousarray(iq, dtype=np.complex64)
            assert iq.shape[1:] == (self.N_RX, self.rx_stride()), (iq.shape, self.rx_stride())
        out = np.zeros(n, dtype=SYNC_RESULT_DTYPE).view(np.recarray) if fetch else None
        _check(lib().dnrp_sync_batch(self._ctx, n, _ptr(iq, np.complex64), int(search_length),
                                     None if out is None else out.ctypes.data_as(C.c_void_p)))
        return out

    def sync_sc16(self, iq16, scale=1.0 / 32768.0, search_length=0):
        """dnrp_sync_batch_sc16: iq16 [n][N_RX][rx_stride][2] int16 host windows; rx_pcc(None, n=n, sync="device") continues"""
        n = iq16.shape[0]
        assert iq16.dtype == np.int16 and iq16.flags["C_CONTIGUOUS"] and iq16.shape[1:] == (self.N_RX, self.rx_stride(), 2)
        out = np.zeros(n, dtype=SYNC_RESULT_DTYPE).view(np.recarray)
        f = lib().dnrp_sync_batch_sc16
        f.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_float, C.c_uint32, C.c_void_p]
        _check(f(self._ctx, n, iq16.ctypes.data_as(C.c_void_p), float(scale), int(search_length), out.ctypes.data_as(C.c_void_p)))
        return out

    # ---- coarse_peak_f_domain_t between coarse and fine peak: beta search and integer CFO (off by default)
    def sync_set_fdomain(self, beta_threshold_dB=0.0, cfo_search_range_subc=0):
        """dnrp_sync_set_fdomain: beta_threshold_dB < 0 switches the beta search on (reference: -10 dB),
        cfo_search_range_subc >= 4 the integer CFO search (reference: 4)"""
        f = lib().dnrp_sync_set_fdomain
        f.argtypes = [C.c_void_p, C.c_float, C.c_uint32]
        _check(f(self._ctx, float(beta_threshold_dB), int(cfo_search_range_subc)))

    def sync_fdomain_read(self, n):
        """dnrp_sync_fdomain_read: (estimated beta [n], integer CFO of the signal in subcarriers [n]) of the last search"""
        b = np.zeros(n, dtype=np.uint32)
        cfo = np.zeros(n, dtype=np.int32)
        _check(lib().dnrp_sync_fdomain_read(self._ctx, n, _ptr(b, np.uint32), _ptr(cfo, np.int32)))
        return b, cfo

    # ---- residual CFO / STO tracking on the DRS (off by default; rx_synced_param.hpp:156-181)
    def rx_set_tracking(self, residual_cfo=True, residual_sto=True):
        """dnrp_rx_set_tracking"""
        f = lib().dnrp_rx_set_tracking
        f.argtypes = [C.c_void_p, C.c_int, C.c_int]
        _check(f(self._ctx, int(bool(residual_cfo)), int(bool(residual_sto))))

    def rx_tracking_read(self, n):
        """dnrp_rx_tracking_read: [n][2] float32 (dphi, dpsi): phase advance per N_step symbols, common and per subcarrier"""
        out = np.zeros((n, 2), dtype=np.float32)
        f = lib().dnrp_rx_tracking_read
        f.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        _check(f(self._ctx, n, out.ctypes.data_as(C.c_void_p)))
        return out

    # ---- streaming front door (worker_sync_t / sync_chunk_t / baton_t): packet search over one continuous stream
    def stream_sync(self, iq, chunk_hop=0, max_out=None):
        """iq [N_RX][n_samples] complex64 host stream; returns the unique packets in time order (STREAM_PACKET_DTYPE)"""
        iq = np.ascontiguousarray(iq, dtype=np.complex64)
        assert iq.ndim == 2 and iq.shape[0] == self.N_RX, iq.shape
        n_samples = iq.shape[1]
        if max_out is None:
            max_out = n_samples // max(self.ps["N_samples_STF"], 1) + 1
        out = np.zeros(max_out, dtype=STREAM_PACKET_DTYPE)
        n_out = C.c_uint32(0)
        _check(lib().dnrp_stream_sync(self._ctx, C.c_uint64(n_samples), _ptr(iq, np.complex64), int(chunk_hop), int(max_out),
                                      out.ctypes.data_as(C.c_void_p), C.byref(n_out)))
        return out[:n_out.value].view(np.recarray)

    def stream_open(self, max_push_samples, chunk_hop=0):
        _check(lib().dnrp_stream_open(self._ctx, int(max_push_samples), int(chunk_hop)))

    def stream_push(self, iq, max_out=256):
        """iq [N_RX][n] complex64 hardware-rate samples (n may be 0) -> the packets complete after this push"""
        iq = np.ascontiguousarray(iq, dtype=np.complex64).reshape(self.N_RX, -1)
        out = np.zeros(max_out, dtype=STREAM_PACKET_DTYPE)
        n_out = C.c_uint32(0)
        _check(lib().dnrp_stream_push(self._ctx, iq.shape[1], _ptr(iq, np.complex64) if iq.shape[1] else None, int(max_out),
                                      out.ctypes.data_as(C.c_void_p), C.byref(n_out)))
        return out[:n_out.value].view(np.recarray)

    def stream_push_sc16(self, iq16, scale=1.0 / 32768.0, max_out=256):
        """iq16 [N_RX][n][2] int16 hardware-rate samples in the radio's wire format -> the packets complete after this push"""
        iq16 = np.ascontiguousarray(iq16, dtype=np.int16).reshape(self.N_RX, -1, 2)
        out = np.zeros(max_out, dtype=STREAM_PACKET_DTYPE)
        n_out = C.c_uint32(0)
        f = lib().dnrp_stream_push_sc16
        f.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_float, C.c_uint32, C.c_void_p, C.c_void_p]
        _check(f(self._ctx, iq16.shape[1], iq16.ctypes.data_as(C.c_void_p) if iq16.shape[1] else None, float(scale), int(max_out),
                 out.ctypes.data_as(C.c_void_p), C.byref(n_out)))
        return out[:n_out.value].view(np.recarray)

    def stream_close(self):
        _check(lib().dnrp_stream_close(self._ctx))

    def stream_extract(self, start_64, t0):
        """windows of the synced receiver <- stream[start - t0, start - t0 + rx_stride) for every start (device copy)"""
        start = np.ascontiguousarray(start_64, dtype=np.int64)
        _check(lib().dnrp_stream_extract(self._ctx, len(start), start.ctypes.data_as(C.c_void_p), int(t0)))

    def rx_pdc(self, n, harq_slot=None, rv=None, harq_reset=None, tb_out=None):
        tb = np.zeros((n, self.ps["N_TB_byte"]), dtype=np.uint8) if tb_out is None else tb_out
        nit = np.zeros((n, self.ps["C"]), dtype=np.uint8)
        slot = None if harq_slot is None else np.ascontiguousarray(harq_slot, dtype=np.uint32)
        rvv = None if rv is None else np.ascontiguousarray(rv, dtype=np.uint8)
        rst = None if harq_reset is None else np.ascontiguousarray(harq_reset, dtype=np.uint8)
        out = np.zeros(n, dtype=PDC_DTYPE)
        _check(lib().dnrp_rx_pdc_batch(self._ctx, n, _ptr(slot, np.uint32), _ptr(rvv, np.uint8), _ptr(rst, np.uint8),
                                       _ptr(tb, np.uint8), out.ctypes.data_as(C.c_void_p), _ptr(nit, np.uint8)))
        self.last_pdc_report = out.view(np.recarray)   # incl. the MIMO report fields
        return out["crc_ok"].astype(bool), tb, out["snr_dB"].copy(), nit

    def rx(self, iq, harq_slot=None, rv=None, harq_reset=None, tb_out=None, search_length=0, want_sync=True,
           want_pcc=True, sc16_scale=1.0 / 32768.0):
        """dnrp_rx_batch: host IQ windows -> (sync results, PCC reports, crc_ok, transport blocks, PDC reports), copies
        pipelined against the kernels"""
        n = iq.shape[0]
        sc16 = iq.dtype == np.int16   # [n][N_RX][rx_stride][2] int16 I/Q: dnrp_rx_batch_sc16
        sc8 = iq.dtype == np.int8     # the same as int8: dnrp_rx_batch_sc8 (pass sc16_scale = 1 / 128 or the gain used)
        if sc16 or sc8:
            assert iq.flags["C_CONTIGUOUS"] and iq.shape[1:] == (self.N_RX, self.rx_stride(), 2)
        else:
            assert iq.dtype == np.complex64 and iq.flags["C_CONTIGUOUS"] and iq.shape[1:] == (self.N_RX, self.rx_stride())
        tb = np.zeros((n, self.ps["N_TB_byte"]), dtype=np.uint8) if tb_out is None else tb_out
        slot = None if harq_slot is None else np.ascontiguousarray(harq_slot, dtype=np.uint32)
        rvv = None if rv is None else np.ascontiguousarray(rv, dtype=np.uint8)
        rst = None if harq_reset is None else np.ascontiguousarray(harq_reset, dtype=np.uint8)
        sres = np.zeros(n, dtype=SYNC_RESULT_DTYPE).view(np.recarray) if want_sync else None
        pcc = np.zeros(n, dtype=PCC_DTYPE).view(np.recarray) if want_pcc else None
        pdc = np.zeros(n, dtype=PDC_DTYPE).view(np.recarray)
        tail = (int(search_length), _ptr(slot, np.uint32), _ptr(rvv, np.uint8), _ptr(rst, np.uint8),
                None if sres is None else sres.ctypes.data_as(C.c_void_p),
                None if pcc is None else pcc.ctypes.data_as(C.c_void_p), _ptr(tb, np.uint8),
                pdc.ctypes.data_as(C.c_void_p), None)
        if sc8:
            _check(lib().dnrp_rx_batch_sc8(self._ctx, n, iq.ctypes.data_as(C.c_void_p), C.c_float(sc16_scale), *tail))
        elif sc16:
            _check(lib().dnrp_rx_batch_sc16(self._ctx, n, iq.ctypes.data_as(C.c_void_p), C.c_float(sc16_scale), *tail))
        else:
            _check(lib().dnrp_rx_batch(self._ctx, n, _ptr(iq, np.complex64), *tail))
        self.last_pdc_report = pdc
        return sres, pcc, pdc.crc_ok.astype(bool), tb, pdc

    def tap(self, which, n):
        lib().dnrp_tap_bytes.restype = C.c_size_t
        nbytes = lib().dnrp_tap_bytes(self._ctx, which, n)
        buf = np.zeros(nbytes, dtype=np.uint8)
        _check(lib().dnrp_tap_read(self._ctx, which, n, _ptr(buf, np.uint8), C.c_size_t(nbytes)))
        if which in (TAP_Y, TAP_DRS_ZF, TAP_STF_ZF, TAP_CHESTIM):
            return buf.view(np.complex64).reshape(n, -1)
        if which in (TAP_LLR_PCC, TAP_LLR_PDC):
            return buf.view(np.int16).reshape(n, -1)
        return buf.reshape(n, -1)

    def loopback_tx_to_rx(self, n, t0=0):
        _check(lib().dnrp_loopback_tx_to_rx(self._ctx, n, t0))

    # ---- simulation channel (reference simulation/vspace.cpp:437-512, wireless/channel_{awgn,flat}.cpp)
    def channel(self, n, snr_dB=100.0, t0=0, mode=0, seed=1, net_bandwidth_norm=None, fetch=False, pdp_idx=0,
                tau_rms_ns=0.0, fD_Hz=0.0, time_64=0, large_scale=0.0, clip_limit=0.0, dac_bits=0, adc_bits=0):
        if net_bandwidth_norm is None:
            net_bandwidth_norm = self.resample[1] / (self.resample[0] * self.oversampling)
        descs = np.zeros(n, dtype=CHANNEL_DTYPE)
        descs["snr_dB"], descs["net_bandwidth_norm"], descs["t0"], descs["mode"] = snr_dB, net_bandwidth_norm, t0, mode
        descs["seed"] = seed
        descs["pdp_idx"], descs["tau_rms_ns"], descs["fD_Hz"], descs["time_64"] = pdp_idx, tau_rms_ns, fD_Hz, time_64
        descs["large_scale"], descs["clip_limit"], descs["dac_bits"], descs["adc_bits"] = large_scale, clip_limit, dac_bits, adc_bits
        out = np.zeros((n, self.N_RX, self.rx_stride()), dtype=np.complex64) if fetch else None
        _check(lib().dnrp_channel_batch(self._ctx, n, descs.ctypes.data_as(C.c_void_p), _ptr(out, np.complex64)))
        return out

    # ---- measurement
    def kernel_time_enable(self, on=True):
        _check(lib().dnrp_kernel_time_enable(self._ctx, int(on)))

    def kernel_time_read(self, kernel):
        ms, cnt = C.c_float(), C.c_uint32()
        _check(lib().dnrp_kernel_time_read(self._ctx, kernel, C.byref(ms), C.byref(cnt)))
        return ms.value, cnt.value

    def launch_count(self):
        lib().dnrp_launch_count.restype = C.c_uint64
        return lib().dnrp_launch_count(self._ctx)

    def time_stage(self, stage, n, reps=10):
        ms = C.c_float()
        _check(lib().dnrp_time_stage(self._ctx, stage, n, reps, C.byref(ms)))
        return ms.value


STAGE_FEC_DECODE_TB, STAGE_FEC_ENCODE, STAGE_FEC_DECODE_PLCF = 1, 2, 3
STAGE_RX_CHAIN, STAGE_TX_CHAIN, STAGE_CHANNEL, STAGE_SYNC, STAGE_SYNC_RX_CHAIN = 4, 5, 6, 7, 8
