"""oracle/ref.py -- ctypes access to oracle/_ref/libdectnrp_ref.so: the reference's OWN PHY sources (tx_t, rx_synced_t,
sync_chunk_t, resampler_t, sections_part3) compiled unmodified from /root/reference by oracle/ref_recipe/Makefile, with
stand-ins for the third-party libraries that are absent on this image (oracle/ref_recipe/shim*).

TEST INFRASTRUCTURE: imported by tests/, __graft_entry__.smoke() and bench.py's reference arm only.  /root/reference does
not exist on the GPU box; there the prebuilt library (it travels with the repository snapshot) is used as it is.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
RECIPE = os.path.join(_HERE, "ref_recipe")
LIB_PATH = os.path.join(_HERE, "_ref", "libdectnrp_ref.so")
REFERENCE_ROOT = "/root/reference"

_PS_FIELDS = ("N_eff_TX", "N_SS", "N_TS", "N_TX", "closed_loop", "N_bps", "N_b_DFT", "N_b_CP", "N_b_OCC", "N_PACKET_symb",
              "N_DF_symb", "N_PDC_subc", "N_DRS_subc", "G", "N_PDC_bits", "N_TB_bits", "N_TB_byte", "C", "K_plus", "K_minus",
              "C_plus", "C_minus", "N_samples_STF", "N_samples_STF_CP_only", "N_samples_DF", "N_samples_GI",
              "N_samples_packet_no_GI", "N_samples_packet")
_PSDEF = ("u", "b", "PacketLengthType", "PacketLength", "tm_mode_index", "mcs_index", "Z")


def can_build():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "lib", "src"))


def build(force=False):
    """compile the reference's sources where they lie (needs /root/reference: this container only)"""
    if not can_build():
        return os.path.exists(LIB_PATH)
    if force:
        subprocess.check_call(["make", "-C", RECIPE, "clean"], stdout=subprocess.DEVNULL)
    subprocess.check_call(["make", "-C", RECIPE, "-j%d" % (os.cpu_count() or 4)], stdout=subprocess.DEVNULL)
    return True


def available():
    return os.path.exists(LIB_PATH)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not available():
            build()
        _lib = C.CDLL(LIB_PATH)
        _lib.ref_tx_create.restype = C.c_void_p
        _lib.ref_rx_create.restype = C.c_void_p
        _lib.ref_resample_n_after.restype = C.c_uint32
        _lib.ref_rx_n_rx.restype = C.c_uint32
        _lib.ref_pool_create.restype = C.c_void_p
        _lib.ref_pool_threads.restype = C.c_uint32
    return _lib


def _u32(v):
    return np.ascontiguousarray(v, dtype=np.uint32)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def get_packet_sizes(*psdef):
    """sp3::get_packet_sizes (+ srsran_cbsegm_FIX) of the reference; dict with the names of dnrp_packet_sizes_t, or None"""
    out = np.zeros(40, dtype=np.uint32)
    n = lib().ref_get_packet_sizes(_p(_u32(psdef)), _p(out))
    if n < 0:
        return None
    d = dict(zip(_PS_FIELDS, (int(x) for x in out[:n])))
    d.update(dict(zip(_PSDEF, (int(x) for x in psdef))))
    return d


def k_i_linear(kind, b, N_TS):
    """pcc_t::get_k_i_linear ("pcc") / pdc_t::get_k_i_linear ("pdc")"""
    k = {"pcc": 0, "pdc": 1}[kind]
    n = lib().ref_k_i_linear(k, b, N_TS, None, 0)
    out = np.zeros(n, dtype=np.uint32)
    lib().ref_k_i_linear(k, b, N_TS, _p(out), n)
    return out


def stf(b, N_eff_TX):
    out = np.zeros(2 * (56 * b + 8), dtype=np.float32)
    n = lib().ref_stf(b, N_eff_TX, _p(out), out.size)
    return out[:2 * n].view(np.complex64).copy()


def W(N_TS, N_TX, idx):
    out = np.zeros(2 * N_TS * N_TX, dtype=np.float32)
    n = lib().ref_W(N_TS, N_TX, idx, _p(out), out.size)
    return out.view(np.complex64).reshape(N_TX, N_TS).copy(), n


def resample(x, L, M, os=1, user=0, unit=0, flush=True):
    """resampler_t (user 0 TX, 1 SYNC, 2 RX_SYNCED filter definitions), x fed in `unit`-sized calls"""
    x = np.ascontiguousarray(x, dtype=np.complex64)
    cap = int(len(x) * L / M) + 4 * L + 64
    out = np.zeros(cap, dtype=np.complex64)
    n = lib().ref_resample(L, M, os, user, _p(x), len(x), unit, int(flush), _p(out), cap)
    assert n >= 0, n
    return out[:n].copy()


def resample_n_after(L, M, n_in, os=1, user=0):
    return lib().ref_resample_n_after(L, M, os, user, n_in)


class Tx:
    """tx_t of the reference for a pool whose maximum packet is psdef_max"""

    def __init__(self, psdef_max, os=1, resample=(1, 1), buffer_samples=None):
        self.ps_max = get_packet_sizes(*psdef_max)
        assert self.ps_max is not None
        if buffer_samples is None:
            buffer_samples = int(self.ps_max["N_samples_packet"] * os * resample[0] / resample[1] * 1.3) + 4096
        self.buffer_samples = buffer_samples
        self._h = C.c_void_p(lib().ref_tx_create(_p(_u32(psdef_max)), os, resample[0], resample[1], buffer_samples))
        assert self._h.value
        self._nets = set()

    def close(self):
        if self._h:
            lib().ref_tx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def packet(self, psdef, plcf, tb, network_id=123456789, plcf_type=1, codebook_index=0, rv=0, new_tb=True, DAC_scale=1.0,
               iq_phase_rad=0.0, iq_phase_increment=0.0, GI_percentage=5, optimal_scaling_DAC=False):
        """generate_tx_packet; returns (iq [N_TX][n_samples], pcc_d [25], pdc_d [ceil(G/8)])"""
        ps = get_packet_sizes(*psdef)
        if network_id not in self._nets:
            lib().ref_tx_add_network_id(self._h, C.c_uint32(network_id))
            self._nets.add(network_id)
        plcf10 = np.zeros(10, dtype=np.uint8)
        plcf10[:len(plcf)] = plcf
        tb = np.ascontiguousarray(tb, dtype=np.uint8)
        assert tb.size == ps["N_TB_byte"]
        meta = np.array([float(optimal_scaling_DAC), DAC_scale, iq_phase_rad, iq_phase_increment, GI_percentage], dtype=np.float32)
        iq = np.zeros((ps["N_TX"], self.buffer_samples), dtype=np.complex64)
        ns = C.c_uint32(0)
        pcc_d = np.zeros(25, dtype=np.uint8)
        pdc_d = np.zeros((ps["G"] + 7) // 8, dtype=np.uint8)
        rc = lib().ref_tx_packet(self._h, _p(_u32(psdef)), plcf_type, C.c_uint32(network_id), _p(plcf10), _p(tb), codebook_index,
                                 _p(meta), rv, int(new_tb), _p(iq), self.buffer_samples, C.byref(ns), _p(pcc_d), _p(pdc_d))
        assert rc == 0, rc
        return iq[:, :ns.value].copy(), pcc_d, pdc_d


class Rx:
    """rx_synced_t of the reference; N_RX = N_TX of psdef_max (the reference's own rule)"""

    def __init__(self, psdef_max, n_window, os=1, resample=(1, 1), chestim_lr=True, stride=2):
        self.ps_max = get_packet_sizes(*psdef_max)
        assert self.ps_max is not None
        self.n_window = n_window
        self.os = os
        self._h = C.c_void_p(lib().ref_rx_create(_p(_u32(psdef_max)), os, resample[0], resample[1], n_window, int(chestim_lr), stride))
        assert self._h.value
        self.N_RX = lib().ref_rx_n_rx(self._h)
        self._nets = set()

    def close(self):
        if self._h:
            lib().ref_rx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def pcc(self, iq, t0, u, b, N_eff_TX, cfo_fractional_rad=0.0, cfo_integer_rad=0.0, rms=None):
        """rms: per-antenna RMS from synchronisation (the reference asserts that at least one antenna has one; zeros are
        estimated by its STF stage).  None: antenna 0 gets the RMS of the STF samples of the window (contexts without
        resampling), the value the STF stage itself would estimate"""
        iq = np.ascontiguousarray(iq, dtype=np.complex64)
        assert iq.shape[0] == self.N_RX and iq.shape[1] <= self.n_window, (iq.shape, self.N_RX, self.n_window)
        sync = np.zeros(14, dtype=np.float64)
        sync[:6] = [t0, u, b, N_eff_TX, cfo_fractional_rad, cfo_integer_rad]
        if rms is None:
            n_stf = (64 + 48 if u == 1 else 64 + 80) * b * (iq.shape[1] and 1)
            n_stf *= self.os
            seg = iq[0, t0:t0 + n_stf]
            sync[6] = np.sqrt(np.float32((np.abs(seg.astype(np.complex64)) ** 2).sum(dtype=np.float32)) / np.float32(n_stf))
        else:
            r = np.asarray(rms, dtype=np.float64).ravel()
            sync[6:6 + r.size] = r
        ok = np.zeros(2, dtype=np.uint8)
        plcf = np.zeros((2, 10), dtype=np.uint8)
        clbf = np.zeros(2, dtype=np.uint8)
        llr = np.zeros(196, dtype=np.int16)
        rep = np.zeros(11, dtype=np.float32)
        has = np.zeros(2, dtype=np.uint8)
        rc = lib().ref_rx_pcc(self._h, _p(iq), iq.shape[1], iq.shape[1], _p(sync), _p(ok), _p(plcf), _p(clbf), _p(llr), _p(rep), _p(has))
        assert rc == 0, rc
        return {"plcf_ok": ok.astype(bool), "plcf": plcf, "cl_bf": clbf, "llr_pcc": llr, "snr_dB": float(rep[0]),
                "cfo_fractional_rad": float(rep[1]), "sto_fractional": float(rep[2]), "rms": rep[3:11].copy(),
                "has_plcf": has.astype(bool)}

    def pdc(self, psdef, network_id=123456789, plcf_type=1, rv=0, new_tb=True):
        ps = get_packet_sizes(*psdef)
        if network_id not in self._nets:
            lib().ref_rx_add_network_id(self._h, C.c_uint32(network_id))
            self._nets.add(network_id)
        tb = np.zeros(ps["N_TB_byte"], dtype=np.uint8)
        llr = np.zeros(ps["G"], dtype=np.int16)
        rep = np.zeros(4, dtype=np.float32)
        rc = lib().ref_rx_pdc(self._h, _p(_u32(psdef)), plcf_type, C.c_uint32(network_id), rv, int(new_tb), _p(tb), _p(llr), _p(rep))
        assert rc >= 0, rc
        return {"crc_ok": bool(rc), "tb": tb, "llr_pdc_descrambled": llr, "snr_dB": float(rep[1]),
                "tm_3_7_beamforming_idx": int(rep[2]), "tm_3_7_beamforming_reciprocal_idx": int(rep[3])}


def sync_stream(iq, psdef_max, os=1, resample=(1, 1), chunk_length=0, max_out=64):
    """sync_chunk_t::search over one stream iq [N_RX][n] until its chunk is used up; list of dicts"""
    iq = np.ascontiguousarray(iq, dtype=np.complex64)
    out = np.zeros((max_out, 12), dtype=np.float64)
    coarse = np.zeros((max_out, 8), dtype=np.float32)
    rms = np.zeros((max_out, 8), dtype=np.float32)
    n = lib().ref_sync_stream(_p(_u32(psdef_max)), os, resample[0], resample[1], iq.shape[0], _p(iq), iq.shape[1], iq.shape[1],
                              chunk_length, max_out, _p(out), _p(coarse), _p(rms))
    assert n >= 0, n
    names = ("fine_peak_time_64", "coarse_peak_time_64", "N_eff_TX", "detection_time_local", "detection_ant_idx", "detection_metric",
             "detection_rms", "coarse_peak_time_local", "cfo_fractional_rad", "fine_peak_time_local",
             "detection_time_with_jump_back_local", "u")
    res = []
    for i in range(n):
        d = dict(zip(names, out[i]))
        for k in ("fine_peak_time_64", "coarse_peak_time_64", "N_eff_TX", "detection_time_local", "detection_ant_idx",
                  "coarse_peak_time_local", "fine_peak_time_local", "detection_time_with_jump_back_local", "u"):
            d[k] = int(d[k])
        d["coarse_peak_array"] = coarse[i].copy()
        d["rms_array"] = rms[i].copy()
        res.append(d)
    return res


class Pool:
    """n_threads workers (ring buffer + sync_chunk_t + rx_synced_t each), the reference's own scaling model; used by
    bench.py's reference arm and cpu_baseline"""

    def __init__(self, psdef_max, n_window, threads, os=1, resample=(1, 1), chestim_lr=True, stride=2, network_id=123456789):
        self.ps_max = get_packet_sizes(*psdef_max)
        self.n_window, self.network_id = n_window, network_id
        self._h = C.c_void_p(lib().ref_pool_create(_p(_u32(psdef_max)), os, resample[0], resample[1], n_window, int(chestim_lr), stride,
                                                   threads, C.c_uint32(network_id)))
        assert self._h.value
        self.N_RX = self.ps_max["N_TX"]
        self.threads = lib().ref_pool_threads(self._h)

    def close(self):
        if self._h:
            lib().ref_pool_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def rx_windows(self, iq, psdef, plcf_type=1):
        """iq [n][N_RX][n_window] complex64 -> dict(crc_ok, tb, fine_peak, plcf_ok, stage_ms)"""
        iq = np.ascontiguousarray(iq, dtype=np.complex64)
        n = iq.shape[0]
        assert iq.shape[1] == self.N_RX and iq.shape[2] <= self.n_window
        ps = get_packet_sizes(*psdef)
        ok = np.zeros(n, dtype=np.uint8)
        tb = np.zeros((n, ps["N_TB_byte"]), dtype=np.uint8)
        fine = np.zeros(n, dtype=np.int32)
        pok = np.zeros(n, dtype=np.uint8)
        ms = np.zeros(3, dtype=np.float64)
        rc = lib().ref_pool_rx_windows(self._h, n, _p(iq), iq.shape[2], iq.shape[2], _p(_u32(psdef)), plcf_type,
                                       C.c_uint32(self.network_id), _p(ok), _p(tb), _p(fine), _p(pok), _p(ms))
        assert rc >= 0, rc
        return {"crc_ok": ok.astype(bool), "tb": tb, "fine_peak": fine, "plcf_ok": pok.astype(bool),
                "stage_ms": {"sync": float(ms[0]), "pcc": float(ms[1]), "pdc": float(ms[2])}}
