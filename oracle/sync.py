"""ORACLE (test infrastructure, not product code) -- Python front of oracle/sync.c (packet synchronisation:
reference lib/src/phy/rx/sync/{sync_chunk,autocorrelator_detection,autocorrelator_peak,crosscorrelator,
stf_template}.cpp).  Parity unpinned (see sync.c header)."""
import ctypes as C
import functools

import numpy as np

from . import lib, spec


class Report(C.Structure):
    _fields_ = [("detected", C.c_int32), ("detection_ant_idx", C.c_int32), ("detection_time_local", C.c_int32),
                ("detection_time_with_jump_back_local", C.c_int32), ("detection_rms", C.c_float),
                ("detection_metric", C.c_float), ("coarse_peak_time_local", C.c_int32),
                ("coarse_peak_array", C.c_float * 8), ("rms_array", C.c_float * 8), ("cfo_fractional_rad", C.c_float),
                ("cfo_integer_rad", C.c_float), ("N_eff_TX", C.c_int32), ("fine_peak_time_local", C.c_int32),
                ("n_detections", C.c_int32), ("coarse_peak_time_hw", C.c_int32)]


def stf_template(u, b, N_eff_TX, os=1):
    """stf_template_t::generate_stf_time_domain (stf_template.cpp:81-204) for L = M = 1: the transmitter's STF of a
    single stream, scale 1/sqrt(N_OCC/4), unnormalised IFFT, long CP, cover sequence"""
    N_OCC, N_os = 56 * b, 64 * b * os
    cp = (N_os // 4) * (3 if u == 1 else 5)
    y = spec.stf_filled(b, N_eff_TX) * np.float32(1.0 / np.sqrt(np.float32(N_OCC // 4)))
    X = np.zeros(N_os, dtype=np.complex128)
    X[:N_OCC // 2 + 1] = y[N_OCC // 2:]
    X[N_os - N_OCC // 2:] = y[:N_OCC // 2]
    x = (np.fft.ifft(X) * N_os).astype(np.complex64)
    t = x[(np.arange(cp + N_os) - cp) % N_os].copy()
    npat = 7 if u == 1 else 9
    plen = t.size // npat
    for p in range(npat):
        t[p * plen:(p + 1) * plen] *= np.float32(spec.STF_COVER[p])
    return t


@functools.lru_cache(maxsize=None)
def templates(u, b, N_eff_TX_max, os=1, L=1, M=1):
    """stf_template_t for every N_eff_TX; with L != M each template goes through the TX resampler (reset, resample,
    flush) and is truncated to stf_len * L / M samples (stf_template.cpp:169-194)"""
    n = [1, 2, 4, 8]
    t = [stf_template(u, b, v, os) for v in n if v <= N_eff_TX_max]
    if L != M:
        from . import resample as rs
        t = [rs.stream(x, L, M, os)[:x.size * L // M] for x in t]
    return np.stack(t)


def search(iq, u, b, os=1, N_eff_TX_max=1, search_length=None, resample=(1, 1)):
    """iq [n_ant][len] complex64 at the hardware rate -> Report (first packet found in the window).
    resample = (L, M) of the TX direction: the window is resampled with M/L for detection and coarse peak (rx_pacer,
    zero history at the window start, no flush); the cross-correlation runs on the hardware-rate samples."""
    lb = lib()
    iq = np.ascontiguousarray(iq, dtype=np.complex64)
    L_, M_ = resample
    iq_hw = iq
    if L_ != M_:
        from . import resample as rs
        iq = np.ascontiguousarray(np.stack([rs.stream(x, M_, L_, os, flush=False) for x in iq_hw]))
    n_ant, n = iq.shape
    tpl = np.ascontiguousarray(templates(u, b, N_eff_TX_max, os, L_, M_))
    stf_len = (112 if u == 1 else 144) * b * os
    if search_length is None:
        search_length = n - stf_len - stf_len // (7 if u == 1 else 9)
    rep = Report()
    lb.orc_sync_search2.restype = C.c_int
    lb.orc_sync_search2.argtypes = [C.c_void_p, C.c_long, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                    C.c_int, C.c_int, C.c_double, C.c_void_p, C.c_long, C.c_int, C.c_int, C.c_int,
                                    C.c_int, C.c_void_p]
    rc = lb.orc_sync_search2(iq.ctypes.data, n, n, n_ant, u, b, os, tpl.ctypes.data, tpl.shape[0], int(search_length),
                             float(u * b * os * 1.728e6), iq_hw.ctypes.data, iq_hw.shape[1], iq_hw.shape[1], L_, M_,
                             tpl.shape[1], C.addressof(rep))
    if rc == -2:
        # the fine search window leaves the buffer (cannot happen in the reference's ring buffer): the product
        # reports such a window as "not detected", and so does this front
        rep.detected = 0
    elif rc != 0:
        raise RuntimeError("oracle sync failed (rc=%d)" % rc)
    return rep


def search_rs(iq_rs, iq_hw, u, b, os, N_eff_TX_max, search_length, resample, ignore_before=0):
    """search() on a window that is already resampled (iq_rs at the DECT rate, iq_hw the same span at the hardware rate,
    iq_rs sample r <-> iq_hw sample r L / M), detections ending before ignore_before suppressed"""
    lb = lib()
    L_, M_ = resample
    iq_rs = np.ascontiguousarray(iq_rs, dtype=np.complex64)
    iq_hw = np.ascontiguousarray(iq_hw, dtype=np.complex64)
    tpl = np.ascontiguousarray(templates(u, b, N_eff_TX_max, os, L_, M_))
    rep = Report()
    lb.orc_sync_search3.restype = C.c_int
    lb.orc_sync_search3.argtypes = [C.c_void_p, C.c_long, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                    C.c_int, C.c_int, C.c_double, C.c_void_p, C.c_long, C.c_int, C.c_int, C.c_int,
                                    C.c_int, C.c_int, C.c_void_p]
    rc = lb.orc_sync_search3(iq_rs.ctypes.data, iq_rs.shape[1], iq_rs.shape[1], iq_rs.shape[0], u, b, os, tpl.ctypes.data,
                             tpl.shape[0], int(search_length), float(u * b * os * 1.728e6), iq_hw.ctypes.data,
                             iq_hw.shape[1], iq_hw.shape[1], L_, M_, tpl.shape[1], int(ignore_before), C.addressof(rep))
    if rc == -2:
        rep.detected = 0
    elif rc != 0:
        raise RuntimeError("oracle sync failed (rc=%d)" % rc)
    return rep


def stream_search_resampled(iq_hw, u, b, os=1, N_eff_TX_max=1, resample=(10, 9), window=None):
    """stream_search for a radio at L / M times the DECT rate: the WHOLE stream goes through the RX resampler once
    (rx_pacer_t with carried history == one pass, no flush), one sync_chunk_t walks the resampled stream; the
    cross-correlation runs on the hardware-rate stream around round(coarse L / M) (sync_chunk.cpp:204-205).  Windows
    start on multiples of lcm(step, 2 M) so that their first sample is an integer hardware-rate sample; the position to
    resume from is handed to the detector as ignore_before_index.  Returns [(fine_hw, coarse_hw, Report)]."""
    import math
    from . import resample as rs
    L_, M_ = resample
    iq_hw = np.ascontiguousarray(iq_hw, dtype=np.complex64)
    iq = np.ascontiguousarray(np.stack([rs.stream(x, M_, L_, os, flush=False) for x in iq_hw]))
    n = iq.shape[1]
    npat = 7 if u == 1 else 9
    stf = (112 if u == 1 else 144) * b * os
    pat = stf // npat
    step = pat // 4
    Q = math.lcm(step, 2 * M_)
    lead, tail = stf + pat, 2 * stf + 3 * pat
    if window is None:
        window = 64 * stf
    found = []
    pos = lead
    n_pkt_guard = stf * L_ // M_
    while pos + stf < n:
        s = max(pos - lead, 0) // Q * Q                 # floor to the window grid
        w = iq[:, s:s + (pos - s) + window + tail]
        need = (pos - s) + stf + tail
        if w.shape[1] < need:
            w = np.concatenate([w, np.zeros((iq.shape[0], need - w.shape[1]), dtype=np.complex64)], axis=1)
        sl = w.shape[1] - tail
        sh = s * L_ // M_
        wh = iq_hw[:, sh:sh + w.shape[1] * L_ // M_ + 8]
        if wh.shape[1] < w.shape[1] * L_ // M_ + 8:
            wh = np.concatenate([wh, np.zeros((wh.shape[0], w.shape[1] * L_ // M_ + 8 - wh.shape[1]), dtype=np.complex64)], axis=1)
        rep = search_rs(w, wh, u, b, os, N_eff_TX_max, sl, resample, ignore_before=pos - s)
        if rep.detected:
            fine, coarse = sh + rep.fine_peak_time_local, sh + rep.coarse_peak_time_hw
            if fine + n_pkt_guard <= iq_hw.shape[1]:
                found.append((fine, coarse, rep))
            pos = s + rep.coarse_peak_time_local + 2 * stf
        else:
            pos = s + sl
    return found


def stream_search(iq, u, b, os=1, N_eff_TX_max=1, window=None):
    """Sequential packet search over one continuous stream iq [n_ant][n], the way ONE sync_chunk_t walks an arbitrarily
    long chunk (sync_chunk.cpp:146-283): search, on a packet move ignore_before_index to coarse peak + 2 STFs
    (autocorrelator_detection.cpp:137-146, RX_SYNC_PARAM_AUTOCORRELATOR_DETECTION_SKIP_AFTER_PEAK_IN_STFS_DP = 2),
    continue.  Implemented with search() on slices that start on the global detection step grid one STF + one
    pattern (the detector's history) before the position to resume from.  Returns [(fine, coarse, Report)] with global
    sample indices."""
    iq = np.ascontiguousarray(iq, dtype=np.complex64)
    n = iq.shape[1]
    npat = 7 if u == 1 else 9
    stf = (112 if u == 1 else 144) * b * os
    pat = stf // npat
    step = pat // 4
    lead, tail = stf + pat, 2 * stf + 3 * pat
    if window is None:
        window = 64 * stf
    found = []
    pos = lead                       # first sample a detection may end on
    while pos + stf < n:
        s = -(-(pos - lead) // step) * step          # ceil to the step grid
        w = iq[:, s:s + lead + window + tail]
        if w.shape[1] < lead + stf + tail:
            w = np.concatenate([w, np.zeros((iq.shape[0], lead + stf + tail - w.shape[1]), dtype=np.complex64)], axis=1)
        sl = w.shape[1] - tail
        rep = search(w, u, b, os=os, N_eff_TX_max=N_eff_TX_max, search_length=sl)
        if rep.detected:
            fine, coarse = s + rep.fine_peak_time_local, s + rep.coarse_peak_time_local
            if fine + stf <= n:
                found.append((fine, coarse, rep))
            pos = coarse + 2 * stf
        else:
            pos = s + sl
    return found
