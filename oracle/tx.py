"""ORACLE (test infrastructure): numpy restatement of the reference transmitter
(lib/src/phy/tx/tx.cpp:165-314 generate_tx_packet and its run_* helpers), symbol by symbol like the
reference, at the DECT NR+ sample rate times the oversampling factor, without resampling (L = M = 1;
the resampler has its own oracle in oracle/resample.py).

FFT: the reference calls FFTW3f through srsran_dft (unnormalised c2c, no mirroring, ofdm.cpp:62-79);
here numpy's double precision FFT is used and the result cast to complex64 -> "math-pinned", fp32
tolerance applies to every IQ comparison.
"""
import functools
import math
import numpy as np
from . import fec, spec

# ---- LTE modulation tables (TS 36.211 7.1; srsRAN modem_table.c/lte_tables.c are not in tree),
# index = bits MSB first; used by reference tx.cpp:55-64 / fix/mod.cpp:137-174


@functools.lru_cache(maxsize=None)
def mod_table(N_bps):
    # N_bps = 10 (1024-QAM, MCS 10/11) is outside the reference (tx.cpp:53-60 stops at 256-QAM): the same recursion one
    # level deeper, normalised by sqrt(682) -- own extension, DESIGN.md "1024-QAM"
    if N_bps == 1:
        return np.array([(1 + 1j), (-1 - 1j)], dtype=np.complex128) / math.sqrt(2)
    half = N_bps // 2
    norm = {1: 2.0, 2: 10.0, 3: 42.0, 4: 170.0, 5: 682.0}[half]
    tab = np.zeros(1 << N_bps, dtype=np.complex128)
    for idx in range(1 << N_bps):
        bits = [(idx >> (N_bps - 1 - i)) & 1 for i in range(N_bps)]

        def axis(bs):
            # bs = (b0, b2, b4, b6)[:half]: amplitude = (1-2b0)*(2^(h-1) - (1-2b2)*(2^(h-2) - ...))
            acc = 0.0
            for i in range(half - 1, 0, -1):
                acc = (1 - 2 * bs[i]) * ((1 << (half - 1 - i)) - acc)
            return (1 - 2 * bs[0]) * ((1 << (half - 1)) - acc)

        tab[idx] = complex(axis(bits[0::2]), axis(bits[1::2])) / math.sqrt(norm)
    return tab


def modulate(bits, N_bps):
    """fix/mod.cpp:137-174: bits (one per element) -> complex64 symbols"""
    bits = np.asarray(bits, dtype=np.int64).reshape(-1, N_bps)
    idx = np.zeros(len(bits), dtype=np.int64)
    for i in range(N_bps):
        idx = (idx << 1) | bits[:, i]
    return mod_table(N_bps)[idx].astype(np.complex64)


# ---- codebooks (beamforming_and_antenna_port_mapping.cpp:251-281), j coded as 2
_j = 2
W_TABLES = {
    (1, 1): [[1]],
    (1, 2): [[1, 0], [0, 1], [1, 1], [1, -1], [1, _j], [1, -_j]],
    (1, 4): [[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0], [0, 0, 0, 1], [1, 0, 1, 0], [1, 0, -1, 0], [1, 0, _j, 0],
             [1, 0, -_j, 0], [0, 1, 0, 1], [0, 1, 0, -1], [0, 1, 0, _j], [0, 1, 0, -_j], [1, 1, 1, 1], [1, 1, _j, _j],
             [1, 1, -1, -1], [1, 1, -_j, -_j], [1, _j, 1, _j], [1, _j, _j, -1], [1, _j, -1, -_j], [1, _j, -_j, 1],
             [1, -1, 1, -1], [1, -1, _j, -_j], [1, -1, -1, 1], [1, -1, -_j, _j], [1, -_j, 1, -_j], [1, -_j, _j, 1],
             [1, -_j, -1, _j], [1, -_j, -_j, -1]],
    (2, 2): [[1, 0, 0, 1], [1, 1, 1, -1], [1, 1, _j, -_j]],
    (2, 4): [[1, 0, 0, 1, 0, 0, 0, 0], [1, 0, 0, 0, 0, 1, 0, 0], [1, 0, 0, 0, 0, 0, 0, 1], [0, 0, 1, 0, 0, 1, 0, 0],
             [0, 0, 1, 0, 0, 0, 0, 1], [0, 0, 0, 0, 1, 0, 0, 1], [1, 0, 0, 1, 1, 0, 0, -_j], [1, 0, 0, 1, 1, 0, 0, _j],
             [1, 0, 0, 1, -_j, 0, 0, 1], [1, 0, 0, 1, -_j, 0, 0, -1], [1, 0, 0, 1, -1, 0, 0, -_j],
             [1, 0, 0, 1, -1, 0, 0, _j], [1, 0, 0, 1, _j, 0, 0, 1], [1, 0, 0, 1, _j, 0, 0, -1],
             [1, 1, 1, 1, 1, -1, 1, -1], [1, 1, 1, 1, _j, -_j, _j, -_j], [1, 1, _j, _j, 1, -1, _j, -_j],
             [1, 1, _j, _j, _j, -_j, -1, 1], [1, 1, -1, -1, 1, -1, -1, 1], [1, 1, -1, -1, _j, -_j, -_j, _j],
             [1, 1, -_j, -_j, 1, -1, -_j, _j], [1, 1, -_j, -_j, _j, -_j, 1, -1]],
    (4, 4): [[1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1], [1, 1, 0, 0, 0, 0, 1, 1, 1, -1, 0, 0, 0, 0, 1, -1],
             [1, 1, 0, 0, 0, 0, 1, 1, _j, -_j, 0, 0, 0, 0, _j, -_j],
             [1, 1, 1, 1, 1, -1, 1, -1, 1, 1, -1, -1, 1, -1, -1, 1],
             [1, 1, 1, 1, 1, -1, 1, -1, _j, _j, -_j, -_j, _j, -_j, -_j, _j]],
    (8, 8): [[1 if r == c else 0 for r in range(8) for c in range(8)]],
}


def codebook(N_TS, N_TX, idx):
    code = {0: 0, 1: 1, -1: -1, 2: 1j, -2: -1j}
    W = np.array([code[v] for v in W_TABLES[(N_TS, N_TX)][idx]], dtype=np.complex64).reshape(N_TX, N_TS)
    scale = np.float32(1.0) / np.sqrt(np.float32(np.count_nonzero(W)))
    return W, scale


def tx_packet(ps, plcf_bytes, tb_bytes, network_id=123456789, plcf_type=1, rv=0, codebook_index=0, os=1,
              DAC_scale=1.0, iq_phase_rad=0.0, iq_phase_increment=0.0, GI_percentage=5, debug=None,
              resample=(1, 1)):
    """returns complex64 [N_TX][N_samples_transmit] (tx.cpp:165-314).  resample = (L, M): every OFDM symbol goes
    through the streaming resampler (tx.cpp:913-934), the history is flushed at the end (:679-704), the mixer runs at
    the hardware rate."""
    u, b = ps["u"], ps["b"]
    N_TS, N_TX, N_SS, N_eff_TX = ps["N_TS"], ps["N_TX"], ps["N_SS"], ps["N_eff_TX"]
    N_OCC = ps["N_b_OCC"]
    N_os = ps["N_b_DFT"] * os
    cp_os = ps["N_b_CP"] * os
    stf_cp_os = ps["N_samples_STF_CP_only"] * os
    txdiv = ps["tm_mode_index"] in (1, 5, 10)
    W, Wscale = codebook(N_TS, N_TX, codebook_index)
    cl, bf = bool(ps["cl"]), codebook_index > 0
    scale_common = np.float32(DAC_scale) * Wscale
    final_scale_STF = np.float32(1.0) / np.sqrt(np.float32(N_OCC // 4)) * scale_common
    final_scale = np.float32(1.0) / np.sqrt(np.float32(N_OCC)) * scale_common

    # FEC (tx.cpp:199-202)
    nb = 5 if plcf_type == 1 else 10
    pcc_bits = fec.unpack_bits(fec.pcc_encode(np.asarray(plcf_bytes, dtype=np.uint8)[:nb], plcf_type, cl, bf), 196)
    gold = fec.pdc_gold(network_id, plcf_type, ps["G"])
    pdc_bits = fec.unpack_bits(fec.pdc_encode(tb_bytes, ps, rv, gold), ps["G"])
    pcc_sym = modulate(pcc_bits, 2)
    pdc_sym = modulate(pdc_bits, ps["N_bps"])

    def flip(sym):  # tx.cpp:654-676 / 1068-1082: swap pairs, then (-re,+im), (+re,-im)
        f = np.empty_like(sym)
        f[0::2] = sym[1::2]
        f[1::2] = sym[0::2]
        f[0::2] = -np.conj(f[0::2])
        f[1::2] = np.conj(f[1::2])
        return f

    pcc_flip = flip(pcc_sym) if N_TS > 1 else None
    pdc_flip = flip(pdc_sym) if txdiv else None

    # resource element lists
    pcc_l, pcc_k = spec.lin2lk(spec.pcc_k_i_linear(b, N_TS), b)
    pdc_l, pdc_k = spec.lin2lk(spec.pdc_k_i_linear(b, N_TS, u, ps["N_PACKET_symb"]), b)
    N_step = spec.get_N_step(N_TS)
    n_drs = spec.nof_drs_symbols_per_TS(ps["N_PACKET_symb"], N_TS)

    n_no_gi = ps["N_samples_packet_no_GI"] * os
    n_tx = n_no_gi + (ps["N_samples_packet"] * os - n_no_gi) * GI_percentage // 100
    out = np.zeros((N_TX, n_tx), dtype=np.complex64)

    def ofdm(streams, n_nonzero, cp, scale):
        # run_beamforming (tx.cpp:729-860) + run_ifft_cp_scale (:862-911)
        ant = (W[:, :n_nonzero].astype(np.complex64) @ streams[:n_nonzero]).astype(np.complex64) * scale
        X = np.zeros((N_TX, N_os), dtype=np.complex64)
        X[:, :N_OCC // 2 + 1] = ant[:, N_OCC // 2:]
        X[:, N_os - N_OCC // 2:] = ant[:, :N_OCC // 2]
        x = (np.fft.ifft(X.astype(np.complex128), axis=1) * N_os).astype(np.complex64)
        idx = (np.arange(cp + N_os) - cp) % N_os
        return x[:, idx], X

    pos = 0
    # STF (tx.cpp:216-237)
    streams = np.zeros((N_TS, N_OCC + 1), dtype=np.complex64)
    streams[0] = spec.stf_filled(b, N_eff_TX)
    sym, X = ofdm(streams, 1, stf_cp_os, final_scale_STF)
    npat = 7 if u == 1 else 9
    plen = 16 * b * os
    for p in range(npat):
        sym[:, p * plen:(p + 1) * plen] *= np.float32(spec.STF_COVER[p])
    out[:, pos:pos + sym.shape[1]] = sym
    pos += sym.shape[1]
    if debug is not None:
        debug["X"] = [X]

    PCC_idx = PDC_idx = 0
    for l in range(1, ps["N_DF_symb"] + 1):
        streams = np.zeros((N_TS, N_OCC + 1), dtype=np.complex64)
        # PCC (tx.cpp:963-988)
        sel = np.nonzero(pcc_l == l)[0]
        for q in sel:
            k = pcc_k[q]
            if N_TS == 1:
                streams[0, k] = pcc_sym[q]
            else:
                A, B = spec.TXDIV_PAIRS[N_TS][(q // 2) % len(spec.TXDIV_PAIRS[N_TS])]
                streams[A, k] = pcc_sym[q]
                streams[B, k] = pcc_flip[q]
        # DRS (tx.cpp:990-1002)
        for t in range(N_TS):
            for n in range(n_drs):
                if l == 1 + t // 4 + n * N_step:
                    streams[t, spec.drs_k_i_single(b, t, n)] = spec.drs_y_single(b, t)
        # PDC (tx.cpp:1004-1116)
        sel = np.nonzero(pdc_l == l)[0]
        if not txdiv:
            for q in sel:
                for s in range(N_SS):
                    streams[s, pdc_k[q]] = pdc_sym[q * N_SS + s]
        else:
            for q in sel:
                A, B = spec.TXDIV_PAIRS[N_TS][(q // 2) % len(spec.TXDIV_PAIRS[N_TS])]
                streams[A, pdc_k[q]] = pdc_sym[q]
                streams[B, pdc_k[q]] = pdc_flip[q]
        sym, X = ofdm(streams, N_TS, cp_os, final_scale)
        out[:, pos:pos + sym.shape[1]] = sym
        pos += sym.shape[1]
        if debug is not None:
            debug["X"].append(X)
    assert pos == n_no_gi
    L, M = resample
    if L != M:
        from . import resample as rs
        blocks = [stf_cp_os + N_os] + [cp_os + N_os] * ps["N_DF_symb"]
        n_no_gi_rs = -(-n_no_gi * L // M)                       # tx.cpp:527-528
        n_packet_rs = ps["N_samples_packet"] * os // M * L      # tx.cpp:541-542
        n_tx = n_no_gi_rs + (n_packet_rs - n_no_gi_rs) * GI_percentage // 100
        res = np.zeros((N_TX, n_tx), dtype=np.complex64)
        for a in range(N_TX):
            y = rs.stream(out[a, :n_no_gi], L, M, os, blocks=blocks)
            assert y.size >= n_no_gi_rs
            res[a, :n_no_gi_rs] = y[:n_no_gi_rs]
        out = res
    # mixer (tx.cpp:913-934), phase continuous from the first sample at the output rate
    if iq_phase_rad != 0.0 or iq_phase_increment != 0.0:
        n = np.arange(n_tx, dtype=np.float64)
        out = (out * np.exp(1j * (iq_phase_rad + iq_phase_increment * n))).astype(np.complex64)
    if debug is not None:
        debug.update(pcc_bits=pcc_bits, pdc_bits=pdc_bits, pcc_sym=pcc_sym, pdc_sym=pdc_sym)
    return out
