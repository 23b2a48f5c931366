"""ORACLE (test infrastructure, not product code) -- numpy restatement of the reference's
`sections_part3` tables.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this package.

Every function cites the reference file:line it follows (paths relative to /root/reference/lib/src).
The structure deliberately mirrors the reference (virtual frame + per-symbol lists), NOT the product's
closed-form classification in dect-nr-plus-sdr_b200/csrc/spec_re.hpp, so the two are independent.

Parity status: the reference ships no golden vectors.  Integer tables here are pinned by
the reference's own internal asserts (counts 98 / N_DRS / G / C), restated as checks in
tests/test_spec.py; the 3GPP tables (QPP f1/f2, CRC polynomials, Gold sequence) are "parity unpinned"
against srsRAN release_23_11 (not in tree) and pinned only by their published definitions.
"""
import math
import numpy as np

PCC_CELLS = 98

# ---- sections_part3/tm_mode.cpp:27-131 (Table 7.2-1): N_eff_TX, N_SS, cl, N_TS, N_TX
TM_MODE = [(1, 1, 0, 1, 1), (2, 1, 0, 2, 2), (2, 2, 0, 2, 2), (1, 1, 1, 1, 2), (2, 2, 1, 2, 2),
           (4, 1, 0, 4, 4), (4, 4, 0, 4, 4), (1, 1, 1, 1, 4), (2, 2, 1, 2, 4), (4, 4, 1, 4, 4),
           (8, 1, 0, 8, 8), (8, 8, 0, 8, 8)]
# ---- sections_part3/mcs.cpp:27-107: N_bps, R_num, R_den
MCS = [(1, 1, 2), (2, 1, 2), (2, 3, 4), (4, 1, 2), (4, 3, 4), (6, 2, 3), (6, 3, 4), (6, 5, 6),
       (8, 3, 4), (8, 5, 6), (10, 3, 4), (10, 5, 6)]

# ---- sections_part3/fix/cbsegm.cpp:33-45 (TS 36.212 Table 5.1.3-3 sizes)
TC_CB_SIZES = ([40 + 8 * i for i in range(60)] + [528 + 16 * i for i in range(32)] +
               [1056 + 32 * i for i in range(32)] + [2112 + 64 * i for i in range(64)])

# ---- TS 36.212 Table 5.1.3-3 (f1, f2); srsRAN tc_interl_lte.c is not in tree ("parity unpinned")
_QPP = """3 10|7 12|19 42|7 16|7 18|11 20|5 22|11 24|7 26|41 84|103 90|15 32|9 34|17 108|9 38|21 120|101 84|21 44|57 46|23 48|13 50|27 52|11 36|27 56|85 58|29 60|33 62|15 32|17 198|33 68|103 210|19 36|19 74|37 76|19 78|21 120|21 82|115 84|193 86|21 44|133 90|81 46|45 94|23 48|243 98|151 40|155 102|25 52|51 106|47 72|91 110|29 168|29 114|247 58|29 118|89 180|91 122|157 62|55 84|31 64|17 66|35 68|227 420|65 96|19 74|37 76|41 234|39 80|185 82|43 252|21 86|155 44|79 120|139 92|23 94|217 48|25 98|17 80|127 102|25 52|239 106|17 48|137 110|215 112|29 114|15 58|147 118|29 60|59 122|65 124|55 84|31 64|17 66|171 204|67 140|35 72|19 74|39 76|19 78|199 240|21 82|211 252|21 86|43 88|149 60|45 92|49 846|71 48|13 28|17 80|25 102|183 104|55 954|127 96|27 110|29 112|29 114|57 116|45 354|31 120|59 610|185 124|113 420|31 64|17 66|171 136|209 420|253 216|367 444|265 456|181 468|39 80|27 164|127 504|143 172|43 88|29 300|45 92|157 188|47 96|13 28|111 240|443 204|51 104|51 212|451 192|257 220|57 336|313 228|271 232|179 236|331 120|363 244|375 248|127 168|31 64|33 130|43 264|33 134|477 408|35 138|233 280|357 142|337 480|37 146|71 444|71 120|37 152|39 462|127 234|39 158|39 80|31 96|113 902|41 166|251 336|43 170|21 86|43 174|45 176|45 178|161 120|89 182|323 184|47 186|23 94|47 190|263 480"""
QPP_F1F2 = [tuple(int(v) for v in r.split()) for r in _QPP.split("|")]
assert len(QPP_F1F2) == 188 == len(TC_CB_SIZES)


def qpp(K):
    f1, f2 = QPP_F1F2[TC_CB_SIZES.index(K)]
    i = np.arange(K, dtype=np.int64)
    return ((f1 * i + f2 * i * i) % K).astype(np.int64)


def cbsegm(tbs, Z):
    """fix/cbsegm.cpp:65-123 (srsran_cbsegm_FIX)."""
    B = tbs + 24
    if B <= Z:
        C, Bp = 1, B
    else:
        C = -(-B // (Z - 24))
        Bp = B + 24 * C
    long_cb = (Bp - 1) // C + 1
    idx1 = next((j for j, s in enumerate(TC_CB_SIZES) if s >= long_cb), None)
    if idx1 is None:
        return None
    K1 = TC_CB_SIZES[idx1]
    if C == 1:
        K2, C2, C1, K2_idx = 0, 0, 1, 0
    else:
        K2 = TC_CB_SIZES[idx1 - 1]
        K2_idx = idx1 - 1
        C2 = (C * K1 - Bp) // (K1 - K2)
        C1 = C - C2
    F = C1 * K1 + C2 * K2 - Bp
    return dict(tbs=tbs, Z=Z, C=C, K1=K1, K2=K2, K1_idx=idx1, K2_idx=K2_idx, C1=C1, C2=C2, F=F)


def get_N_DF_symb(u, N_PACKET_symb):
    """pdc.cpp:155-165."""
    return N_PACKET_symb - {1: 2, 2: 3, 4: 3, 8: 4}[u]


def get_N_step(N_eff_TX):
    return 5 if N_eff_TX <= 2 else 10


def nof_drs_symbols_per_TS(N_PACKET_symb, N_eff_TX):
    """pdc.cpp:167-196."""
    N_step = get_N_step(N_eff_TX)
    n = N_PACKET_symb // N_step
    if N_step == 10 and N_PACKET_symb % 10 != 0:
        n += 1
    return n


def get_N_TB_bits(N_PDC_bits, Z):
    """transport_block_size.cpp:39-79."""
    L = 24
    M = 8 if N_PDC_bits <= 512 else 16 if N_PDC_bits <= 1024 else 32 if N_PDC_bits <= 2048 else 64
    N_M = (N_PDC_bits // M) * M
    if N_M <= L:
        return 0
    if N_M <= Z:
        return N_M - L
    C = -(-(N_M - L) // Z)
    return N_M - (C + 1) * L


def get_packet_sizes(u, b, PacketLengthType, PacketLength, tm_mode_index, mcs_index, Z):
    """derivative/packet_sizes.cpp:97-227; returns None where the reference returns nullopt."""
    N_eff_TX, N_SS, cl, N_TS, N_TX = TM_MODE[tm_mode_index]
    N_bps, Rn, Rd = MCS[mcs_index]
    N_b_DFT, N_b_CP, N_b_OCC = 64 * b, 8 * b, 56 * b
    N_PACKET_symb = PacketLength * (u * 10) // (u * 2) if PacketLengthType == 0 else PacketLength * u * 10
    if N_eff_TX == 4 and N_PACKET_symb < 15:
        return None
    if u == 8 and N_eff_TX == 8 and (N_PACKET_symb < 20 or N_PACKET_symb % 10 != 0):
        return None
    N_DF_symb = get_N_DF_symb(u, N_PACKET_symb)
    N_DRS_subc = N_eff_TX * N_b_OCC // 4 * nof_drs_symbols_per_TS(N_PACKET_symb, N_eff_TX)
    if N_DF_symb * N_b_OCC <= N_DRS_subc + PCC_CELLS:
        return None
    N_PDC_subc = N_DF_symb * N_b_OCC - N_DRS_subc - PCC_CELLS
    G = N_SS * N_PDC_subc * N_bps
    N_PDC_bits = (G * Rn) // Rd
    N_TB_bits = get_N_TB_bits(N_PDC_bits, Z)
    if N_TB_bits == 0:
        return None
    seg = cbsegm(N_TB_bits, Z)
    if seg is None or seg["F"] > 0:
        return None
    sym = 72 * b
    N_samples_STF = sym * 14 // 9 if u == 1 else sym * 2
    N_samples_GI = sym * 4 // 9 if u == 1 else (sym * 2 if u == 8 else sym)
    return dict(u=u, b=b, PacketLengthType=PacketLengthType, PacketLength=PacketLength,
                tm_mode_index=tm_mode_index, mcs_index=mcs_index, Z=Z,
                N_eff_TX=N_eff_TX, N_SS=N_SS, cl=cl, N_TS=N_TS, N_TX=N_TX, N_bps=N_bps,
                N_b_DFT=N_b_DFT, N_b_CP=N_b_CP, N_b_OCC=N_b_OCC,
                N_PACKET_symb=N_PACKET_symb, N_DF_symb=N_DF_symb, N_DRS_subc=N_DRS_subc,
                N_PDC_subc=N_PDC_subc, G=G, N_PDC_bits=N_PDC_bits, N_TB_bits=N_TB_bits,
                N_TB_byte=-(-N_TB_bits // 8), C=seg["C"], seg=seg,
                N_samples_OFDM_symbol=sym, N_samples_STF=N_samples_STF,
                N_samples_STF_CP_only=N_samples_STF - 64 * b, N_samples_DF=sym * N_DF_symb,
                N_samples_GI=N_samples_GI, N_samples_packet_no_GI=N_samples_STF + sym * N_DF_symb,
                N_samples_packet=N_samples_STF + sym * N_DF_symb + N_samples_GI)


# ------------------------------------------------------------------------------------------------
# physical resources: physical_resources.cpp:25-35
def k_b_OCC(b):
    N = 56 * b
    return np.array(list(range(-N // 2, 0)) + list(range(1, N // 2 + 1)), dtype=np.int64)


# ---- STF: stf.cpp:159-287
_Y_B_1 = [1, -1, 1, 1, -1, 1, 1, -1, 1, 1, 1, -1, -1, -1]
_Y_B_2 = [-1, 1, -1, 1, 1, -1, 1, 1, -1, 1, 1, 1, -1, 1, -1, -1, -1, 1, -1, -1, -1, 1, 1, 1, -1, -1, -1, -1]
_Y_B_4 = [-1, -1, -1, 1, -1, 1, -1, -1, 1, 1, 1, 1, -1, 1, -1, -1, -1, 1, -1, 1, 1, -1, -1, -1, -1, -1, 1, -1,
          1, 1, 1, -1, 1, -1, 1, 1, -1, -1, -1, -1, 1, -1, -1, -1, -1, 1, -1, 1, 1, -1, -1, -1, -1, -1, 1, -1]
STF_COVER = [1.0, -1.0, 1.0, 1.0, -1.0, -1.0, -1.0, -1.0, -1.0]  # stf.hpp:146-151 (active)


def _fliplr_pm(v):
    v = list(v)[::-1]
    return [x * (1 if i % 2 == 0 else -1) for i, x in enumerate(v)]


def stf_filled(b, N_eff_TX, scale=1.0):
    """stf.cpp:34-82,159-287: complex vector of length N_b_OCC+1 (index 0 = lowest subcarrier)."""
    N = 56 * b
    if b == 1:
        pol = _Y_B_1
    elif b == 2:
        pol = _Y_B_2
    elif b == 4:
        pol = _Y_B_4
    else:
        y8 = _Y_B_4 + _fliplr_pm(_Y_B_4)
        if b == 8:
            pol = y8
        else:
            y16 = y8 + _fliplr_pm(y8)
            pol = y16[28:28 + 168] if b == 12 else y16
    occ = k_b_OCC(b)
    k_i = [occ[i * 4] for i in range(N // 8)] + [occ[N // 2 + 3 + (i - N // 8) * 4] for i in range(N // 8, N // 4)]
    lg = {1: 0, 2: 1, 4: 2, 8: 3}[N_eff_TX]
    fac = np.complex64(scale) * np.complex64(complex(np.float32(math.cos(math.pi / 4)), np.float32(math.sin(math.pi / 4))))
    out = np.zeros(N + 1, dtype=np.complex64)
    for i in range(N // 4):
        out[k_i[i] + N // 2] = np.complex64(pol[(i + 2 * lg) % (N // 4)]) * fac
    return out


# ---- DRS: drs.cpp:124-258, drs.hpp:124-132
_DRS_Y_B_1 = [1, 1, 1, 1, -1, 1, 1, -1, -1, 1, 1, 1, 1, -1, 1, -1, 1, 1, -1, 1, -1, 1, -1, 1, 1, 1, 1, 1, -1, 1,
              -1, -1, 1, 1, -1, -1, -1, -1, 1, -1, -1, -1, -1, -1, 1, 1, 1, -1, 1, 1, -1, -1, 1, -1, -1, -1]


def drs_k_i_single(b, t, n):
    """drs.cpp:218-238: subcarrier indices (0 = lowest occupied, N_b_OCC/2 = DC)."""
    N = 56 * b
    occ = k_b_OCC(b)
    return np.array([occ[i * 4 + (t + (n % 2) * 2) % 4] + N // 2 for i in range(N // 4)], dtype=np.int64)


def drs_y_single(b, t):
    """drs.cpp:240-256."""
    N = 56 * b
    s = 1 if t < 4 else -1
    return np.array([s * _DRS_Y_B_1[(4 * i + t % 4) % 56] for i in range(N // 4)], dtype=np.float32)


def drs_k_i_linear(u, b, N_PACKET_symb, N_TS):
    """drs.cpp:124-177: linear indices into a frame of N_b_DFT columns per symbol."""
    N_b_DFT, N = 64 * b, 56 * b
    occ = k_b_OCC(b)
    N_step = get_N_step(N_TS)
    n_symb = nof_drs_symbols_per_TS(N_PACKET_symb, N_TS)
    out = []
    for t in range(N_TS):
        row = []
        for n in range(n_symb):
            l = 1 + t // 4 + n * N_step
            off = N_b_DFT // 2 + N_b_DFT * l
            row += [occ[i * 4 + (t + (n % 2) * 2) % 4] + off for i in range(N // 4)]
        out.append(row)
    return out


def _virtual_frame(b, n_symb):
    N_b_DFT = 64 * b
    g_top = (N_b_DFT - 56 * b) // 2 - 1
    g_bot = g_top + 1
    vf = np.arange(n_symb * N_b_DFT, dtype=np.int64).reshape(n_symb, N_b_DFT)
    vf[:, N_b_DFT // 2] = -1
    vf[:, :g_bot] = -2
    vf[:, N_b_DFT - g_top:] = -2
    return vf, g_bot


def pcc_k_i_linear(b, N_TS):
    """pcc.cpp:110-246 (TS 103 636-3 clause 5.2.4), sorted linear indices of the 98 PCC cells."""
    N_b_DFT = 64 * b
    vf, _ = _virtual_frame(b, 20)
    flat = vf.reshape(-1)
    for row in drs_k_i_linear(8, b, 20, N_TS):
        flat[np.array(row)] = -3
    out = []
    l, unalloc = 1, PCC_CELLS
    while True:
        k0 = [int(v) for v in vf[l] if v >= 0]
        U = len(k0)
        if U < unalloc:
            out += k0
            l += 1
            unalloc -= U
            continue
        R = 7
        Cc = U // R
        assert U % R == 0
        done = False
        for c in range(Cc):
            for r in range(R):
                out.append(k0[r * Cc + c])
                unalloc -= 1
                if unalloc == 0:
                    done = True
                    break
            if done:
                break
        break
    assert len(out) == PCC_CELLS
    return sorted(out)


def pdc_k_i_linear(b, N_TS, u, N_PACKET_symb):
    """pdc.cpp:221-310, but for the actual packet length instead of the 30-symbol template + repetition
    (pdc.cpp:40-153 compresses the same set; tx.cpp:299-313 asserts the counts)."""
    N_b_DFT = 64 * b
    n_rows = max(N_PACKET_symb, 21)
    vf, _ = _virtual_frame(b, n_rows)
    flat = vf.reshape(-1)
    for row in drs_k_i_linear(u, b, N_PACKET_symb, N_TS):
        flat[np.array(row)] = -3
    flat[np.array(pcc_k_i_linear(b, N_TS))] = -4
    N_DF = get_N_DF_symb(u, N_PACKET_symb)
    out = [int(v) for l in range(1, 1 + N_DF) for v in vf[l] if v >= 0]
    return out


def lin2lk(lin, b):
    """pcc.cpp:248-315 idx2sub_*: linear index -> (symbol l, subcarrier k with 0 = lowest occupied)."""
    N_b_DFT = 64 * b
    g_bot = (N_b_DFT - 56 * b) // 2
    lin = np.asarray(lin, dtype=np.int64)
    return lin // N_b_DFT, lin % N_b_DFT - g_bot


# ---- transmit diversity: transmit_diversity_precoding.cpp:45-85
TXDIV_PAIRS = {2: [(0, 1)],
               4: [(0, 1), (2, 3), (0, 2), (1, 3), (0, 3), (1, 2)],
               8: [(0, 1), (2, 3), (4, 5), (6, 7), (0, 4), (1, 5), (2, 6), (3, 7), (0, 2), (1, 3), (4, 6), (5, 7)]}


# ---- Gold sequence, TS 36.211 7.2 (srsran_sequence_LTE_pr; call sites scrambling_pdc.cpp:43-56,
# phy/fec/pcc_enc.cpp:104-107)
def gold_sequence(c_init, length):
    Nc = 1600
    n = Nc + length + 31
    x1 = np.zeros(n, dtype=np.uint8)
    x2 = np.zeros(n, dtype=np.uint8)
    x1[0] = 1
    for i in range(31):
        x2[i] = (c_init >> i) & 1
    for i in range(n - 31):
        x1[i + 31] = x1[i + 3] ^ x1[i]
        x2[i + 31] = x2[i + 3] ^ x2[i + 2] ^ x2[i + 1] ^ x2[i]
    return (x1[Nc:Nc + length] ^ x2[Nc:Nc + length]).astype(np.uint8)
