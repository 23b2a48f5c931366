"""TEST INFRASTRUCTURE (oracle).  Residual CFO / STO tracking on the DRS: float64 numpy specification of k_rx_track
(dect-nr-plus-sdr_b200/csrc/phy_rx.cu) and of the rotation the detection applies to every cell.

The reference plans this stage and leaves it switched off ("not functional yet"):
  lib/include/dectnrp/phy/rx/rx_synced/rx_synced_param.hpp:156-181   the two compile-time switches
  lib/src/phy/rx/rx_synced/offsets/estimator_cfo.cpp:31-70            process_drs: sum of products of neighbouring pilots
  lib/src/phy/rx/rx_synced/rx_synced.cpp:839-851                      feedback into the mixer, symbol by symbol
  lib/src/phy/rx/rx_synced/rx_synced.cpp:1165-1169                    post-correction in mode lr: empty todo
so the specification is this repo's own and parity is CUDA == this file, not CUDA == reference: PARITY UNPINNED.

Batched form: ONE estimate per packet from all its DRS symbols.  hzf[a][ts][n][i] = zero-forced pilot i of DRS occurrence n
of transmit stream ts at antenna a; the pilot sits at occupied-subcarrier position q = 4 i + (ts + 2 (n % 2)) % 4, i.e. at
k = q (q < N_OCC/2) or q + 1 (DC skipped), in OFDM symbol l = 1 + n N_step (+ 1 for ts >= 4).
  pilot i of occurrence n + 1 lies midway between two pilots of occurrence n (the DRS are interlaced by two subcarriers):
        P_i = h_{n+1}[i] conj((h_n[i'] + h_n[i'+1]) / 2): the channel cancels, the phase left is dphi + (k_i - N_OCC/2) dpsi
  dpsi = arg(sum P_{i+1} conj(P_i)) / 4    (pilots 4 subcarriers apart; the pair across DC is 5 apart and left out)
  dphi = arg(sum P_i e^{-j (k_i - N_OCC/2) dpsi})       sums over antennas, streams, occurrences n, pilots i
Both are phase advances per N_step OFDM symbols.  Cell (l, k) is multiplied by
  e^{-j (l - 1) / N_step (dphi + (k - N_OCC/2) dpsi)}."""
import numpy as np


def pilot_k(i, ts, n, N_OCC):
    q = 4 * np.asarray(i) + ((ts & 3) + 2 * (n & 1)) % 4
    return np.where(q < N_OCC // 2, q, q + 1)


def estimate(hzf, N_OCC, cfo=True, sto=True):
    """hzf complex [N_RX][N_eff_TX][n_drs][NP] -> (dphi, dpsi) in rad per N_step symbols"""
    h = np.asarray(hzf, dtype=np.complex128)
    N_RX, N_TS, n_drs, NP = h.shape
    dphi = dpsi = 0.0
    if n_drs < 2:
        return dphi, dpsi
    terms = []                                                    # (P over the valid pilots, their k, DC flags)
    for n in range(n_drs - 1):
        for ts in range(N_TS):
            pc = ((ts & 3) + 2 * (n & 1)) % 4
            lo = 0 if pc < 2 else 1
            idx = np.arange(lo, NP - 1 + lo)                      # pilots of occurrence n + 1 that have both neighbours
            ref = 0.5 * (h[:, ts, n, idx - lo] + h[:, ts, n, idx - lo + 1])
            q = 4 * idx + (pc + 2) % 4
            terms.append((h[:, ts, n + 1, idx] * np.conj(ref), pilot_k(idx, ts, n + 1, N_OCC), q))
    if sto:
        Zs = 0.0 + 0.0j
        for P, k, q in terms:
            keep = ~((q[:-1] < N_OCC // 2) & (q[:-1] + 4 >= N_OCC // 2))
            Zs += np.sum((P[:, 1:] * np.conj(P[:, :-1]))[:, keep])
        dpsi = float(np.angle(Zs)) / 4 if Zs != 0 else 0.0
    if cfo:
        Z = 0.0 + 0.0j
        for P, k, q in terms:
            Z += np.sum(P * np.exp(-1j * (k - N_OCC // 2) * dpsi)[None, :])
        dphi = float(np.angle(Z)) if Z != 0 else 0.0
    return dphi, dpsi


def phasor(l, k, dphi, dpsi, N_OCC, N_step):
    """factor of cell (OFDM symbol l = 1.., subcarrier index k = 0..N_OCC with DC at N_OCC/2)"""
    return np.exp(-1j * (np.asarray(l) - 1) / N_step * (dphi + (np.asarray(k) - N_OCC // 2) * dpsi))


def derotate_pilots(hzf, dphi, dpsi, N_OCC, N_step):
    h = np.array(hzf, dtype=np.complex128)
    N_RX, N_TS, n_drs, NP = h.shape
    i = np.arange(NP)
    for ts in range(N_TS):
        for n in range(n_drs):
            l = 1 + n * N_step + (1 if ts >= 4 else 0)
            h[:, ts, n, :] *= phasor(l, pilot_k(i, ts, n, N_OCC), dphi, dpsi, N_OCC, N_step)[None, :]
    return h
