"""TEST INFRASTRUCTURE (oracle): frequency-domain stage of the packet search between coarse and fine peak --
coarse_peak_f_domain_t::process (reference lib/src/phy/rx/sync/coarse_peak_f_domain.cpp:70-201).  The reference compiles both
parts out (sync_param.hpp:244, 252) and leaves the integer CFO as "#error not implemented", so this file is the written
specification the CUDA kernel k_sync_fdomain is held against, in float64 with numpy's FFT: parity unpinned
against the reference by necessity.

  spectrum: the N_os samples behind the STF's cyclic prefix at the coarse peak, times exp(j cfo_fractional_rad n) (the
      correction increment of the sync report), patterns 3..6 of the STF times their cover sequence (1, -1, -1, -1);
  beta search (estimate_beta, :94-190): P[c] = power of subcarrier c summed over the antennas; b = 1; for the next larger
      bt of (1, 2, 4, 8, 12, 16) <= b_max: side = sum of P over c in [32 b, 32 bt) and [-32 bt, -32 b), accepted while
      (side / ((bt - b) 64)) / (central / (b 64)) > 10^(threshold_dB / 20), central += side; bands symmetric (the reference
      moves its right-hand index by the new width, :150-151 -- a slip in code that is never compiled);
  integer CFO (own): for s in [-S, S], S = range // 4, and cyclic shifts 2 lg (N_eff_TX = 2^lg <= N_eff_TX_max):
      D = sum_a sum_i X_a[c_i + 4 s] conj(X_a[c_(i+1) + 4 s]) y_i y_(i+1), c_i = the STF's subcarriers (multiples of 4
      from -28 b to 28 b without 0; the pair across DC is left out), y = the STF sequence of b; arg max of abs(D).
"""
import numpy as np

from . import spec

TAB = (1, 2, 4, 8, 12, 16)


def stf_sequence(b):
    """y_i of the STF cells in comb order (lower half first), N_eff_TX = 1"""
    N_OCC = 56 * b
    full = np.asarray(spec.stf_filled(b, 1))            # [N_OCC + 1] complex, zero where empty
    c = np.concatenate([np.arange(-N_OCC // 2, 0, 4), np.arange(4, N_OCC // 2 + 1, 4)])
    v = full[c + N_OCC // 2]
    return np.sign(v.real).astype(np.int64), c


def fdomain(win, coarse, cfo_fractional_rad, b_max, os, stf_cp_os, beta_threshold_dB=0.0, cfo_range=0, N_eff_TX_max=1):
    """win [n_ant][len] complex; returns (b, integer CFO of the signal in subcarriers)"""
    N = 64 * b_max * os
    start = int(coarse) + int(stf_cp_os)
    n = np.arange(N)
    cover = np.where(n < N // 4, 1.0, -1.0)
    X = []
    for a in range(win.shape[0]):
        x = win[a, start:start + N].astype(np.complex128) * np.exp(1j * float(cfo_fractional_rad) * n) * cover
        X.append(np.fft.fft(x))
    X = np.array(X)
    P = (np.abs(X) ** 2).sum(0)
    b = b_max
    if beta_threshold_dB < 0 and b_max > 1:
        thr = 10.0 ** (beta_threshold_dB / 20.0)
        band = lambda lo, hi: sum(P[c % N] + P[(-c - 1) % N] for c in range(lo, hi))
        b = 1
        central = band(0, 32)
        for bt in TAB[1:]:
            if bt > b_max:
                break
            side = band(32 * b, 32 * bt)
            if (side / ((bt - b) * 64)) / (central / (b * 64)) > thr:
                b, central = bt, central + side
            else:
                break
    s_best = 0
    S = min(cfo_range // 4, 8)
    if S:
        y, c = stf_sequence(b)
        NS = len(y)
        best = -1.0
        lg_max = int(np.log2(N_eff_TX_max))
        for lg in range(lg_max + 1):
            ys = np.roll(y, -2 * lg)
            for s in range(-S, S + 1):
                D = 0j
                for i in range(NS - 1):
                    if i + 1 == NS // 2:
                        continue
                    c0 = c[i] + 4 * s
                    D += (X[:, c0 % N] * np.conj(X[:, (c0 + 4) % N])).sum() * ys[i] * ys[i + 1]
                if abs(D) ** 2 > best:
                    best, s_best = abs(D) ** 2, s
    return b, 4 * s_best
