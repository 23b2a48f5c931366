"""ORACLE (test infrastructure, not product code) -- spatial multiplexing receive, transmission modes 2, 4, 6, 8, 9, 11.

The reference has no implementation to follow here: rx_synced_t::run_pdc_mode_AxA_MIMO is an empty stub
(rx_synced.cpp:1331-1333, "todo") and the PCC of such packets is even discarded while
RX_SYNCED_PARAM_BLOCK_N_SS_TX_LARGER_1_AT_PCC is defined (rx_synced.cpp:307-318).  Parity for this stage is therefore
UNPINNED by construction; what is pinned is everything around it (layer order of tx.cpp:1054-1060, channel estimates,
demapper, FEC).  The specification is this repo's own (DESIGN.md "Spatial multiplexing receive"):

  per PDC cell (l, k), with H[a][s] the Wiener estimate of stream s at antenna a as held when symbol l is collected
  and y[a] the received cell:
      A = H^H H,  b = H^H y,  x = A^-1 b                       zero forcing
      rho_s = min(1, 1 / ([A^-1]_ss A_ss))                     loss of stream s against its matched filter bound
  per packet: rho_ref = mean of rho over the cells q = 0, 8, 16, ... (all streams)
      llr(q, s, :) = trunc(min(rho_s / rho_ref, 4) * demap(x_s))   demap = the reference's int16 demapper (srsRAN scales)
  LLRs of cell q, stream s at (q N_SS + s) N_bps.  rho = 1 for orthogonal columns: the arithmetic then equals the
  reference's MRC (rx_synced.cpp:1204-1273); a flat channel keeps the demapper's scale (decoder input range +-1023).

Two restatements exist on purpose: oracle/rx.c detect_sm (fp32, Cholesky, the order the CUDA kernel uses, inside the
sequential receiver so that the processing-stage bookkeeping is shared) and this file (float64, numpy.linalg, from the
Y / H snapshots), checked against each other in tests/test_rx_mimo_cpu.py.
"""
import numpy as np

from . import rx


def detect(Y, H, pdc_l, pdc_k, N_bps):
    """Y [N_RX][N_DF][Nk], H [N_RX][N_SS][N_DF][Nk] (oracle/rx.py rx_packet(want_Y, want_H)), cell positions in
    collection order -> (llr int16 [n_cells * N_SS * N_bps], rho float64 [n_cells][N_SS])"""
    Y = np.asarray(Y, dtype=np.complex128)
    H = np.asarray(H, dtype=np.complex128)
    N_SS = H.shape[1]
    l = np.asarray(pdc_l) - 1
    k = np.asarray(pdc_k)
    Hc = np.transpose(H[:, :, l, k], (2, 0, 1))           # [cell][rx][ss]
    yc = np.transpose(Y[:, l, k], (1, 0))[:, :, None]      # [cell][rx][1]
    Hh = np.conj(np.transpose(Hc, (0, 2, 1)))
    A = Hh @ Hc
    b = Hh @ yc
    Ainv = np.linalg.inv(A)
    x = (Ainv @ b)[:, :, 0]                                # [cell][ss]
    g = np.real(np.einsum("cii->ci", A))
    v = np.real(np.einsum("cii->ci", Ainv))
    rho = np.minimum(1.0, 1.0 / (v * g))
    w = np.minimum(rho / max(float(np.mean(rho[0::8])), 1e-12), 4.0)
    raw = rx.demap(x.reshape(-1).astype(np.complex64), N_bps).reshape(len(l), N_SS, N_bps)
    llr = np.trunc(w[:, :, None] * raw.astype(np.float64))
    return np.clip(llr, -32768, 32767).astype(np.int16).reshape(-1), rho
