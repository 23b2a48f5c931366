"""CPU checks of the two oracle restatements of spatial multiplexing receive against each other (oracle/rx.c
detect_sm in fp32 / Cholesky inside the sequential receiver, oracle/rx_mimo.py in float64 / numpy.linalg from the
Y and H snapshots) and of the 1024-QAM extension.  The reference has neither (rx_synced.cpp:1331-1333 stub,
tx.cpp:53-60), so nothing here can be pinned against it; see oracle/rx_mimo.py."""
import numpy as np
import pytest

from oracle import fec, rx, rx_mimo, spec, tx
from test_rx_gpu import make_windows

NET = 123456789


@pytest.mark.parametrize("cfg,N_RX,snr", [((1, 1, 1, 2, 2, 3, 2048), 2, 25.0), ((1, 2, 1, 2, 8, 5, 2048), 4, 30.0),
                                           ((1, 1, 1, 4, 9, 1, 2048), 4, 25.0), ((1, 1, 1, 6, 11, 1, 2048), 8, 30.0)])
def test_sm_restatements_agree_and_decode(cfg, N_RX, snr):
    ps = spec.get_packet_sizes(*cfg)
    rng = np.random.default_rng(11)
    case = dict(cfg=cfg, snr_db=snr, N_RX=N_RX, t0=7)
    win, tbs, plcfs = make_windows(ps, case, 2, ps["N_samples_packet"] + 64, rng, 1)
    orc = rx.RxConfig(ps, N_RX=N_RX)
    gold = fec.pdc_gold(NET, 1, ps["G"])
    n_ok = 0
    for i in range(2):
        r = orc.rx_packet(win[i], t0=7, want_H=True)
        llr, rho = rx_mimo.detect(r["Y"], r["H"], orc.pdc_l, orc.pdc_k, ps["N_bps"])
        assert rho.min() > 0 and rho.max() <= 1.0
        d = np.abs(llr.astype(int) - r["llr_pdc"].astype(int))
        # fp32 Cholesky against float64 inverse: rounding grows with the condition number of the cell
        assert np.mean(d <= 1) > 0.99 and d.max() <= 2 + 0.02 * np.abs(llr).max(), (d.max(), np.mean(d <= 1))
        ok, tb, nit = fec.pdc_decode(r["llr_pdc"], ps, 0, gold)
        if ok:   # an unlucky random channel may lose a packet; a decoded one must be the sent one
            n_ok += 1
            assert np.array_equal(tb, tbs[i])
        assert fec.pcc_decode(r["llr_pcc"], 1)[0]
    assert n_ok >= 1


def test_sm_orthogonal_columns_equal_mrc():
    """rho = 1 and x = the per-stream MRC when the columns of H are orthogonal: the arithmetic of the reference's single
    stream path (rx_synced.cpp:1204-1273)"""
    rng = np.random.default_rng(5)
    n_cells, N = 64, 2
    Nk = n_cells
    q, _ = np.linalg.qr(rng.standard_normal((n_cells, 4, 4)) + 1j * rng.standard_normal((n_cells, 4, 4)))
    Hc = q[:, :, :N] * rng.uniform(0.5, 2.0, (n_cells, 1, N))            # [cell][rx][ss], orthogonal columns
    x = tx.modulate(rng.integers(0, 2, n_cells * N * 4), 4).reshape(n_cells, N)
    y = np.einsum("cas,cs->ca", Hc, x)
    H = np.transpose(Hc, (1, 2, 0))[:, :, None, :]                        # [rx][ss][N_DF = 1][Nk]
    Y = np.transpose(y, (1, 0))[:, None, :]
    llr, rho = rx_mimo.detect(Y, H, np.ones(n_cells, dtype=int), np.arange(Nk), 4)
    assert np.allclose(rho, 1.0, atol=1e-9)
    mrc = np.einsum("cas,ca->cs", np.conj(Hc), y) / np.einsum("cas,cas->cs", np.conj(Hc), Hc).real
    ref = rx.demap(mrc.reshape(-1).astype(np.complex64), 4)
    assert np.abs(llr.astype(int) - ref.astype(int)).max() <= 1


@pytest.mark.parametrize("N_bps", [1, 2, 4, 6, 8, 10])
def test_modulate_demap_roundtrip(N_bps):
    """every constellation point demaps to its own bits (LLR > 0 <=> bit 1), 1024-QAM included; unit average power"""
    tab = tx.mod_table(N_bps)
    assert abs(np.mean(np.abs(tab) ** 2) - 1.0) < 1e-12
    idx = np.arange(1 << N_bps)
    bits = ((idx[:, None] >> (N_bps - 1 - np.arange(N_bps))) & 1).reshape(-1)
    sym = tx.modulate(bits, N_bps)
    llr = rx.demap(sym, N_bps)
    assert np.array_equal((llr > 0).astype(int), bits)
    if N_bps >= 4:   # Gray mapping: nearest neighbours on an axis differ in exactly one bit
        re = np.round(tab.real * np.sqrt({4: 10, 6: 42, 8: 170, 10: 682}[N_bps])).astype(int)
        im = np.round(tab.imag * np.sqrt({4: 10, 6: 42, 8: 170, 10: 682}[N_bps])).astype(int)
        pos = {(a, b): i for i, (a, b) in enumerate(zip(re, im))}
        for (a, b), i in pos.items():
            if (a + 2, b) in pos:
                assert bin(i ^ pos[(a + 2, b)]).count("1") == 1


def test_packet_sizes_1024qam():
    """MCS 10 / 11 geometry is the reference's (mcs.cpp:90-101: N_bps 10, R = 3/4, 5/6)"""
    for mcs, R in ((10, (3, 4)), (11, (5, 6))):
        ps = spec.get_packet_sizes(1, 1, 1, 2, 0, mcs, 2048)
        assert ps["N_bps"] == 10 and ps["G"] == ps["N_PDC_subc"] * 10
