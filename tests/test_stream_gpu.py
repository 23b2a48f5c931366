"""Streaming front door (dnrp_stream_sync / dnrp_stream_extract: the job of worker_sync_t, sync_chunk_t and baton_t)
against the oracle's sequential search over the same stream (oracle/sync.py::stream_search): packets at random gaps,
some back to back, chunk hops chosen so that STFs straddle chunk borders; then the found packets go through the batched
receiver and must decode to the transmitted transport blocks."""
import numpy as np
import pytest

from oracle import sync as osync

pytestmark = pytest.mark.gpu
NET = 123456789


def _make_stream(product, cfg, n_pkt, snr_db, seed, N_RX=1):
    rng = np.random.default_rng(seed)
    phy = product.Phy(cfg, max_batch=n_pkt, N_RX=N_RX, network_id=NET, plcf_type=1)
    ps = phy.ps
    tb = rng.integers(0, 256, (n_pkt, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n_pkt, 10), dtype=np.uint8)
    out = phy.tx(plcf, tb)
    iq = out[0] if isinstance(out, tuple) else out
    S = ps["N_samples_packet_no_GI"]
    stf = ps["N_samples_STF"]
    gaps = rng.integers(stf // 4, 2 * S, n_pkt)
    gaps[::5] = ps["N_samples_GI"]                     # every fifth packet follows its predecessor after the guard interval
    starts = np.zeros(n_pkt, dtype=np.int64)
    t = 3 * stf
    for i in range(n_pkt):
        starts[i] = t
        t += S + int(gaps[i])
    n = int(t + 4 * stf)
    sigma = 10 ** (-snr_db / 20)
    stream = (sigma / np.sqrt(2)) * (rng.standard_normal((N_RX, n)) + 1j * rng.standard_normal((N_RX, n)))
    for i in range(n_pkt):
        for a in range(N_RX):
            stream[a, starts[i]:starts[i] + S] += iq[i, 0, :S] * np.exp(1j * 0.7 * a)
    return phy, stream.astype(np.complex64), starts, tb, plcf


@pytest.mark.parametrize("cfg,hops", [((1, 2, 1, 1, 0, 3, 2048), (0, 1000, 2344)), ((1, 4, 0, 3, 0, 2, 2048), (1500,))])
def test_stream_search_matches_sequential_oracle(product, cfg, hops):
    n_pkt = 40
    phy, stream, starts, tb, plcf = _make_stream(product, cfg, n_pkt, snr_db=18.0, seed=3)
    ref = osync.stream_search(stream, cfg[0], cfg[1], N_eff_TX_max=1)
    ref_fine = np.array([f for f, _, _ in ref], dtype=np.int64)
    # the oracle itself finds every transmitted packet at its start
    assert np.array_equal(ref_fine, starts), (ref_fine[:5], starts[:5])
    for hop in hops:
        got = phy.stream_sync(stream, chunk_hop=hop)
        fine = got.fine_peak_time_64
        assert np.all(np.diff(fine) > 0)
        # every packet of the sequential search is found at the same sample; chunk borders may add detections on the
        # falling edge of an STF (the reference's overlap regions do the same), never closer than one pattern
        assert np.isin(ref_fine, fine).all(), (hop, np.setdiff1d(ref_fine, fine))
        extra = np.setdiff1d(fine, ref_fine)
        assert len(extra) <= 2, (hop, extra)
        sel = np.isin(fine, ref_fine)
        by_time = {f: r for f, _, r in ref}
        for g in got[sel]:
            r = by_time[int(g.fine_peak_time_64)]
            assert g.coarse_peak_time_64 - g.fine_peak_time_64 == r.coarse_peak_time_local - r.fine_peak_time_local
            assert g.sync.N_eff_TX == r.N_eff_TX
            assert abs(g.sync.cfo_fractional_rad - r.cfo_fractional_rad) < 2e-5
    phy.close()


def test_stream_packets_decode_through_the_batched_receiver(product):
    cfg = (1, 2, 1, 1, 0, 3, 2048)
    n_pkt = 32
    phy, stream, starts, tb, plcf = _make_stream(product, cfg, n_pkt, snr_db=22.0, seed=11, N_RX=2)
    # pinned host memory: the upload is truly asynchronous (regression: the zero fill of the freshly allocated device
    # stream buffer used to race with it on the first call)
    import torch
    pinned = torch.empty(stream.shape, dtype=torch.complex64, pin_memory=True)
    pinned.numpy()[...] = stream
    got = phy.stream_sync(pinned.numpy(), chunk_hop=3000)
    assert np.array_equal(got.fine_peak_time_64, starts)
    t0 = 64
    phy.stream_extract(got.fine_peak_time_64, t0)
    sync = phy.make_sync(n_pkt, t0=t0, cfo_fractional_rad=got.sync.cfo_fractional_rad, rms=got.sync.rms_array)
    pcc = phy.rx_pcc(None, n=n_pkt, sync=sync)
    ok, tb_out, snr, nit = phy.rx_pdc(n_pkt)
    assert pcc.plcf_ok[:, 0].all() and np.array_equal(pcc.plcf[:, 0, :5], plcf[:, :5])
    assert ok.all() and np.array_equal(tb_out, tb)
    phy.close()


def test_stream_edge_cases(product):
    """noise only: nothing found; a stream shorter than one STF plus its history is refused; one packet right at the
    earliest detectable position and one cut off at the end of the stream"""
    cfg = (1, 2, 1, 1, 0, 3, 2048)
    phy, stream, starts, tb, plcf = _make_stream(product, cfg, 3, snr_db=20.0, seed=5)
    rng = np.random.default_rng(1)
    noise = (0.05 * (rng.standard_normal((1, 20000)) + 1j * rng.standard_normal((1, 20000)))).astype(np.complex64)
    assert len(phy.stream_sync(noise)) == 0
    with pytest.raises(product.DnrpError):
        phy.stream_sync(noise[:, :300])
    # cut the stream in the middle of the last packet: its STF is still there, so it is reported (the receiver will
    # fail its CRC); cut inside its STF: not reported
    S = phy.ps["N_samples_packet_no_GI"]
    cut = stream[:, :starts[2] + S // 2]
    got = phy.stream_sync(np.ascontiguousarray(cut))
    assert np.array_equal(got.fine_peak_time_64, starts)
    cut = stream[:, :starts[2] + phy.ps["N_samples_STF"] // 2]
    got = phy.stream_sync(np.ascontiguousarray(cut))
    assert np.array_equal(got.fine_peak_time_64, starts[:2])
    phy.close()
