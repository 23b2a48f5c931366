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


# ---- stateful front door: dnrp_stream_open / dnrp_stream_push ------------------------------------------------------------
def _make_stream_hw(product, cfg, n_pkt, snr_db, seed, N_RX=1, os_=1, resample=(1, 1)):
    """like _make_stream, at the hardware rate of a context with oversampling / resampling"""
    rng = np.random.default_rng(seed)
    phy = product.Phy(cfg, max_batch=n_pkt, N_RX=N_RX, network_id=NET, plcf_type=1, oversampling=os_, resample=resample)
    ps = phy.ps
    tb = rng.integers(0, 256, (n_pkt, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n_pkt, 10), dtype=np.uint8)
    iq, ns = phy.tx(plcf, tb)
    S = int(ns[0])
    stf = ps["N_samples_STF"] * os_ * resample[0] // resample[1]
    gaps = rng.integers(stf // 4, 2 * S, n_pkt)
    gaps[::5] = S // 20                                # every fifth packet follows its predecessor almost back to back
    starts = np.zeros(n_pkt, dtype=np.int64)
    t = 3 * stf
    for i in range(n_pkt):
        starts[i] = t
        t += S + int(gaps[i])
    n = int(t + 6 * stf)
    sigma = 10 ** (-snr_db / 20)
    stream = (sigma / np.sqrt(2)) * (rng.standard_normal((N_RX, n)) + 1j * rng.standard_normal((N_RX, n)))
    for i in range(n_pkt):
        for a in range(N_RX):
            stream[a, starts[i]:starts[i] + S] += iq[i, 0, :S] * np.exp(1j * 0.7 * a)
    return phy, stream.astype(np.complex64), starts, tb, plcf, S


def _push_all(phy, stream, cuts, max_push, hop, decode=None):
    """feed stream[:, cuts[i]:cuts[i+1]] push by push; returns the records handed out, in order"""
    phy.stream_open(max_push, chunk_hop=hop)
    got = []
    for a, b in zip(cuts[:-1], cuts[1:]):
        r = phy.stream_push(np.ascontiguousarray(stream[:, a:b]))
        if len(r):
            assert (r.fine_peak_time_64 + 1 <= b).all()           # nothing from the future
            if decode is not None:
                decode(r)
            got.append(r.copy())
    phy.stream_close()
    return np.concatenate(got).view(np.recarray) if got else np.zeros(0)


@pytest.mark.parametrize("cfg,os_,resample,hop", [((1, 2, 1, 1, 0, 3, 2048), 1, (1, 1), 0), ((1, 2, 1, 1, 0, 3, 2048), 1, (1, 1), 1100),
                                                  ((1, 1, 1, 1, 0, 2, 2048), 2, (10, 9), 0),
                                                  ((1, 2, 1, 1, 0, 3, 2048), 2, (10, 9), 2000)])
def test_stream_push_matches_sequential_oracle(product, cfg, os_, resample, hop):
    """one long stream in 7 unequal pushes (packets and STFs straddle the cuts), without and with resampling 10/9 (C1,
    the reference's loopback configuration, is the third case): the oracle's packets at the oracle's samples, the same
    for a different cut of the same stream, and every packet handed out decodes"""
    n_pkt = 24
    phy, stream, starts, tb, plcf, S = _make_stream_hw(product, cfg, n_pkt, snr_db=20.0, seed=8, os_=os_, resample=resample)
    n = stream.shape[1]
    if resample == (1, 1):
        ref = osync.stream_search(stream, cfg[0], cfg[1], os=os_, N_eff_TX_max=1)
    else:
        ref = osync.stream_search_resampled(stream, cfg[0], cfg[1], os=os_, N_eff_TX_max=1, resample=resample)
    ref_fine = np.array([f for f, _, _ in ref], dtype=np.int64)
    assert np.array_equal(ref_fine, starts), (ref_fine[:6], starts[:6])
    rng = np.random.default_rng(5)
    # 7 pushes: cuts inside an STF, inside a packet body, one tiny push, one large
    cuts = np.array([0, starts[2] + 40, starts[5] + S // 2, starts[5] + S // 2 + 64, starts[11] - 3, starts[17] + 7, starts[20] + S - 1, n])
    assert np.all(np.diff(cuts) > 0)
    max_push = int(np.diff(cuts).max())
    t0 = 64
    decoded = {}

    def decode(r):
        m = len(r)
        phy.stream_extract(r.fine_peak_time_64, t0)
        sync = phy.make_sync(m, t0=t0, cfo_fractional_rad=r.sync.cfo_fractional_rad, rms=r.sync.rms_array)
        phy.rx_pcc(None, n=m, sync=sync)
        ok, tb_out, _, _ = phy.rx_pdc(m)
        for i in range(m):
            decoded[int(r.fine_peak_time_64[i])] = (bool(ok[i]), tb_out[i].copy())

    got = _push_all(phy, stream, cuts, max_push, hop, decode)
    fine = got.fine_peak_time_64
    assert np.all(np.diff(fine) > 0)
    assert np.isin(ref_fine, fine).all(), np.setdiff1d(ref_fine, fine)
    assert len(np.setdiff1d(fine, ref_fine)) <= 2
    by_time = {f: (c, r) for f, c, r in ref}
    for g in got[np.isin(fine, ref_fine)]:
        c, r = by_time[int(g.fine_peak_time_64)]
        assert g.coarse_peak_time_64 == c
        assert g.sync.N_eff_TX == r.N_eff_TX
        assert abs(g.sync.cfo_fractional_rad - r.cfo_fractional_rad) < 2e-5
    for i, s in enumerate(starts):
        ok, out = decoded[int(s)]
        assert ok and np.array_equal(out, tb[i]), i
    # a different cut of the same stream (equal pushes): the same packets
    cuts2 = np.linspace(0, n, 12).astype(np.int64)
    got2 = _push_all(phy, stream, cuts2, int(np.diff(cuts2).max()), hop)
    assert np.array_equal(np.intersect1d(got2.fine_peak_time_64, ref_fine), ref_fine)
    phy.close()


def test_stream_push_limits(product):
    cfg = (1, 2, 1, 1, 0, 3, 2048)
    phy = product.Phy(cfg, max_batch=4, network_id=NET, plcf_type=1)
    with pytest.raises(product.DnrpError, match="dnrp_stream_open first"):
        phy.stream_push(np.zeros((1, 16), dtype=np.complex64))
    phy.stream_open(1000)
    with pytest.raises(product.DnrpError, match="max_push_samples"):
        phy.stream_push(np.zeros((1, 1001), dtype=np.complex64))
    assert len(phy.stream_push(np.zeros((1, 1000), dtype=np.complex64))) == 0
    assert len(phy.stream_push(np.zeros((1, 0), dtype=np.complex64))) == 0
    phy.stream_close()
    # a chunk hop whose detection prefix sums cannot fit the shared memory of a block is refused with a message
    with pytest.raises(product.DnrpError, match="shared memory"):
        phy.stream_open(1000, chunk_hop=10_000_000)
    with pytest.raises(product.DnrpError, match="shared memory"):
        phy.stream_sync(np.zeros((1, 4000), dtype=np.complex64), chunk_hop=10_000_000)
    phy.close()


@pytest.mark.parametrize("cfg,os_,resample", [((1, 2, 1, 1, 0, 3, 2048), 1, (1, 1)), ((1, 1, 1, 1, 0, 2, 2048), 2, (10, 9))])
def test_stream_push_sc16_equals_push_of_the_dequantised_stream(product, cfg, os_, resample):
    """dnrp_stream_push_sc16 (the radio's wire format, converted on the device at whatever ring offset the push lands)
    hands out record for record what dnrp_stream_push hands out for the float stream int16 * scale, cut at odd positions"""
    phy, stream, starts, tb, plcf, S = _make_stream_hw(product, cfg, 16, snr_db=20.0, seed=11, os_=os_, resample=resample)
    n = stream.shape[1]
    gain = np.float32(0.1 / np.sqrt(np.mean(np.abs(stream) ** 2)) * 32768.0)
    s16 = np.clip(np.rint(stream.view(np.float32).reshape(stream.shape[0], n, 2) * gain), -32768, 32767).astype(np.int16)
    scale = np.float32(1.0) / gain
    deq = np.ascontiguousarray(s16.astype(np.float32) * scale).view(np.complex64).reshape(stream.shape[0], n)
    cuts = np.array([0, starts[1] + 37, starts[4] + S // 2 + 1, starts[4] + S // 2 + 66, starts[9] - 3, starts[13] + 7, n])
    max_push = int(np.diff(cuts).max())
    ref = _push_all(phy, deq, cuts, max_push, 0)
    phy.stream_open(max_push, chunk_hop=0)
    got = []
    for a, b in zip(cuts[:-1], cuts[1:]):
        r = phy.stream_push_sc16(np.ascontiguousarray(s16[:, a:b]), scale=float(scale))
        if len(r):
            got.append(r.copy())
    phy.stream_close()
    got = np.concatenate(got).view(np.recarray)
    assert len(got) == len(ref) and np.isin(starts, got.fine_peak_time_64).all()
    assert np.array_equal(got.fine_peak_time_64, ref.fine_peak_time_64)
    assert np.array_equal(got.coarse_peak_time_64, ref.coarse_peak_time_64)
    assert np.array_equal(got.sync.cfo_fractional_rad, ref.sync.cfo_fractional_rad)
    assert np.array_equal(got.sync.rms_array, ref.sync.rms_array)
    phy.close()
