"""throughput of the streaming front door on a C2 stream: n back-to-back-ish packets in one continuous stream,
dnrp_stream_sync (H2D of the stream included) + extract + PCC + PDC; prints samples/s and packets/s"""
import sys, time
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
from conftest import load_product
product = load_product()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
cfg = (1, 16, 1, 1, 0, 7, 6144)
phy = product.Phy(cfg, max_batch=n, N_RX=1, network_id=123456789)
ps = phy.ps
rng = np.random.default_rng(1)
tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
out = phy.tx(plcf, tb)
iq = out[0] if isinstance(out, tuple) else out
S, gi, stf = ps["N_samples_packet_no_GI"], ps["N_samples_GI"], ps["N_samples_STF"]
starts = 3 * stf + np.arange(n, dtype=np.int64) * (S + gi) + rng.integers(0, 200, n).cumsum()
N = int(starts[-1] + S + 4 * stf)
sigma = 10 ** (-25 / 20)
stream = ((sigma / np.sqrt(2)) * (rng.standard_normal(N, dtype=np.float32) + 1j * rng.standard_normal(N, dtype=np.float32))).astype(np.complex64)[None, :]
for i in range(n):
    stream[0, starts[i]:starts[i] + S] += iq[i, 0, :S]
import torch
pinned = torch.empty(stream.shape, dtype=torch.complex64, pin_memory=True)
pinned.numpy()[...] = stream
stream = pinned.numpy()
for rep in range(3):
    t0 = time.perf_counter()
    got = phy.stream_sync(stream)
    t1 = time.perf_counter()
    phy.stream_extract(got.fine_peak_time_64, 64)
    sync = phy.make_sync(len(got), t0=64, cfo_fractional_rad=got.sync.cfo_fractional_rad, rms=got.sync.rms_array)
    pcc = phy.rx_pcc(None, n=len(got), sync=sync)
    ok, tb_out, _, _ = phy.rx_pdc(len(got))
    t2 = time.perf_counter()
    print("stream %d samples (%.1f MB), %d packets: found %d (exact %d), rounds max %d, sync %.1f ms = %.2f GS/s, +rx %.1f ms, decoded %d, %.0f packets/s end to end" % (
        N, N * 8 / 1e6, n, len(got), int(np.isin(got.fine_peak_time_64, starts).sum()), int(got["round"].max()), (t1 - t0) * 1e3, N / (t1 - t0) / 1e9,
        (t2 - t1) * 1e3, int(ok.sum()), len(got) / (t2 - t0)))
