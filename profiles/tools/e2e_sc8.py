"""host-buffer call of the north-star workload with UHD's 8-bit wire format (dnrp_rx_batch_sc8) next to the int16 one: P
windows in pinned host memory, the whole call timed on the host clock (copies inside), best of a few repetitions"""
import os
import sys
import time
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import load_product
product = load_product()
P = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
cfg = (1, 16, 1, 2, 5, 7, 6144)
phy = product.Phy(cfg, max_batch=P, N_RX=4, network_id=123456789)
ps = phy.ps
rng = np.random.default_rng(1)
tb = rng.integers(0, 256, (P, ps["N_TB_byte"]), dtype=np.uint8)
plcf = rng.integers(0, 256, (P, 10), dtype=np.uint8)
phy.tx(plcf, tb, fetch=False)
stride, stf, pat = phy.rx_stride(), ps["N_samples_STF"], 16 * 16
t0 = rng.integers(2 * pat, 2 * stf - 4, P).astype(np.uint32)
win = phy.channel(P, snr_dB=25.0, t0=t0, mode=2, seed=3, pdp_idx=1, tau_rms_ns=100.0, fD_Hz=50.0, fetch=True)
rms = float(np.sqrt(np.mean(np.abs(win[:64]) ** 2)))
search_len = int(stride - ps["N_samples_packet"] + stf + 2 * pat)
tb_pinned = torch.empty((P, ps["N_TB_byte"]), dtype=torch.uint8, pin_memory=True)
for name, dt, full, backoff in (("sc16", torch.int16, 32768.0, 0.1), ("sc8", torch.int8, 128.0, 0.25)):
    gain = np.float32(backoff / rms * full)
    host = torch.empty((P, 4, stride, 2), dtype=dt, pin_memory=True)
    w = host.numpy()
    for i0 in range(0, P, 256):
        v = win[i0:i0 + 256].view(np.float32).reshape(-1, 4, stride, 2) * gain
        w[i0:i0 + 256] = np.clip(np.rint(v), -full, full - 1).astype(w.dtype)
    best, ok = 1e9, None
    for rep in range(6):
        torch.cuda.synchronize()
        t = time.perf_counter()
        _, _, ok, tbo, _ = phy.rx(w, tb_out=tb_pinned.numpy(), sc16_scale=float(1.0 / gain), search_length=search_len)
        dt_s = time.perf_counter() - t
        if rep:
            best = min(best, dt_s)
    assert np.array_equal(tbo[ok], tb[ok])
    print("%-5s %d windows, %.1f MB in, %.3f ms, %.1f Mbit/s decoded, crc ok %d" %
          (name, P, w.nbytes / 1e6, best * 1e3, ok.sum() * ps["N_TB_bits"] / best / 1e6, int(ok.sum())))
