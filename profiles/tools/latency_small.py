"""latency of the host-buffer receive call for small batches (a live radio hands over one or a few packets at a time):
profiles/tools/latency_small.py [C2 | C3] -- median wall time of dnrp_rx_batch_sc16 over 50 calls per batch size"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import load_product
import torch
product = load_product()
C3 = len(sys.argv) > 1 and sys.argv[1] == "C3"
cfg = (1, 16, 1, 2, 5, 7, 6144) if C3 else (1, 16, 1, 1, 0, 7, 6144)
N_RX = 4 if C3 else 1
P = 256
phy = product.Phy(cfg, max_batch=P, N_RX=N_RX, network_id=123456789)
rng = np.random.default_rng(1)
tb = rng.integers(0, 256, (P, phy.ps["N_TB_byte"]), dtype=np.uint8)
plcf = rng.integers(0, 256, (P, 10), dtype=np.uint8)
phy.tx(plcf, tb, fetch=False)
t0 = rng.integers(512, 3000, P).astype(np.uint32)
win = phy.channel(P, snr_dB=25.0, t0=t0, mode=0, seed=3, fetch=True)
scale = 1.0 / 8192
w16 = np.clip(np.round(win.view(np.float32).reshape(P, N_RX, -1, 2) / scale), -32768, 32767).astype(np.int16)
host = torch.from_numpy(w16).pin_memory().numpy()
out = {}
for n in (1, 2, 4, 8, 16, 32, 64, 128, 256):
    ts = []
    for rep in range(60):
        a = time.perf_counter()
        sres, pcc, ok, tbo, pdc = phy.rx(host[:n], sc16_scale=scale)
        ts.append(time.perf_counter() - a)
    ts = np.array(ts[10:]) * 1e3
    out[n] = (float(np.median(ts)), float(ts.min()), int(ok.sum()))
    print("n=%4d median %.3f ms min %.3f ms  (%.1f us/packet)  crc ok %d" % (n, out[n][0], out[n][1], out[n][0] * 1e3 / n, out[n][2]))
