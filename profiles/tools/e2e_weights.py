"""e2e timing of dnrp_rx_batch / _sc16 for chunk weight schedules (C2, 10k packets)"""
import os, sys, time
import numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import torch
from conftest import load_product
product = load_product()
P = 10000
cfg = (1, 16, 1, 1, 0, 7, 6144)
phy = product.Phy(cfg, max_batch=P, N_RX=1, network_id=123456789)
rng = np.random.default_rng(1)
tb = rng.integers(0, 256, (P, phy.ps["N_TB_byte"]), dtype=np.uint8)
plcf = rng.integers(0, 256, (P, 10), dtype=np.uint8)
phy.tx(plcf, tb, fetch=False)
t0 = rng.integers(512, 3000, P).astype(np.uint32)
host = torch.empty((P, 1, phy.rx_stride()), dtype=torch.complex64, pin_memory=True)
win = host.numpy(); win[...] = phy.channel(P, snr_dB=25.0, t0=t0, mode=0, seed=3, fetch=True)
tbp = torch.empty((P, phy.ps["N_TB_byte"]), dtype=torch.uint8, pin_memory=True)
gain = np.float32(0.1 / np.sqrt(np.mean(np.abs(win[:64]) ** 2)) * 32768.0)
h16 = torch.empty((P, 1, phy.rx_stride(), 2), dtype=torch.int16, pin_memory=True)
w16 = h16.numpy()
for i in range(0, P, 512):
    w16[i:i + 512] = np.clip(np.rint(win[i:i + 512].view(np.float32).reshape(-1, 1, phy.rx_stride(), 2) * gain), -32768, 32767).astype(np.int16)
for ws in sys.argv[1:]:
    os.environ["DNRP_RX_WEIGHTS"] = ws
    out = []
    for buf, kw in ((win, {}), (w16, {"sc16_scale": float(1 / gain)})):
        for _ in range(2): phy.rx(buf, tb_out=tbp.numpy(), **kw)
        t = time.perf_counter()
        for _ in range(6): s, p, ok, o, d = phy.rx(buf, tb_out=tbp.numpy(), **kw)
        out.append((time.perf_counter() - t) / 6 * 1e3)
    print("%-40s cf32 %.2f ms  sc16 %.2f ms  ok %d" % (ws, out[0], out[1], int(ok.sum())))
