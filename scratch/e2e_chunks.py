"""e2e timing of dnrp_rx_batch for different chunk counts + raw H2D rate (C2, 10k packets)"""
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
dev = torch.empty_like(host, device="cuda")
for _ in range(2): dev.copy_(host, non_blocking=True); torch.cuda.synchronize()
t = time.perf_counter()
for _ in range(5): dev.copy_(host, non_blocking=True)
torch.cuda.synchronize(); dt = (time.perf_counter() - t) / 5
print("raw H2D %.2f ms = %.1f GB/s" % (dt * 1e3, host.numel() * 8 / dt / 1e9))
for ch in (1, 4, 8, 16, 32, 64):
    os.environ["DNRP_RX_CHUNKS"] = str(ch)
    for _ in range(2): phy.rx(win, tb_out=tbp.numpy())
    t = time.perf_counter()
    for _ in range(5): s, p, ok, o, d = phy.rx(win, tb_out=tbp.numpy())
    dt = (time.perf_counter() - t) / 5
    print("chunks %3d: %.2f ms  ok %d" % (ch, dt * 1e3, int(ok.sum())))
