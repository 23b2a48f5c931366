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
# ---- sc16 ingest
gain = np.float32(0.1 / np.sqrt(np.mean(np.abs(win[:64]) ** 2)) * 32768.0)
h16 = torch.empty((P, 1, phy.rx_stride(), 2), dtype=torch.int16, pin_memory=True)
w16 = h16.numpy()
for i in range(0, P, 512):
    w16[i:i + 512] = np.clip(np.rint(win[i:i + 512].view(np.float32).reshape(-1, 1, phy.rx_stride(), 2) * gain), -32768, 32767).astype(np.int16)
for ch in (1, 2, 4, 8, 16, 32):
    os.environ["DNRP_RX_CHUNKS"] = str(ch)
    for _ in range(2): phy.rx(w16, tb_out=tbp.numpy(), sc16_scale=float(1 / gain))
    t = time.perf_counter()
    for _ in range(5): s, p, ok, o, d = phy.rx(w16, tb_out=tbp.numpy(), sc16_scale=float(1 / gain))
    dt = (time.perf_counter() - t) / 5
    print("sc16 chunks %3d: %.2f ms  ok %d" % (ch, dt * 1e3, int(ok.sum())))
# where does the time go: resident chain in one piece vs python/report overhead
ms = phy.time_stage(product.STAGE_SYNC_RX_CHAIN, P, reps=3)
print("resident step %.2f ms" % ms)
t = time.perf_counter()
for _ in range(5):
    np.zeros(P, dtype=product.PDC_DTYPE); np.zeros(P, dtype=product.PCC_DTYPE); np.zeros(P, dtype=product.SYNC_RESULT_DTYPE)
print("numpy output allocation %.2f ms" % ((time.perf_counter() - t) / 5 * 1e3))
