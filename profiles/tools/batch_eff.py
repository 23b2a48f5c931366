"""resident rx chain time vs batch size (C2): per-packet cost of small launches"""
import sys
import numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
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
phy.channel(P, snr_dB=25.0, t0=t0, mode=0, seed=3)
for n in (148, 296, 312, 444, 500, 592, 625, 740, 1000, 1250, 1480, 2500, 5000, 10000):
    phy.time_stage(product.STAGE_SYNC_RX_CHAIN, n, reps=2)
    ms = phy.time_stage(product.STAGE_SYNC_RX_CHAIN, n, reps=10)
    print("n %5d: %.3f ms  %.3f us/packet" % (n, ms, ms / n * 1e3))
