"""profiling driver: one resident receive-chain step (sync + rx chain) on P packets, run twice (the second repetition is
the warm one).  prof_rx.py P [C2 | C3 | C3TDL]: C2 over AWGN, C3 over flat Rayleigh, C3TDL = the bench's north-star
workload (C3 through the doubly selective TDL channel, profile 1, 100 ns, 50 Hz)"""
import os
import sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import load_product
product = load_product()
P = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
TM6 = len(sys.argv) > 2 and sys.argv[2] == "TM6"      # 4x4 spatial multiplexing, QPSK R=1/2, TDL channel at 30 dB
C3 = len(sys.argv) > 2 and sys.argv[2] in ("C3", "C3TDL", "TM6")
TDL = len(sys.argv) > 2 and sys.argv[2] in ("C3TDL", "TM6")
cfg = (1, 16, 1, 2, 6, 1, 6144) if TM6 else ((1, 16, 1, 2, 5, 7, 6144) if C3 else (1, 16, 1, 1, 0, 7, 6144))
phy = product.Phy(cfg, max_batch=P, N_RX=4 if C3 else 1, network_id=123456789)
rng = np.random.default_rng(1)
tb = rng.integers(0, 256, (P, phy.ps["N_TB_byte"]), dtype=np.uint8)
plcf = rng.integers(0, 256, (P, 10), dtype=np.uint8)
phy.tx(plcf, tb, fetch=False)
t0 = rng.integers(512, 3000, P).astype(np.uint32)
if TDL:
    phy.channel(P, snr_dB=30.0 if TM6 else 25.0, t0=t0, mode=2, seed=3, pdp_idx=1, tau_rms_ns=100.0, fD_Hz=50.0)
else:
    phy.channel(P, snr_dB=25.0, t0=t0, mode=1 if C3 else 0, seed=3)
for _ in range(2):
    ms = phy.time_stage(product.STAGE_SYNC_RX_CHAIN, P, reps=1)
ok = phy.tap(product.TAP_TB_OK, P).reshape(-1)
print("ms", ms, "ok", int(ok.sum()), "of", P)
