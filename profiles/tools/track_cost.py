"""cost of the residual CFO / STO stage on the north-star workload: the resident receive chain (sync + receiver) on P
packets with the stage off, with the CFO part, with both; per-kernel times from the context's kernel timers"""
import os
import sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import load_product
product = load_product()
P = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
cfg = (1, 16, 1, 2, 5, 7, 6144)
phy = product.Phy(cfg, max_batch=P, N_RX=4, network_id=123456789)
rng = np.random.default_rng(1)
tb = rng.integers(0, 256, (P, phy.ps["N_TB_byte"]), dtype=np.uint8)
plcf = rng.integers(0, 256, (P, 10), dtype=np.uint8)
phy.tx(plcf, tb, fetch=False)
t0 = rng.integers(512, 3000, P).astype(np.uint32)
phy.channel(P, snr_dB=25.0, t0=t0, mode=2, seed=3, pdp_idx=1, tau_rms_ns=100.0, fD_Hz=50.0)
for name, sw in (("off", (False, False)), ("cfo", (True, False)), ("cfo+sto", (True, True)), ("off", (False, False))):
    phy.rx_set_tracking(*sw)
    ms = min(phy.time_stage(product.STAGE_SYNC_RX_CHAIN, P, reps=3) for _ in range(3))
    ok = int(phy.tap(product.TAP_TB_OK, P).reshape(-1).sum())
    est = phy.rx_tracking_read(P) if any(sw) else np.zeros((P, 2))
    print("%-8s %.3f ms per %d packets, crc ok %d, |dphi| median %.4f rad, |dpsi| median %.6f rad" %
          (name, ms, P, ok, float(np.median(np.abs(est[:, 0]))), float(np.median(np.abs(est[:, 1])))))
