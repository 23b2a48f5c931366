import os, sys
import numpy as np
ROOT = "/root/repo"
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import load_product
product = load_product()
for name, cfg, nrx in (("C3", (1, 16, 1, 2, 5, 7, 6144), 4), ("C2", (1, 16, 1, 1, 0, 7, 6144), 1)):
    P = 10000
    phy = product.Phy(cfg, max_batch=P, N_RX=nrx, network_id=123456789)
    rng = np.random.default_rng(1)
    tb = rng.integers(0, 256, (P, phy.ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (P, 10), dtype=np.uint8)
    phy.tx(plcf, tb, fetch=False)
    print(name, "fec_encode ms", phy.time_stage(product.STAGE_FEC_ENCODE, P, reps=3), "tx_chain ms", phy.time_stage(product.STAGE_TX_CHAIN, P, reps=3))
    phy.close()
