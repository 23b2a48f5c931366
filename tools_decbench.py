"""scratch: decoder micro-benchmark (resident LLRs)"""
import sys, numpy as np
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
from conftest import load_product, CONFIGS
m = load_product()
name = sys.argv[1] if len(sys.argv) > 1 else "C3_b16_txdiv4_mcs7"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
phy = m.Phy(CONFIGS[name], max_batch=n)
ps = phy.ps
rng = np.random.default_rng(0)
tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
pcc, pdc = phy.fec_encode(plcf, tb)
bits = np.unpackbits(pdc, axis=1)[:, :ps["G"]].astype(np.float32)
for snr in (20.0, 7.0, 6.0):
    sigma = 10 ** (-snr / 20)
    llr = np.clip(np.round(((2 * bits - 1) + sigma * rng.standard_normal(bits.shape, dtype=np.float32)) * 2 / sigma ** 2 * 16), -1000, 1000).astype(np.int16)
    ok, out, nit = phy.fec_decode_tb(llr)
    ms = phy.time_stage(m.STAGE_FEC_DECODE_TB, n, reps=5)
    print(name, "n", n, "snr", snr, "ok", ok.mean(), "avg it", nit.mean(), "ms", ms, "pkt/s", n / ms * 1e3, "Mbit/s", n * ps["N_TB_bits"] / ms / 1e3)
ms = phy.time_stage(m.STAGE_FEC_ENCODE, n, reps=5)
print("encode ms", ms, "pkt/s", n / ms * 1e3)
