import sys; sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
from conftest import load_product
from oracle import spec, rx
import test_rx_gpu as T
product = load_product()
name = sys.argv[1]
case = dict(T.RX_CASES[name]); cfg = case["cfg"]; os_ = case.get("os", 1); N_RX = case.get("N_RX", 1)
lr, st = case.get("chestim_lr", True), case.get("stride", 2)
ps = spec.get_packet_sizes(*cfg); n = 3
phy = product.Phy(cfg, max_batch=n, N_RX=N_RX, network_id=T.NET, plcf_type=1, oversampling=os_, chestim_mode_lr=lr, chestim_stride=st)
rng = np.random.default_rng(sum(map(ord, name)))
win, tbs, plcfs = T.make_windows(ps, case, n, phy.rx_stride(), rng, os_)
t0 = case.get("t0", 0); cfo = case.get("cfo", 0.0) + case.get("cfo_err", 0.0)
pcc = phy.rx_pcc(win, t0=t0, cfo_fractional_rad=cfo)
ok, tb, snr, nit = phy.rx_pdc(n)
llr = phy.tap(product.TAP_LLR_PDC, n)
orc = rx.RxConfig(ps, N_RX=N_RX, os=os_, chestim_lr=lr, stride=st)
dl = orc.pdc_l; nb = ps["N_bps"]
for i in range(n):
    r = orc.rx_packet(win[i], t0=t0, cfo_fractional_rad=cfo)
    d = np.abs(llr[i].astype(int) - r["llr_pdc"].astype(int)).reshape(-1, nb).max(axis=1)
    bad = np.nonzero(d > 2)[0]
    print("pkt", i, "n bad cells", len(bad), "of", len(d), "max", d.max())
    ls = sorted(set(dl[bad].tolist()))
    print("  symbols with bad cells:", ls[:60])
    for q in bad[:8]:
        print("   cell", q, "l", dl[q], "k", orc.pdc_k[q], "got", llr[i][q*nb:(q+1)*nb], "ref", r["llr_pdc"][q*nb:(q+1)*nb])
