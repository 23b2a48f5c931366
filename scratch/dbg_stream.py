import sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
from conftest import load_product
import test_stream_gpu as T
product = load_product()
cfg = (1, 2, 1, 1, 0, 3, 2048)
phy, stream, starts, tb, plcf = T._make_stream(product, cfg, 32, snr_db=22.0, seed=11, N_RX=2)
got = phy.stream_sync(stream, chunk_hop=3000)
for g in got:
    print(g.fine_peak_time_64, g.coarse_peak_time_64, "chunk", g.chunk, "round", g["round"], "det", g.sync.detection_time_local, "ant", g.sync.detection_ant_idx, "ndet", g.sync.n_detections)
print("missing", np.setdiff1d(starts, got.fine_peak_time_64))
from oracle import sync as osync
ref = osync.stream_search(stream, 1, 2)
print("oracle", [f for f, _, _ in ref][:14])
w = stream[:, 30000:30000 + 256 + 3000 + 544]
for pos in (476,):
    s = -(-(pos - 256) // 8) * 8
    r = osync.search(np.ascontiguousarray(w[:, s:]), 1, 2, search_length=w.shape[1] - s - 544)
    print("oracle from", pos, "->", r.detected, s + r.detection_time_local, s + r.coarse_peak_time_local, s + r.fine_peak_time_local, r.n_detections)
