"""PER sweep tool (per_sweep.py, BASELINE.json configs[4]).  CPU: cell partition and the counter-matrix reduction over two
gloo ranks, result files in the reference's rx_loopback_MCS_xxxx layout.  GPU: a small sweep on one device -- every
CRC-passing transport block equals the transmitted one, PER falls with SNR, low-SNR points lose packets."""
import json
import os
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _fake_row(ps, ci, n):
    """deterministic counters of cell ci (stands in for the device pipeline in the CPU test)"""
    ok = n - (ci * 7) % n
    row = np.zeros(len(ps.COUNTERS), dtype=np.int64)
    row[:ps.SUMMED] = [n, n, n - 1, ok, ok, ok - 1 if ok else 0, ok - 1 if ok else 0, 12 * n, 1000 * ci * max(ok - 1, 0)]
    row[ps.SUMMED], row[ps.SUMMED + 1] = (1000 * ci - 5, 1000 * ci + 5) if ok > 1 else (ps.BIG, -ps.BIG)
    return row


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import per_sweep as ps
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cells = ps.sweep_cells([0, 1], [1, 2, 3], [0.0, 5.0, 10.0, 15.0, 20.0])
    mine = ps.cells_of_rank(len(cells), rank, world)
    mat = ps.neutral_matrix(len(cells))
    for ci in mine:
        mat[ci] = _fake_row(ps, ci, 100)
    out = ps.reduce_matrix(mat, world, "cpu")
    q.put((rank, mine, out.tolist()))
    dist.barrier()
    dist.destroy_process_group()


def test_cells_are_partitioned_and_reduced_over_two_ranks():
    import per_sweep as ps
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, mine0, m0), (_, mine1, m1) = res
    n_cells = 2 * 3 * 5
    assert sorted(mine0 + mine1) == list(range(n_cells)) and not set(mine0) & set(mine1)
    expect = np.stack([_fake_row(ps, ci, 100) for ci in range(n_cells)])
    assert np.array_equal(np.array(m0), expect) and np.array_equal(np.array(m1), expect)


def test_result_files_follow_the_reference_layout():
    import per_sweep as ps
    snr = [0.0, 5.0, 10.0]
    cells = ps.sweep_cells([0], [2, 4], snr)
    mat = np.stack([_fake_row(ps, ci, 50) for ci in range(len(cells))])
    files = ps.result_files(cells, mat, 50, [0], [2, 4], snr)
    assert sorted(files) == ["rx_loopback_MCS_0002", "rx_loopback_MCS_0004"]      # tfw_loopback_snr.cpp:163
    j = files["rx_loopback_MCS_0004"]
    assert j["experiment_range"] == {"snr_vec": snr, "nof_experiment_per_snr": 50} and j["parameter"]["mcs"] == 4
    for k in ("snr_max_vec", "snr_min_vec", "PER_pcc_crc", "PER_pcc_crc_and_plcf", "PER_pdc_crc"):   # :174-178
        assert len(j["result"][k]) == 3
    ci = 3   # first cell of MCS 4
    assert j["result"]["PER_pdc_crc"][0] == pytest.approx(1.0 - mat[ci, 5] / 50)
    json.dumps(files)


@pytest.mark.gpu
def test_small_sweep_on_one_gpu(tmp_path):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "per_sweep.py"), "--beta", "1", "--mcs", "1", "3", "--snr", "-4", "16", "4",
                        "--packets", "200", "--out", str(tmp_path)], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    s = json.loads(r.stdout.strip().splitlines()[-1])
    assert s["cells"] == 12 and s["packets"] == 2400 and s["tb_mismatch_after_crc_ok"] == 0
    for mcs in (1, 3):
        j = json.load(open(tmp_path / ("rx_loopback_MCS_%04d.json" % mcs)))
        per = j["result"]["PER_pdc_crc"]
        assert per[0] > 0.5 and per[-1] < 0.02, per                       # -4 dB loses packets, 16 dB does not
        assert all(per[i + 1] <= per[i] + 0.05 for i in range(len(per) - 1)), per
        assert all(a <= b + 1e-9 for a, b in zip(j["result"]["PER_pcc_crc"], j["result"]["PER_pcc_crc_and_plcf"]))


def test_ratio_result_files_follow_the_reference_layout():
    import per_sweep as ps
    snr = [0.0, 5.0, 10.0]
    cells = ps.ratio_cells([30, 60], snr)
    mat = np.stack([_fake_row(ps, ci, 100) for ci in range(len(cells))])
    files = ps.ratio_result_files(cells, mat, 100, [30, 60], snr)
    assert sorted(files) == ["rx_loopback_ratio_000_30", "rx_loopback_ratio_001_60"]      # tfw_loopback_ratio.cpp:176-178
    j = files["rx_loopback_ratio_001_60"]
    assert j["experiment_range"] == {"snr_vec": snr, "nof_experiment_per_snr": 100} and j["parameter"] == {"ratio": 60}
    assert sorted(j["result"]) == sorted(["snr_max_vec", "snr_min_vec", "PER_pcc_crc", "PER_pcc_crc_and_plcf", "PER_pdc_crc"])
    json.dumps(files)


@pytest.mark.gpu
def test_ratio_experiment_on_one_gpu(tmp_path):
    """tfw_loopback_ratio through the streaming front door: with both packets equally strong (0 dB) the wanted packet
    decodes at 15 dB and the PCC counter sees both packets of every experiment (error rate -1, like the reference's
    counting); the error rate of the wanted packet does not fall when the interferer grows by 40 dB"""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "per_sweep.py"), "--experiment", "ratio", "--beta", "1", "--mcs", "2",
                        "--ratio", "0", "40", "--snr", "5", "15", "10", "--packets", "40", "--out", str(tmp_path)],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    s = json.loads(r.stdout.strip().splitlines()[-1])
    assert s["cells"] == 4 and s["tb_mismatch_after_crc_ok"] == 0
    j0 = json.load(open(tmp_path / "rx_loopback_ratio_000_0.json"))
    j40 = json.load(open(tmp_path / "rx_loopback_ratio_001_40.json"))
    assert j0["experiment_range"]["snr_vec"] == [5.0, 15.0] and j0["parameter"]["ratio"] == 0
    assert j0["result"]["PER_pdc_crc"][1] <= 0.05 and j0["result"]["PER_pcc_crc"][1] <= -0.9
    assert j40["result"]["PER_pdc_crc"][1] >= j0["result"]["PER_pdc_crc"][1] - 1e-9


def test_mmie_result_files_follow_the_reference_layout():
    import per_sweep as ps
    snr = [0.0, 5.0, 10.0]
    cells = [(m, s) for m in (3, 7) for s in snr]
    mat = np.stack([_fake_row(ps, ci, 100) for ci in range(len(cells))])
    files = ps.mmie_result_files(cells, mat, 100, [3, 7])
    assert sorted(files) == ["rx_loopback_mmie_0003", "rx_loopback_mmie_0007"]          # tfw_loopback_mmie.cpp:152-153
    j = files["rx_loopback_mmie_0003"]
    assert j["experiment_range"] == {"snr_vec": snr, "nof_experiment_per_snr": 100} and j["parameter"] == {"mmie_idx": 3}
    assert list(j["result"]) == ["snr_max_vec", "snr_min_vec", "PER_pcc_crc", "PER_pcc_crc_and_plcf", "PER_pdc_crc"]
    json.dumps(files)


@pytest.mark.gpu
def test_mmie_experiment_on_one_gpu(tmp_path):
    """tfw_loopback_mmie: PER over SNR per MMIE index (random MAC PDU bytes, like the reference's set_mac_pdu)"""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "per_sweep.py"), "--experiment", "mmie", "--beta", "1", "--mcs", "2",
                        "--snr", "-4", "16", "10", "--packets", "60", "--out", str(tmp_path)],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    s = json.loads(r.stdout.strip().splitlines()[-1])
    assert s["cells"] == 3 and s["tb_mismatch_after_crc_ok"] == 0 and s["files"] == ["rx_loopback_mmie_0003"]
    j = json.load(open(tmp_path / "rx_loopback_mmie_0003.json"))
    assert j["result"]["PER_pdc_crc"][0] > 0.5 and j["result"]["PER_pdc_crc"][-1] < 0.05
