#!/usr/bin/env python
"""per_sweep.py -- PER-vs-SNR sweep of the DECT NR+ PHY hot path (BASELINE.json configs[4]) sharded over the GPUs of one box.

The reference runs this experiment packet by packet in its loopback firmware (lib/src/upper/loopback/tfw_loopback_snr.cpp:
MCS 1-6, SNR -2...20 dB, 100 packets per point; result.cpp:47-58 for the three error rates) and in the FEC test
tb2pdc_awgn.cpp.  Here a sweep cell = (transmission mode, MCS, SNR); every cell is one batch through
TX -> AWGN channel -> packet synchronisation -> PCC -> PDC on the device.  Cells are dealt to the ranks round robin
(no data-path collective); ONE all-reduce (SUM, int64) of the [cells x counters] matrix at the end collects the result
(NCCL over NVLink under torchrun, nothing at world size 1).  Rank 0 writes one JSON file per (mode, MCS) in the
reference's rx_loopback_MCS_xxxx layout (tfw_loopback_snr.cpp:158-185) and prints a summary line.

  python per_sweep.py --mcs 1 2 3 4 --snr -2 20 1 --packets 1000
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 per_sweep.py ...
"""
import argparse
import importlib.util
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# counters per cell; PER definitions follow result_t::set_PERs (result.cpp:47-58): a packet counts for "pcc" when
# work_pcc was reached (sync found it and a PLCF CRC passed), for "pcc_and_plcf" when the PLCF content is the
# transmitted one, for "pdc" when the transport block CRC passed
COUNTERS = ("packets", "sync_detected", "sync_exact", "pcc_crc", "pcc_crc_and_plcf", "pdc_crc", "tb_equal", "turbo_iterations",
            "snr_milli_dB_sum", "snr_milli_dB_min", "snr_milli_dB_max")
SUMMED = len(COUNTERS) - 2   # the last two are reduced with MIN / MAX


def load_product():
    path = os.path.join(ROOT, "dect-nr-plus-sdr_b200", "__init__.py")
    spec = importlib.util.spec_from_file_location("dectnrp_b200", path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules["dectnrp_b200"] = mod
    spec.loader.exec_module(mod)
    return mod


def sweep_cells(tm_modes, mcs_list, snr_list):
    """cell order: mode-major, then MCS, then SNR (the reference's loop order: parameter outside, SNR inside)"""
    return [(tm, mcs, snr) for tm in tm_modes for mcs in mcs_list for snr in snr_list]


def cells_of_rank(n_cells, rank, world):
    """static round-robin partition: neighbouring SNR points (similar decoder work) land on different ranks"""
    return list(range(rank, n_cells, world))


def reduce_matrix(mat, world, device):
    """the sweep's only collective.  mat [cells][len(COUNTERS)] int64, rows of cells this rank did not run are neutral
    (0 for sums, +/- big for min / max)."""
    import torch
    import torch.distributed as dist
    t = torch.as_tensor(mat, dtype=torch.int64).to(device)
    if world > 1:
        s, mn, mx = t[:, :SUMMED].contiguous(), t[:, SUMMED].contiguous(), t[:, SUMMED + 1].contiguous()
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
        dist.all_reduce(mn, op=dist.ReduceOp.MIN)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        t = torch.cat([s, mn[:, None], mx[:, None]], dim=1)
    return t.cpu().numpy()


BIG = 10 ** 12


def neutral_matrix(n_cells):
    m = np.zeros((n_cells, len(COUNTERS)), dtype=np.int64)
    m[:, SUMMED] = BIG
    m[:, SUMMED + 1] = -BIG
    return m


def result_files(cells, mat, n_packets, tm_modes, mcs_list, snr_list):
    """{file name: json dict} in the reference's layout, one per (mode, MCS)"""
    files = {}
    for tm in tm_modes:
        for mcs in mcs_list:
            rows = [i for i, c in enumerate(cells) if c[0] == tm and c[1] == mcs]
            per = lambda k: [float(1.0 - mat[i, COUNTERS.index(k)] / max(mat[i, 0], 1)) for i in rows]
            name = "rx_loopback_MCS_%04d" % mcs + ("" if len(tm_modes) == 1 else "_TM_%02d" % tm)
            files[name] = {
                "experiment_range": {"snr_vec": [float(cells[i][2]) for i in rows], "nof_experiment_per_snr": int(n_packets)},
                "parameter": {"mcs": int(mcs), "tm_mode": int(tm)},
                "result": {
                    "snr_max_vec": [mat[i, SUMMED + 1] / 1000.0 if mat[i, SUMMED + 1] > -BIG else -100.0e6 for i in rows],
                    "snr_min_vec": [mat[i, SUMMED] / 1000.0 if mat[i, SUMMED] < BIG else 100.0e6 for i in rows],
                    "PER_pcc_crc": per("pcc_crc"), "PER_pcc_crc_and_plcf": per("pcc_crc_and_plcf"), "PER_pdc_crc": per("pdc_crc"),
                    # beyond the reference's fields
                    "sync_detection_rate": [float(mat[i, 1] / max(mat[i, 0], 1)) for i in rows],
                    "turbo_iterations_per_cb": [float(mat[i, COUNTERS.index("turbo_iterations")]) for i in rows]}}
    return files


def run_cell(product, phy, cell, n_packets, seed, plcf_type=1):
    """one batch through the device pipeline; returns the counter row"""
    tm, mcs, snr = cell
    ps = phy.ps
    rng = np.random.default_rng(seed)
    tb = rng.integers(0, 256, (n_packets, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n_packets, 10), dtype=np.uint8)
    stf_len, pat = ps["N_samples_STF"], 16 * ps["b"]
    t0 = rng.integers(2 * pat, 2 * stf_len - 4, n_packets).astype(np.uint32)
    phy.tx(plcf, tb, fetch=False)
    phy.channel(n_packets, snr_dB=float(snr), t0=t0, mode=0, seed=seed & 0x7fffffff, fetch=False)
    sres = phy.sync(None, n=n_packets)
    pcc = phy.rx_pcc(None, n=n_packets, sync="device")
    ok, tb_out, snr_est, nit = phy.rx_pdc(n_packets)
    det = sres.detected == 1
    nb = 5 if plcf_type == 1 else 10
    pcc_ok = det & (pcc.plcf_ok[:, plcf_type - 1] != 0)
    plcf_eq = pcc_ok & (pcc.plcf[:, plcf_type - 1, :nb] == plcf[:, :nb]).all(axis=1)
    pdc_ok = pcc_ok & ok
    tb_eq = pdc_ok & (tb_out == tb).all(axis=1)
    row = np.zeros(len(COUNTERS), dtype=np.int64)
    snr_m = np.round(snr_est[pdc_ok].astype(np.float64) * 1000.0).astype(np.int64)
    row[:SUMMED] = [n_packets, int(det.sum()), int((det & (sres.fine_peak_time_local == t0)).sum()), int(pcc_ok.sum()),
                    int(plcf_eq.sum()), int(pdc_ok.sum()), int(tb_eq.sum()), int(nit.sum()), int(snr_m.sum())]
    row[SUMMED] = int(snr_m.min()) if snr_m.size else BIG
    row[SUMMED + 1] = int(snr_m.max()) if snr_m.size else -BIG
    return row


def mmie_result_files(cells, mat, n_packets, mmie_list):
    """tfw_loopback_mmie::save_all_results_to_file (tfw_loopback_mmie.cpp:148-171): one file per MMIE index of the
    reference's mmie_pool_tx_t that derives from serdes::testing_t"""
    files = {}
    for m in mmie_list:
        rows = [i for i, c in enumerate(cells) if c[0] == m]
        per = lambda k: [float(1.0 - mat[i, COUNTERS.index(k)] / max(mat[i, 0], 1)) for i in rows]
        files["rx_loopback_mmie_%04d" % m] = {
            "experiment_range": {"snr_vec": [float(cells[i][1]) for i in rows], "nof_experiment_per_snr": int(n_packets)},
            "parameter": {"mmie_idx": int(m)},
            "result": {
                "snr_max_vec": [mat[i, SUMMED + 1] / 1000.0 if mat[i, SUMMED + 1] > -BIG else -100.0e6 for i in rows],
                "snr_min_vec": [mat[i, SUMMED] / 1000.0 if mat[i, SUMMED] < BIG else 100.0e6 for i in rows],
                "PER_pcc_crc": per("pcc_crc"), "PER_pcc_crc_and_plcf": per("pcc_crc_and_plcf"), "PER_pdc_crc": per("pdc_crc")}}
    return files


def ratio_cells(ratio_list, snr_list):
    """tfw_loopback_ratio: parameter = power ratio of the interferer in dB (outside), SNR inside"""
    return [(ratio, snr) for ratio in ratio_list for snr in snr_list]


def run_ratio_cell(product, phy, cell, n_exp, seed):
    """The near-far experiment of tfw_loopback_ratio.cpp:125-147: packet A (identity A, amplitude db2mag(ratio)) and,
    starting exactly one packet length later, packet B (identity B, amplitude 1, the wanted one); the noise level is set
    against B.  All n_exp experiments sit in ONE stream that goes through the streaming front door (packet search ->
    receiver windows -> PCC -> PDC).  Counters as in the firmware: n_pcc = every packet that reached work_pcc (A and B:
    the error rate can be negative, like the reference's), n_pcc_and_plcf = B's PLCF, n_pdc = B's transport block."""
    ratio, snr = cell
    ps = phy.ps
    rng = np.random.default_rng(seed)
    n = 2 * n_exp
    tb = rng.integers(0, 256, (n, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    plcf[0::2, 1] = 0xA5                         # stands for the transmitter identity: A on even, B on odd packets
    plcf[1::2, 1] = 0x5B
    amp = np.ones(n, dtype=np.float32)
    amp[0::2] = 10.0 ** (float(ratio) / 20.0)
    iq, ns = phy.tx(plcf, tb, DAC_scale=amp)
    S, stf = int(ps["N_samples_packet"]), int(ps["N_samples_STF"])
    gap = 6 * stf + int(rng.integers(0, 64))
    period = 2 * S + gap
    lead = 4 * stf
    total = lead + n_exp * period
    sigma = 10.0 ** (-float(snr) / 20.0)         # B has unit power per antenna (DAC_scale 1)
    stream = (sigma / np.sqrt(2.0)) * (rng.standard_normal((phy.N_RX, total)) + 1j * rng.standard_normal((phy.N_RX, total)))
    startB = np.zeros(n_exp, dtype=np.int64)
    for i in range(n_exp):
        a0 = lead + i * period
        startB[i] = a0 + S
        for a in range(phy.N_RX):
            stream[a, a0:a0 + int(ns[2 * i])] += iq[2 * i, a % iq.shape[1], :int(ns[2 * i])]
            stream[a, a0 + S:a0 + S + int(ns[2 * i + 1])] += iq[2 * i + 1, a % iq.shape[1], :int(ns[2 * i + 1])]
    found = phy.stream_sync(stream.astype(np.complex64), max_out=4 * n)
    row = np.zeros(len(COUNTERS), dtype=np.int64)
    row[0] = n_exp
    row[SUMMED], row[SUMMED + 1] = BIG, -BIG
    if len(found) == 0:
        return row
    found = found[:phy.max_batch]
    m, t0 = len(found), stf // 2
    phy.stream_extract(found.fine_peak_time_64, t0)
    sync = phy.make_sync(m, t0=t0, cfo_fractional_rad=found.sync.cfo_fractional_rad, rms=found.sync.rms_array)
    pcc = phy.rx_pcc(None, n=m, sync=sync)
    ok, tb_out, snr_est, nit = phy.rx_pdc(m)
    pcc_ok = pcc.plcf_ok[:, 0] != 0
    # which experiment's B is it: PLCF content (the firmware compares the transmitter identity), not the position
    is_b = np.zeros(m, dtype=bool)
    pdc_b = np.zeros(m, dtype=bool)
    seen = set()
    for q in range(m):
        if not pcc_ok[q]:
            continue
        hit = np.nonzero((plcf[1::2, :5] == pcc.plcf[q, 0, :5]).all(axis=1))[0]
        if hit.size and int(hit[0]) not in seen:
            seen.add(int(hit[0]))
            is_b[q] = True
            pdc_b[q] = bool(ok[q]) and np.array_equal(tb_out[q], tb[2 * int(hit[0]) + 1])
    snr_m = np.round(snr_est[pdc_b].astype(np.float64) * 1000.0).astype(np.int64)
    exact = int(np.isin(found.fine_peak_time_64, startB).sum())
    row[:SUMMED] = [n_exp, m, exact, int(pcc_ok.sum()), int(is_b.sum()), int(pdc_b.sum()), int(pdc_b.sum()), int(nit[pdc_b].sum()),
                    int(snr_m.sum())]
    if snr_m.size:
        row[SUMMED], row[SUMMED + 1] = int(snr_m.min()), int(snr_m.max())
    return row


def ratio_result_files(cells, mat, n_exp, ratio_list, snr_list):
    """rx_loopback_ratio_<idx>_<ratio> in the layout of tfw_loopback_ratio_t::save_all_results_to_file (:173-201)"""
    files = {}
    for idx, ratio in enumerate(ratio_list):
        rows = [i for i, c in enumerate(cells) if c[0] == ratio]
        per = lambda k: [float(1.0 - mat[i, COUNTERS.index(k)] / max(mat[i, 0], 1)) for i in rows]
        files["rx_loopback_ratio_%03d_%d" % (idx, int(ratio))] = {
            "experiment_range": {"snr_vec": [float(cells[i][1]) for i in rows], "nof_experiment_per_snr": int(n_exp)},
            "parameter": {"ratio": int(ratio)},
            "result": {"snr_max_vec": [mat[i, SUMMED + 1] / 1000.0 if mat[i, SUMMED + 1] > -BIG else -100.0e6 for i in rows],
                       "snr_min_vec": [mat[i, SUMMED] / 1000.0 if mat[i, SUMMED] < BIG else 100.0e6 for i in rows],
                       "PER_pcc_crc": per("pcc_crc"), "PER_pcc_crc_and_plcf": per("pcc_crc_and_plcf"), "PER_pdc_crc": per("pdc_crc")}}
    return files


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mu", type=int, default=1)
    ap.add_argument("--beta", type=int, default=1)
    ap.add_argument("--packet-length", type=int, default=2, help="slots (PacketLengthType 1)")
    ap.add_argument("--tm", type=int, nargs="+", default=[0], help="transmission mode indices (0 SISO, 1 TxDiv 2, 5 TxDiv 4, 10 TxDiv 8)")
    ap.add_argument("--mcs", type=int, nargs="+", default=[1, 2, 3, 4, 5, 6])
    ap.add_argument("--snr", type=float, nargs=3, default=[-2.0, 20.0, 1.0], metavar=("FROM", "TO", "STEP"))
    ap.add_argument("--packets", type=int, default=1000, help="packets per (mode, MCS, SNR) cell")
    ap.add_argument("--n-rx", type=int, default=1)
    ap.add_argument("--out", default="", help="directory for the rx_loopback_MCS_xxxx JSON files (rank 0)")
    ap.add_argument("--experiment", choices=("snr", "ratio", "mmie"), default="snr",
                    help="snr: tfw_loopback_snr (PER over MCS and SNR); ratio: tfw_loopback_ratio (a strong packet right in "
                         "front of the wanted one, PER over power ratio and SNR); mmie: tfw_loopback_mmie (PER over SNR per "
                         "MMIE index -- the reference fills the MAC PDU with random bytes whatever the MMIE, "
                         "tfw_loopback_mmie.cpp:104-113, so a cell is an SNR cell at the first MCS)")
    ap.add_argument("--mmie", type=int, nargs="+", default=[3],
                    help="--experiment mmie: indices into the reference's mmie_pool_tx_t (3 = cluster beacon message, the "
                         "only entry that derives from serdes::testing_t, mmie_pool_tx.cpp:47-70)")
    ap.add_argument("--ratio", type=int, nargs="+", default=[30, 40, 50, 60, 70, 80], help="--experiment ratio: dB")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("per_sweep.py: no CUDA device; the product has no CPU fallback")
    torch.cuda.set_device(local)
    # stdout carries the one JSON line only: library chatter written to file descriptor 1 (NCCL prints its version
    # there) goes to stderr until the result is printed
    sys.stdout.flush()
    stdout_fd = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    product = load_product()
    snr_list = [float(v) for v in np.arange(args.snr[0], args.snr[1] + 0.5 * args.snr[2], args.snr[2])]
    ratio = args.experiment == "ratio"
    mmie = args.experiment == "mmie"
    cells = ratio_cells(args.ratio, snr_list) if ratio else sweep_cells(args.tm, args.mcs, snr_list)
    if mmie:
        cells = [(m, snr) for m in args.mmie for snr in snr_list]
    mine = cells_of_rank(len(cells), rank, world)
    mat = neutral_matrix(len(cells))
    t_start = time.perf_counter()
    phys = {}
    if ratio:   # one geometry (tfw_loopback.cpp:93-99: the firmware's MCS), every cell = one stream of n experiments
        psdef = (args.mu, args.beta, 1, args.packet_length, args.tm[0], args.mcs[0], 2048 if args.beta < 4 else 6144)
        phy = product.Phy(psdef, max_batch=4 * args.packets, N_RX=args.n_rx, network_id=123456789, plcf_type=1, device=local)
        for ci in mine:
            mat[ci] = run_ratio_cell(product, phy, cells[ci], args.packets, seed=0x5EED1000 + ci)
        phy.close()
        mine = []
    for ci in mine:
        tm, mcs, snr_c = (args.tm[0], args.mcs[0], cells[ci][1]) if mmie else cells[ci]
        if (tm, mcs) not in phys:      # one context per packet geometry, reused over the SNR points
            for p in phys.values():
                p.close()
            phys.clear()
            psdef = (args.mu, args.beta, 1, args.packet_length, tm, mcs, 2048 if args.beta < 4 else 6144)
            phys[(tm, mcs)] = product.Phy(psdef, max_batch=args.packets, N_RX=args.n_rx, network_id=123456789,
                                          plcf_type=1, device=local)
        mat[ci] = run_cell(product, phys[(tm, mcs)], (tm, mcs, snr_c), args.packets, seed=0x5EED0000 + ci)
    for p in phys.values():
        p.close()
    torch.cuda.synchronize()
    mat = reduce_matrix(mat, world, "cuda")
    if world > 1:
        dist.barrier()
    dt = time.perf_counter() - t_start
    if rank == 0:
        files = (ratio_result_files(cells, mat, args.packets, args.ratio, snr_list) if ratio else
                 mmie_result_files(cells, mat, args.packets, args.mmie) if mmie else
                 result_files(cells, mat, args.packets, args.tm, args.mcs, snr_list))
        if args.out:
            os.makedirs(args.out, exist_ok=True)
            for name, j in files.items():
                with open(os.path.join(args.out, name + ".json"), "w") as f:
                    json.dump(j, f, indent=4)
        tot = mat[:, :SUMMED].sum(axis=0)
        sys.stdout.flush()
        os.dup2(stdout_fd, 1)
        print(json.dumps({"cells": len(cells), "n_gpus": world, "packets": int(tot[0]), "seconds": dt,
                          "packets_per_s": float(tot[0] / dt), "pdc_crc_ok": int(tot[5]), "tb_mismatch_after_crc_ok": int(tot[5] - tot[6]),
                          "files": sorted(files)}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
