#!/usr/bin/env python
"""bench.py -- decoded PDC throughput of the DECT NR+ receive hot path on B200.

One "step" = one pass of the receiver from raw IQ windows to transport blocks (packet synchronisation, STF stage,
FFTs, DRS zero forcing, Wiener interpolation, equalisation, demapping, blind PLCF decode, de-rate-matching/HARQ,
turbo decode, CRC) over one batch of synthetic packets of BASELINE.json configs[1] (mu=1 beta=16 N_TX=1 64-QAM, AWGN 25 dB).

  value   decoded PDC Mbit/s (N_TB_bits of CRC-passing packets / device time), IQ windows resident in HBM
  e2e     same metric through the C ABI call dnrp_rx_batch with HOST buffers (pinned IQ in, sync/PCC/PDC reports and
          transport blocks out; copies chunked and overlapped with the kernels)
  roofline  the dominant kernel (k_turbo_decode) against the measured HBM copy bandwidth
  cpu_baseline  the oracle (scalar CPU port of the same algorithms) on the host cores, bounded sample

Multi-GPU: packets are sharded over ranks (no data-path collective); one NCCL all-reduce sums the PER
counters, one takes the max of the step time.   `--impl reference` times the CPU oracle only.
"""
import argparse
import importlib.util
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (psdef (u, b, PacketLengthType, PacketLength, tm_mode, mcs, Z), N_RX, snr_dB, channel mode)
    "C2_mu1_b16_siso_64qam_r56_awgn25": ((1, 16, 1, 1, 0, 7, 6144), 1, 25.0, 0),
    "C3_mu1_b16_txdiv4_64qam_r56_flat25": ((1, 16, 1, 2, 5, 7, 6144), 4, 25.0, 1),
    "C1_mu1_b1_siso_qpsk_awgn10": ((1, 1, 1, 1, 0, 2, 2048), 1, 10.0, 0),
}
NETWORK_ID = 123456789


def load_product():
    path = os.path.join(ROOT, "dect-nr-plus-sdr_b200", "__init__.py")
    spec = importlib.util.spec_from_file_location("dectnrp_b200", path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules["dectnrp_b200"] = mod
    spec.loader.exec_module(mod)
    return mod


class ClockSampler:
    """nvidia-smi sampling during the timed region (B200_PROFILING.md 'clocks' recipe)"""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([v.strip() for v in line.split(",")])

    def stop(self, first=0):
        """first = number of rows that had arrived when the timed region started: the rows from there on were taken
        during the timed region; if the region was shorter than the sampling period, the rows of the loaded warm-up
        right before it are used and samples_in_timed_region says so"""
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.thread.join(timeout=2)
        timed = self.rows[first:]
        rows = timed if timed else self.rows
        sm = [float(r[0]) for r in rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in rows if len(r) >= 6 for i in range(4) if r[2 + i] == "Active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "samples_in_timed_region": len(timed)}


# ---- CPU arm: the oracle (scalar C port of the reference algorithms) on the host cores ---------------
def cpu_windows(ps, N_RX, snr_db, mode, n_distinct, seed):
    """oracle TX + numpy channel: n_distinct IQ windows [n][N_RX][L] of the workload"""
    from oracle import tx
    rng = np.random.default_rng(seed)
    wins, tbs = [], []
    for _ in range(n_distinct):
        tb = rng.integers(0, 256, ps["N_TB_byte"], dtype=np.uint8)
        plcf = rng.integers(0, 256, 10, dtype=np.uint8)
        iq = tx.tx_packet(ps, plcf, tb, network_id=NETWORK_ID, plcf_type=1, rv=0)
        N_TX, L = iq.shape
        g = np.ones((N_RX, N_TX), dtype=np.complex64)
        if mode == 1:
            g = ((rng.standard_normal((N_RX, N_TX)) + 1j * rng.standard_normal((N_RX, N_TX))) / np.sqrt(2))
        stf_len = ps["N_samples_STF"]
        t0 = int(rng.integers(stf_len // 2, 2 * stf_len - 4))
        y = np.zeros((N_RX, L + 2 * stf_len), dtype=np.complex64)
        y[:, t0:t0 + L] = g @ iq
        sigma = 10 ** (-snr_db / 20)
        y = y + (sigma / np.sqrt(2)) * (rng.standard_normal(y.shape) + 1j * rng.standard_normal(y.shape))
        wins.append(np.ascontiguousarray(y, dtype=np.complex64))
        tbs.append(tb)
    return wins, tbs


def cpu_rx_rate(workload, n_packets, threads):
    """decoded PDC Mbit/s of the oracle receiver (rx_synced restatement + PLCF blind decode + PDC decode)"""
    from concurrent.futures import ThreadPoolExecutor
    from oracle import fec, rx, spec, sync as osync
    psdef, N_RX, snr_db, mode = WORKLOADS[workload]
    ps = spec.get_packet_sizes(*psdef)
    wins, tbs = cpu_windows(ps, N_RX, snr_db, mode, 8, seed=7)
    cfg = rx.RxConfig(ps, N_RX=N_RX)
    gold = fec.pdc_gold(NETWORK_ID, 1, ps["G"])

    def one(i):
        w = wins[i % len(wins)]
        s = osync.search(w, psdef[0], psdef[1], N_eff_TX_max=ps["N_eff_TX"])
        if not s.detected:
            return False
        r = cfg.rx_packet(w, t0=s.fine_peak_time_local, cfo_fractional_rad=s.cfo_fractional_rad,
                          cfo_integer_rad=s.cfo_integer_rad, rms=np.array(s.rms_array[:], dtype=np.float32),
                          want_Y=False)
        fec.pcc_decode(r["llr_pcc"], 1)
        fec.pcc_decode(r["llr_pcc"], 2)
        ok, out, _ = fec.pdc_decode(r["llr_pdc"], ps, 0, gold)
        return bool(ok) and np.array_equal(out, tbs[i % len(wins)])

    one(0)  # page in the library and the LUTs
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:  # ctypes releases the GIL inside the C oracle
        oks = list(ex.map(one, range(n_packets)))
    dt = time.perf_counter() - t0
    n_ok = int(sum(oks))
    return n_ok * ps["N_TB_bits"] / dt / 1e6, n_packets / dt, n_ok, dt, ps


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle
    oracle.build()
    threads = os.cpu_count() or 1
    per_step = args.cpu_packets
    if not per_step:  # about 10 s of CPU work per step
        _, pps0, _, _, _ = cpu_rx_rate(args.workload, threads * 4, threads)
        per_step = int(min(max(pps0 * 10.0, threads * 8), 100000))
    for _ in range(min(args.warmup, 1)):
        cpu_rx_rate(args.workload, threads, threads)
    rates, pkts, t_all = [], [], 0.0
    for _ in range(args.steps):
        mbps, pps, n_ok, dt, ps = cpu_rx_rate(args.workload, per_step, threads)
        rates.append(mbps)
        pkts.append(pps)
        t_all += dt
    value = float(np.mean(rates))
    sample = "%d packets per step x %d steps of %s, %d threads" % (per_step, args.steps, args.workload, threads)
    print(json.dumps({
        "impl": "reference", "metric": "decoded_pdc_throughput", "value": value, "unit": "Mbit/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_all / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int16+f32", "data": "synthetic",
        "config": {"workload": args.workload, "packets_per_step": per_step, "packets_per_s": float(np.mean(pkts)),
                   "note": "CPU restatement (oracle/, scalar C) of the reference algorithms; the reference build "
                           "(srsRAN/VOLK/FFTW/Eigen) is unavailable on this box"},
        "cpu_baseline": {"value": value, "unit": "Mbit/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "Mbit/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def rank_seed(base, rank):
    """disjoint synthetic packet streams per rank (packets are sharded, no rank sees another's data)"""
    return base + 1000003 * rank


def reduce_over_ranks(times_ms, counters, world, device):
    """the path's only collective: MAX over ranks of the device-timed step times, SUM of the PER / throughput counters
    (NCCL over NVLink on the GPUs; gloo in the CPU test).  Returns (times, counters) valid on every rank."""
    import torch
    import torch.distributed as dist
    t = torch.tensor(list(times_ms), dtype=torch.float64, device=device)
    c = torch.tensor(list(counters), dtype=torch.int64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
    return [float(v) for v in t.cpu()], [int(v) for v in c.cpu()]


def whole_job_value(total_bits, ms_step_max):
    """decoded Mbit/s of the whole job: bits of all ranks / slowest rank's step time"""
    return total_bits / (ms_step_max * 1e-3) / 1e6


# ---- GPU arm ------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local)
    # stdout carries the one JSON line only: library chatter written to file descriptor 1 (NCCL prints its version
    # there) goes to stderr until the result is printed
    sys.stdout.flush()
    stdout_fd = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    product = load_product()
    if not os.path.exists(product.LIB_PATH):
        product.build()
    psdef, N_RX, snr_db, mode = WORKLOADS[args.workload]
    P = args.packets
    phy = product.Phy(psdef, max_batch=P, N_RX=N_RX, network_id=NETWORK_ID, plcf_type=1, device=local)
    ps = phy.ps
    stride = phy.rx_stride()

    # ---- synthetic batch of this rank: TX and channel run on the device, IQ windows fetched once for e2e
    rng = np.random.default_rng(rank_seed(1000, rank))
    tb = rng.integers(0, 256, (P, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (P, 10), dtype=np.uint8)
    stf_len, pat = ps["N_samples_STF"], 16 * psdef[1]
    t0 = rng.integers(2 * pat, 2 * stf_len - 4, P).astype(np.uint32)   # unknown to the receiver: found by sync
    phy.tx(plcf, tb, fetch=False)
    host_iq = torch.empty((P, N_RX, stride), dtype=torch.complex64, pin_memory=True)
    win = host_iq.numpy()
    win[...] = phy.channel(P, snr_dB=snr_db, t0=t0, mode=mode, seed=0xD0C0 + rank, fetch=True)
    tb_pinned = torch.empty((P, ps["N_TB_byte"]), dtype=torch.uint8, pin_memory=True)

    def e2e_step():
        # dnrp_rx_batch: pinned host IQ windows in, sync results + PCC reports + transport blocks + PDC reports out;
        # the H2D copy is chunked and overlaps the kernels
        sres, pcc, ok, tb_out, pdc = phy.rx(win, tb_out=tb_pinned.numpy())
        return pcc, ok, tb_out, sres

    # correctness gate before any timing: every CRC-passing TB must equal what was sent
    pcc, ok, tb_out, sres = e2e_step()
    n_ok = int(ok.sum())
    n_sync_ok = int(((sres.detected == 1) & (sres.fine_peak_time_local == t0)).sum())
    if not np.array_equal(tb_out[ok], tb[ok]):
        raise SystemExit("bench.py: decoded transport blocks differ from the transmitted ones")
    n_plcf_ok = int(pcc.plcf_ok[:, 0].sum())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- resident steps: IQ windows + sync reports already in HBM (left there by e2e_step)
    clocks = ClockSampler(local)
    clocks.start()
    for _ in range(args.warmup):
        phy.time_stage(product.STAGE_SYNC_RX_CHAIN, P, reps=1)
    t_wait = time.perf_counter()   # nvidia-smi needs a moment to deliver its first row: keep the device loaded meanwhile
    while clocks.proc is not None and not clocks.rows and time.perf_counter() - t_wait < 3.0:
        phy.time_stage(product.STAGE_SYNC_RX_CHAIN, P, reps=1)
    phy.kernel_time_enable(True)
    barrier()
    rows_before = len(clocks.rows)
    l0 = phy.launch_count()
    ms_step = phy.time_stage(product.STAGE_SYNC_RX_CHAIN, P, reps=args.steps)   # CUDA events on the kernels' stream
    launches = phy.launch_count() - l0
    barrier()
    clk = clocks.stop(rows_before)
    k_ms, k_cnt = phy.kernel_time_read(product.KERNEL_TURBO_DECODE_PDC)
    others = {name: phy.kernel_time_read(k) for name, k in (("k_rx_fft", product.KERNEL_RX_FFT),
                                                            ("k_rx_demap_pdc", product.KERNEL_RX_DEMAP_PDC),
                                                            ("k_rx_stf", product.KERNEL_RX_STF),
                                                            ("k_derm_pdc", product.KERNEL_DERM_PDC),
                                                            ("k_sync_x4", product.KERNEL_SYNC))}
    phy.kernel_time_enable(False)
    ok_res = phy.tap(product.TAP_TB_OK, P).reshape(-1).astype(bool)
    n_ok_res = int(ok_res.sum())

    # ---- e2e steps: host IQ -> C ABI -> host TBs
    for _ in range(min(args.warmup, 3)):
        e2e_step()
    barrier()
    t_start = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    e2e_s = (time.perf_counter() - t_start) / args.steps

    # ---- the same call fed with the radio's wire format (int16 I/Q, dnrp_rx_batch_sc16): half the H2D bytes.
    # Reported beside e2e, not instead of it: the reference's buffers hold cf32 (UHD converts on the host).
    rms = float(np.sqrt(np.mean(np.abs(win[: min(P, 64)]) ** 2)))
    q_gain = 0.1 / max(rms, 1e-12) * 32768.0            # windows at -20 dBFS rms of the int16 range
    host_sc16 = torch.empty((P, N_RX, stride, 2), dtype=torch.int16, pin_memory=True)
    w16 = host_sc16.numpy()
    for i0 in range(0, P, 512):                          # chunked: no second full-size float temporary
        v = win[i0:i0 + 512].view(np.float32).reshape(-1, N_RX, stride, 2) * np.float32(q_gain)
        w16[i0:i0 + 512] = np.clip(np.rint(v), -32768, 32767).astype(np.int16)

    def e2e16_step():
        return phy.rx(w16, tb_out=tb_pinned.numpy(), sc16_scale=1.0 / q_gain)

    _, _, ok16, tb16, _ = e2e16_step()
    n_ok16 = int(ok16.sum())
    if not np.array_equal(tb16[ok16], tb[ok16]):
        raise SystemExit("bench.py: sc16 path decoded transport blocks differ from the transmitted ones")
    for _ in range(2):
        e2e16_step()
    barrier()
    t_start = time.perf_counter()
    for _ in range(args.steps):
        e2e16_step()
    barrier()
    e2e16_s = (time.perf_counter() - t_start) / args.steps

    # ---- the other half of the path, outside the timed step: transmit chain (FEC encode + OFDM) and simulation
    # channel on the resident batch (device time per pass over the P packets)
    tx_ms = phy.time_stage(product.STAGE_TX_CHAIN, P, reps=3)
    ch_ms = phy.time_stage(product.STAGE_CHANNEL, P, reps=3)

    # ---- reduce over ranks: max of the times, sum of the counters (the one collective of this path)
    (ms_step_max, e2e_ms_max, e2e16_ms_max), (tot_P, tot_ok, tot_bits, tot_ok_e2e, tot_plcf, tot_sync, tot_ok16) = \
        reduce_over_ranks([ms_step, e2e_s * 1e3, e2e16_s * 1e3],
                          [P, n_ok_res, n_ok_res * ps["N_TB_bits"], n_ok, n_plcf_ok, n_sync_ok, n_ok16], world, "cuda")
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    value = whole_job_value(tot_bits, ms_step_max)
    e2e_value = tot_ok_e2e * ps["N_TB_bits"] / (e2e_ms_max * 1e-3) / 1e6
    # roofline of the dominant kernel: algorithmic bytes per packet of the fused de-rate-match + HARQ +
    # turbo decode + CRC kernel (DESIGN.md "k_turbo_decode"): LLRs in, soft buffer out, TB + flags out
    seg_K = [ps["K_plus"]] * ps["C_plus"] + [ps["K_minus"]] * ps["C_minus"]
    alg_bytes = ps["G"] * 2 + sum((3 * K + 12) * 2 for K in seg_K) + ps["N_TB_byte"] + 3 + ps["C"]
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    else:
        peak, peak_src = 6500.0, "fallback"
    k_ms_per_launch = k_ms / max(k_cnt, 1)
    achieved = alg_bytes * P / (k_ms_per_launch * 1e-3) / 1e9 if k_cnt else None
    traffic, traffic_src, issue = None, None, None
    tpath = os.path.join(ROOT, "profiles", "turbo_decode_traffic.json")
    if os.path.exists(tpath):
        t = json.load(open(tpath)).get(args.workload)
        if t:  # ncu capture of a smaller launch, scaled linearly to this launch's packets (the kernel is per-packet)
            traffic = t["dram_bytes_per_launch"] / t["packets_per_launch"] * P
            traffic_src = "%s, scaled from %d to %d packets per launch" % (t["source"], t["packets_per_launch"], P)
            if t.get("warp_inst_per_launch") and k_cnt:
                # the bound that does describe this kernel: warp instructions issued per second against the issue
                # slots of the chip (148 SMs x 4 schedulers x SM clock); instruction count from the same ncu capture
                winst = t["warp_inst_per_launch"] / t["packets_per_launch"] * P
                issue = {"warp_inst_per_launch": winst, "achieved_Ginst_per_s": winst / (k_ms_per_launch * 1e-3) / 1e9,
                         "sm_count": 148, "schedulers_per_sm": 4}

    # per-kernel roofline of the other streaming kernels (algorithmic bytes per packet)
    S_win = stride                                  # samples per antenna window (packet + slack for the offset)
    Nk, N_DF = ps["N_b_OCC"] + 1, ps["N_DF_symb"]
    soft_bytes = sum((3 * K + 12) * 2 for K in seg_K)
    kernel_alg = {
        "k_sync_x4": N_RX * S_win * 8 + 128,                                        # whole window once, sync record out
        "k_rx_fft": N_RX * N_DF * ps["N_b_DFT"] * 8 + N_RX * N_DF * Nk * 8,          # symbols in (no CP), Y out
        "k_rx_demap_pdc": N_RX * N_DF * Nk * 8 + ps["G"] * 2,                        # Y in, LLRs out
        "k_derm_pdc": ps["G"] * 2 + soft_bytes,                                     # LLRs in, soft buffer out (first tx)
        "k_rx_stf": N_RX * ps["N_samples_STF"] * 8,
    }
    kernels = {}
    for name, (ms_sum, cnt) in others.items():
        ms1 = ms_sum / max(cnt, 1)
        a = kernel_alg.get(name)
        kernels[name] = {"ms_per_launch": ms1, "alg_bytes_per_packet": a,
                         "achieved_GBps": (a * P / (ms1 * 1e-3) / 1e9) if (a and ms1 > 0) else None}
        if kernels[name]["achieved_GBps"]:
            kernels[name]["frac"] = kernels[name]["achieved_GBps"] / peak

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        import oracle
        oracle.build()
        threads = os.cpu_count() or 1
        n_cpu = args.cpu_packets
        if not n_cpu:  # calibrate on a short run, then size the sample to about 15 s of CPU work
            _, pps0, _, _, _ = cpu_rx_rate(args.workload, threads * 4, threads)
            n_cpu = int(min(max(pps0 * 15.0, threads * 8), 200000))
        mbps, pps, n_ok_cpu, dt, _ = cpu_rx_rate(args.workload, n_cpu, threads)
        cpu = {"value": mbps, "unit": "Mbit/s", "cores": threads, "kind": "port",
               "sample": "%d packets of %s (%.1f s, %d CRC ok), scalar C oracle" % (n_cpu, args.workload, dt, n_ok_cpu)}

    h2d = P * N_RX * stride * 8
    d2h = P * ps["N_TB_byte"] + P * (ps["C"] + 8 + 76 + 112)
    if issue and clk.get("sm_mhz"):
        issue["peak_Ginst_per_s"] = 148 * 4 * clk["sm_mhz"] * 1e6 / 1e9
        issue["frac"] = issue["achieved_Ginst_per_s"] / issue["peak_Ginst_per_s"]
    out = {
        "metric": "decoded_pdc_throughput", "value": value, "unit": "Mbit/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step_max, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int16+f32", "data": "synthetic",
        "config": {"workload": args.workload, "packets_per_step_per_gpu": P, "packets_per_s": tot_P / (ms_step_max * 1e-3),
                   "crc_ok": tot_ok, "packets": tot_P, "plcf_ok": tot_plcf, "sync_exact": tot_sync,
                   "N_TB_bits": ps["N_TB_bits"],
                   "stages": "raw IQ windows -> TBs: sync (detection, coarse peak, CFO, xcorr fine peak), STF stage, "
                             "FFT+DRS ZF, SNR class, PCC demap+PLCF decode, PDC demap, de-rate-match+HARQ, turbo "
                             "decode+CRC",
                   "l2": "inputs larger than L2 (%.0f MB of IQ per step)" % (P * N_RX * stride * 8 / 1e6)},
        "clocks": clk,
        "e2e": {"value": e2e_value, "unit": "Mbit/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": e2e_ms_max},
        "e2e_sc16": {"value": tot_ok16 * ps["N_TB_bits"] / (e2e16_ms_max * 1e-3) / 1e6, "unit": "Mbit/s",
                     "h2d_bytes_per_step": h2d // 2, "d2h_bytes_per_step": d2h, "ms_per_step": e2e16_ms_max,
                     "crc_ok": tot_ok16,
                     "note": "dnrp_rx_batch_sc16: host IQ as int16 I/Q (the radio's wire format), converted on the device"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": "k_turbo_decode", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": (achieved / peak) if achieved else None, "traffic": traffic, "traffic_source": traffic_src,
                     "peak_source": peak_src,
                     "alg_bytes_per_packet": alg_bytes, "ms_per_launch": k_ms_per_launch, "launches": k_cnt,
                     "share_of_step": k_ms_per_launch / ms_step_max if ms_step_max else None,
                     "note": "compute/shared-memory bound kernel: issue_slots is the bound that describes it",
                     "issue_slots": issue,
                     "other_kernels": kernels},
        "cpu_baseline": cpu,
        "other_stages": {"tx_chain_ms": tx_ms, "tx_packets_per_s": P / (tx_ms * 1e-3), "channel_ms": ch_ms,
                         "channel_packets_per_s": P / (ch_ms * 1e-3),
                         "note": "per rank, %d packets per pass; not part of the timed step" % P},
    }
    sys.stdout.flush()
    os.dup2(stdout_fd, 1)
    print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C2_mu1_b16_siso_64qam_r56_awgn25", choices=list(WORKLOADS))
    ap.add_argument("--packets", type=int, default=10000, help="packets per step and GPU")
    ap.add_argument("--cpu-packets", type=int, default=0, help="CPU sample size (0 = cores x 48)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
