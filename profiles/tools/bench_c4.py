"""BASELINE.json configs[3] as a measurement (not the driver's bench line, which stays on the north-star C3 workload):
mu = 8, beta = 16, eight transmit streams (TM 10), 256-QAM R = 5/6 (59 code blocks of 6080 bits), polyphase resampling
10/9 to the device rate, HARQ: every packet is received twice, rv 0 at an SNR where it does not decode, rv 2 combined in
the same soft-buffer slot.  One step = both receptions of P packets, device-resident windows, packet synchronisation
included.  Prints one JSON record with the per-kernel times; k_resample is quoted against the measured HBM peak.

    python profiles/tools/bench_c4.py [--packets 1000] [--steps 5] [--n-rx 8] > profiles/r02_bench_c4.json
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--packets", type=int, default=1000)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--n-rx", type=int, default=8)
    ap.add_argument("--snr", type=float, default=19.0)
    args = ap.parse_args()
    product = bench.load_product()
    cfg, LM = (8, 16, 1, 1, 10, 9, 6144), (10, 9)
    P, N_RX = args.packets, args.n_rx
    phy = product.Phy(cfg, max_batch=P, N_RX=N_RX, network_id=bench.NETWORK_ID, resample=LM)
    ps = phy.ps
    rng = np.random.default_rng(5)
    tb = rng.integers(0, 256, (P, ps["N_TB_byte"]), dtype=np.uint8)
    plcf = rng.integers(0, 256, (P, 10), dtype=np.uint8)
    stf_rs = ps["N_samples_STF"] * LM[0] // LM[1]
    slots = np.arange(P, dtype=np.uint32)
    kernels = (("k_turbo_decode_pdc", product.KERNEL_TURBO_DECODE_PDC), ("k_rx_fft", product.KERNEL_RX_FFT),
               ("k_rx_demap_pdc", product.KERNEL_RX_DEMAP_PDC), ("k_rx_stf", product.KERNEL_RX_STF),
               ("k_derm_pdc", product.KERNEL_DERM_PDC), ("k_sync_x4", product.KERNEL_SYNC),
               ("k_resample", product.KERNEL_RESAMPLE))
    t_rx, ok_rounds, n_sync = 0.0, [0, 0], 0
    acc = {name: [0.0, 0] for name, _ in kernels}
    for step in range(args.warmup + args.steps):
        timed = step >= args.warmup
        for rnd, rv in enumerate((0, 2)):
            phy.tx(plcf, tb, rv=rv, fetch=False)
            t0 = rng.integers(stf_rs // 2, 2 * stf_rs - 8, P).astype(np.uint32)
            phy.channel(P, snr_dB=args.snr, t0=t0, mode=0, seed=100 + 2 * step + rnd, fetch=False)
            if timed:
                phy.kernel_time_enable(True)
            t = time.perf_counter()
            res = phy.sync(n=P)
            phy.rx_pcc(None, n=P, sync="device")
            ok, tb_out, _, _ = phy.rx_pdc(P, harq_slot=slots, rv=np.full(P, rv, np.uint8),
                                          harq_reset=np.full(P, 1 if rnd == 0 else 0, np.uint8))
            dt = time.perf_counter() - t
            if timed:   # enable() starts from zero: collect this round's kernel times before switching the timers off
                for name, k in kernels:
                    ms, cnt = phy.kernel_time_read(k)
                    acc[name][0] += ms
                    acc[name][1] += cnt
            phy.kernel_time_enable(False)   # TX and channel of the next round stay outside the kernel timers
            if not np.array_equal(tb_out[ok], tb[ok]):
                raise SystemExit("bench_c4: decoded transport blocks differ from the transmitted ones")
            if timed:
                t_rx += dt
                ok_rounds[rnd] += int(ok.sum())
                n_sync += int(((res.detected == 1) & (np.abs(res.fine_peak_time_local.astype(np.int64) - t0) <= 1)).sum())
    ms_step = t_rx / args.steps * 1e3
    try:
        hbm = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except (OSError, KeyError, ValueError):
        hbm = 6553.6
    # resampling per packet and round: the whole window hardware -> DECT rate for detection, then the packet from its
    # fine peak; cf32 in and out
    rx_stride = phy.rx_stride()
    n_no_gi_hw = -(-ps["N_samples_packet_no_GI"] * LM[0] // LM[1])
    rs_bytes = 8 * N_RX * (rx_stride + rx_stride * LM[1] // LM[0] + n_no_gi_hw + ps["N_samples_packet_no_GI"])
    out_k = {}
    for name, _ in kernels:
        out_k[name] = {"ms_per_step": acc[name][0] / args.steps, "launches_per_step": acc[name][1] / args.steps}
    rs_ms = out_k["k_resample"]["ms_per_step"]
    out_k["k_resample"].update(alg_bytes_per_packet_and_round=rs_bytes,
                               achieved_GBps=2 * P * rs_bytes / (rs_ms * 1e-3) / 1e9 if rs_ms else None,
                               frac_of_hbm_peak=(2 * P * rs_bytes / (rs_ms * 1e-3) / 1e9 / hbm) if rs_ms else None)
    decoded = ok_rounds[1] / args.steps
    print(json.dumps({
        "workload": "C4_mu8_b16_txdiv8_%drx_256qam_r56_resample_10_9_harq_rv0_rv2_awgn%g" % (N_RX, args.snr),
        "packets_per_step": P, "steps": args.steps, "receptions_per_step": 2 * P,
        "ms_per_step": ms_step, "timing": "host clock around sync + rx_pcc + rx_pdc of both rounds (each call returns "
        "after its reports are on the host); TX, channel outside",
        "decoded_pdc_Mbit_s": decoded * ps["N_TB_bits"] / (ms_step * 1e-3) / 1e6,
        "receptions_per_s": 2 * P / (ms_step * 1e-3),
        "crc_ok_after_rv0": ok_rounds[0] / args.steps, "crc_ok_after_rv2": decoded, "sync_within_1_sample": n_sync / args.steps,
        "N_TB_bits": ps["N_TB_bits"], "C": ps["C"], "G": ps["G"], "rx_stride": rx_stride,
        "hbm_peak_GBps": hbm, "kernels": out_k}))
    phy.close()


if __name__ == "__main__":
    main()
