# round-2 closing measurement job after the last widenings (tracking stage, limited soft buffer, multi-mu adapter): full GPU
# test suite, bench both arms and the other workloads, ncu launch list, ncu --set full of the receive chain (10 000 packets,
# warm repetition; the detection kernels carry the tracking branch now)
set -x
python -c "import __graft_entry__ as g; g.smoke(); print(\"smoke ok\")" > gpurun_out/r02w_smoke.log 2>&1; tail -1 gpurun_out/r02w_smoke.log
python -m pytest tests -m gpu -x -q > gpurun_out/r02w_gputests.log 2>&1; tail -2 gpurun_out/r02w_gputests.log
python bench.py > gpurun_out/r02w_bench_c3_tdl_n1.json 2> gpurun_out/r02w_bench_c3.err; tail -2 gpurun_out/r02w_bench_c3.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02w_bench_reference_arm.json 2> gpurun_out/r02w_ref.err; tail -2 gpurun_out/r02w_ref.err
python bench.py --workload TM6_mu1_b16_sm4x4_qpsk_r12_tdl_doubly_1_100_50_snr30 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02w_bench_tm6_qpsk.json 2> gpurun_out/r02w_tm6.err; tail -2 gpurun_out/r02w_tm6.err
python bench.py --workload C2_mu1_b16_siso_64qam_r56_awgn25 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r02w_bench_c2.json 2> gpurun_out/r02w_c2.err; tail -2 gpurun_out/r02w_c2.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02w_launches_bench_c3.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02w_ncu_bench.log 2>&1
timeout 1200 ncu --set full --import-source on --clock-control none -k 'regex:k_turbo_decode|k_rx_demap|k_rx_fft|k_derm|k_sync|k_rx_stf|k_rx_track' --launch-skip 13 -c 13 -f -o gpurun_out/r02w_c3tdl_10k python profiles/tools/prof_rx.py 10000 C3TDL > gpurun_out/r02w_ncu_full.log 2>&1; tail -3 gpurun_out/r02w_ncu_full.log
du -sh gpurun_out
