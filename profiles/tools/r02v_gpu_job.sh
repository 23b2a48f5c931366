# round-2 closing measurement job (after the last kernel changes): full GPU test suite, bench both arms, other workloads,
# ncu launch list, ncu --set full of the receive chain (10 000 packets, warm repetition), of the transmit chain + channel
# (2000 packets) and of the C4 resampler
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r02v_gputests.log 2>&1; tail -2 gpurun_out/r02v_gputests.log
python bench.py > gpurun_out/r02v_bench_c3_tdl_n1.json 2> gpurun_out/r02v_bench_c3.err; tail -2 gpurun_out/r02v_bench_c3.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02v_bench_reference_arm.json 2> gpurun_out/r02v_ref.err; tail -2 gpurun_out/r02v_ref.err
python bench.py --workload TM6_mu1_b16_sm4x4_qpsk_r12_tdl_doubly_1_100_50_snr30 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02v_bench_tm6_qpsk.json 2> gpurun_out/r02v_tm6.err; tail -2 gpurun_out/r02v_tm6.err
python bench.py --workload C2_mu1_b16_siso_64qam_r56_awgn25 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r02v_bench_c2.json 2> gpurun_out/r02v_c2.err; tail -2 gpurun_out/r02v_c2.err
timeout 600 python profiles/tools/bench_c4.py --packets 1000 --steps 3 > gpurun_out/r02v_bench_c4.json 2> gpurun_out/r02v_c4.err; tail -2 gpurun_out/r02v_c4.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02v_launches_bench_c3.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02v_ncu_bench.log 2>&1
timeout 1200 ncu --set full --import-source on --clock-control none -k 'regex:k_turbo_decode|k_rx_demap|k_rx_fft|k_derm|k_sync|k_rx_stf' --launch-skip 13 -c 13 -f -o gpurun_out/r02v_c3tdl_10k python profiles/tools/prof_rx.py 10000 C3TDL > gpurun_out/r02v_ncu_full.log 2>&1; tail -3 gpurun_out/r02v_ncu_full.log
timeout 900 ncu --set full --clock-control none -k 'regex:k_tx_ofdm|k_tx_fec|k_channel' -c 6 -f -o gpurun_out/r02v_tx_2k python profiles/tools/prof_rx.py 2000 C3TDL > gpurun_out/r02v_tx_ncu.log 2>&1; tail -2 gpurun_out/r02v_tx_ncu.log
timeout 900 ncu --set full --clock-control none -k 'regex:k_resample' --launch-skip 4 -c 4 -f -o gpurun_out/r02v_c4_resample python profiles/tools/bench_c4.py --packets 200 --steps 1 > gpurun_out/r02v_rs_ncu.log 2>&1; tail -2 gpurun_out/r02v_rs_ncu.log
python profiles/tools/latency_small.py C3 > gpurun_out/r02v_latency_c3.txt 2>&1
du -sh gpurun_out
