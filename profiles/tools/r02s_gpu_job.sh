# round-2 final measurement job: full GPU test suite, bench both arms, TM6, C4, C2, ncu launch list, one ncu --set full
# capture of the warm repetition of the whole chain at 10 000 packets
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r02s_gputests.log 2>&1; tail -2 gpurun_out/r02s_gputests.log
python bench.py > gpurun_out/r02s_bench_c3_tdl_n1.json 2> gpurun_out/r02s_bench_c3.err; tail -2 gpurun_out/r02s_bench_c3.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02s_bench_reference_arm.json 2> gpurun_out/r02s_ref.err; tail -2 gpurun_out/r02s_ref.err
python bench.py --workload TM6_mu1_b16_sm4x4_qpsk_r12_tdl_doubly_1_100_50_snr30 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02s_bench_tm6_qpsk.json 2> gpurun_out/r02s_tm6.err; tail -2 gpurun_out/r02s_tm6.err
python bench.py --workload C2_mu1_b16_siso_64qam_r56_awgn25 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r02s_bench_c2.json 2> gpurun_out/r02s_c2.err; tail -2 gpurun_out/r02s_c2.err
timeout 600 python profiles/tools/bench_c4.py --packets 1000 --steps 3 > gpurun_out/r02s_bench_c4.json 2> gpurun_out/r02s_c4.err; tail -2 gpurun_out/r02s_c4.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02s_launches_bench_c3.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02s_ncu_bench.log 2>&1
timeout 1200 ncu --set full --import-source on --clock-control none -k 'regex:k_turbo_decode|k_rx_demap|k_rx_fft|k_derm|k_sync|k_rx_stf' --launch-skip 13 -c 13 -f -o gpurun_out/r02s_c3tdl_10k python profiles/tools/prof_rx.py 10000 C3TDL > gpurun_out/r02s_ncu_full.log 2>&1; tail -3 gpurun_out/r02s_ncu_full.log
du -sh gpurun_out; ls -la gpurun_out/ | tail -12
