python -m pytest tests/test_comm_gpu.py tests/test_host_cpp.py -m gpu -x -q 2>&1 | tail -4
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02p_bench_n2.json 2> gpurun_out/r02p_bench_n2.err; tail -c 600 gpurun_out/r02p_bench_n2.json | head -c 400; echo
