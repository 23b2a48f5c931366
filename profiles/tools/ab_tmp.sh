python -m pytest tests/test_sync_gpu.py tests/test_stream_gpu.py tests/test_golden_gpu.py tests/test_ref_gpu.py tests/test_resample_gpu.py tests/test_rx_gpu.py tests/test_tx_gpu.py tests/test_rx_mimo_gpu.py -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 10 --warmup 3 > gpurun_out/ab_sync.json 2>/dev/null; python - <<PY
import json
d=json.loads(open('gpurun_out/ab_sync.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], [(k['kernel'],round(k['ms_per_step'],3)) for k in d['roofline']['kernels']], d['config']['crc_ok'], d['config']['sync_exact_e2e'], d['other_stages']['tx_chain_ms'])
PY
