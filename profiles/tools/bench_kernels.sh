#!/bin/bash
# per-kernel times of the C2 and C3 bench workloads (no CPU baseline)
for w in C2_mu1_b16_siso_64qam_r56_awgn25 C3_mu1_b16_txdiv4_64qam_r56_flat25; do
python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['config']['workload'][:2], round(d['ms_per_step'],3), 'tdec', round(d['roofline']['ms_per_launch'],3), {k:round(v['ms_per_launch'],3) for k,v in d['roofline']['other_kernels'].items()}, 'crc_ok', d['config']['crc_ok'])"
done
