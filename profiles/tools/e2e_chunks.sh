# host-buffer call (e2e) under different chunk schedules of dnrp_rx_batch_sc16 (DNRP_RX_WEIGHTS), north-star workload and C2
for wl in C3_mu1_b16_txdiv4_4x4_64qam_r56_tdl_doubly_1_100_50_snr25 C2_mu1_b16_siso_64qam_r56_awgn25; do
for w in "" "1,2,3,4,4,4,3,2,1,1" "4,8,8,8,8,6,4,2,1,1" "2,4,6,8,8,8,6,4,2,1,1" "4,8,8,8,8,8,4,2,1"; do
  DNRP_RX_WEIGHTS="$w" python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$wl'[:2], 'weights [%s]' % '$w', 'e2e ms', round(d['e2e']['ms_per_step'],3), 'Mbit/s', round(d['e2e']['value'],1), 'h2d GB/s', round(d['e2e']['h2d_GBps_per_rank_concurrent_min'],2), 'crc', d['e2e']['crc_ok'])
"
done
done
