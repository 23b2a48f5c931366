for w in "" "1,2,3,4,4,4,4,4,4,4,4,3,2,1,1" "2,3,4,4,4,4,4,4,3,2,1,1"; do
  DNRP_RX_WEIGHTS="$w" python bench.py --e2e-packets 9000 --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('weights [%s]' % '$w', 'e2e ms', round(d['e2e']['ms_per_step'],3), 'pkts', d['e2e']['packets_per_step_per_gpu'], 'Mbit/s', round(d['e2e']['value'],1), 'crc', d['e2e']['crc_ok'])
"
done
