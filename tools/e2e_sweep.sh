for cfg in "1024 3 2" "512 3 2" "768 3 2" "1536 3 2" "1024 4 2" "1024 4 3" "512 4 3" "2048 3 2" "683 3 2"; do
  set -- $cfg
  python bench.py --no-cpu-baseline --steps 6 --warmup 3 --e2e-chunk $1 --e2e-workers $2 --e2e-concurrent $3 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.readlines()[-1]); print('$cfg', round(d['ms_per_step'],2), round(d['e2e']['ms_per_step'],2), d['e2e']['d2h_bytes_per_step'])"
done
