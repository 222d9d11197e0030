for s in 6 8 12; do for t in 8 10 14; do
  B2_TRACE_SLICE=$s B2_TRI_THRESH=$t python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-prb 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('slice $s thresh $t', round(d['value'],1), {k:round(v,2) for k,v in d['roofline']['whole_step']['ms_per_step_by_kind'].items()})"
done; done
