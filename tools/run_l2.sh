for m in 0 1 2; do echo "== B2_L2_PERSIST=$m"; B2_L2_PERSIST=$m python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-prb 2> gpurun_out/l2_$m.err | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['roofline']['whole_step']['ms_per_step_by_kind'])"; grep "L2 persist" gpurun_out/l2_$m.err; done
