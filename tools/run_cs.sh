run() { python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value'],1), round(d['e2e']['value'],1), {k:round(v,2) for k,v in d['roofline']['whole_step']['ms_per_step_by_kind'].items()}, 'prb', round(d['prb']['value'],2))"; }
cd mitsuba3_experiments_b200/csrc; cp libb2path.so /tmp/lib_cs.so; cd ../..
echo "== streaming state + L2 node pin"; run
echo "== streaming state, no pin"; B2_L2_PERSIST=0 run
cp mitsuba3_experiments_b200/csrc/alt/libb2path_nocs.so mitsuba3_experiments_b200/csrc/libb2path.so
echo "== plain state + L2 node pin"; run
echo "== plain state, no pin"; B2_L2_PERSIST=0 run
cp /tmp/lib_cs.so mitsuba3_experiments_b200/csrc/libb2path.so
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
