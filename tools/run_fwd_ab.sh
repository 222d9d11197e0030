prb() { python bench.py --workload cornell --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('cornell', round(d['value'],1), 'prb ms', round(d['prb']['value'],2))"; }
cd mitsuba3_experiments_b200/csrc; cp libb2path.so /tmp/lib_cur.so; cd ../..
echo "== current"; prb; prb
cp mitsuba3_experiments_b200/csrc/alt/libb2path_nofwd.so mitsuba3_experiments_b200/csrc/libb2path.so
echo "== forward branches compiled out"; prb; prb
cp /tmp/lib_cur.so mitsuba3_experiments_b200/csrc/libb2path.so
python tools/profile_prb.py 2>&1 | tail -12
