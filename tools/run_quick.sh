echo "== default grid"; python tools/exp_overlap.py 2>&1 | tail -7
echo "== trace blocks 2"; B2_TRACE_BLOCKS=2 python tools/exp_overlap.py 2>&1 | tail -7
echo "== trace blocks 3"; B2_TRACE_BLOCKS=3 python tools/exp_overlap.py 2>&1 | tail -7
