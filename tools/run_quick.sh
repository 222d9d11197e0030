python tools/exp_kinds.py $PWD/tuning_libs/r1 soup 128 2>&1 | tail -1
python tools/exp_kinds.py $PWD soup 128 2>&1 | tail -1
python tools/exp_kinds.py $PWD/tuning_libs/r1 soup 32 2>&1 | tail -1
python tools/exp_kinds.py $PWD soup 32 2>&1 | tail -1
