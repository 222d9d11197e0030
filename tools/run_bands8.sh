N=${1:-8}
for b in 1 4 8; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$b bench.py --gpus $N --steps 5 --warmup 3 --bands $b > gpurun_out/bench_${N}gpu_b$b.log 2>&1; echo "bands=$b"; tail -1 gpurun_out/bench_${N}gpu_b$b.log | cut -c1-160
done
