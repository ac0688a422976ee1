# the scaling record: bench.py at N = 8, 4, 2 on one box (run with gpurun --gpus 8); extra bench.py flags as arguments
for n in 8 4 2; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) bench.py --gpus $n --steps 50 --warmup 10 "$@" > gpurun_out/r02_bench_n$n.json 2> gpurun_out/r02_bench_n$n.err
  tail -1 gpurun_out/r02_bench_n$n.json | cut -c1-250
done
