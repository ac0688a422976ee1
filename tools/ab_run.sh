# developer A/B run on the GPU box: tensor-core tests, a short bench, and the phase-clock profile build
timeout 300 python -m pytest tests/test_gpu_tcw.py tests/test_gpu_tc_shapes.py -x -q -m gpu 2>&1 | tail -2
timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-sa 2>&1 | tail -1 > gpurun_out/b_g.json; cat gpurun_out/b_g.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['parity']['rel'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"
if [ -f tools/probe/libtdq_prof.so ]; then
TDQ_LIB=tools/probe/libtdq_prof.so timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-sa 2>&1 | grep -a "tcp" > gpurun_out/tcp2.txt; for k in "tcp ctl" "tcp w0 a" "tcp w0 b" "tcp w5 a" "tcp w5 b"; do grep -a "$k" gpurun_out/tcp2.txt | tail -1; done
fi
