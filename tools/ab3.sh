# developer A/B on one box: the same bench against several builds of the library
for L in "$@"; do
  for rep in 1 2; do
    TDQ_LIB=$L timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-sa 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$L', d['ms_per_step'], d['roofline']['kernel_ms'], d['clocks']['sm_mhz'])"
  done
done
