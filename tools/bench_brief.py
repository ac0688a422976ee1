"""Run bench.py and print the few numbers watched during tuning (developer tool)."""
import json, subprocess, sys
out = subprocess.run([sys.executable, "bench.py", "--no-cpu-baseline"] + sys.argv[1:], stdout=subprocess.PIPE, text=True).stdout
for l in out.splitlines():
    if l.startswith("{"):
        d = json.loads(l)
        r = d["roofline"]
        print("ms/step %.3f  Mpts/s %.2f  kernel_ms %.3f  TFLOP/s %.2f  frac %.4f  e2e Mpts/s %.2f  clocks %s" % (
            d["ms_per_step"], d["value"] / 1e6, r["kernel_ms"], r["achieved"], r["frac"], d["e2e"]["value"] / 1e6, d["clocks"]))
    else:
        print(l)
