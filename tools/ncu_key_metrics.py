"""Key raw metrics of the first kernel in an ncu --set full report, as text (profiles/r02_ncu_*.txt) and optionally as the
JSON bench.py reads (profiles/r02_traffic.json).  usage: ncu_key_metrics.py <report.ncu-rep> [n_f json_out]"""
import csv, json, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units, r = rows[0], rows[1], rows[2]
d, u = dict(zip(hdr, r)), dict(zip(hdr, units))
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_tc_wavefronts_mem_shared.sum", "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum",
        "derived__memory_l1_wavefronts_shared_excessive", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "smsp__sass_inst_executed_op_tmem_ldt.sum", "smsp__sass_inst_executed_op_tmem_stt.sum"]
print("--- " + d["Kernel Name"])
for w in want:
    if w in d:
        print("  %-92s %s %s" % (w, d[w], u[w]))
if len(sys.argv) > 3:
    num = lambda k: float(d[k].replace(",", ""))
    by = lambda k: num(k) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3}.get(u[k], 1)
    json.dump({"n_f": int(sys.argv[2]), "kernel": d["Kernel Name"], "dram_bytes_read": by("dram__bytes_read.sum"),
               "dram_bytes_write": by("dram__bytes_write.sum"),
               "tensor_pipe_active_pct": num("TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed"),
               "smem_pipe_tc_pct": num("l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed"),
               "smem_pipe_lsu_pct": num("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed"),
               "source": "ncu --set full, one launch (tools/ncu_key_metrics.py)"}, open(sys.argv[3], "w"), indent=1)
