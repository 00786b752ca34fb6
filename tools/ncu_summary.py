"""Extract the metrics we report from an .ncu-rep (run where ncu is installed; no GPU needed).
usage: python tools/ncu_summary.py rep.ncu-rep > profiles/xxx.txt"""
import csv, io, re, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
KEEP = re.compile(r"^(Kernel Name|Grid Size|Block Size|gpu__time_duration\.sum|sm__cycles_elapsed\.avg(\.per_second)?|launch__registers_per_thread|launch__shared_mem_per_block_dynamic|launch__occupancy_limit_\w+|launch__waves_per_multiprocessor|"
                  r"sm__throughput\.avg\.pct_of_peak_sustained_elapsed|smsp__issue_active\.avg\.pct_of_peak_sustained_active|smsp__inst_executed\.sum|smsp__thread_inst_executed_per_inst_executed\.ratio|"
                  r"sm__inst_executed_pipe_(fma|alu|fp64|xu|lsu|adu|cbu|uniform)\.avg\.pct_of_peak_sustained_active|sm__pipe_(fma|fmaheavy|alu|fp64)_cycles_active\.avg\.pct_of_peak_sustained_(active|elapsed)|"
                  r"sm__warps_active\.avg\.pct_of_peak_sustained_active|sm__sass_thread_inst_executed_op_(ffma|fadd|fmul|dfma|dadd|dmul)_pred_on\.sum|smsp__sass_thread_inst_executed_op_(ffma|fadd|fmul|dfma|dadd|dmul)_pred_on\.sum|"
                  r"dram__bytes_(read|write)\.sum(\.per_second)?|lts__t_bytes\.sum(\.per_second)?|lts__t_sectors_op_read\.sum|l1tex__t_bytes\.sum(\.per_second)?|l1tex__data_pipe_lsu_wavefronts_mem_shared\.sum|"
                  r"smsp__average_warps_issue_stalled_\w+_per_issue_active\.ratio|smsp__inst_executed_op_branch\.sum|derived__smsp__inst_executed_op_branch_pct)$")
for r in rows[2:]:
    print("== launch", r[0])
    for h, u, v in zip(hdr, units, r):
        if KEEP.match(h):
            print(f"{h} [{u}] = {v}")
