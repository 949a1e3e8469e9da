"""Flatten `ncu -i REP --page details --csv` into the text summary kept under profiles/:
    python tools/ncu_details.py REP.ncu-rep OUT.txt"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "details", "--csv"], capture_output=True, text=True).stdout
rows = [r for r in csv.reader(out.splitlines()) if len(r) > 10]
hdr, rows = rows[0], rows[1:]
ix = {k: i for i, k in enumerate(hdr)}
with open(sys.argv[2], "w") as f:
    f.write(f"# ncu --set full --clock-control none; kernel: {rows[0][ix['Kernel Name']]}\n")
    for r in rows:
        if r[ix["Metric Name"]]:
            f.write(f"{r[ix['Section Name']]:40s} | {r[ix['Metric Name']]:50s} | {r[ix['Metric Unit']]:14s} | {r[ix['Metric Value']]}\n")
print(sys.argv[2])

# raw-page metrics the details page leaves out: tensor / FMA pipe activity, issue slots, TMEM instructions, DRAM bytes
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(raw.splitlines()))
if len(rr) >= 3:
    keys = ("sm__pipe_tensor_cycles_active.avg.pct", "sm__pipe_tensor_subpipe_imma_cycles_active.avg.pct",
            "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct", "sm__pipe_fma_cycles_active.avg.pct",
            "sm__pipe_alu_cycles_active.avg.pct", "smsp__issue_active.avg.pct", "sm__inst_executed_pipe_tmem.avg.pct",
            "sm__inst_executed_pipe_uniform.avg.pct", "sm__throughput.avg.pct", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "dram__throughput.avg.pct", "gpu__time_duration.sum", "lts__t_bytes.sum", "sm__cycles_active.avg",
            "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__inst_executed.sum")
    with open(sys.argv[2], "a") as f:
        f.write("# --page raw (selected)\n")
        for h, u, v in zip(rr[0], rr[1], rr[-1]):
            if any(h.startswith(k) for k in keys) and "per_second" not in h and ".max" not in h and ".min" not in h:
                f.write(f"{'raw':40s} | {h:90s} | {u:14s} | {v}\n")
