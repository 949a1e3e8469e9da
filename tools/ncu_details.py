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
