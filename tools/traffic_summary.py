"""DRAM traffic per launch from an ncu launch list with dram__bytes_read.sum / dram__bytes_write.sum:
    python tools/traffic_summary.py CSV LAST_N OUT.json
LAST_N = launches of one bench step (the tail of the list), so that pack-time launches are left out.  bench.py reads
the newest profiles/*_traffic.json to fill roofline.traffic."""
import csv, json, sys, collections
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr, rows = rows[0], rows[1:]
ix = {k: i for i, k in enumerate(hdr)}
by_id = collections.OrderedDict()
for r in rows:
    d = by_id.setdefault(int(r[ix["ID"]]), {"kernel": r[ix["Kernel Name"]]})
    v = float(r[ix["Metric Value"]].replace(",", ""))
    u = r[ix["Metric Unit"]].lower()
    v *= {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
    d[r[ix["Metric Name"]]] = v
launches = list(by_id.values())[-int(sys.argv[2]):]
out = {"source": f"ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum over `bench.py --layers 2` "
                 f"(last {len(launches)} launches = one step)", "kernels": {}}
for key in ("mixed_gemm_kernel", "quant_rows_h2", "quant_rows_fast", "decode_gemm_kernel"):
    sel = [l for l in launches if key in l["kernel"]]
    if sel:
        rd = sum(l.get("dram__bytes_read.sum", 0) for l in sel) / len(sel)
        wr = sum(l.get("dram__bytes_write.sum", 0) for l in sel) / len(sel)
        out["kernels"][key] = {"launches": len(sel), "dram_read_bytes_per_launch": rd, "dram_write_bytes_per_launch": wr,
                               "dram_bytes_per_launch": rd + wr}
json.dump(out, open(sys.argv[3], "w"), indent=1)
print(json.dumps(out))
