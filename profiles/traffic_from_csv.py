"""Turns an `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv` log of
profiles/phase_times.py (full bench workload) into profiles/r1_traffic.json: DRAM bytes per launch for each phase of a
commit. bench.py quotes the dominant kernel's entry as roofline.traffic.
usage: python profiles/traffic_from_csv.py gpurun_out/r2_traffic.csv <commit> > profiles/r2_traffic.json
(also keeps smsp__inst_executed.sum and the per-pipe instruction counts when the log has them)"""
import csv
import json
import sys

rows = list(csv.reader(open(sys.argv[1])))
i0 = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
hdr = rows[i0]
K, M, V, G = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Grid Size")
phase_of = {"11, 0, 1": "interp_strided", "11, 1, 1": "interp_contig", "11, 0, 0": "lde_strided", "11, 1, 0": "lde_contig"}
out = {}
seen = {}
for r in rows[i0 + 1:]:
    name = r[K]
    ph = None
    for k, v in phase_of.items():
        if "ntt_pass_kernel<Gl64, " + k in name:
            ph = v
    if "hash_rows" in name:
        ph = "hash_rows"
    if "merkle_level" in name:
        ph = "merkle_level_%d" % seen.setdefault(r[0], len([s for s in seen if s != r[0]]))
    if ph is None:
        continue
    d = out.setdefault(ph, {"kernel": name, "grid": r[G]})
    d[r[M]] = int(float(r[V].replace(",", "")))
INT_PEAK = 592 * 1.965e9  # warp-instructions / s at 1.0 per clock per SMSP (profiles/r2_int_peak_summary.md)
for d in out.values():
    d["dram_bytes"] = d.get("dram__bytes_read.sum", 0) + d.get("dram__bytes_write.sum", 0)
    if "smsp__inst_executed.sum" in d:
        d["warp_instructions"] = d["smsp__inst_executed.sum"]
commit = sys.argv[2] if len(sys.argv) > 2 else "?"
json.dump({"workload": "2^22 x 100 f64, blowup 8, Blake3_256 (one commit)", "source": "ncu --clock-control none, one launch each",
           "commit": commit, "int_peak_warp_inst_per_s": INT_PEAK, "phases": out}, sys.stdout, indent=1)
