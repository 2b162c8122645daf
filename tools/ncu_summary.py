"""Summarises an `ncu --set full` capture (raw page) per kernel: duration, DRAM bytes, instructions,
issue activity, instruction-cache hit rate, top stall reasons.  usage: ncu_summary.py <raw.csv> [json_out]"""
import csv, json, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
def g(r, k, d=0.0):
    try: return float(r[idx[k]])
    except Exception: return d
out = {}
for r in rows[2:]:
    name = r[idx["Kernel Name"]].replace("void ", "").split("(")[0].replace("pcseg::", "")
    tu = units[idx["gpu__time_duration.sum"]]
    t_ms = g(r, "gpu__time_duration.sum") * {"ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3}.get(tu, 1.0)
    bu = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}
    rd = g(r, "dram__bytes_read.sum") * bu.get(units[idx["dram__bytes_read.sum"]], 1.0)
    wr = g(r, "dram__bytes_write.sum") * bu.get(units[idx["dram__bytes_write.sum"]], 1.0)
    stalls = {h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""): g(r, h)
              for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")}
    top = sorted(stalls.items(), key=lambda kv: -kv[1])[:5]
    out[name] = {"duration_ms_under_ncu": round(t_ms, 3), "dram_bytes_per_launch": int(rd + wr),
                 "dram_read": int(rd), "dram_write": int(wr),
                 "warp_instructions": int(g(r, "smsp__inst_executed.sum")),
                 "issue_active_pct": round(g(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"), 1),
                 "icc_hit_rate_pct": round(g(r, "sm__icc_request_hit_rate.pct"), 1),
                 "registers": int(g(r, "launch__registers_per_thread")),
                 "grid": int(g(r, "launch__grid_size")), "block": int(g(r, "launch__block_size")),
                 "l1_hit_pct": round(g(r, "l1tex__t_sector_hit_rate.pct"), 1),
                 "l2_hit_pct": round(g(r, "lts__t_sector_hit_rate.pct"), 1),
                 "top_stalls_per_issue": {k: round(v, 2) for k, v in top}}
print(json.dumps(out, indent=1))
if len(sys.argv) > 2:
    json.dump(out, open(sys.argv[2], "w"), indent=1)
