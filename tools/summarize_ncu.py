"""Summarises ncu output brought back from the GPU box into small committed files under profiles/.
usage: python tools/summarize_ncu.py <tag>    (reads gpurun_out/launches_<tag>.csv, prof_<tag>.ncu-rep, prof_aql_<tag>.ncu-rep)"""
import collections
import csv
import os
import statistics
import subprocess
import sys

tag = sys.argv[1]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
KEEP = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__cluster_dim_x",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum",
        "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__ops_path_tensor_op_utchmma_src_tf32_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed",
        "sm__cycles_elapsed.max"]

out = []
lst = os.path.join(G, "launches_%s.csv" % tag)
if os.path.exists(lst):
    rows = [r for r in csv.reader(open(lst)) if len(r) > 5]
    hdr = rows[0]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    d = collections.defaultdict(list)
    for r in rows[1:]:
        try:
            d[r[ki].split("(")[0]].append(float(r[vi].replace(",", "")))
        except ValueError:
            pass
    tot = sum(statistics.median(v) for v in d.values())
    out.append("# launch list (ncu --metrics gpu__time_duration.sum --clock-control none; cold-cache, serialised: compare SHARES)")
    out.append("kernel,launches,median_ns,min_ns,max_ns,share_of_step_pct")
    for k, v in d.items():
        out.append("%s,%d,%.0f,%.0f,%.0f,%.1f" % (k, len(v), statistics.median(v), min(v), max(v), 100 * statistics.median(v) / tot))
for rep in ("prof_%s.ncu-rep" % tag, "prof_aql_%s.ncu-rep" % tag):
    path = os.path.join(G, rep)
    if not os.path.exists(path):
        continue
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = [hdr.index(k) for k in KEEP if k in hdr]
    out.append("")
    out.append("# %s (ncu --set full --clock-control none)" % rep)
    out.append(",".join(hdr[i] for i in idx))
    out.append(",".join(units[i] for i in idx))
    for r in rows[2:]:
        out.append(",".join('"%s"' % r[i] if "," in r[i] else r[i] for i in idx))
    if rep.startswith("prof_%s" % tag) and "dram__bytes_read.sum" in hdr:  # what bench.py reports as roofline.traffic
        import json
        ki, ri, wi = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        for r in rows[2:]:
            if "k_gather" in r[ki]:
                tot = float(r[ri].replace(",", "")) * scale[units[ri]] + float(r[wi].replace(",", "")) * scale[units[wi]]
                name = "k_gather_bal" if "k_gather_bal" in r[ki] else ("k_gather_tma" if "tma" in r[ki] else "k_gather_ldg<16>")
                json.dump({"kernel": name, "dram_bytes_per_launch": int(tot), "source": "ncu --set full, %s" % rep,
                           "dram__bytes_read.sum": r[ri] + " " + units[ri], "dram__bytes_write.sum": r[wi] + " " + units[wi]},
                          open(os.path.join(P, "%s_gather_traffic.json" % tag), "w"))
                break
    if rep.startswith("prof_aql") and "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active" in hdr:
        import json
        ti = hdr.index("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active")
        for r in rows[2:]:
            if "k_aql_q" in r[hdr.index("Kernel Name")]:
                json.dump({"sm__pipe_tensor_cycles_active_pct": float(r[ti]), "source": "ncu --set full, %s" % rep},
                          open(os.path.join(P, "%s_aql_tensor_pipe.json" % tag), "w"))
                break
dst = os.path.join(P, "%s_ncu_summary.csv" % tag)
open(dst, "w").write("\n".join(out) + "\n")
print(open(dst).read())
