"""ncu report -> profiles/<tag>_ncu_summary.md (one block per kernel: duration, occupancy, issue, lanes, caches, DRAM, pipes, top stalls) and
profiles/ncu_traffic.json (per kernel: DRAM bytes, warp instructions, issue-active %, lanes per instruction of ONE launch — what bench.py's
`roofline.traffic` and instruction-issue ceiling read).      python tools/ncu_to_profiles.py gpurun_out/x.ncu-rep r2 [layout]"""
import csv, io, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, tag = sys.argv[1], sys.argv[2]
layout = sys.argv[3] if len(sys.argv) > 3 else "wide"
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
def val(r, k, d=0.0):
    try: return float(r[hdr.index(k)].replace(",", ""))
    except Exception: return d
def byt(r, k):
    u = units[hdr.index(k)]; v = val(r, k)
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
def usec(r):
    u = units[hdr.index("gpu__time_duration.sum")]; v = val(r, "gpu__time_duration.sum")
    return v * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(u, 1)   # ncu prints "us" / "ms" ... 
stalls = [h for h in hdr if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued")]
best = {}
for r in data:
    name = r[hdr.index("Kernel Name")].replace("void ", "").split("(")[0].replace("rtrt::", "")
    if name not in best or usec(r) > usec(best[name]): best[name] = r            # the longest launch of each kernel (bounce round 1 for k_pt_trace)
traffic_path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
traffic = json.load(open(traffic_path)) if os.path.exists(traffic_path) else {}
md = [f"# ncu --set full summary, {tag} ({os.path.basename(rep)}; layout {layout}; longest launch of each kernel; --clock-control none)", ""]
for name, r in sorted(best.items(), key=lambda kv: -usec(kv[1])):
    st = sorted(((val(r, h), h.replace("smsp__pcsamp_warps_issue_stalled_", "")) for h in stalls), reverse=True)
    tot = sum(v for v, _ in st) or 1
    rec = {"dram_bytes": int(byt(r, "dram__bytes_read.sum") + byt(r, "dram__bytes_write.sum")), "dram_read_bytes": int(byt(r, "dram__bytes_read.sum")),
           "dram_write_bytes": int(byt(r, "dram__bytes_write.sum")), "ncu_duration_us": usec(r), "warp_instructions": int(val(r, "smsp__inst_executed.sum")),
           "issue_active_pct": val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"), "lanes_per_instruction": val(r, "smsp__thread_inst_executed_per_inst_executed.ratio"),
           "source": f"{tag}: {os.path.basename(rep)}"}
    key = name
    if name.startswith("k_pt_trace"): key = "k_pt_trace_round1_wide" if "Wide" in name else "k_pt_trace_round1"
    if name.startswith("k_pt_primary"): key = "k_pt_primary_wide" if "Wide" in name else "k_pt_primary"
    traffic[key] = rec
    md += [f"## {name}", "",
           f"- duration {usec(r):.1f} us, {int(val(r, 'launch__registers_per_thread'))} registers, grid {r[hdr.index('launch__grid_size')]} x {r[hdr.index('launch__block_size')]}, "
           f"warps active {val(r, 'sm__warps_active.avg.pct_of_peak_sustained_active'):.1f} %",
           f"- warp instructions {int(val(r, 'smsp__inst_executed.sum')):,}, issue active {rec['issue_active_pct']:.1f} %, lanes per instruction {rec['lanes_per_instruction']:.2f} of 32, "
           f"eligible warps per cycle {val(r, 'smsp__warps_eligible.avg.per_cycle_active'):.2f}",
           f"- L1 hit {val(r, 'l1tex__t_sector_hit_rate.pct'):.1f} %, L2 hit {val(r, 'lts__t_sector_hit_rate.pct'):.1f} %, l1tex throughput {val(r, 'l1tex__throughput.avg.pct_of_peak_sustained_active'):.1f} %, "
           f"L2 throughput {val(r, 'lts__throughput.avg.pct_of_peak_sustained_elapsed'):.1f} %, DRAM throughput {val(r, 'dram__throughput.avg.pct_of_peak_sustained_elapsed'):.1f} %",
           f"- DRAM read {rec['dram_read_bytes'] / 1e6:.1f} MB, written {rec['dram_write_bytes'] / 1e6:.1f} MB",
           f"- pipes: alu {val(r, 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active'):.1f} %, fma {val(r, 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active'):.1f} %, "
           f"lsu {val(r, 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active'):.1f} %, xu {val(r, 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active'):.1f} %",
           "- stalls (share of the samples): " + ", ".join(f"{n} {100 * v / tot:.0f} %" for v, n in st[:6]), ""]
open(os.path.join(ROOT, "profiles", f"{tag}_ncu_summary.md"), "w").write("\n".join(md))
json.dump(traffic, open(traffic_path, "w"), indent=1)
print("wrote", f"profiles/{tag}_ncu_summary.md", "and profiles/ncu_traffic.json:", ", ".join(sorted(best)))
