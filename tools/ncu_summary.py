"""Summaries for profiles/ from one `ncu --set full` report and one launch-list csv.
    python tools/ncu_summary.py <report.ncu-rep> <launches.csv> <out-prefix> [points]
writes <out-prefix>_raw_metrics.csv, <out-prefix>_source_lines.txt, profiles/knn_traffic.json and prints the launch shares."""
import collections, csv, json, os, subprocess, sys
rep, launches, prefix = sys.argv[1:4]
points = int(sys.argv[4]) if len(sys.argv) > 4 else 10_000_000
here = os.path.dirname(os.path.abspath(__file__))
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h, u, v = rows[0], rows[1], rows[2]
keep = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
        'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_dynamic']
d = {}
with open(prefix + "_raw_metrics.csv", "w") as f:
    f.write("metric,unit,value\n")
    for k, uu, x in zip(h, u, v):
        if k in keep or (k.startswith('smsp__average_warps_issue_stalled') and k.endswith('per_issue_active.ratio')):
            f.write(f"{k},{uu},{x}\n"); d[k] = (x, uu)
def tobytes(k):
    x, uu = d[k]
    return float(x) * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[uu]
tr = tobytes('dram__bytes_read.sum') + tobytes('dram__bytes_write.sum')
json.dump({"kernel": "knn_lane_kernel<0>", "points": points, "dram_bytes_per_launch": tr, "dram_read": tobytes('dram__bytes_read.sum'),
           "dram_write": tobytes('dram__bytes_write.sum'), "source": f"ncu --set full --clock-control none, {os.path.basename(prefix)}_raw_metrics.csv"},
          open(os.path.join(os.path.dirname(prefix), "knn_traffic.json"), "w"), indent=1)
print("traffic", tr, "time", d['gpu__time_duration.sum'], "inst", d['smsp__inst_executed.sum'], "issue", d['smsp__issue_active.avg.pct_of_peak_sustained_active'],
      "warps", d['sm__warps_active.avg.pct_of_peak_sustained_active'])
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
open("/tmp/_src.csv", "w").write(src)
open(prefix + "_source_lines.txt", "w").write(subprocess.run([sys.executable, os.path.join(here, "ncu_lines.py"), "/tmp/_src.csv", "40"], capture_output=True, text=True).stdout)
rows = list(csv.reader(open(launches)))
i = [k for k, r in enumerate(rows) if r and r[0] == 'ID'][0]
hh = rows[i]; kn, mv, mu = hh.index('Kernel Name'), hh.index('Metric Value'), hh.index('Metric Unit')
t, c = collections.Counter(), collections.Counter()
for r in rows[i + 1:]:
    if len(r) <= mv: continue
    val = float(r[mv].replace(',', '')) * {'ns': 1, 'us': 1e3, 'ms': 1e6}.get(r[mu], 1)
    name = r[kn].split('(')[0]; t[name] += val; c[name] += 1
tot = sum(t.values())
for k, vv in t.most_common(10): print(f"{100*vv/tot:6.2f}% {vv/1e6:9.3f} ms {c[k]:4d}  {k}")
print('total ms', tot / 1e6)
