"""Turn the scratch outputs of one gpurun round (gpurun_out/*_<tag>.*) into the tracked summaries under profiles/.
usage: python tools/summarise_profile.py <tag> [<round-name>]   e.g.  r01a r01"""
import csv
import io
import json
import re
import shutil
import subprocess
import sys
from collections import defaultdict
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
tag = sys.argv[1]
rnd = sys.argv[2] if len(sys.argv) > 2 else tag
G, P = ROOT / "gpurun_out", ROOT / "profiles"
P.mkdir(exist_ok=True)

# 1. launch list (ncu --metrics gpu__time_duration.sum) of `python bench.py --steps 5 --warmup 3 --cpu-seconds 0`
src = G / f"launches_{tag}.csv"
if src.exists():
    shutil.copy(src, P / f"{rnd}_launches_bench.csv")
    rows = [r for r in csv.reader(open(src)) if len(r) > 10]
    hdr = rows[0]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    d = defaultdict(list)
    for r in rows[1:]:
        d[re.sub(r"\(.*", "", r[ki])[:70]].append(float(r[vi].replace(",", "")))
    tot = sum(sum(v) for v in d.values())
    with open(P / f"{rnd}_launches_bench_summary.txt", "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none  python bench.py --steps 5 --warmup 3 --cpu-seconds 0\n")
        f.write("# per-launch times are cold-cache and serialised: compare SHARES, not absolutes\n")
        for k, v in sorted(d.items(), key=lambda kv: -sum(kv[1])):
            f.write(f"{k:72s} launches={len(v):3d} total_us={sum(v)/1e3:9.1f} mean_us={sum(v)/len(v)/1e3:8.2f} share={sum(v)/tot:.3f}\n")
    print(open(P / f"{rnd}_launches_bench_summary.txt").read())

# 2. full capture: raw metrics + per-line stalls
rep = G / f"prof_{tag}.ncu-rep"
if rep.exists():
    out = subprocess.run(["ncu", "-i", str(rep), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    keep = re.compile(r"^(Kernel Name|Grid Size|Block Size|gpu__time_duration.sum|dram__bytes_(read|write).sum$|"
                      r"gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed|launch__registers_per_thread$|"
                      r"launch__occupancy_limit_(registers|shared_mem|warps)|sm__warps_active.avg.pct_of_peak_sustained_active|"
                      r"smsp__issue_active.avg.pct_of_peak_sustained_active|smsp__inst_executed.sum$|"
                      r"l1tex__t_sector_hit_rate.pct|lts__t_sector_hit_rate.pct|lts__t_bytes.sum$|"
                      r"smsp__inst_executed_op_tma_st.sum$|sm__cycles_elapsed.avg$|"
                      r"smsp__average_warps_issue_stalled_(long_scoreboard|short_scoreboard|wait|barrier|mio_throttle|lg_throttle|math_pipe_throttle|not_selected)_per_issue_active.ratio)")
    with open(P / f"{rnd}_ncu_full_raw.txt", "w") as f:
        f.write(f"# ncu --set full --clock-control none --import-source on -k regex:pk_decide_kernel  python tools/prof_decide.py (last round)\n")
        f.write("# columns: launch 1 = pk_encode (features+mask only), 2 = pk_decide with features"
                + (", 3 = pk_decide without features" if len(data) > 2 else "") + "; B = 4096\n")
        for i, h in enumerate(hdr):
            if keep.search(h):
                f.write(f"{h:90s} {units[i]:16s} " + "  ".join(f"{r[i][:40]:>16s}" for r in data) + "\n")
    print(open(P / f"{rnd}_ncu_full_raw.txt").read()[:3000])
    hi = {h: i for i, h in enumerate(hdr)}
    tr = [float(r[hi["dram__bytes_read.sum"]].replace(",", "")) * {"Kbyte": 1e3, "Mbyte": 1e6, "byte": 1, "Gbyte": 1e9}[units[hi["dram__bytes_read.sum"]]] for r in data]
    tw = [float(r[hi["dram__bytes_write.sum"]].replace(",", "")) * {"Kbyte": 1e3, "Mbyte": 1e6, "byte": 1, "Gbyte": 1e9}[units[hi["dram__bytes_write.sum"]]] for r in data]
    traffic = {"pk_decide_kernel@B=4096": tr[1] + tw[1], "pk_encode@B=4096": tr[0] + tw[0],
               "note": "dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full, one isolated launch: the 92 MB "
                       "of output of one 4096-state launch fits the 126 MB L2, so part of it is still dirty in L2 when the "
                       "kernel ends and is not counted; the steady-state (rotating buffers) DRAM write traffic is the full 92 MB",
               "source": f"profiles/{rnd}_ncu_full_raw.txt"}
    if len(data) > 2:
        traffic["pk_decide_nofeat@B=4096"] = tr[2] + tw[2]
    (P / "traffic.json").write_text(json.dumps(traffic, indent=1) + "\n")
    for which, name in ((1, "decide_feat"),):
        o = subprocess.run([sys.executable, str(ROOT / "tools" / "ncu_lines.py"), str(rep), str(which), "45"],
                           capture_output=True, text=True).stdout
        (P / f"{rnd}_ncu_lines_{name}.txt").write_text(o)

# 3. dense-input forward: full capture of pk_dense_tc_kernel at 65 536 states
rep = G / f"prof_dense_{tag}.ncu-rep"
if rep.exists():
    out = subprocess.run(["ncu", "-i", str(rep), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    keep = re.compile(r"^(Kernel Name|Grid Size|Block Size|gpu__time_duration.sum|dram__bytes_(read|write).sum$|"
                      r"gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed|launch__registers_per_thread$|"
                      r"sm__warps_active.avg.pct_of_peak_sustained_active|smsp__issue_active.avg.pct_of_peak_sustained_active|"
                      r"smsp__inst_executed.sum$|lts__t_sector_hit_rate.pct|lts__t_bytes.sum$|sm__cycles_elapsed.avg$|"
                      r"sm__pipe_tensor_subpipe.*cycles_active.avg.pct_of_peak_sustained_active|sm__inst_executed_pipe_tensor.*sum$|"
                      r"l1tex__data_pipe_lsu_wavefronts_mem_shared.sum$|l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum$|"
                      r"smsp__average_warps_issue_stalled_(long_scoreboard|short_scoreboard|wait|barrier|mio_throttle|lg_throttle|membar|sleeping)_per_issue_active.ratio)")
    with open(P / f"{rnd}_ncu_dense_summary.txt", "w") as f:
        f.write("# ncu --set full --clock-control none --import-source on -k regex:pk_dense_tc_kernel  python tools/prof_dense.py\n")
        f.write("# pk_mlp_forward on dense (65536, 617, 9) fp32 features, default MLP 2x64 (one isolated launch, cold L2)\n")
        for i, h in enumerate(hdr):
            if keep.search(h):
                f.write(f"{h:100s} {units[i]:16s} " + "  ".join(f"{r[i][:40]:>16s}" for r in data) + "\n")
    print(open(P / f"{rnd}_ncu_dense_summary.txt").read()[:3000])
if (G / f"{rnd}_trace_dense.txt").exists():
    shutil.copy(G / f"{rnd}_trace_dense.txt", P / f"{rnd}_trace_dense.txt")

for name in (f"bench_{tag}.json", f"bench_ref_{tag}.json", f"quick_{tag}.log", f"pytest_{tag}.log", f"smoke_{tag}.log", f"gpu_{tag}.txt"):
    if (G / name).exists():
        shutil.copy(G / name, P / name.replace(f"_{tag}", f"_{rnd}"))
