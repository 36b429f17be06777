"""Print selected raw metrics of every launch in an ncu report. usage: ncu_raw.py report.ncu-rep [regex]"""
import csv, io, re, subprocess, sys
rep = sys.argv[1]
pat = re.compile(sys.argv[2]) if len(sys.argv) > 2 else None
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units, data = rows[0], rows[1], rows[2:]
DEFAULT = r"^(gpu__time_duration.sum|dram__bytes_(read|write).sum|gpu__dram_throughput.avg.pct|launch__(registers_per_thread|grid_size|block_size|occupancy_limit.*)|smsp__inst_executed.sum$|smsp__issue_active.avg.pct|sm__warps_active.avg.pct|smsp__average_warp.*stall.*|smsp__warps_issue_stalled_.*_per_warp_active.pct|l1tex__data_pipe_lsu_wavefronts(_mem_shared)?.sum$|lts__t_sectors_op_(read|write).sum$|l1tex__t_sector_hit_rate.pct|sm__inst_executed_pipe_(fma|alu|lsu|fmaheavy|fmalite|xu|uniform).*sum$|smsp__cycles_active.avg$|sm__cycles_elapsed.avg$|lts__t_bytes.sum$|smsp__inst_executed_op_.*)"
pat = pat or re.compile(DEFAULT)
for i, h in enumerate(hdr):
    if pat.search(h):
        print(f"{h:95s} {units[i]:10s} " + "  ".join(f"{r[i]:>14s}" for r in data))
