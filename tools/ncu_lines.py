"""Summarise `ncu --page source --csv --print-source cuda,sass` per CUDA source line.
usage: python tools/ncu_lines.py report.ncu-rep [kernel_index] [top_n]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
SORT = sys.argv[4] if len(sys.argv) > 4 else "samples"
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
kern = -1
hdr = None
cur_file = None
lines = {}
name = None
wanted = None
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        if r[1] != name:
            kern += 1
            name = r[1]
            if kern == which:
                wanted = name
        continue
    if r[0] in ("Line No", "Address") or (len(r) > 2 and r[0] == "Line No"):
        hdr = r
        continue
    if kern != which or hdr is None or len(r) < 10:
        continue
    if r[0].isdigit():
        d = dict(zip(hdr, r))
        try:
            samples = int(d.get("# Samples", d.get("Warp Stall Sampling (All Samples)", "0")) or 0)
            inst = int(d.get("Instructions Executed", "0") or 0)
        except ValueError:
            continue
        key = (cur_file, int(r[0]))
        e = lines.setdefault(key, {"src": r[1].strip()[:90], "samples": 0, "inst": 0, "stalls": {}})
        e["samples"] += samples
        e["inst"] += inst
        for k, v in d.items():
            if k.startswith("stall_") and "Not Issued" not in k:
                try:
                    e["stalls"][k] = e["stalls"].get(k, 0) + int(v or 0)
                except ValueError:
                    pass
print("kernel:", wanted)
tot_s = sum(e["samples"] for e in lines.values()) or 1
tot_i = sum(e["inst"] for e in lines.values()) or 1
print(f"total samples {tot_s}  total warp-instructions {tot_i}")
for (f, ln), e in sorted(lines.items(), key=lambda kv: -kv[1][SORT])[:top]:
    st = sorted(e["stalls"].items(), key=lambda kv: -kv[1])[:3]
    st = " ".join(f"{k[6:]}={v}" for k, v in st if v)
    print(f"{f}:{ln:4d} samp {100*e['samples']/tot_s:5.1f}% inst {100*e['inst']/tot_i:5.1f}% | {st:45s} | {e['src']}")
