"""Registers / stack (spill) / static shared memory per kernel of the built library (cuobjdump --dump-resource-usage),
names demangled.  usage: python tools/resource_usage.py > profiles/r02_resource_usage.txt"""
import re
import subprocess
from pathlib import Path

LIB = Path(__file__).resolve().parents[1] / "pokeai_b200" / "_lib" / "libpokeai_b200.so"


def main():
    txt = subprocess.run(["cuobjdump", "--dump-resource-usage", str(LIB)], capture_output=True, text=True, check=True).stdout
    rows, name = [], None
    for line in txt.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            name = m.group(1)
            continue
        m = re.search(r"REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", line)
        if m and name:
            rows.append((name, *map(int, m.groups())))
            name = None
    dem = subprocess.run(["c++filt"], input="\n".join(r[0] for r in rows), capture_output=True, text=True).stdout.splitlines()
    print("# cuobjdump --dump-resource-usage libpokeai_b200.so (sm_100a): registers per thread, stack bytes (spills),")
    print("# static shared bytes (dynamic shared memory is set at launch and not listed), local bytes")
    print(f"# {len(rows)} kernels; {sum(1 for r in rows if r[2])} with a stack frame")
    print(f"{'reg':>4} {'stack':>5} {'smem':>6} {'local':>5}  kernel")
    for (n, reg, st, sh, lo), d in zip(rows, dem):
        d = re.sub(r"\(.*$", "", d).replace("void ", "")
        print(f"{reg:>4} {st:>5} {sh:>6} {lo:>5}  {d}")


if __name__ == "__main__":
    main()
