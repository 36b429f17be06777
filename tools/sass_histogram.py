"""Opcode histogram per kernel of the built library (cuobjdump -sass), the evidence that the kernels are what
DESIGN.md says they are: FFMA2 / UBLKCP (TMA bulk) in the decision kernels, UTCHMMA / UTMALDG / LDTM in the
tensor-core ones.  usage: python tools/sass_histogram.py > profiles/r02_sass_histogram.txt"""
import collections
import re
import subprocess
import sys
from pathlib import Path

LIB = Path(__file__).resolve().parents[1] / "pokeai_b200" / "_lib" / "libpokeai_b200.so"
INTEREST = ["FFMA2", "FFMA", "FADD", "LDS", "STS", "LDG", "STG", "UBLKCP", "UTMALDG", "UTMASTG", "UTCHMMA", "UTCQMMA",
            "LDTM", "STTM", "UTCBAR", "SYNCS", "LDGSTS", "HMMA", "SHFL", "BAR", "RED", "ATOM", "MUFU", "ACQBULK"]
KEEP = re.compile(r"pk_decide_kernelILi2ELi9ELi512ELb0|pk_decide2_kernelILi4ELi9ELb[01]|pk_decide_tc_kernelILi9|"
                  r"pk_dense_forward_kernelILi2ELi9|pk_dense_tc_kernel|pk_dqn_backward_kernelILi2ELi2ELb1|pk_wide_gemm_kernel|"
                  r"pk_wide_encode|pk_peer_clamp_adam|pk_clamp_adam_kernel|pk_dqn_reduce_flat|pk_select_kernel|pk_wide_head")


def main():
    txt = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    name, hist = None, collections.OrderedDict()
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1) if KEEP.search(m.group(1)) else None
            if name:
                hist[name] = collections.Counter()
            continue
        if name is None:
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m:
            op = m.group(1)
            hist[name]["_total"] += 1
            hist[name][op.split(".")[0]] += 1
            if op.startswith(("UTMALDG", "UTCHMMA", "UBLKCP", "LDTM", "UTMASTG")):
                hist[name][op] += 1
    print(f"# cuobjdump -sass {LIB.name}: instructions per kernel (static counts), selected opcodes")
    for name, h in hist.items():
        demangled = subprocess.run(["cu++filt", name], capture_output=True, text=True).stdout.strip() or name
        print(f"\n{demangled[:150]}")
        print(f"  total {h['_total']}")
        shown = [(k, v) for k, v in h.items() if k != "_total" and (k.split(".")[0] in INTEREST)]
        print("  " + "  ".join(f"{k} {v}" for k, v in sorted(shown, key=lambda kv: -kv[1])))


main()
