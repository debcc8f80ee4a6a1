"""Opcode summary per kernel from `cuobjdump -sass` of the built library: the instructions that prove which
hardware path a kernel is on (tcgen05 MMA / TMEM loads / bulk copies / cp.async / MUFU) next to its size.

    python tools/sass_summary.py [scvi_tools_b200/libscvi_b200.so] > profiles/r02_sass_opcodes.txt
"""
import collections
import re
import subprocess
import sys

LIB = sys.argv[1] if len(sys.argv) > 1 else "scvi_tools_b200/libscvi_b200.so"
WATCH = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTCATOMSWS", "UBLKCP", "UTMALDG", "LDGSTS", "SYNCS", "MUFU", "FFMA", "HFMA2",
         "F2FP", "LDG", "STG", "LDS", "STS", "SHFL", "BAR", "ELECT", "R2UR", "ATOM", "RED", "ACQBULK", "FENCE", "MEMBAR"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    name = None
    counts = collections.OrderedDict()
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
            counts[name] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and name:
            op = m.group(1)
            counts[name]["_total"] += 1
            for w in WATCH:
                if op == w or op.startswith(w + "."):
                    counts[name][w] += 1
    cols = [w for w in WATCH if any(c[w] for c in counts.values())]
    print(f"# cuobjdump -sass {LIB}: instruction counts per kernel (static SASS, sm_100a)")
    print(f"{'kernel':58s} {'total':>6s} " + " ".join(f"{w:>7s}" for w in cols))
    for k, c in counts.items():
        print(f"{k[-58:]:58s} {c['_total']:6d} " + " ".join(f"{c[w]:7d}" for w in cols))


if __name__ == "__main__":
    main()
