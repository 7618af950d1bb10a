"""Opcode histogram of the shipped library per kernel (cuobjdump -sass), the proof of what the hot path is built from:
UTCHMMA / UTCIMMA (tcgen05.mma kind::f16 / kind::i8), LDTM / STTM (tcgen05.ld / st), UTMALDG / UTMASTG / UTMAREDG (tiled TMA
load / store / reduce), UBLKCP (bulk copy), UTCBAR (tcgen05.commit), SYNCS (mbarrier), HFMA2 / PRMT / LOP3 (operand widening).
    python tests/sass_histogram.py [lib.so] [out.txt]
Called by __graft_entry__.build(); the output is committed as profiles/sass_r02.txt."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEY = ["UTCHMMA", "UTCIMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP", "SYNCS", "IMMA", "IDP",
       "HFMA2", "PRMT", "LOP3", "FFMA2", "FMUL2", "I2F", "STS", "LDS", "LDG", "STG", "REDUX", "SHFL", "MEMBAR", "FENCE", "LDL", "STL"]


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "llama.cpp-rrs_b200", "libtcq4_b200.so")
    out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "profiles", "sass_r02.txt")
    txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    per = collections.OrderedDict()
    cur = None
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            per[cur] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and cur:
            per[cur][m.group(1)] += 1
            per[cur]["_total"] += 1
    demangle = subprocess.run(["c++filt"], input="\n".join(per), capture_output=True, text=True).stdout.splitlines()
    with open(out, "w") as f:
        f.write(f"# {os.path.basename(lib)}: SASS opcode counts per kernel (static instruction counts; cuobjdump -sass)\n")
        tot = collections.Counter()
        for (name, c), dn in zip(per.items(), demangle):
            short = re.sub(r"\(.*", "", dn)
            keys = [k for k in KEY if c[k]]
            f.write(f"{short}: total {c['_total']}; " + ", ".join(f"{k} {c[k]}" for k in keys) + "\n")
            tot.update(c)
        f.write("# library total: " + ", ".join(f"{k} {tot[k]}" for k in KEY if tot[k]) + "\n")
    print(f"wrote {out}: {len(per)} kernels")


if __name__ == "__main__":
    main()
