#!/usr/bin/env python
"""Per-kernel SASS opcode histogram of libw4a16_b200.so (cuobjdump -sass): the mnemonics that prove what the hot path
is made of (UTCHMMA / LDTM / STTM = tcgen05 + TMEM, UTMALDG / UBLKCP = TMA, IMMA / HMMA = mma.sync, SYNCS = mbarrier,
LDGSTS = cp.async, REDUX).      python tools/sass_ops.py > profiles/r2_sass_ops.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "vllm-rust_b200", "lib", "libw4a16_b200.so")
KEY = ["UTCHMMA", "UTCIMMA", "LDTM", "STTM", "UTCBAR", "UTMALDG", "UTMASTG", "UBLKCP", "UBLKPF", "IMMA", "HMMA", "LDGSTS",
       "SYNCS", "REDUX", "CREDUX", "LDS", "STS", "LDG", "STG", "RED", "ATOM", "BAR", "LOP3", "PRMT", "IDP", "MEMBAR", "ACQBULK"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m and cur:
            kernels[cur][m.group(1)] += 1
            kernels[cur]["#instructions"] += 1
    print(f"# SASS opcode histogram of {os.path.relpath(LIB, ROOT)} (sm_100a), by kernel; columns: instructions, then the")
    print("# counts of the mnemonic families that matter (prefix match), zero counts omitted.\n")
    for name, c in kernels.items():
        short = re.sub(r"\(.*", "", name)
        fam = []
        for k in KEY:
            n = sum(v for op, v in c.items() if op.startswith(k) and not (k == "RED" and op.startswith("REDUX")))
            if n:
                fam.append(f"{k}={n}")
        print(f"{short:70s} {c['#instructions']:6d}  " + " ".join(fam))


if __name__ == "__main__":
    main()
