#!/usr/bin/env python3
"""SASS opcode histogram per kernel of liblgasr.so (cuobjdump -sass): the mnemonics that prove the Blackwell paths
(UBLKCP = 1-D TMA bulk copy, UBLKPF = bulk L2 prefetch, SYNCS = mbarrier, UCGABAR = cluster barrier, MUFU, REDUX, ATOMS ...).
usage: sass_histogram.py [out.txt]"""
import collections, os, re, subprocess, sys
so = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "llm_guided_asr_b200", "csrc", "liblgasr.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
kern, hist, order = None, {}, []
for l in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", l)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        if kern not in hist:
            hist[kern] = collections.Counter()
            order.append(kern)
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", l)
    if m and kern:
        hist[kern][m.group(1).split(".")[0]] += 1
KEY = ["UBLKCP", "UBLKPF", "UTMALDG", "SYNCS", "UCGABAR_ARV", "UCGABAR_WAIT", "ACQBULK", "MUFU", "REDUX", "ATOMS", "ATOMG", "LDGSTS",
       "SHFL", "BAR", "LDG", "STG", "LDS", "STS", "HMMA", "UTCHMMA"]
lines = ["SASS opcode histogram of llm_guided_asr_b200/csrc/liblgasr.so (cuobjdump -sass, sm_100a); no tensor-core opcodes by design (no contraction on the path)", ""]
for k in order:
    h = hist[k]
    tot = sum(h.values())
    lines.append(f"{k}: {tot} instructions")
    lines.append("   " + "  ".join(f"{o}={h[o]}" for o in KEY if h.get(o)))
    lines.append("   top: " + "  ".join(f"{o}={c}" for o, c in h.most_common(8)))
txt = "\n".join(lines) + "\n"
if len(sys.argv) > 1:
    open(sys.argv[1], "w").write(txt)
print(txt[:3000])
