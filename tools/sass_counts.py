#!/usr/bin/env python
"""SASS instruction counts per kernel of libwf_bitnet.so (cuobjdump -sass | this table): python tools/sass_counts.py > profiles/<name>.txt"""
import re, subprocess, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
so = ROOT / "wrinklefree-inference-engine_b200" / "libwf_bitnet.so"
sass = subprocess.run(["cuobjdump", "-sass", str(so)], capture_output=True, text=True).stdout
names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True, text=True).stdout.split("\n")
cols = ["UTCIMMA", "UTCHMMA", "UTCBAR", "STTM", "LDTM", "UTMALDG", "UBLKCP", "SYNCS", "IMMA", "HMMA", "IDP", "LDGSTS", "LDSM", "REDG", "ATOMG", "RED.E", "UCGABAR", "ELECT"]
print("SASS instruction counts per kernel of libwf_bitnet.so (cuobjdump -sass, sm_100a).  Blackwell-native evidence: UTCIMMA = tcgen05.mma kind::i8,")
print("STTM/LDTM = tcgen05.st/ld (TMEM), UTMALDG = cp.async.bulk.tensor (TMA), UBLKCP = cp.async.bulk, SYNCS = mbarrier, UCGABAR = barrier.cluster, ELECT = elect.sync;")
print("legacy paths: IMMA/HMMA = mma.sync, IDP = dp4a, LDGSTS = cp.async, LDSM = ldmatrix, REDG/ATOMG = global reductions.\n")
print("kernel | " + " | ".join(cols) + " | total")
blocks = re.split(r"\n\s*Function : ", sass)[1:]
for blk, nm in zip(blocks, names):
    body = blk.split("\n", 1)[1] if "\n" in blk else ""
    ops = re.findall(r"^\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", body, flags=re.M)
    short = re.sub(r"^void\s+", "", nm)
    short = re.sub(r"\(.*$", "", short).replace("wf::", "wf::", 1)
    row = [sum(1 for o in ops if o.startswith(c)) for c in cols]
    print(f"{short} | " + " | ".join(map(str, row)) + f" | {len(ops)}")
