#!/usr/bin/env python
"""Top SASS instructions by stall samples: python tools/ncu_hot.py rep [n]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
lines = out.splitlines()
# first kernel only
start = [i for i, l in enumerate(lines) if l.startswith('"Kernel Name"')]
blk = lines[start[0] + 1:(start[1] if len(start) > 1 else len(lines))]
rows = list(csv.reader(io.StringIO("\n".join(blk))))
hdr = rows[0]
ix = {h: i for i, h in enumerate(hdr)}
data = []
for r in rows[1:]:
    if len(r) < len(hdr): continue
    try: s = int(r[ix["# Samples"]]); e = int(r[ix["Instructions Executed"]])
    except ValueError: continue
    data.append((s, e, r))
tot = sum(d[0] for d in data); tote = sum(d[1] for d in data)
print("total samples", tot, "instructions", tote)
for s, e, r in sorted(data, key=lambda x: -x[0])[:n]:
    st = {k: r[ix[k]] for k in ("stall_long_sb", "stall_short_sb", "stall_wait", "stall_mio", "stall_not_selected", "stall_math", "stall_lg", "stall_branch_resolving") if r[ix[k]] not in ("0", "")}
    print("%5.1f%% %5.1f%%exec %s  %s  %s" % (100.0 * s / tot, 100.0 * e / tote, r[ix["Address"]][-5:], r[ix["Source"]][:60], st))
