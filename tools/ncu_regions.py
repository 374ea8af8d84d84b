#!/usr/bin/env python
"""Stall samples and executed instructions of a kernel by SASS address bucket (to see which phase of a long kernel
costs what): python tools/ncu_regions.py rep [bucket_bytes=1024]"""
import csv, subprocess, sys, io, collections
rep = sys.argv[1]; bucket = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
lines = out.splitlines()
start = [i for i, l in enumerate(lines) if l.startswith('"Kernel Name"')]
blk = lines[start[0] + 1:(start[1] if len(start) > 1 else len(lines))]
rows = list(csv.reader(io.StringIO("\n".join(blk))))
hdr = rows[0]; ix = {h: i for i, h in enumerate(hdr)}
agg = collections.OrderedDict(); first = None
for r in rows[1:]:
    if len(r) < len(hdr): continue
    try: s = int(r[ix["# Samples"]]); e = int(r[ix["Instructions Executed"]]); a = int(r[ix["Address"]], 16)
    except ValueError: continue
    if first is None: first = a
    b = (a - first) // bucket
    t = agg.setdefault(b, [0, 0, ""])
    t[0] += s; t[1] += e
    if not t[2]: t[2] = r[ix["Source"]][:50]
ts = sum(v[0] for v in agg.values()); te = sum(v[1] for v in agg.values())
for b, (s, e, src) in agg.items():
    print("+0x%05x  samples %5.1f%%  exec %5.1f%%   %s" % (b * bucket, 100.0 * s / ts, 100.0 * e / te, src))
