#!/usr/bin/env python
"""BASELINE config 5: L1 / L2 hit rates, L2 sectors and DRAM bytes per sweep point from the ncu metrics pass of
    ncu --metrics gpu__time_duration.sum,lts__t_sector_hit_rate.pct,lts__t_sectors.sum,l1tex__m_xbar2l1tex_read_bytes.sum,\
dram__bytes_read.sum,dram__bytes_write.sum,l1tex__t_sector_hit_rate.pct --clock-control none -k regex:'k_scan_only64|k_raycast32' \
        --csv --log-file ncu.csv python bench.py --workload config5 --steps 2 --warmup 1
joined with the un-profiled run of the same command (its JSON line):   python tools/config5_ncu.py ncu.csv sweep.json > out.json
A sweep point = its plain launches (3 warm-up + 2 timed) followed by one launch of the lookup-counting instantiation."""
import collections
import csv
import json
import sys


def main():
    lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
    sweep = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    launches = collections.OrderedDict()
    for row in csv.DictReader(lines):
        k = int(row["ID"])
        d = launches.setdefault(k, {"kernel": row["Kernel Name"]})
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"].lower()
        v *= {"kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "us": 1e-3, "ns": 1e-6, "usecond": 1e-3, "nsecond": 1e-6, "msecond": 1.0, "ms": 1.0}.get(u, 1.0)
        d[row["Metric Name"]] = v
    points, cur = [], []
    for d in launches.values():
        name = d["kernel"]
        # the counting instantiation: COUNT is the first template argument of k_raycast32*, the third of k_scan_only64
        args = name[name.index("<") + 1:name.index(">")].replace("(bool)", "").replace(" ", "").split(",")
        counting = args[0] == "1" if "k_raycast32" in name else args[2] == "1"
        if counting:
            if cur:
                points.append(cur)
            cur = []
        else:
            cur.append(d)
    if cur:
        points.append(cur)
    pts = sweep["sweep"]
    assert len(points) == len(pts), (len(points), len(pts))
    out = []
    med = lambda xs: sorted(xs)[len(xs) // 2]
    l2 = sweep.get("l2_roofline", {}).get("26MB", {})
    for p, ls in zip(pts, points):
        ms = med([l["gpu__time_duration.sum"] for l in ls])
        sect = med([l["lts__t_sectors.sum"] for l in ls])
        rd = med([l["l1tex__m_xbar2l1tex_read_bytes.sum"] for l in ls])
        e = dict(edt=p["edt"], beams=p["beams"], dtype=p["dtype"], edt_MB=round(p["edt_MB"], 1), scans_per_s_events=p["scans_per_s"] / sweep["n_gpus"],
                 ncu_ms=ms, l1_hit_pct=med([l["l1tex__t_sector_hit_rate.pct"] for l in ls]), l2_hit_pct=med([l["lts__t_sector_hit_rate.pct"] for l in ls]),
                 l2_sector_GBps=sect * 32 / (ms * 1e-3) / 1e9, l2_to_l1_read_GBps=rd / (ms * 1e-3) / 1e9,
                 dram_read_MB=med([l["dram__bytes_read.sum"] for l in ls]) / 1e6, dram_write_MB=med([l["dram__bytes_write.sum"] for l in ls]) / 1e6)
        e["hbm_GBps"] = (e["dram_read_MB"] + e["dram_write_MB"]) * 1e6 / (ms * 1e-3) / 1e9
        if l2:
            e["l2_read_frac_of_gather_peak"] = e["l2_to_l1_read_GBps"] / l2["gather_sector_GBps"]
            e["l2_read_frac_of_stream_peak"] = e["l2_to_l1_read_GBps"] / l2["stream_GBps"]
        out.append(e)
    print(json.dumps({"how": __doc__.split("\n\n")[0], "l2_peaks_measured": l2, "points": out}, indent=1))


main()
