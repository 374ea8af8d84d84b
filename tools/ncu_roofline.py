#!/usr/bin/env python
"""Per-scan counters of the ray-cast kernel from one `ncu --set full` capture of the bench-size launch, for the roofline
object of bench.py:  python tools/ncu_roofline.py gpurun_out/x.ncu-rep fp32 <scans in the launch> [kernel-lookups per scan]
Prints the JSON entry for profiles/roofline_traffic.json."""
import csv
import io
import json
import subprocess
import sys


def main():
    rep, dtype, scans = sys.argv[1], sys.argv[2], float(sys.argv[3])
    lookups = float(sys.argv[4]) if len(sys.argv) > 4 else None
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, r = rows[0], rows[2]

    def g(name):
        return float(r[hdr.index(name)].replace(",", ""))
    unit = rows[1]
    def in_bytes(name):
        v, u = g(name), unit[hdr.index(name)].lower()
        return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
    inst = g("smsp__inst_executed.sum")
    tpi = g("smsp__thread_inst_executed_per_inst_executed.ratio")
    d = {"kernel": r[hdr.index("Kernel Name")], "scans_in_capture": scans,
         "dram_bytes_per_scan": (in_bytes("dram__bytes_read.sum") + in_bytes("dram__bytes_write.sum")) / scans,
         "l2_read_sectors_per_scan": in_bytes("l1tex__m_xbar2l1tex_read_bytes.sum") / 32.0 / scans,      # what the L1s fetched from L2
         "warp_inst_per_scan": inst / scans, "threads_per_warp_inst": tpi,
         "issue_active_pct": g("smsp__issue_active.avg.pct_of_peak_sustained_active"),
         "l1_hit_pct": g("l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": g("lts__t_sector_hit_rate.pct"),
         "l2_throughput_pct": g("lts__throughput.avg.pct_of_peak_sustained_elapsed"),
         "dram_throughput_pct": g("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
         "duration_ms_under_ncu": g("gpu__time_duration.sum") / 1e6 if unit[hdr.index("gpu__time_duration.sum")] in ("ns", "nsecond") else g("gpu__time_duration.sum"),
         "registers": g("launch__registers_per_thread"), "source": "ncu --set full --clock-control none, " + rep.split("/")[-1]}
    if lookups:
        d["thread_inst_per_lookup"] = inst * tpi / scans / lookups
        # the marching loop is 10 instructions: what share of the lane-slots it was issued for did useful lookups
        d["loop_lane_efficiency"] = None
    print(json.dumps({dtype: d}, indent=1))


main()
