#!/usr/bin/env python
"""Turn ncu outputs into the small text summaries kept under profiles/.

  python tools/summarize_ncu.py launches gpurun_out/launches_c3.csv           # per-kernel share of the step
  python tools/summarize_ncu.py report   gpurun_out/prof_rec_c3.ncu-rep       # key metrics per captured launch
"""
import collections
import csv
import io
import re
import subprocess
import sys


def launches(path):
    lines = [l for l in open(path).read().splitlines() if not l.startswith("==")]
    agg = collections.OrderedDict()
    n = 0
    for row in csv.DictReader(io.StringIO("\n".join(lines))):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("void ", "").replace("crnn::", "")
        v = float(row["Metric Value"].replace(",", ""))
        v = {"ns": v / 1e3, "ms": v * 1e3}.get(row["Metric Unit"], v)
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
        n += 1
    tot = sum(a[1] for a in agg.values())
    print("launches %d   total %.1f us (serialised, cold cache: compare SHARES)" % (n, tot))
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-60s n=%4d  sum=%10.1f us  avg=%9.1f us  share=%5.1f%%" % (k[:60], a[0], a[1], a[1] / a[0], 100 * a[1] / tot))


WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu.sum",
        "smsp__inst_executed.sum"]


def report(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    idx = [hdr.index(w) for w in WANT if w in hdr]
    for r in rows[2:]:
        print(re.sub(r"\(.*", "", r[hdr.index("Kernel Name")]))
        for i in idx:
            print("   %-70s %s %s" % (hdr[i], r[i], units[i]))


if __name__ == "__main__":
    {"launches": launches, "report": report}[sys.argv[1]](sys.argv[2])
