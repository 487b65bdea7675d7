#!/usr/bin/env python
"""Turn ncu outputs into the small text summaries kept under profiles/.

  python tools/summarize_ncu.py launches gpurun_out/launches_c3.csv           # per-kernel share of the step
  python tools/summarize_ncu.py report   gpurun_out/prof_rec_c3.ncu-rep       # key metrics per captured launch
  python tools/summarize_ncu.py rawcsv   gpurun_out/r02_rec_c3.raw.csv        # the same from `ncu -i X --page raw --csv`
  python tools/summarize_ncu.py traffic  gpurun_out/r02_rec_c3.raw.csv ...    # {kernel: dram bytes per launch} as JSON
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


HBM_PEAK_GBS = 6455.9      # MEASURED_PEAKS.json (copy bandwidth)


def _rows(path):
    rows = list(csv.reader(open(path)))
    start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    return rows[start], rows[start + 1], rows[start + 2:]


def _num(x):
    try:
        return float(x.replace(",", ""))
    except Exception:
        return None


def rawcsv(path):
    hdr, units, rows = _rows(path)
    idx = [hdr.index(w) for w in WANT if w in hdr]
    for r in rows:
        if len(r) <= max(idx):
            continue
        print(re.sub(r"\(.*", "", r[hdr.index("Kernel Name")]).replace("void ", "").replace("crnn::", ""),
              " grid", r[hdr.index("Grid Size")] if "Grid Size" in hdr else "")
        for i in idx:
            print("   %-70s %s %s" % (hdr[i], r[i], units[i]))
        t, rd, wr = (_num(r[hdr.index(k)]) if k in hdr else None for k in ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum"))
        if t and rd is not None and wr is not None:
            tu, bu = units[hdr.index("gpu__time_duration.sum")], units[hdr.index("dram__bytes_read.sum")]
            sec = t * {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}.get(tu, 1e-9)
            scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(bu, 1.0)
            wscale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(units[hdr.index("dram__bytes_write.sum")], 1.0)
            tot = rd * scale + wr * wscale
            print("   %-70s %.3f MB -> %.1f GB/s = %.3f of the measured copy peak" % ("DRAM traffic (read + write) / duration", tot / 1e6, tot / sec / 1e9,
                                                                                     tot / sec / 1e9 / HBM_PEAK_GBS))


def traffic(*paths):
    import json
    out = {}
    for path in paths:
        hdr, units, rows = _rows(path)
        for r in rows:
            try:
                name = re.sub(r"[<(].*", "", r[hdr.index("Kernel Name")]).replace("void ", "").replace("crnn::", "").replace("tc::", "")
                ri, wi = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
                sc = lambda i: {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(units[i], 1.0)
                out.setdefault(name, []).append(_num(r[ri]) * sc(ri) + _num(r[wi]) * sc(wi))
            except Exception:
                continue
    print(json.dumps({k: sum(v) / len(v) for k, v in out.items()}, indent=1))


if __name__ == "__main__":
    fn = {"launches": launches, "report": report, "rawcsv": rawcsv, "traffic": traffic}[sys.argv[1]]
    fn(*sys.argv[2:]) if sys.argv[1] == "traffic" else fn(sys.argv[2])
