"""Summarises ncu outputs brought back in gpurun_out/ into the tracked profiles/ directory.

    python tools/summarize_ncu.py launches gpurun_out/launches.csv profiles/r01_launches.md
    python tools/summarize_ncu.py full     gpurun_out/prof.ncu-rep  profiles/r01_top_kernels.md
"""
import collections
import csv
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
    "smsp__cycles_active.avg", "sm__cycles_elapsed.max",
]


def launches(src, dst):
    lines = [l for l in open(src) if not l.startswith("==")]
    rows = list(csv.DictReader(lines))
    agg = collections.OrderedDict()
    for x in rows:
        k = x["Kernel Name"]
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += float(x["Metric Value"].replace(",", ""))
    tot = sum(a[1] for a in agg.values())
    with open(dst, "w") as f:
        f.write("# ncu launch list (gpu__time_duration.sum, --clock-control none; cold-cache, serialised: read SHARES)\n\n")
        f.write("source: %s, %d launches, total %.1f us\n\n" % (src, len(rows), tot / 1e3))
        f.write("| launches | total us | share | avg us | kernel |\n|---:|---:|---:|---:|---|\n")
        for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write("| %d | %.1f | %.1f%% | %.2f | `%s` |\n" % (c, t / 1e3, 100 * t / tot, t / c / 1e3, k[:150]))


def full(src, dst):
    raw = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    with open(dst, "w") as f:
        f.write("# ncu --set full --clock-control none (per launch; numbers under the profiler are not bench values)\n\n")
        f.write("source: %s\n\n" % src)
        for r in rows[2:]:
            f.write("## `%s`\n\n" % r[col["Kernel Name"]][:160])
            for k in KEYS:
                if k in col:
                    f.write("- %s = %s %s\n" % (k, r[col[k]], units[col[k]]))
            f.write("\n")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
