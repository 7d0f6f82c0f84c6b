"""Markdown table of an `ncu --set full` report of profiles/prof_hbm.py: per launch duration, dram bytes, throughput.
usage: python profiles/ncu_hbm_table.py report.ncu-rep"""
import csv
import io
import subprocess
import sys

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[0]
col = {k: hdr.index(k) for k in ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
                                 "FBSP.TriageCompute.dram__throughput.avg.pct_of_peak_sustained_elapsed", "Grid Size", "Block Size",
                                 "launch__registers_per_thread") if k in hdr}
units = rows[1]


def num(r, k):
    v = float(r[col[k]].replace(",", "") or "nan")
    u = units[col[k]]
    return v * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "%": 1.0}.get(u, 1.0)


print("| kernel | grid x block | us | dram read MB | dram write MB | (read + write) / time, GB/s | of the measured 6,540 GB/s |")
print("|---|---|---|---|---|---|---|")
for r in rows[2:]:
    name = r[col["Kernel Name"]]
    name = name[:name.find("(")] if "(" in name else name
    t, rd, wr = num(r, "gpu__time_duration.sum"), num(r, "dram__bytes_read.sum"), num(r, "dram__bytes_write.sum")
    print("| `%s` | %s x %s | %.1f | %.1f | %.1f | %.0f | %.1f %% |" % (
        name.replace("void ", ""), r[col["Grid Size"]], r[col["Block Size"]], t, rd / 1e6, wr / 1e6, (rd + wr) / t / 1e3,
        100.0 * (rd + wr) / t / 1e3 / 6540.5))
