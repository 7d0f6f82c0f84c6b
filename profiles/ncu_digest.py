"""Digest of an ncu report: per-launch headline metrics (raw page) and the hottest source lines (source page).
usage: python profiles/ncu_digest.py report.ncu-rep [launch_index_for_source]"""
import collections
import csv
import io
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "launch__registers_per_thread",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio"]


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print("==", r[hdr.index("Kernel Name")][:110])
        for w in WANT:
            if w in hdr:
                print("   %-85s %s %s" % (w, r[hdr.index(w)], units[hdr.index(w)]))


def source(rep, skip, top=30):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-skip",
                          str(skip), "--launch-count", "1"], capture_output=True, text=True).stdout
    cur, agg, src, total = None, collections.Counter(), {}, 0
    for r in csv.reader(io.StringIO(out)):
        if len(r) == 2 and r[0] == "File Path":
            cur = r[1].split("/")[-1]
        elif len(r) == 2 and r[0] == "Function Name":
            print("== source of", r[1][:110])
        elif len(r) > 5 and r[0].isdigit():
            try:
                n = int(r[4])
            except ValueError:
                continue
            agg[(cur, int(r[0]))] += n
            src[(cur, int(r[0]))] = r[1].strip()[:105]
            total += n
    for k, n in agg.most_common(top):
        print("%5.1f%% %s:%d  %s" % (100.0 * n / max(total, 1), k[0], k[1], src[k]))


if __name__ == "__main__":
    raw(sys.argv[1])
    for a in sys.argv[2:]:
        source(sys.argv[1], int(a))
