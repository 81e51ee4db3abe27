"""Summaries of ncu CSV logs for profiles/: per-kernel totals of a launch list, and DRAM bytes per launch of a metrics pass.
usage: python profiles/summarize_ncu.py launches <ncu --csv log> | dram <ncu --csv log> | full <report.ncu-rep>"""
import csv
import sys
from collections import defaultdict


def rows_of(path):
    lines = [l for l in open(path, errors="replace") if l.startswith('"')]
    r = list(csv.reader(lines))
    return r[0], r[1:]


def main():
    mode, path = sys.argv[1], sys.argv[2]
    hdr, rows = rows_of(path)
    ik, im, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    iid = hdr.index("ID")
    if mode == "launches":
        tot, cnt = defaultdict(float), defaultdict(int)
        for r in rows:
            if r[im] != "gpu__time_duration.sum":
                continue
            v = float(r[iv].replace(",", ""))
            ms = v / 1e6 if r[iu] in ("ns", "nsecond") else v / 1e3 if r[iu] in ("us", "usecond") else v
            tot[r[ik]] += ms; cnt[r[ik]] += 1
        allms = sum(tot.values())
        for k in sorted(tot, key=lambda k: -tot[k]):
            print("%-72s n=%4d ms=%9.3f %5.1f%%" % (k[:72], cnt[k], tot[k], 100 * tot[k] / allms))
        print("total %.3f ms in %d launches" % (allms, sum(cnt.values())))
    else:
        per = defaultdict(dict)
        for r in rows:
            v = float(r[iv].replace(",", ""))
            u = r[iu]
            scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "nsecond": 1e-9,
                     "usecond": 1e-6, "msecond": 1e-3, "second": 1.0}.get(u, 1.0)
            per[r[iid]][r[im]] = v * scale
            per[r[iid]]["kernel"] = r[ik][:40]
        for i, d in per.items():
            b = d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
            t = d.get("gpu__time_duration.sum", 0.0)
            print("launch %s %s: dram read %.3f GB + write %.3f GB = %.3f GB in %.3f ms (%.0f GB/s under ncu)" % (
                i, d["kernel"], d.get("dram__bytes_read.sum", 0) / 1e9, d.get("dram__bytes_write.sum", 0) / 1e9, b / 1e9, t * 1e3,
                b / t / 1e9 if t else 0))




def full_summary(rep):
    """one line + stall breakdown per captured launch of an `ncu --set full` report (reads it through `ncu -i --page raw --csv`)."""
    import subprocess
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader([l for l in out.splitlines() if l.startswith('"')]))
    hdr, rows = r[0], r[2:]
    col = {h: i for i, h in enumerate(hdr)}
    g = lambda row, k: float(row[col[k]].replace(",", "")) if k in col and row[col[k]] not in ("", "n/a") else float("nan")
    print("%-34s %8s %10s %10s %6s %13s %7s %7s %7s %7s %8s %10s" % ("kernel", "time_ms", "dram_rd_GB", "dram_wr_MB", "regs",
                                                                   "grid x block", "L1tex%", "L2%", "dram%", "fp64%", "warps%", "inst"))
    units = {h: r[1][i] for i, h in enumerate(hdr)}
    sc = lambda k: {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "nsecond": 1e-9,
                    "usecond": 1e-6, "msecond": 1e-3}.get(units.get(k, ""), 1.0)
    for row in rows:
        name = row[col["Kernel Name"]].split("(")[0].replace("void ", "")
        t = g(row, "gpu__time_duration.sum") * sc("gpu__time_duration.sum")
        print("%-34s %8.3f %10.2f %10.2f %6d %13s %7.1f %7.1f %7.1f %7.1f %8.1f %10.2e" % (
            name[:34], t * 1e3, g(row, "dram__bytes_read.sum") * sc("dram__bytes_read.sum") / 1e9,
            g(row, "dram__bytes_write.sum") * sc("dram__bytes_write.sum") / 1e6, int(g(row, "launch__registers_per_thread")),
            "%dx%d" % (int(g(row, "launch__grid_size")), int(g(row, "launch__block_size"))),
            g(row, "l1tex__throughput.avg.pct_of_peak_sustained_active"), g(row, "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
            g(row, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
            g(row, "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"),
            g(row, "sm__warps_active.avg.pct_of_peak_sustained_active"), g(row, "smsp__inst_executed.sum")))
        st = []
        for h in hdr:
            if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio"):
                v = g(row, h)
                if v == v and v >= 0.25:
                    st.append((v, h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
        print("    stalls per issue: " + ", ".join("%s %.2f" % (n, v) for v, n in sorted(st, reverse=True)))


if __name__ == "__main__":
    if sys.argv[1] == "full":
        full_summary(sys.argv[2])
    else:
        main()
