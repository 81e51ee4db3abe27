"""SASS opcode counts per kernel of the built library (cuobjdump -sass): the proof that the TMA / mbarrier / RED / FP64 paths are in
the shipped code.  usage: python profiles/sass_summary.py > profiles/r2_sass_summary.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "fempar_b200", "libfempar_b200.so")
KEEP = ["UBLKCP", "SYNCS", "REDG", "RED.", "DFMA", "DMUL", "DADD", "LDS.128", "LDS.64", "LDS.U16", "STS.64", "STG.E.64", "SHFL",
        "BAR.SYNC", "LDG.E.64", "LDG.E.CONSTANT", "ATOMG", "ST.E", "MEMBAR", "LDC"]
WANT = re.compile(r"k_spmv|k_assemble|k_cg_|k_dot|k_row_fill|k_peer|k_rows_|k_tile_|k_class|k_sum_partials|k_error_norms|k_integrate")


def main():
    out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
    print("SASS opcode counts per kernel of fempar_b200/libfempar_b200.so (cuobjdump -sass, sm_100a); UBLKCP = cp.async.bulk (TMA), "
          "SYNCS.* = mbarrier")
    name, cnt, total = None, None, 0

    def flush():
        if name and WANT.search(name):
            parts = ["%s=%d" % (k, v) for k, v in cnt.items()]
            print("%-110s total %5d  %s" % (name[:110], total, "  ".join(parts)))

    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            flush()
            name, cnt, total = m.group(1), collections.OrderedDict(), 0
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and name:
            total += 1
            op = m.group(1)
            if any(op.startswith(k) for k in KEEP):
                key = op if op.startswith(("UBLKCP", "SYNCS", "REDG", "ATOMG")) else ".".join(op.split(".")[:2]) if op.startswith(("LDS", "STS", "LDG", "STG", "BAR", "SHFL")) else op.split(".")[0]
                cnt[key] = cnt.get(key, 0) + 1
    flush()


if __name__ == "__main__":
    main()
