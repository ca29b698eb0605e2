"""Aggregate the per-instruction warp-stall samples of an .ncu-rep source page.
usage: python tools/ncu_stalls.py REP KERNEL_SUBSTR [--top N] [--seq]
Prints per-opcode totals of (not-issued) stall samples by reason, and the top-N hottest
instructions; --seq prints the running sample profile along the program (64-instr buckets)."""
import csv
import collections
import io
import re
import subprocess
import sys


def main():
    rep, sub = sys.argv[1], sys.argv[2]
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 25
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    # the dump is a sequence of per-kernel tables, each preceded by a "Kernel Name" row
    blocks, cur = [], None
    for row in csv.reader(io.StringIO(raw)):
        if row and row[0] == "Kernel Name":
            cur = {"name": row[1], "rows": []}
            blocks.append(cur)
        elif cur is not None:
            cur["rows"].append(row)
    blk = [b for b in blocks if sub in b["name"].replace("(int)", "")][-1]
    hdr, rows = blk["rows"][0], blk["rows"][1:]
    col = {h: i for i, h in enumerate(hdr)}
    reasons = [h for h in hdr if h.startswith("stall_") and h.endswith("(Not Issued)")]
    by_op = collections.defaultdict(lambda: collections.Counter())
    tot = collections.Counter()
    hot = []
    execd = collections.Counter()
    for n, r in enumerate(rows):
        src = r[col["Source"]].strip()
        m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_]+)", src)
        op = m.group(2) if m else "?"
        ns = int(r[col["Warp Stall Sampling (Not-issued Samples)"]] or 0)
        execd[op] += int(r[col["Instructions Executed"]] or 0)
        for h in reasons:
            v = int(r[col[h]] or 0)
            if v:
                by_op[op][h[6:-13]] += v
                tot[h[6:-13]] += v
        hot.append((ns, n, src, {h[6:-13]: int(r[col[h]] or 0) for h in reasons if int(r[col[h]] or 0)}))
    print("kernel:", blk["name"].replace("(int)", ""))
    allv = sum(tot.values())
    print("not-issued samples by reason:", ", ".join("%s %.1f%%" % (k, 100.0 * v / allv) for k, v in tot.most_common()))
    print("\nper opcode (share of not-issued samples; executed warp-instr share):")
    allx = sum(execd.values())
    for op, c in sorted(by_op.items(), key=lambda kv: -sum(kv[1].values()))[:18]:
        s = sum(c.values())
        print("  %-8s %5.1f%%  exec %5.1f%%  %s" % (op, 100.0 * s / allv, 100.0 * execd[op] / allx,
                                                  ", ".join("%s %d" % (k, v) for k, v in c.most_common(4))))
    print("\nhottest instructions:")
    for ns, n, src, d in sorted(hot, reverse=True)[:top]:
        print("  #%-5d %6d  %-60s %s" % (n, ns, src[:60], d))
    if "--seq" in sys.argv:
        print("\nprofile along the program (bucket of 64 instructions: samples):")
        for b in range(0, len(hot), 64):
            print("  %5d %7d" % (b, sum(h[0] for h in hot[b:b + 64])))


if __name__ == "__main__":
    main()
