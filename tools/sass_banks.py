"""Register-bank statistics of the FFMA/FMUL/FADD instructions of one SASS function:
how many read three (two) distinct source registers of the same bank without a .reuse hit.
usage: python tools/sass_banks.py file.sass 'substring of demangled name'"""
import re, sys, subprocess, collections

txt = open(sys.argv[1]).read()
flt = sys.argv[2]
for f in re.split(r"\n\s*Function : ", txt)[1:]:
    name = subprocess.run(["c++filt", f.split("\n")[0].strip()], capture_output=True, text=True).stdout.strip()
    if flt not in name:
        continue
    tot = collections.Counter(); c2 = collections.Counter(); c4 = collections.Counter(); reuse = 0; nsrc = collections.Counter()
    for l in f.split("\n"):
        m = re.search(r"/\*[0-9a-f]{4,6}\*/\s+(@!?U?P\d+\s+)?(FFMA|FMUL|FADD)(\.\S+)?\s+(.*?);", l)
        if not m:
            continue
        op = m.group(2)
        ops = [o.strip() for o in m.group(4).split(",")]
        srcs = ops[1:]
        regs = []
        for o in srcs:
            r = re.match(r"[-|]?R(\d+)(\.reuse)?", o)
            if r and not o.startswith("RZ"):
                if r.group(2):
                    reuse += 1
                regs.append((int(r.group(1)), bool(r.group(2))))
        tot[op] += 1
        distinct = {r for r, _ in regs}
        nsrc[(op, len(distinct))] += 1
        for mod, cnt in ((2, c2), (4, c4)):
            banks = collections.Counter(r % mod for r in distinct)
            worst = max(banks.values()) if banks else 0
            cnt[(op, worst)] += 1
    print(name)
    print("  totals", dict(tot), "reuse-flags", reuse)
    print("  distinct reg sources", dict(sorted(nsrc.items())))
    print("  mod2 worst-bank occupancy", dict(sorted(c2.items())))
    print("  mod4 worst-bank occupancy", dict(sorted(c4.items())))
