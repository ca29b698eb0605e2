"""Static report for one per-D object file: registers / spills per pair kernel (from
`nvcc -Xptxas -v` output) and the opcode mix of the per-point loop (largest innermost
backward-branch region containing MUFU) from `cuobjdump -sass`.

usage: python tools/sass_report.py D [filter-substring]
"""
import collections
import os
import re
import subprocess
import sys

CSRC = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "density_deconvolution_b200", "csrc")


def demangle(n):
    return subprocess.run(["c++filt", n.strip()], capture_output=True, text=True).stdout.strip()


def main():
    D = int(sys.argv[1])
    rest = [a for i, a in enumerate(sys.argv[2:]) if not a.startswith("--") and sys.argv[i + 1] != "--obj"]
    flt = rest[0] if rest else ""
    obj = os.path.join(CSRC, "build", "dgmm_inst_d%02d.o" % D)
    if "--obj" in sys.argv:
        obj = sys.argv[sys.argv.index("--obj") + 1]
    flags = "-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -Xptxas -v"
    if "--no-build" not in sys.argv:
        r = subprocess.run("nvcc %s -c dgmm_inst_d%02d.cu -o %s" % (flags, D, obj), shell=True, cwd=CSRC,
                           capture_output=True, text=True)
        cur = None
        for line in r.stderr.splitlines():
            m = re.search(r"Compiling entry function '(\S+)'", line)
            if m:
                cur = demangle(m.group(1))
            m = re.search(r"Used (\d+) registers", line)
            if m and cur and flt in cur:
                print("%-75s regs %s" % (cur.replace("dgmm::", "").replace("(PairParams)", ""), m.group(1)), end="")
            m = re.search(r"(\d+) bytes spill stores, (\d+) bytes spill loads", line)
            if m and cur and flt in cur:
                spill = "  spill %s/%s" % (m.group(1), m.group(2))
                print(spill if m.group(1) != "0" else "", end="\n" if False else "")
            if "Used" in line and cur and flt in cur:
                print()
    txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    for f in re.split(r"\n\s*Function : ", txt)[1:]:
        name = demangle(f.split("\n")[0])
        if "pair_kernel" not in name or flt not in name:
            continue
        ins = []
        for l in f.split("\n"):
            m = re.search(r"/\*([0-9a-f]{4,5})\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)(.*?);", l)
            if m:
                ins.append((int(m.group(1), 16), m.group(3), m.group(4)))
        best = None
        for addr, op, rest in ins:
            if op.startswith("BRA"):
                t = re.search(r"0x([0-9a-f]+)", rest)
                if t and int(t.group(1), 16) < addr:
                    tgt = int(t.group(1), 16)
                    body = [o for a, o, _ in ins if tgt <= a <= addr]
                    if sum(o.startswith("FFMA") for o in body) >= 20 and \
                            any(o.startswith("MUFU.RCP") or o.startswith("MUFU.RSQ") for o in body):
                        if best is None or len(body) < len(best):
                            best = body
        if best:
            c = collections.Counter(o.split(".")[0] for o in best)
            fp = c["FFMA"] + c["FMUL"] + c["FADD"]
            print("%s\n   per-point loop: %d instr  (FFMA %d FMUL %d FADD %d MUFU %d LDS %d STS %d LDG %d SHFL %d | other %d)"
                  % (name.replace("dgmm::", "").replace("(PairParams)", ""), len(best), c["FFMA"], c["FMUL"], c["FADD"],
                     c["MUFU"], c["LDS"], c["STS"], c["LDG"], c["SHFL"],
                     len(best) - fp - c["MUFU"] - c["LDS"] - c["STS"] - c["LDG"] - c["SHFL"]))
            other = {k: v for k, v in c.items() if k not in ("FFMA", "FMUL", "FADD", "MUFU", "LDS", "STS", "LDG", "SHFL")}
            print("   other:", sorted(other.items(), key=lambda x: -x[1])[:14])


if __name__ == "__main__":
    main()
