"""Print the per-point loop of one pair kernel (see sass_report.py); optional: hide FP ops.
usage: python tools/sass_loop.py D '<10, 1, 1, 0, 128, 3>' [--nofp]"""
import re, subprocess, sys, os
CSRC = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "density_deconvolution_b200", "csrc")
D = int(sys.argv[1]); flt = sys.argv[2]; nofp = "--nofp" in sys.argv
obj = os.path.join(CSRC, "build", "dgmm_inst_d%02d.o" % D)
if "--obj" in sys.argv: obj = sys.argv[sys.argv.index("--obj") + 1]
txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
for f in re.split(r"\n\s*Function : ", txt)[1:]:
    name = subprocess.run(["c++filt", f.split("\n")[0].strip()], capture_output=True, text=True).stdout.strip()
    if flt not in name: continue
    ins = []
    for l in f.split("\n"):
        m = re.search(r"/\*([0-9a-f]{4,5})\*/\s+(.*?);", l)
        if m: ins.append((int(m.group(1), 16), m.group(2).strip()))
    best = None
    for addr, t in ins:
        m = re.match(r"(@!?U?P\d+\s+)?BRA\S*\s.*0x([0-9a-f]+)", t)
        if m and int(m.group(2), 16) < addr:
            tgt = int(m.group(2), 16)
            body = [(a, x) for a, x in ins if tgt <= a <= addr]
            if sum("FFMA" in x for _, x in body) >= 20 and any("MUFU.RCP" in x or "MUFU.RSQ" in x for _, x in body):
                if best is None or len(body) < len(best): best = body
    for a, x in best:
        if nofp and re.match(r"(@!?P\d+\s+)?(FFMA|FMUL|FADD)\b", x): continue
        print("%05x  %s" % (a, x))
