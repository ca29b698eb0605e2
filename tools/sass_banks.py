"""Register-file read model of one SASS function (per-point loop of a pair kernel): how many FFMA / FMUL / FADD
need more than one cycle of operand fetch because two or three of their register sources that are NOT served
by the operand-reuse cache sit in the same register bank.  Reported for the two bank hypotheses
(bank = reg % 2, bank = reg % 4); the predicted issue efficiency is instr / (instr + extra fetch cycles).

usage: python tools/sass_banks.py file.sass 'substring of the demangled kernel name'"""
import collections
import re
import subprocess
import sys


def main():
    txt = open(sys.argv[1]).read()
    flt = sys.argv[2]
    for f in re.split(r"\n\s*Function : ", txt)[1:]:
        name = subprocess.run(["c++filt", f.split("\n")[0].strip()], capture_output=True, text=True).stdout.strip()
        if flt not in name:
            continue
        ins = []
        for l in f.split("\n"):
            m = re.search(r"/\*([0-9a-f]{4,6})\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)\s*(.*?);", l)
            if m:
                ins.append((int(m.group(1), 16), m.group(3), m.group(4)))
        # the per-point loop: the largest backward-branch region that contains MUFU.RCP
        best = None
        for k, (addr, op, rest) in enumerate(ins):
            if op.startswith("BRA"):
                t = re.search(r"0x([0-9a-f]+)", rest)
                if t and int(t.group(1), 16) < addr:
                    tgt = int(t.group(1), 16)
                    body = [x for x in ins if tgt <= x[0] <= addr]
                    if sum(o.startswith("FFMA") for _, o, _ in body) >= 20 and any(o.startswith("MUFU.RCP") for _, o, _ in body):
                        if best is None or len(body) < len(best):
                            best = body
        body = best or ins
        reuse_slot = {}   # slot index -> register held by the reuse cache
        extra = {2: 0, 4: 0}
        n_fp = 0
        hist = collections.Counter()
        for addr, op, rest in body:
            base = op.split(".")[0]
            ops = [o.strip() for o in rest.split(",")]
            srcs = ops[1:] if base in ("FFMA", "FMUL", "FADD") else []
            new_reuse = {}
            reads = []
            for slot, o in enumerate(srcs):
                r = re.match(r"[-|]?R(\d+)(\.reuse)?", o)
                if not r or o.startswith("RZ") or o.startswith("-RZ"):
                    continue
                reg = int(r.group(1))
                if reuse_slot.get(slot) != reg:
                    reads.append(reg)
                if r.group(2):
                    new_reuse[slot] = reg
            if base in ("FFMA", "FMUL", "FADD"):
                n_fp += 1
                distinct = set(reads)
                hist[len(distinct)] += 1
                for mod in (2, 4):
                    banks = collections.Counter(x % mod for x in distinct)
                    extra[mod] += max(banks.values()) - 1 if banks else 0
            reuse_slot = new_reuse if base in ("FFMA", "FMUL", "FADD") else {}
        n = len(body)
        print(name)
        print("  loop: %d instr, %d FFMA/FMUL/FADD; register-file reads per FP instr (after the reuse cache): %s"
              % (n, n_fp, dict(sorted(hist.items()))))
        for mod in (2, 4):
            print("  bank = reg %% %d: %d extra operand-fetch cycles -> issue efficiency <= %.3f"
                  % (mod, extra[mod], n / (n + extra[mod])))


if __name__ == "__main__":
    main()
