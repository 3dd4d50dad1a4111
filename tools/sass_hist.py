#!/usr/bin/env python3
"""Static SASS opcode histogram of a kernel's main loop: instructions per outer-loop iteration,
inner loops (backward branches nested inside) weighted by their trip count (default 2 for the
j-loop of the blind rotation).  usage: sass_hist.py file.cubin [kernel-substring] [inner_trips]"""
import re, subprocess, sys, collections

def main():
    cubin = sys.argv[1]
    want = sys.argv[2] if len(sys.argv) > 2 else ""
    trips = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    txt = subprocess.check_output(["cuobjdump", "-sass", cubin], text=True)
    kernels = re.split(r"\n\s*Function : ", txt)[1:]
    for k in kernels:
        name = k.split("\n", 1)[0]
        if want not in name:
            continue
        ins = []
        for m in re.finditer(r"/\*([0-9a-f]{4,6})\*/\s+(@!?U?P\d\s+)?([A-Z0-9_.]+)(.*?);", k):
            ins.append((int(m.group(1), 16), m.group(3), m.group(4)))
        addr_idx = {a: i for i, (a, _, _) in enumerate(ins)}
        loops = []
        for i, (a, op, rest) in enumerate(ins):
            if op.startswith("BRA"):
                t = re.search(r"0x([0-9a-f]+)", rest)
                if t and int(t.group(1), 16) <= a and int(t.group(1), 16) in addr_idx:
                    loops.append((addr_idx[int(t.group(1), 16)], i))
        if not loops:
            print(name, "no loop"); continue
        outer = max(loops, key=lambda l: l[1] - l[0])
        weight = [0] * len(ins)
        for i in range(outer[0], outer[1] + 1):
            weight[i] = 1
        for l in loops:
            if l != outer and l[0] >= outer[0] and l[1] <= outer[1]:
                for i in range(l[0], l[1] + 1):
                    weight[i] = trips
        hist = collections.Counter()
        for (a, op, _), w in zip(ins, weight):
            if w:
                hist[op] += w
        total = sum(hist.values())
        def cls(op):
            b = op.split(".")[0]
            if b == "IMAD":
                if ".HI" in op or ".WIDE" in op: return "fma2"
                return "fma1"
            if b in ("FFMA", "FMUL", "FADD", "IDP", "IDP4A"): return "fma1"
            if b in ("LDS", "STS", "LDG", "STG", "LDC", "LDCU", "LD", "ST", "ATOMS", "LDSM"): return "lsu"
            if b in ("BAR", "BRA", "WARPSYNC", "BSSY", "BSYNC", "NOP", "EXIT", "CALL", "RET", "YIELD", "S2R", "S2UR", "CS2R", "R2UR", "DEPBAR", "ERRBAR", "MEMBAR"): return "ctl"
            if b.startswith("U"): return "uni"
            return "alu"
        c = collections.Counter()
        for op, n in hist.items():
            c[cls(op)] += n
        slots = c["fma1"] + 2 * c["fma2"]
        print(f"{name[:100]}")
        print(f"  loop body {outer[1]-outer[0]+1} static instr, {total} dynamic instr/iter: fma-pipe instr {c['fma1']+c['fma2']} "
              f"(slots {slots} = {2*slots} cycles), alu {c['alu']} (= {2*c['alu']} cycles), lsu {c['lsu']}, ctl {c['ctl']}, uniform {c['uni']}")
        for op, n in hist.most_common(40):
            print(f"    {op:28s} {n:6d}")

main()
