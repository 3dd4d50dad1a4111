#!/usr/bin/env python3
"""Static single-warp issue model of a kernel's main loop from the SASS control words:
sum of the per-instruction stall fields (bits [105:109)) over one outer-loop iteration (inner
loops weighted by `inner_trips`), by opcode.  usage: sass_stalls.py file.cubin [kernel-substring] [inner_trips]"""
import re, subprocess, sys, collections

def parse(cubin, want):
    txt = subprocess.check_output(["cuobjdump", "-sass", cubin], text=True)
    for k in re.split(r"\n\s*Function : ", txt)[1:]:
        name = k.split("\n", 1)[0]
        if want not in name:
            continue
        ins = []
        lines = k.split("\n")
        i = 0
        while i < len(lines):
            m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/\s+(@!?U?P\d\s+)?([A-Z0-9_.]+)(.*?);\s*/\* 0x([0-9a-f]{16}) \*/", lines[i])
            if m and i + 1 < len(lines):
                m2 = re.search(r"/\* 0x([0-9a-f]{16}) \*/", lines[i + 1])
                hi = int(m2.group(1), 16) if m2 else 0
                ins.append(dict(addr=int(m.group(1), 16), op=m.group(3), rest=m.group(4), stall=(hi >> 41) & 0xF,
                                yld=(hi >> 45) & 1, wmask=(hi >> 52) & 0x3F))
                i += 2
            else:
                i += 1
        yield name, ins

def main():
    cubin = sys.argv[1]
    want = sys.argv[2] if len(sys.argv) > 2 else ""
    trips = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    for name, ins in parse(cubin, want):
        idx = {x["addr"]: i for i, x in enumerate(ins)}
        loops = []
        for i, x in enumerate(ins):
            if x["op"].startswith("BRA"):
                t = re.search(r"0x([0-9a-f]+)", x["rest"])
                if t and int(t.group(1), 16) <= x["addr"] and int(t.group(1), 16) in idx:
                    loops.append((idx[int(t.group(1), 16)], i))
        if not loops:
            continue
        outer = max(loops, key=lambda l: l[1] - l[0])
        w = [0] * len(ins)
        for i in range(outer[0], outer[1] + 1):
            w[i] = 1
        for l in loops:
            if l != outer and l[0] >= outer[0] and l[1] <= outer[1]:
                for i in range(l[0], l[1] + 1):
                    w[i] = trips
        n = sum(w)
        tot = sum(x["stall"] * wi for x, wi in zip(ins, w))
        by = collections.Counter(); cnt = collections.Counter(); hist = collections.Counter()
        for x, wi in zip(ins, w):
            if wi:
                by[x["op"]] += x["stall"] * wi; cnt[x["op"]] += wi; hist[x["stall"]] += wi
        print(name[:90])
        print(f"  {n} instr/iter, sum of stall fields {tot} cycles (avg {tot / n:.2f}), stall histogram {dict(sorted(hist.items()))}")
        for op, c in by.most_common(12):
            print(f"    {op:24s} n={cnt[op]:5d} stall-sum={c:6d} avg={c / cnt[op]:.2f}")

main()
