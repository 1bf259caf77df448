"""SASS of transit_kernel<S=1, WPT=8, 256 threads>'s chord loop (both sigma-window variants), with the issue-slot count
the FP64-pipe roofline in DESIGN.md rests on.

    python profiles/sass_excerpt.py > profiles/sass_transit_inner_loop.txt      (build container: cuobjdump only)
    python profiles/sass_excerpt.py moment > profiles/sass_moment_item_loop.txt (kernel (m)'s item loop, same accounting)
    python profiles/sass_excerpt.py cells  > profiles/sass_cell_moment_loop.txt  (sigma cells: the chord loop of cell_moment_kernel)

The chord loop is the innermost backward branch whose body holds 6 FP64 instructions per (chord, wavelength) sample:
DFMA t, DADD t - magic, DFMA r, DFMA q1, DFMA q0, DFMA acc (pb200_transit.cu: transit_chords)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "prometheus_b200", "libprometheus_b200.so")
FUN = "_ZN5pb20014transit_kernelILi1ELi8ELi256ELb%dEEEvPKNS_8DevSceneENS_13TransitParamsE"
WPT = 8


def loops(fun):
    txt = subprocess.run(["cuobjdump", "-sass", "-fun", fun, LIB], capture_output=True, text=True).stdout
    ins = [(int(m.group(1), 16), m.group(2).strip()) for m in re.finditer(r"/\*([0-9a-f]{4,})\*/\s+(.*?);", txt)]
    out = []
    for a, t in ins:
        m = re.search(r"\bBRA\b.*?(0x[0-9a-f]+)", t)
        if m and int(m.group(1), 16) < a:
            out.append([x for x in ins if int(m.group(1), 16) <= x[0] <= a])
    return out


def opcode(t):
    return re.sub(r"^@!?U?P\d+\s+", "", t).split()[0].split(".")[0]


def main():
    for hi in (0, 1):
        fun = FUN % hi
        print("=" * 110)
        print("transit_kernel<S=1, WPT=8, THREADS=256, HI=%s>   %s" % ("true" if hi else "false", fun))
        for body in loops(fun):
            h = collections.Counter(opcode(t) for _, t in body)
            fp64 = sum(v for k, v in h.items() if k in ("DFMA", "DADD", "DMUL", "DSETP", "DMNMX"))
            if len(body) > 150 or fp64 < 6 * WPT:
                continue
            kind = "plain (tau < 693.8 guaranteed by compact_kernel's window class)" if h.get("DMUL", 0) == 0 else \
                   "clamped (windows that hold a chord with tau >= 693.8: v = min(v, clamp) before the table)"
            print("-" * 110)
            print("chord loop, %s: 0x%04x..0x%04x, one iteration = 1 chord record x %d wavelengths per thread" % (kind, body[0][0], body[-1][0], WPT))
            print("  issue slots %d = %d FP64 + %d other   -> %.2f slots, %.2f FP64 per (chord, wavelength) sample" %
                  (len(body), fp64, len(body) - fp64, len(body) / WPT, fp64 / WPT))
            print("  opcodes: " + ", ".join("%s %d" % kv for kv in sorted(h.items(), key=lambda kv: -kv[1])))
            for a, t in body:
                print("    /*%04x*/  %s" % (a, t))
    print("=" * 110)
    print("Reading: an sm_100a SM sub-partition (SMSP) issues one warp-wide DFMA/DADD/DMUL every 2 cycles (16 FP64 lanes per SMSP,")
    print("64 per SM; measured: profiles/r01_microbench_issue_model.txt, 2.02 cycles).  The plain loop's 48 FP64 instructions per 8")
    print("samples = 96 cycles per iteration per SMSP are the floor behind bench.py's roofline (6 FP64 instructions = 11 flops per")
    print("sample: 5 DFMA + 1 DADD).  The other 41 slots (8 table LDS.64 + 2 record LDS.128 broadcasts, LOP3 index masks, IMAD index")
    print("scaling and exponent adds, BREV/FLO of the ballot walk, loop control) go to other pipes; 89 slots < 96 cycles, so the loop")
    print("is FP64-bound only if every one of them issues in the shadow of an FP64 instruction.  Measured: 0.555 of the pipe.")


MOMENT = "_ZN5pb20021transit_moment_kernelEPKNS_8DevSceneENS_13TransitParamsEPKdiiiNS_9MolLayoutE"


def moment():
    print("=" * 110)
    print("transit_moment_kernel (kernel (m), consumer threads): the item loop   %s" % MOMENT)
    best = None
    for body in loops(MOMENT):
        h = collections.Counter(opcode(t) for _, t in body)
        fp64 = sum(v for k, v in h.items() if k in ("DFMA", "DADD", "DMUL"))
        if len(body) < 300 and fp64 >= 60 and (best is None or len(body) < len(best[0])):
            best = (body, h, fp64)
    body, h, fp64 = best
    n_items = h["DMUL"]      # one DMUL (r * poly) per item
    print("item loop 0x%04x..0x%04x: %d items per iteration (compiler unroll)" % (body[0][0], body[-1][0], n_items))
    print("  issue slots %d = %d FP64 + %d other  -> per (item, wavelength): %.1f slots, %.1f FP64 instructions" %
          (len(body), fp64, len(body) - fp64, len(body) / n_items, fp64 / n_items))
    print("  FP64 per item: %d DFMA + %d DADD + %d DMUL = %d flops (bench.py FLOPS_MOMENT_ITEM = 33)" %
          (h["DFMA"] // n_items, h["DADD"] // n_items, h["DMUL"] // n_items,
           (2 * h["DFMA"] + h["DADD"] + h["DMUL"]) // n_items))
    print("  opcodes: " + ", ".join("%s %d" % kv for kv in sorted(h.items(), key=lambda kv: -kv[1])))
    for a, t in body:
        print("    /*%04x*/  %s" % (a, t))
    print("=" * 110)
    print("Reading: d = V[j+1] - V[j] (DADD), 8 moment Horner steps + V + t_c d (DFMA), 2^x by a 256-entry table and a cubic (DFMA")
    print("magic, DADD, DFMA reduce, 2 DFMA + DMUL poly, DFMA table), accumulate (DFMA).  FP64 floor 36 cycles per item per SMSP against")
    print("%.1f issue slots: 9 LDS per item (5 broadcast LDS.128 of the item, V pair, table entry) make issue and LSU co-limiters." % (len(body) / n_items))


CELLS = "_ZN5pb20018cell_moment_kernelEPKNS_8DevSceneENS_13TransitParamsEPdS4_i"


def cells():
    print("=" * 110)
    print("cell_moment_kernel (sigma cells): the chord loop   %s" % CELLS)
    best = None
    for body in loops(CELLS):
        h = collections.Counter(opcode(t) for _, t in body)
        fp64 = sum(v for k, v in h.items() if k in ("DFMA", "DADD", "DMUL", "DSETP"))
        if fp64 >= 30 and (best is None or len(body) < len(best[0])):
            best = (body, h, fp64)
    body, h, fp64 = best
    n = max(1, h["LDS"] // 3)        # two broadcast LDS.128 of the record + one table LDS.64 per chord
    print("chord loop 0x%04x..0x%04x: %d chord record(s) per iteration, lane = cell" % (body[0][0], body[-1][0], n))
    print("  issue slots %d = %d FP64 + %d other  -> per (chord, cell): %.1f slots, %.1f FP64 instructions (bench.py FLOPS_CENTRE_PAIR = 49 flops)" %
          (len(body), fp64, len(body) - fp64, len(body) / n, fp64 / n))
    print("  opcodes: " + ", ".join("%s %d" % kv for kv in sorted(h.items(), key=lambda kv: -kv[1])))
    for a, t in body:
        print("    /*%04x*/  %s" % (a, t))
    print("=" * 110)
    print("Reading: opaque test (DMUL + DSETP), table exp at the cell centre (DFMA magic, DADD, DFMA reduce, 2 DFMA poly, DMUL by the")
    print("table entry), u = N h (DMUL), 13 running sums (DADD, DMUL + DADD, 11 x (DMUL + DFMA)).  ncu: FP64 pipe 67 % busy at C5.")


if __name__ == "__main__":
    a = sys.argv[1:]
    moment() if "moment" in a else cells() if "cells" in a else main()
