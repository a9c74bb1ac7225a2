"""Opcode mix of one kernel from an ncu report's source page:
   ncu -i rep.ncu-rep --page source --csv --kernel-name regex:NAME > src.csv;  python tools/sass_mix.py src.csv CELLS"""
import collections
import csv
import sys


def main(path, cells):
    rows = list(csv.reader(open(path)))
    h = next(r for r in rows if "Instructions Executed" in r)
    ci, si, st = h.index("Instructions Executed"), h.index("Source"), h.index("Warp Stall Sampling (All Samples)")
    ops, stalls, tot = collections.Counter(), collections.Counter(), 0
    for r in rows:
        if len(r) <= ci or not r[ci].isdigit():
            continue
        toks = r[si].split()
        if not toks:
            continue
        op = toks[1] if toks[0].startswith("@") else toks[0]
        op = op.split(".")[0]
        n = int(r[ci])
        ops[op] += n
        tot += n
        stalls[op] += int(r[st] or 0)
    print(f"{tot} warp instructions = {tot * 32 / cells:.0f} thread instructions per cell")
    print("| opcode | warp inst | share | per cell | stall samples |\n|---|---:|---:|---:|---:|")
    for op, n in ops.most_common(28):
        print(f"| {op} | {n} | {100 * n / tot:.1f}% | {n * 32 / cells:.1f} | {stalls[op]} |")


if __name__ == "__main__":
    main(sys.argv[1], float(sys.argv[2]))
