"""Executed warp instructions per opcode from an ncu report's source page.

    ncu -i report.ncu-rep --page source --csv > src.csv
    python tools/ncu_opcodes.py src.csv [pairs]      (pairs: divide by this many warp-level units)
"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
unit = float(sys.argv[2]) if len(sys.argv) > 2 else None
start = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[start]
i_src, i_exec, i_samp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
ops, samp = collections.Counter(), collections.Counter()
for r in rows[start + 1:]:
    if len(r) <= i_exec or not r[i_exec].isdigit():
        continue
    toks = r[i_src].split()
    op = toks[1] if toks[0].startswith("@") else toks[0]
    op = op.split(".")[0] if not op.startswith("MUFU") else op
    ops[op] += int(r[i_exec])
    samp[op] += int(r[i_samp])
total = sum(ops.values())
print("total warp instructions: %d" % total + (" = %.2f per unit" % (total / unit) if unit else ""))
for op, n in ops.most_common(25):
    print("%-14s %12d  %5.1f %%  %s  samples %d" % (op, n, 100.0 * n / total,
          ("%.3f/unit" % (n / unit)) if unit else "", samp[op]))
