"""Summarise `ncu -i X.ncu-rep --page source --csv`: instruction mix by opcode
and the hottest SASS lines (executed warp-instructions and stall samples)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
H = rows[1]
ix = {h: i for i, h in enumerate(H)}
S, E, P = ix["Source"], ix["Instructions Executed"], ix["# Samples"]
lines = []
for r in rows[2:]:
    try:
        lines.append((int(r[E]), int(r[P]), r[S].strip()))
    except Exception:
        pass
tot = sum(l[0] for l in lines); tots = sum(l[1] for l in lines)
print("total warp-instructions", tot, "samples", tots, "lines", len(lines))
agg = {}
for e, s, t in lines:
    parts = t.split()
    op = parts[1] if parts[0].startswith("@") else parts[0]
    op = op.split(".")[0]
    a = agg.setdefault(op, [0, 0]); a[0] += e; a[1] += s
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%-10s exec %10d (%5.1f%%) samples %6d (%5.1f%%)" % (k, v[0], 100 * v[0] / tot, v[1], 100 * v[1] / max(tots, 1)))
print("--- hottest lines by samples")
for i, (e, s, t) in sorted(enumerate(lines), key=lambda x: -x[1][1])[:top]:
    print("%5d exec %9d samples %6d  %s" % (i, e, s, t[:90]))
