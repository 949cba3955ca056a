"""Instruction mix / hottest SASS of the first kernel in `ncu -i rep --page source --csv --print-source sass` output (stdin)."""
import collections, csv, sys
rows = list(csv.reader(sys.stdin))
hdr = rows[1]
ci, ie, sm = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
end = next((i for i in range(2, len(rows)) if rows[i] and rows[i][0] == "Kernel Name"), len(rows))
print(rows[0][1][:100])
ops, samp, tot, tots = collections.Counter(), collections.Counter(), 0, 0
body = []
for r in rows[2:end]:
    try:
        n, s = int(r[ie]), int(r[sm] or 0)
    except (ValueError, IndexError):
        continue
    toks = r[ci].split()
    op = toks[1] if toks and toks[0].startswith("@") and len(toks) > 1 else (toks[0] if toks else "")
    ops[op.split(".")[0]] += n; samp[op.split(".")[0]] += s; tot += n; tots += s
    body.append((n, s, r[ci]))
print("total warp-instructions", tot, "samples", tots)
for k, v in ops.most_common(22):
    print(f"{k:12s} {v:11d} {100*v/tot:5.1f}%   samples {100*samp[k]/max(tots,1):5.1f}%")
if len(sys.argv) > 1:
    print("-- hottest by samples")
    for n, s, src in sorted(body, key=lambda t: -t[1])[:int(sys.argv[1])]:
        print(f"{n:9d} {s:6d}  {src[:110]}")
