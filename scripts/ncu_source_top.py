"""Top stall-sample instructions of a kernel from `ncu -i rep --page source --csv ...` output: python scripts/ncu_source_top.py file.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
isrc, isamp, iexec = col["Source"], col["Warp Stall Sampling (All Samples)"], col["Instructions Executed"]
iw, iwi = col.get("L1 Wavefronts Shared"), col.get("L1 Wavefronts Shared Ideal")
data = []
for i, r in enumerate(rows[hi + 1:]):
    if len(r) <= isamp or not r[0].startswith("0x"):
        continue
    data.append((int(r[isamp] or 0), r[isrc].strip(), int(r[iexec] or 0), r[iw] if iw else "", r[iwi] if iwi else "", i))
tot = sum(d[0] for d in data)
print("total samples", tot, "instructions", len(data), "warp-instr executed", sum(d[2] for d in data))
for d in sorted(data, reverse=True)[:top]:
    print(f"{d[0]:6d} {100*d[0]/max(tot,1):5.1f}%  exec {d[2]:8d}  wf {d[3]:>9s}/{d[4]:>9s}  #{d[5]:4d} {d[1][:100]}")
