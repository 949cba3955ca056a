"""Summarise an `ncu --set full` capture (exported with `ncu -i rep --page raw --csv`) per launch and per kernel family.
Usage: python scripts/ncu_summary.py raw.csv [traffic.json]
Families follow the C-ABI entry points: convT = convt64_kernel + tapgemm_{persistent,pair}_kernel<BN, 1> + tapgemm_kernel launched
with 4 parities (grid.z = 4), conv = tapgemm_{persistent,pair}_kernel<BN, 0> + the other tapgemm_kernel launches, wgrad =
wgrad_kernel; every other kernel is its own family (bandwidth kernels: read the dram% / dramMB columns)."""
import collections
import csv
import json
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
ix = {h: i for i, h in enumerate(hdr)}


def col(prefix):
    for h in hdr:
        if h == prefix or h.startswith(prefix):
            return ix[h]
    return None


C = {"name": col("Kernel Name"), "grid": col("Grid Size"), "dur": col("gpu__time_duration.sum"),
     "rd": col("dram__bytes_read.sum"), "wr": col("dram__bytes_write.sum"),
     "tensor": col("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
     "lts": col("lts__throughput.avg.pct_of_peak_sustained_elapsed"),
     "dram": col("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
     "regs": col("launch__registers_per_thread")}
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3}


def val(row, key):
    i = C[key]
    v = float(row[i].replace(",", ""))
    return v * SCALE.get(units[i], 1.0)


def family(name, grid):
    if "convt64" in name:
        return "b2c_convt4s2_bf16"
    if "wgrad_kernel" in name or "wgrad_pair_kernel" in name:
        return "b2c_wgrad4s2_bf16"
    if "tapgemm_kernel<(int)32" in name or "tapgemm_kernel<32" in name:
        return "b2c_first_dgrad"
    m = re.search(r"tapgemm_(?:persistent|pair)_kernel<\(?(?:int\))?\d+, \(?(?:int\))?(\d)>", name)
    if m:
        return "b2c_convt4s2_bf16" if m.group(1) == "1" else "b2c_conv4s2_bf16"
    if "tapgemm" in name:
        z = int(re.findall(r"\d+", grid)[2])
        return "b2c_convt4s2_bf16" if z == 4 else "b2c_conv4s2_bf16"
    return re.sub(r"\(.*", "", name)


fam = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
print(f"{'kernel':44s} {'grid':>16s} {'us':>8s} {'dramMB':>8s} {'tensor%':>8s} {'lts%':>6s} {'dram%':>6s} {'regs':>5s}")
for r in rows[2:]:
    name, grid = r[C["name"]], r[C["grid"]]
    us, by = val(r, "dur"), val(r, "rd") + val(r, "wr")
    t = float(r[C["tensor"]]) if C["tensor"] is not None and r[C["tensor"]] not in ("", "n/a") else 0.0
    print(f"{re.sub(r'[(]FirstIn.*|[(]CUtensorMap.*|[(]const.*', '', name).replace('(int)', '').replace('(bool)', '')[:44]:44s} {grid:>16s} {us:8.1f} {by/1e6:8.1f} {t:8.1f} "
          f"{float(r[C['lts']]):6.1f} {float(r[C['dram']]):6.1f} {r[C['regs']]:>5s}")
    f = fam[family(name, grid)]
    f[0] += 1; f[1] += us; f[2] += by; f[3] += t * us
print()
out = {}
for k, (n, us, by, tw) in sorted(fam.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:28s} launches {n:4d}  {us/1e3:8.3f} ms  dram {by/n/1e6:8.1f} MB/launch  time-weighted tensor-pipe active {tw/us:5.1f} %")
    out[k] = {"launches": n, "dram_bytes_per_launch": by / n, "us_per_launch": us / n, "tensor_pipe_active_pct": tw / us}
if len(sys.argv) > 2:
    json.dump(out, open(sys.argv[2], "w"), indent=1)
