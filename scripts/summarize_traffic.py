"""Summarise gpurun_out/traffic.csv (ncu: duration + DRAM bytes of every launch) for the last full training step:
writes profiles/<round>_ncu_step_traffic.txt (per-kernel table) and profiles/traffic.json (what bench.py reports as
roofline.traffic: average DRAM bytes per launch of the conv forward/dgrad kernel class)."""
import collections, csv, json, re, sys
src = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/traffic.csv"
tag = sys.argv[2] if len(sys.argv) > 2 else "r01"
rows = list(csv.reader(open(src)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
hdr = rows[hi]
ki, mi, vi, ui, ii = (hdr.index(k) for k in ("Kernel Name", "Metric Name", "Metric Value", "Metric Unit", "ID"))
launch = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) <= vi:
        continue
    d = launch.setdefault(r[ii], {"name": r[ki], "grid": r[hdr.index("Grid Size")].replace(",", " ")})
    v = float(r[vi].replace(",", ""))
    unit = r[ui]
    if r[mi].startswith("dram__bytes"):
        v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]
    elif r[mi].startswith("sm__pipe_tensor"):
        pass   # percent
    else:
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1.0)   # -> us
    d[r[mi]] = v
L = list(launch.values())
# a step ends with step_increment_kernel (once per step; Adam runs once per gradient bucket)
adam = [i for i, l in enumerate(L) if "step_increment_kernel" in l["name"]]
step = L[adam[-2] + 1: adam[-1] + 1]
agg = collections.OrderedDict()
for l in step:
    k = re.sub(r"\(.*", "", l["name"]).split("::")[-1]
    a = agg.setdefault(k, [0, 0.0, 0.0, 0.0, 0.0])
    a[0] += 1
    a[4] += l.get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", 0.0) * l.get("gpu__time_duration.sum", 0.0)
    a[1] += l.get("gpu__time_duration.sum", 0.0)
    a[2] += l.get("dram__bytes_read.sum", 0.0)
    a[3] += l.get("dram__bytes_write.sum", 0.0)
tot = sum(v[1] for v in agg.values())
lines = [f"{'kernel':34s} {'n':>3s} {'us':>8s} {'share':>6s} {'rd MB':>8s} {'wr MB':>8s} {'GB/s':>7s} {'tensor%':>8s}"]
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    lines.append(f"{k:34s} {v[0]:3d} {v[1]:8.1f} {v[1]/tot*100:5.1f}% {v[2]/1e6:8.1f} {v[3]/1e6:8.1f} "
                 f"{(v[2]+v[3])/1e3/max(v[1],1e-9):7.0f} {v[4]/max(v[1],1e-9):8.1f}")
lines.append(f"launches/step {len(step)}  total {tot:.1f} us (cold-cache, serialised under ncu: compare shares, not "
             f"absolutes)  DRAM {sum(v[2]+v[3] for v in agg.values())/1e9:.2f} GB/step")
txt = "\n".join(lines)
print(txt)
open(f"profiles/{tag}_ncu_step_traffic.txt", "w").write(txt + "\n")
with open(f"profiles/{tag}_ncu_launches_step.csv", "w") as f:
    f.write("idx,kernel,grid,us,dram_read_bytes,dram_write_bytes,tensor_pipe_active_pct\n")
    for i, l in enumerate(step):
        k = re.sub(r"\(.*", "", l["name"]).split("::")[-1]
        f.write(f"{i},{k},{l.get('grid','')},{l.get('gpu__time_duration.sum',0):.2f},{l.get('dram__bytes_read.sum',0):.0f},"
                f"{l.get('dram__bytes_write.sum',0):.0f},{l.get('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed',0):.1f}\n")
conv = [agg[k] for k in agg if k.startswith("conv_halo") or k.startswith("igemm_kernel")]
n = sum(c[0] for c in conv)
json.dump({"igemm_dram_bytes_per_launch": sum(c[2] + c[3] for c in conv) / n, "launches": n,
           "source": f"profiles/{tag}_ncu_step_traffic.txt (ncu dram__bytes_read.sum + dram__bytes_write.sum, "
                     "conv_halo_rp_kernel + conv_halo_kernel + igemm_kernel launches of one training step)"},
          open("profiles/traffic.json", "w"), indent=1)
