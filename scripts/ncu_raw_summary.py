"""Key metrics per captured launch from a gzipped `ncu --page raw --csv` export (scripts/gpu_ncu_r02.sh)."""
import csv, gzip, io, sys
path = sys.argv[1]
rows = list(csv.reader(io.TextIOWrapper(gzip.open(path), newline="")))
hdr = rows[0]
KEYS = [
    ("gpu__time_duration.sum", "ns"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
    ("sm__issue_active.avg.pct_of_peak_sustained_elapsed", "issue%"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor%"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
    ("l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smem->tc%"),
    ("dram__bytes_read.sum", "rd"), ("dram__bytes_write.sum", "wr"),
]
STALL = "smsp__average_warps_issue_stalled_"
stall_cols = [h for h in hdr if h.startswith("smsp__average_warp_latency_issue_stalled") or h.startswith(STALL)]
idx = {h: i for i, h in enumerate(hdr)}
for r in rows[2:]:
    name = r[idx["Kernel Name"]].split("(")[0][-60:]
    vals = []
    for k, lab in KEYS:
        if k in idx:
            vals.append(f"{lab}={r[idx[k]]}")
    print(name, " ".join(vals))
    st = []
    for h in stall_cols:
        try:
            v = float(r[idx[h]].replace(",", ""))
        except ValueError:
            continue
        if v > 0:
            st.append((v, h.replace("smsp__average_warp_latency_issue_stalled_", "").replace(STALL, "").replace(".ratio", "").replace("_per_warp_active.pct", "")))
    st.sort(reverse=True)
    if st:
        print("     stalls:", ", ".join(f"{n}={v:.2f}" for v, n in st[:6]))
