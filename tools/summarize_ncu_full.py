"""Condense `ncu -i X.ncu-rep --page raw --csv` into one row per captured launch with the metrics the roofline
claims rest on: duration, DRAM bytes (read+write = `traffic`), DRAM throughput %, tensor-pipe activity, TMEM pipe,
registers, grid.  Usage: summarize_ncu_full.py raw.csv > profiles/rNN_ncu_full_summary.csv"""
import csv
import sys

rd = list(csv.reader(open(sys.argv[1])))
hdr, units, rows = rd[0], rd[1], rd[2:]
idx = {h: i for i, h in enumerate(hdr)}
cols = [
    ("Kernel Name", "kernel"), ("launch__grid_size", "grid"), ("launch__registers_per_thread", "regs"),
    ("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram_read"), ("dram__bytes_write.sum", "dram_write"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
    ("lts__t_bytes.sum", "l2_bytes"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
    ("sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "tensor_hmma_active_pct"),
    ("TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
     "tensor_pipe_elapsed_pct"),
    ("sm__inst_executed_pipe_tmem.avg.pct_of_peak_sustained_active", "tmem_pipe_pct"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu_pipe_pct"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma_pipe_pct"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu_pipe_pct"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy_pct"),
]
cols = [(c, n) for c, n in cols if c in idx]
SCALE = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1.0,
         "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9, "second": 1.0}


def num(r, c):
    return float(r[idx[c]].replace(",", "")) * SCALE.get(units[idx[c]], 1.0)


w = csv.writer(sys.stdout)
w.writerow([n + (f"[{units[idx[c]]}]" if units[idx[c]] else "") for c, n in cols] + ["traffic_MB", "dram_GBps"])
for r in rows:
    out = []
    for c, n in cols:
        v = r[idx[c]]
        if n == "kernel":
            v = v.split("(")[0].replace("void ", "")
        out.append(v)
    traffic = num(r, "dram__bytes_read.sum") + num(r, "dram__bytes_write.sum")
    t = num(r, "gpu__time_duration.sum")
    out += [f"{traffic / 1e6:.2f}", f"{traffic / t / 1e9:.1f}"]
    w.writerow(out)

# optional second argument: write {device kernel: {bytes_per_launch, time_ms, note}} for bench.py's `traffic`
# (last captured launch of each kernel = warm caches)
if len(sys.argv) > 2:
    import json

    tr = {}
    for r in rows:
        name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "").replace("npfk::", "").replace("tc::", "")
        traffic = num(r, "dram__bytes_read.sum") + num(r, "dram__bytes_write.sum")
        tr[name] = {"bytes_per_launch": traffic, "time_ms": num(r, "gpu__time_duration.sum") * 1e3,
                    "dram_read": num(r, "dram__bytes_read.sum"), "dram_write": num(r, "dram__bytes_write.sum"),
                    "note": "dram__bytes_read.sum + dram__bytes_write.sum of one launch under ncu --set full "
                            "(tools/ncu_targets.py, bench shapes); writes still dirty in L2 at kernel end are not counted"}
    with open(sys.argv[2], "w") as f:
        json.dump(tr, f, indent=1, sort_keys=True)
