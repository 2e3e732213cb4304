#!/usr/bin/env python
"""One row per kernel from an `ncu --page raw --csv` export: duration, DRAM bytes and rate, L2 and tensor-pipe
utilisation, occupancy limiters, and the fraction of the measured peaks (MEASURED_PEAKS.json; fall-back 6.65 TB/s).
Usage: tools/ncu_summary.py gpurun_out/ncu_step_raw.csv > profiles/r2/ncu_step_summary.txt"""
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    peaks = {"hbm_gbs": 6650.0}
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            peaks.update({k: v for k, v in json.load(open(p)).items() if isinstance(v, (int, float))})
        except Exception:
            pass
    hbm = float(peaks.get("hbm_copy_gbs", peaks.get("hbm_gbs", 6650.0)))

    def f(r, name, scale=1.0):
        i = col.get(name)
        if i is None or r[i] in ("", "n/a"):
            return float("nan")
        v = float(r[i].replace(",", ""))
        u = units[i]
        if name.startswith("dram__bytes") or name.startswith("lts__t_bytes"):
            v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
        if name == "gpu__time_duration.sum":
            v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}.get(u, 1e-3)
        return v * scale

    def short(name):
        name = re.sub(r"^void\s+", "", name)
        name = re.sub(r"ddpb::", "", name)
        m = re.match(r"([A-Za-z0-9_]+)(<[^(]*>)?", name)
        if not m:
            return name[:48]
        targs = m.group(2) or ""
        targs = re.sub(r"__nv_bfloat16", "bf16", targs).replace("(bool)", "").replace("(int)", "")
        return (m.group(1) + targs)[:58]

    print(f"# ncu --set full, one eager training step (batch 256, bf16), clock-control none; HBM peak used: {hbm:.0f} GB/s (measured copy)")
    print(f"{'kernel':58s} {'grid':>6s} {'blk':>4s} {'regs':>4s} {'us':>7s} {'dram rd MB':>10s} {'dram wr MB':>10s} {'GB/s':>7s} {'of HBM':>6s} "
          f"{'L2 %':>5s} {'tensor %':>8s} {'warps %':>7s} {'occ lim (reg/smem/warps/blk)':>28s}")
    tot = 0.0
    for r in data:
        us = f(r, "gpu__time_duration.sum")
        rd, wr = f(r, "dram__bytes_read.sum"), f(r, "dram__bytes_write.sum")
        gbs = (rd + wr) / (us * 1e-6) / 1e9 if us > 0 else float("nan")
        l2 = f(r, "lts__throughput.avg.pct_of_peak_sustained_elapsed")
        tens = f(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active")
        if tens != tens:
            tens = f(r, "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed")
        warps = f(r, "sm__warps_active.avg.pct_of_peak_sustained_active")
        lim = "/".join(str(int(f(r, "launch__occupancy_limit_" + k))) if f(r, "launch__occupancy_limit_" + k) == f(r, "launch__occupancy_limit_" + k) else "-"
                       for k in ("registers", "shared_mem", "warps", "blocks"))
        tot += us
        print(f"{short(r[col['Kernel Name']]):58s} {int(f(r, 'launch__grid_size')):6d} {int(f(r, 'launch__block_size')):4d} "
              f"{int(f(r, 'launch__registers_per_thread')):4d} {us:7.1f} {rd / 1e6:10.2f} {wr / 1e6:10.2f} {gbs:7.0f} {gbs / hbm:6.2f} "
              f"{l2:5.1f} {tens:8.1f} {warps:7.1f} {lim:>28s}")
    print(f"# sum of kernel durations (serialised under ncu, no overlap between streams): {tot:.1f} us")


if __name__ == "__main__":
    main()
