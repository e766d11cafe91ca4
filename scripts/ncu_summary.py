#!/usr/bin/env python
"""Turn an .ncu-rep (ncu --set full) into the compact text summary committed under profiles/:
per profiled launch the DRAM bytes, duration, throughput, occupancy limits, and per kernel the
warp-stall breakdown + hottest SASS lines.   usage: ncu_summary.py rep.ncu-rep [kernel-regex ...]"""
import csv
import io
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers", "l1tex__t_sector_hit_rate.pct",
        "lts__t_sector_hit_rate.pct", "smsp__inst_executed.sum", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]


def ncu(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    rows = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "raw", "--csv"]))))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    print(f"# ncu summary of {rep}\n")
    names = []
    for r in rows[2:]:
        nm = r[ki].split("(")[0]
        names.append(nm)
        print(f"## launch {r[0]}: {nm}")
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                print(f"  {w:68s} {r[i]} {units[i]}")
        rd, wr = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        t = hdr.index("gpu__time_duration.sum")

        def tob(v, u):
            m = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            return float(v.replace(",", "")) * m.get(u, 1)

        def tos(v, u):
            m = {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9, "second": 1.0}
            return float(v.replace(",", "")) * m.get(u, 1)
        by = tob(r[rd], units[rd]) + tob(r[wr], units[wr])
        print(f"  => dram traffic {by / 1e6:.1f} MB, {by / tos(r[t], units[t]) / 1e9:.0f} GB/s over the launch (cold caches, under ncu)\n")
    for pat in sys.argv[2:] or sorted(set(names)):
        out = ncu(["-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name", "regex:" + pat.split("<")[0].split("::")[-1]])
        rows = list(csv.reader(io.StringIO(out)))
        his = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
        if not his:
            continue
        hi = his[0]
        end = his[1] - 1 if len(his) > 1 else len(rows)
        hdr = rows[hi]
        si, src = hdr.index("# Samples"), hdr.index("Source")
        stalls = [c for c in hdr if c.startswith("stall_") and "Not Issued" not in c]
        tot = {s: 0 for s in stalls}
        data, total = [], 0
        for r in rows[hi + 1:end]:
            if len(r) < len(hdr):
                continue
            try:
                n = int(r[si])
            except ValueError:
                continue
            total += n
            for s in stalls:
                try:
                    tot[s] += int(r[hdr.index(s)])
                except ValueError:
                    pass
            data.append((n, r[src]))
        if not total:
            continue
        print(f"## stall samples, first profiled launch of {pat}: {total}")
        for s, v in sorted(tot.items(), key=lambda x: -x[1])[:6]:
            print(f"  {s:26s} {100 * v / total:5.1f}%")
        print("  hottest SASS (index, share, instruction):")
        for i in sorted(sorted(range(len(data)), key=lambda i: -data[i][0])[:12]):
            print(f"   {i:5d} {100 * data[i][0] / total:5.1f}%  {data[i][1][:100]}")
        print()


if __name__ == "__main__":
    main()
