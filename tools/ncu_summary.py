#!/usr/bin/env python
"""ncu_summary.py — condense ncu outputs into the small text / JSON files kept under profiles/.
  python tools/ncu_summary.py launches <launches.csv> <out.txt>    per-kernel totals and shares of a launch list
  python tools/ncu_summary.py full <report.ncu-rep> <out.json>    key metrics of the first kernel of a --set full capture
(no GPU needed; ncu must be on PATH for `full`)."""
import collections, csv, io, json, subprocess, sys

def launches(src, dst):
    rows = [r for r in csv.reader(open(src)) if len(r) > 5]
    ix = {h: i for i, h in enumerate(rows[0])}
    agg = collections.OrderedDict()
    seq = []
    for r in rows[1:]:
        try:
            v = float(r[ix["Metric Value"]])
        except ValueError:
            continue
        if r[ix["Metric Unit"]] in ("us", "usecond"): v *= 1e3
        if r[ix["Metric Unit"]] in ("ms", "msecond"): v *= 1e6
        name = r[ix["Kernel Name"]].split("(")[0]
        a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v
        seq.append((name, r[ix["Grid Size"]], v))
    tot = sum(a[1] for a in agg.values())
    with open(dst, "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none -k regex:rk_  (cold-cache, serialised launches: compare shares)\n")
        f.write("# source: %s, %d launches, %.3f ms in total\n" % (src, len(seq), tot / 1e6))
        f.write("%-34s %8s %12s %8s\n" % ("kernel", "launches", "total_ms", "share"))
        for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
            f.write("%-34s %8d %12.3f %7.2f%%\n" % (k[:34], a[0], a[1] / 1e6, 100 * a[1] / tot))
        f.write("\n# last 24 launches in order (kernel, grid, ms)\n")
        for name, grid, v in seq[-24:]:
            f.write("%-34s %-14s %10.4f\n" % (name[:34], grid, v / 1e6))

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__m_xbar2l1tex_read_bytes.sum",
        "l1tex__m_xbar2l1tex_read_bytes_mem_global_op_tma_ld.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "smsp__inst_executed.sum", "sm__cycles_elapsed.max", "sm__inst_executed.sum.per_cycle_elapsed"]

def full(src, dst):
    out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(out)))
    hdr, units, vals = r[0], r[1], r[2]
    d = {"kernel": vals[hdr.index("Kernel Name")], "source": src}
    for h, u, v in zip(hdr, units, vals):
        if h in WANT:
            d[h] = {"value": v, "unit": u}
    src_out = subprocess.run(["ncu", "-i", src, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rr = list(csv.reader(io.StringIO(src_out)))
    h2, rows = rr[1], rr[2:]
    ix = {h: i for i, h in enumerate(h2)}
    stalls = [h for h in h2 if h.startswith("stall_") and "Not Issued" not in h]
    num = lambda x: int(x) if x.isdigit() else 0
    tot = {s: sum(num(row[ix[s]]) for row in rows if len(row) == len(h2)) for s in stalls}
    n = sum(num(row[ix["# Samples"]]) for row in rows if len(row) == len(h2))
    d["warp_stall_samples"] = n
    d["warp_stall_share"] = {k[6:]: round(v / max(n, 1), 4) for k, v in sorted(tot.items(), key=lambda x: -x[1])[:8]}
    sass = " ".join(row[ix["Source"]].split()[0] for row in rows if len(row) == len(h2) and row[ix["Source"]].split())
    d["sass_has"] = {m: (m in sass) for m in ("UTCHMMA", "LDTM", "UTMALDG", "UBLKCP", "SYNCS")}
    json.dump(d, open(dst, "w"), indent=1)

if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
