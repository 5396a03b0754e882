#!/usr/bin/env python
"""Turns ncu output into the summaries committed under profiles/ (the recipes behind profiles/r02_v22_*).

  launches  <launches.csv> [out.txt]
      per-kernel totals and shares of an `ncu --metrics gpu__time_duration.sum --clock-control none --csv` launch list
  full      <raw.csv> <source.csv> [<raw.csv> <source.csv> ...] [-o out.txt] [-j out.json]
      key metrics of `ncu -i X.ncu-rep --page raw --csv` plus stall reasons / spill share / hottest SASS lines of
      `ncu -i X.ncu-rep --page source --csv` for one captured launch per pair of files
  ptxas     <nvcc -Xptxas -v log> [out.txt]
      registers, stack, spills, barriers, static shared memory per kernel

GPU side of the recipe (one B200, never a bench value):
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-parity
  ncu --set full --clock-control none --import-source on -k regex:pcg_dual -s 3 -c 1 -f -o gpurun_out/pcg_dual python bench.py --steps 1 --warmup 3 --no-cpu --no-parity
  ncu -i gpurun_out/pcg_dual.ncu-rep --page raw --csv > gpurun_out/pcg_dual_raw.csv        # and --page source --csv
"""
from __future__ import annotations

import collections
import csv
import json
import re
import subprocess
import sys

UNIT_NS = {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9, "nsecond": 1.0, "usecond": 1e3, "msecond": 1e6, "second": 1e9}
RAW_KEYS = [
    "Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes.sum.per_second",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
]
STALLS = ("stall_long_sb", "stall_barrier", "stall_wait", "stall_membar", "stall_short_sb", "stall_not_selected", "stall_selected",
          "stall_math", "stall_branch_resolving", "stall_lg", "stall_mio")


def short_name(kernel: str) -> str:
    return re.sub(r"^void ", "", re.sub(r"\(.*", "", kernel)).replace("feab::", "")


def launches(path: str) -> str:
    lines = [l for l in open(path) if not l.startswith("==")]
    agg: "collections.OrderedDict[str, list]" = collections.OrderedDict()
    total = 0.0
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        t = float(row["Metric Value"].replace(",", "")) * UNIT_NS[row["Metric Unit"]]
        a = agg.setdefault(short_name(row["Kernel Name"]), [0, 0.0])
        a[0] += 1
        a[1] += t
        total += t
    out = [f"# per-kernel totals of the ncu launch list {path}",
           "# --metrics gpu__time_duration.sum --clock-control none; per-launch times are cold-cache and serialised: the SHARE is what counts",
           f"{'kernel':60s} {'launches':>8s} {'total ms':>10s} {'share':>7s}"]
    for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.append(f"{k[:60]:60s} {c:8d} {t / 1e6:10.3f} {100 * t / total:6.1f}%")
    out.append(f"{'total':60s} {sum(c for c, _ in agg.values()):8d} {total / 1e6:10.3f}")
    return "\n".join(out) + "\n"


def raw_page(path: str) -> dict:
    rows = list(csv.reader(open(path)))
    return {h: (v, u) for h, u, v in zip(rows[0], rows[1], rows[2])}


def source_page(path: str) -> list:
    rows = list(csv.reader(open(path)))
    hdr = rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    data = rows[2:]

    def num(r, k):
        return int(r[ix[k]] or 0)

    ti = sum(num(r, "Instructions Executed") for r in data)
    ts = sum(num(r, "# Samples") for r in data)
    loc = [r for r in data if "STL" in r[ix["Source"]] or "LDL" in r[ix["Source"]]]
    li = sum(num(r, "Instructions Executed") for r in loc)
    ls = sum(num(r, "# Samples") for r in loc)
    out = [f"source page: {len(data)} SASS lines, {ti} warp instructions, {ts} stall samples",
           f"local-memory (spill) instructions: {len(loc)} sites, {li} executed = {100 * li / ti:.3f} % of instructions, {100 * ls / ts:.2f} % of samples",
           "top stall sites (samples, share, long-scoreboard, barrier):"]
    for r in sorted(data, key=lambda r: -num(r, "# Samples"))[:12]:
        s = num(r, "# Samples")
        out.append(f"  {r[ix['Source']].strip()[:60]:60s} {s:8d} {100 * s / ts:5.1f}%  long_sb {r[ix['stall_long_sb']]:>7s}  barrier {r[ix['stall_barrier']]:>7s}")
    agg = {k: sum(num(r, k) for r in data) for k in STALLS if k in ix}
    out.append("stall reasons, share of all samples: " +
               ", ".join(f"{k[6:]} {100 * v / ts:.1f}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v / ts > 0.003))
    return out


def full(pairs: list, peak_gbs: float = 6551.7):
    text, js = [], {}
    for raw, src in pairs:
        d = raw_page(raw)
        name = short_name(d["Kernel Name"][0]).split("<")[0]
        text.append(f"== {name}")
        for k in RAW_KEYS:
            if k in d:
                text.append(f"{k:75s} {d[k][0]} {d[k][1]}")
        to_bytes = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
        rd = float(d["dram__bytes_read.sum"][0]) * to_bytes[d["dram__bytes_read.sum"][1]]
        wr = float(d["dram__bytes_write.sum"][0]) * to_bytes[d["dram__bytes_write.sum"][1]]
        ns = float(d["gpu__time_duration.sum"][0]) * UNIT_NS[d["gpu__time_duration.sum"][1]]
        gbs = (rd + wr) / ns
        text.append(f"{'DRAM read + write / duration':75s} {gbs:.0f} GB/s = {gbs / peak_gbs:.3f} of the measured copy peak ({peak_gbs} GB/s)")
        text += source_page(src) + [""]
        js[name] = {"name": d["Kernel Name"][0].strip(), "duration_ms": ns / 1e6, "dram_bytes_read": rd, "dram_bytes_write": wr,
                    "registers": int(d["launch__registers_per_thread"][0]), "grid": int(d["launch__grid_size"][0]),
                    "block": int(d["launch__block_size"][0]), "dynamic_smem_kb": float(d["launch__shared_mem_per_block_dynamic"][0]),
                    "l2_hit_pct": float(d["lts__t_sector_hit_rate.pct"][0]),
                    "dram_pct_of_peak": float(d["gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"][0])}
    return "\n".join(text), js


def ptxas(path: str) -> str:
    txt = open(path).read()
    ents = re.findall(r"Compiling entry function '([^']+)' for 'sm_100a'\nptxas info\s+: Function properties for \S+\n\s+(\d+) bytes stack frame, "
                      r"(\d+) bytes spill stores, (\d+) bytes spill loads\nptxas info\s+: Used (\d+) registers(?:, used (\d+) barriers)?(?:, (\d+) bytes smem)?", txt)
    names = subprocess.run(["c++filt"] + [e[0] for e in ents], capture_output=True, text=True).stdout.strip().split("\n")
    out = ["# nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -Xptxas -v : resources per kernel (static smem only)",
           f"{'kernel':95s} {'regs':>4s} {'stack':>5s} {'spill st/ld':>11s} {'bar':>3s} {'smem':>6s}"]
    spill = 0
    for e, n in sorted(zip(ents, names), key=lambda x: x[1]):
        out.append(f"{short_name(n)[:95]:95s} {e[4]:>4s} {e[1]:>5s} {e[2] + '/' + e[3]:>11s} {e[5] or '0':>3s} {e[6] or '0':>6s}")
        spill += int(e[2]) + int(e[3])
    out.append(f"# {len(ents)} kernels, total spill bytes {spill}")
    return "\n".join(out) + "\n"


def main(argv):
    if len(argv) < 2 or argv[0] not in ("launches", "full", "ptxas"):
        sys.stderr.write(__doc__)
        return 2
    if argv[0] == "launches":
        text = launches(argv[1])
        out = argv[2] if len(argv) > 2 else None
    elif argv[0] == "ptxas":
        text = ptxas(argv[1])
        out = argv[2] if len(argv) > 2 else None
    else:
        args, out, jout = [], None, None
        it = iter(argv[1:])
        for a in it:
            if a == "-o":
                out = next(it)
            elif a == "-j":
                jout = next(it)
            else:
                args.append(a)
        text, js = full(list(zip(args[0::2], args[1::2])))
        if jout:
            json.dump({"kernels": js}, open(jout, "w"), indent=1)
    if out:
        open(out, "w").write(text)
    else:
        sys.stdout.write(text)
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
