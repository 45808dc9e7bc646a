#!/usr/bin/env python3
"""Turns an .ncu-rep (ncu --set full --import-source on) into the markdown summary kept under profiles/: the metrics the
roofline argument uses, DRAM bytes per unit of work, the SASS opcode mix (warp-level instructions per unit) and the stall
reasons of the hottest instructions.  Runs here (no GPU): `ncu -i` only reads the report.

  python tools/ncu_summary.py gpurun_out/x.ncu-rep --units 1e8 --unit-name particle --title "..." > profiles/x.summary.md"""
import argparse
import collections
import csv
import io
import subprocess

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__shared_mem_per_block_static",
        "launch__shared_mem_per_block_dynamic", "lts__t_sector_hit_rate.pct", "smsp__warps_eligible.avg.per_cycle_active",
        "sm__cycles_elapsed.avg.per_second", "launch__grid_size", "launch__block_size"]
TO_BYTES = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
TO_MS = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}


def ncu(args):
    return subprocess.run(["ncu", *args], capture_output=True, text=True).stdout


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("--units", type=float, nargs="*", default=[], help="units of work per captured launch (particles, pairs ...), in capture order")
    ap.add_argument("--unit-name", default="particle")
    ap.add_argument("--title", default="")
    ap.add_argument("--alg-bytes", type=float, default=0.0, help="algorithmic bytes per unit (for the achieved-bandwidth line)")
    ap.add_argument("--peak-gbs", type=float, default=6556.5)
    ap.add_argument("--top", type=int, default=14)
    args = ap.parse_args()
    raw = list(csv.reader(io.StringIO(ncu(["-i", args.rep, "--page", "raw", "--csv"]))))
    head, units, rows = raw[0], raw[1], raw[2:]
    print(f"# {args.title or args.rep}\n")
    print(f"Source: `{args.rep}` (`ncu --set full --clock-control none --import-source on`); read with `tools/ncu_summary.py`.\n")
    kernels = []
    for r in rows:
        d = dict(zip(head, r))
        kernels.append(d)
    print("| metric | " + " | ".join(k["Kernel Name"].split("(")[0].replace("void ", "").replace("opal::fast::", "") for k in kernels) + " | unit |")
    print("|---|" + "---|" * (len(kernels) + 1))
    for key in KEYS:
        if key not in head:
            continue
        print(f"| {key} | " + " | ".join(k[key] for k in kernels) + f" | {units[head.index(key)]} |")
    print()
    for n, k in enumerate(kernels):
        u = args.units[n] if n < len(args.units) else None
        rd = float(k["dram__bytes_read.sum"]) * TO_BYTES[units[head.index("dram__bytes_read.sum")]]
        wr = float(k["dram__bytes_write.sum"]) * TO_BYTES[units[head.index("dram__bytes_write.sum")]]
        ms = float(k["gpu__time_duration.sum"]) * TO_MS[units[head.index("gpu__time_duration.sum")]]
        name = k["Kernel Name"].split("(")[0].replace("void ", "").replace("opal::fast::", "")
        line = f"`{name}`: {ms:.4f} ms, DRAM read {rd / 1e9:.4f} GB + written {wr / 1e9:.4f} GB"
        if u:
            line += f" = **{(rd + wr) / u:.2f} B/{args.unit_name}** (read {rd / u:.2f} + written {wr / u:.2f}); {float(k['smsp__inst_executed.sum']) / u * 32:.1f} warp instructions issued per 32-{args.unit_name} chunk " \
                    f"(= per lane-iteration)"
            if args.alg_bytes:
                ach = args.alg_bytes * u / (ms * 1e-3) / 1e9
                line += f"; achieved {ach:.0f} GB/s on the {args.alg_bytes:.0f}-B algorithmic basis = {ach / args.peak_gbs:.3f} of the measured copy peak ({args.peak_gbs} GB/s) at this capture's clock"
        print(line + "\n")
    # ---- source page: opcode mix and hottest instructions ----
    src = ncu(["-i", args.rep, "--page", "source", "--csv"])
    blocks = src.split('"Kernel Name",')
    for n, blk in enumerate(blocks[1:]):
        lines = list(csv.reader(io.StringIO(blk)))
        kname = lines[0][0].split("(")[0].replace("void ", "").replace("opal::fast::", "")
        h = lines[1]
        iS, iN, iSamp = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
        ops, total, samples, hot = collections.Counter(), 0.0, 0.0, []
        for r in lines[2:]:
            if len(r) <= iN or not r[iN].replace(".", "").isdigit():
                continue
            cnt = float(r[iN])
            toks = r[iS].split()
            toks = [t for t in toks if not t.startswith("@")]
            op = toks[0].split(".")[0] if toks else "?"
            ops[op] += cnt
            total += cnt
            s = float(r[iSamp]) if r[iSamp].replace(".", "").isdigit() else 0.0
            samples += s
            hot.append((s, r[iS].strip()))
        u = args.units[n] if n < len(args.units) else None
        scale = (32.0 / u) if u else 1.0          # warp instructions per lane-iteration (= per unit when one lane serves one unit)
        print(f"## SASS opcode mix of `{kname}` ({'warp instructions per 32-' + args.unit_name + ' chunk = per lane-iteration' if u else 'warp-level instructions, total'})\n")
        print(f"total {total * scale:.1f}\n")
        print("| opcode | count | share |")
        print("|---|---|---|")
        for op, cnt in ops.most_common(args.top):
            print(f"| {op} | {cnt * scale:.2f} | {100 * cnt / total:.1f} % |")
        print()
        hot.sort(reverse=True)
        print(f"Hottest instructions by warp-stall samples ({samples:.0f} samples):\n")
        print("| samples | share | SASS |")
        print("|---|---|---|")
        for s, text in hot[:8]:
            print(f"| {s:.0f} | {100 * s / max(samples, 1):.1f} % | `{text[:90]}` |")
        print()


if __name__ == "__main__":
    main()
