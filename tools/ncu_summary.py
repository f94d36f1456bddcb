#!/usr/bin/env python
"""Text summary of an .ncu-rep (run here, no GPU needed): key raw metrics per kernel, the SASS
opcode mix and the most-stalled instructions.  Output is what gets committed under profiles/.

    python tools/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x.ncu.txt
"""
import collections
import csv
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_write.sum",
    "lts__t_sectors_srcunit_tex_op_red.sum", "lts__t_sectors_srcunit_tex_op_red_lookup_hit.sum",
    "lts__t_sectors_srcunit_tex_op_red_lookup_miss.sum", "lts__d_atomic_input_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "l1tex__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    # the SM -> L2 request port (what bounds the scatter backward)
    "l1tex__m_l1tex2xbar_req_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "l1tex__m_l1tex2xbar_req_cycles_active.max.pct_of_peak_sustained_elapsed",
    "l1tex__m_l1tex2xbar_write_bytes.sum",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
]


def ncu(rep, page):
    return subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], capture_output=True, text=True).stdout


def raw(rep):
    rows = list(csv.reader(ncu(rep, "raw").splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print("=" * 100)
        print("kernel:", r[hdr.index("Kernel Name")])
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print(f"  {k:82s} {r[i]} {units[i]}")


def source(rep, top=16):
    out = ncu(rep, "source")
    blocks, cur = [], []
    for line in out.splitlines():
        if line.startswith('"Kernel Name"'):
            if cur:
                blocks.append(cur)
            cur = [line]
        elif cur:
            cur.append(line)
    if cur:
        blocks.append(cur)
    seen = set()
    for b in blocks:
        name = next(csv.reader([b[0]]))[1]
        if name in seen:
            continue
        seen.add(name)
        rows = list(csv.reader(b[1:]))
        hdr = rows[0]
        if "# Samples" not in hdr or "Instructions Executed" not in hdr:
            continue  # kernels too short to be sampled
        iS, iE, iN = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
        stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        ops, samp, stalls, data = collections.Counter(), collections.Counter(), collections.Counter(), []
        tot = tots = 0
        for r in rows[1:]:
            if len(r) < len(hdr):
                continue
            try:
                e, s = int(r[iE]), int(r[iN])
            except ValueError:
                continue
            toks = r[iS].split()
            op = toks[1] if toks[0].startswith("@") else toks[0]
            parts = op.split(".")
            op = parts[0] + ("." + ".".join(parts[1:3]) if parts[0] in ("LDG", "STG", "RED", "LDS", "STS") else "")
            ops[op] += e
            samp[op] += s
            tot += e
            tots += s
            for i in stall_cols:
                try:
                    stalls[hdr[i]] += int(r[i])
                except ValueError:
                    pass
            data.append((s, e, r[iS]))
        print("=" * 100)
        print("kernel:", name)
        print(f"  warp instructions executed {tot}, stall samples {tots}")
        print("  opcode mix (share of instructions | share of stall samples):")
        for op, c in ops.most_common(14):
            print(f"    {op:16s} {100 * c / max(tot, 1):5.1f}% | {100 * samp[op] / max(tots, 1):5.1f}%")
        print("  stall reasons:", ", ".join(f"{k[6:]} {100 * v / max(tots, 1):.0f}%" for k, v in stalls.most_common(6)))
        print("  most-sampled instructions (samples, executed, SASS):")
        for s, e, src in sorted(data, reverse=True)[:top]:
            print(f"    {s:7d} {e:10d}  {src}")


def traffic(rep, workload, out_txt):
    """Writes the DRAM bytes per launch of the forward / backward kernels of `rep` into profiles/traffic.json under
    `workload` (what bench.py reports as roofline.traffic): dram__bytes_read.sum + dram__bytes_write.sum, units
    taken from the report."""
    import json
    import os

    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    rows = list(csv.reader(ncu(rep, "raw").splitlines()))
    hdr, units = rows[0], rows[1]
    ir, iw, iname = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("Kernel Name")
    res = {}
    port = ["l1tex__m_l1tex2xbar_req_cycles_active.avg.pct_of_peak_sustained_elapsed",
            "l1tex__m_l1tex2xbar_req_cycles_active.max.pct_of_peak_sustained_elapsed",
            "l1tex__m_l1tex2xbar_write_bytes.sum"]
    for r in rows[2:]:
        name = r[iname]
        key = "fwd" if "fwd" in name else ("bwd" if "bwd" in name else None)
        if key and key not in res:
            res[key] = int(float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]])
            if all(m in hdr for m in port):
                # the SM -> L2 request port of the same launch: busy cycles (avg / busiest SM, %) and bytes through it
                ib = hdr.index(port[2])
                res[key + "_sm_l2_port"] = {
                    "req_cycles_active_avg_pct": float(r[hdr.index(port[0])]),
                    "req_cycles_active_max_pct": float(r[hdr.index(port[1])]),
                    "write_bytes": int(float(r[ib]) * scale[units[ib]])}
    res["source"] = f"{out_txt} (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch; written by tools/ncu_summary.py --traffic)"
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "traffic.json")
    try:
        allt = json.load(open(path))
    except Exception:
        allt = {}
    allt[workload] = res
    json.dump(allt, open(path, "w"), indent=1)
    print(f"# traffic.json[{workload}] = {res}", file=sys.stderr)


if __name__ == "__main__":
    # python tools/ncu_summary.py x.ncu-rep [--traffic WORKLOAD profiles/x.ncu.txt] > profiles/x.ncu.txt
    if "--traffic" in sys.argv:
        i = sys.argv.index("--traffic")
        traffic(sys.argv[1], sys.argv[i + 1], sys.argv[i + 2])
    rep = sys.argv[1]
    print(f"# ncu summary of {rep} (ncu --set full --clock-control none --import-source on; cold-cache, serialised)")
    raw(rep)
    source(rep)
