"""Condense ncu outputs into the small, committed files under profiles/.

  python tools/ncu_summary.py raw   <rep.ncu-rep> <out.md> [title]   key metrics of one --set full capture
  python tools/ncu_summary.py list  <launches.csv> <out.csv>         launch list: kernel, grid, block, ns (short names)
  python tools/ncu_summary.py hot   <rep.ncu-rep> <out.txt> [topn]   hottest SASS lines with stall reasons
"""
import csv
import io
import re
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "sm__cycles_elapsed.max",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum",
    "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg",
    "sm__inst_executed_pipe_tensor.sum", "sm__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_fma.sum",
    "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_uniform.sum", "sm__inst_executed.sum",
    "smsp__inst_executed.avg.per_cycle_active", "smsp__issue_active.avg.pct",
    "sm__inst_executed_pipe_xu_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__t_requests_pipe_lsu_mem_local_op_ld.sum",
    "l1tex__t_requests_pipe_lsu_mem_local_op_st.sum", "smsp__cycles_active.avg",
]


def ncu_csv(rep, page):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def raw(rep, dst, title):
    rows = ncu_csv(rep, "raw")
    hdr, units = rows[0], rows[1]
    with open(dst, "w") as f:
        f.write(f"# {title}\n\nSource: `ncu --set full --clock-control none --import-source on` (one launch per kernel row).\n")
        for vals in rows[2:]:
            d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
            name = re.sub(r"\(.*", "", d.get("Kernel Name", ("?", ""))[0])
            f.write(f"\n## {name}\n\n| metric | value | unit |\n|---|---|---|\n")
            for k in KEYS:
                hit = [h for h in hdr if h == k or h.endswith("." + k)]
                for h in hit[:1]:
                    f.write(f"| {k} | {d[h][0]} | {d[h][1]} |\n")


def launches(src, dst):
    lines = [l for l in open(src) if l.startswith('"')]
    rows = list(csv.reader(lines))
    ix = {h: i for i, h in enumerate(rows[0])}
    with open(dst, "w") as f:
        f.write("id,kernel,grid,block,duration_ns\n")
        for r in rows[1:]:
            k = r[ix["Kernel Name"]]
            k = re.sub(r"<.*", "", k).replace("void ", "")[:80]
            f.write(f"{r[ix['ID']]},{k},{r[ix['Grid Size']]!r},{r[ix['Block Size']]!r},{r[ix['Metric Value']]}\n")


def hot(rep, dst, topn):
    rows = ncu_csv(rep, "source")
    i0 = next(i for i, r in enumerate(rows) if "Source" in r and "# Samples" in r)
    hdr, data = rows[i0], rows[i0 + 1:]
    ix = {h: i for i, h in enumerate(hdr)}
    stall = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    num = lambda x: float(x) if re.fullmatch(r"[-0-9.e+]+", x or "") else 0.0  # noqa: E731
    tot = sum(num(r[ix["# Samples"]]) for r in data) or 1.0
    agg = {c: sum(num(r[ix[c]]) for r in data) / tot for c in stall}
    with open(dst, "w") as f:
        f.write(f"total samples {tot:.0f}, {len(data)} SASS instructions\n")
        f.write("stall mix: " + ", ".join(f"{k[6:]} {v:.3f}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v > 0.005) + "\n\n")
        order = sorted(range(len(data)), key=lambda i: -num(data[i][ix["# Samples"]]))[:topn]
        for i in sorted(order):
            r = data[i]
            st = sorted(((num(r[ix[c]]), c[6:]) for c in stall), reverse=True)[:3]
            f.write(f"{i:5d} {num(r[ix['# Samples']]):7.0f} {100 * num(r[ix['# Samples']]) / tot:5.1f}%  {r[ix['Source']][:100]:100s} "
                    f"{[(s, int(v)) for v, s in st if v > 0]}\n")


if __name__ == "__main__":
    mode, src, dst = sys.argv[1:4]
    if mode == "raw":
        raw(src, dst, sys.argv[4] if len(sys.argv) > 4 else src)
    elif mode == "list":
        launches(src, dst)
    else:
        hot(src, dst, int(sys.argv[4]) if len(sys.argv) > 4 else 40)
