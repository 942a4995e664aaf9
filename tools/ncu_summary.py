"""Turn the ncu artefacts a gpurun call brought back into the small text summaries kept under profiles/.

usage:
  python tools/ncu_summary.py launches gpurun_out/launches.csv            > profiles/rNN_launches.txt
  python tools/ncu_summary.py raw      gpurun_out/prof_render.ncu-rep     > profiles/rNN_render_raw.txt
  python tools/ncu_summary.py source   gpurun_out/prof_render.ncu-rep render_bwd [top]  > profiles/rNN_render_bwd_source.txt
  python tools/ncu_summary.py traffic  gpurun_out/prof_render.ncu-rep gpurun_out/prof_render.cmd   (rewrites profiles/traffic.json)
"""
import collections
import csv
import io
import subprocess
import sys

RAW_METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__grid_size", "launch__block_size",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__maximum_warps_per_active_cycle_pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_lsu.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_red.sum",
    "lts__t_sectors_op_red.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
]


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()     # full demangled name -> [launches, time us, dram bytes]
    for row in csv.DictReader(lines):
        name, metric = row["Kernel Name"], row.get("Metric Name")
        if metric not in ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum"):
            continue
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        a = agg.setdefault(name, [0, 0.0, 0.0])
        if metric == "gpu__time_duration.sum":
            a[0] += 1
            a[1] += v / 1000 if u == "ns" else (v * 1000 if u in ("ms", "msecond") else v)
        else:
            a[2] += v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    # kernels that share a 70-character prefix (the reference's forward and backward renderCUDA<3>) stay separate:
    # the label is the prefix plus the parameter count
    tot = sum(a[1] for a in agg.values())
    print(f"# ncu --metrics gpu__time_duration.sum,dram__bytes_*: {sum(a[0] for a in agg.values())} launches, {tot / 1000:.2f} ms")
    print("# cold-cache, serialised launches: compare SHARES, not absolutes; dram = read + write bytes per launch (mean)")
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        label = f"{k[:64]} [{k.count(',') + 1} args]"
        print(f"{label:76s} n={a[0]:4d} total={a[1]:10.1f}us mean={a[1] / a[0]:8.1f}us share={a[1] / tot * 100:5.1f}% dram={a[2] / a[0] / 1e6:8.2f}MB")


def _ncu(rep, page, extra=()):
    return subprocess.run(["ncu", "-i", rep, "--page", page, "--csv", *extra], stdout=subprocess.PIPE,
                          stderr=subprocess.DEVNULL, text=True).stdout


def raw(rep):
    rows = list(csv.reader(io.StringIO(_ncu(rep, "raw"))))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        print("---- " + r[idx["Kernel Name"]])
        for m in RAW_METRICS:
            if m in idx:
                print(f"  {m:84s} {r[idx[m]]:>18s} {units[idx[m]]}")


def traffic(rep, cmd_file):
    """profiles/traffic.json = dram__bytes_read.sum + dram__bytes_write.sum of the captured launches, per kernel (mean
    over the launches in the report), with the ncu command line that produced the report.  bench.py reads
    `render_bwd` / `render_fwd` from it for roofline.traffic; nothing in it is typed by hand."""
    import json
    import os
    rows = list(csv.reader(io.StringIO(_ncu(rep, "raw"))))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    per = collections.OrderedDict()
    for r in rows[2:]:
        name = r[idx["Kernel Name"]]
        key = next((k for k in ("render_bwd", "render_fwd", "preprocess_bwd", "preprocess_fwd", "tile_sort", "emit") if k in name), name[:40])
        rd = float(r[idx["dram__bytes_read.sum"]].replace(",", "")) * scale.get(units[idx["dram__bytes_read.sum"]], 1.0)
        wr = float(r[idx["dram__bytes_write.sum"]].replace(",", "")) * scale.get(units[idx["dram__bytes_write.sum"]], 1.0)
        per.setdefault(key, []).append((rd, wr, name))
    out = {}
    detail = {}
    for k, v in per.items():
        out[k] = int(sum(a + b for a, b, _ in v) / len(v))
        detail[k] = {"kernel": v[0][2], "launches": len(v), "read_bytes": [int(a) for a, _, _ in v], "write_bytes": [int(b) for _, b, _ in v]}
    out["_detail"] = detail
    out["_command"] = open(cmd_file).read().strip() if cmd_file and os.path.exists(cmd_file) else None
    out["_report"] = os.path.basename(rep)
    out["_source"] = "generated by tools/ncu_summary.py traffic from the ncu --set full report of the command above; bytes per launch"
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with open(os.path.join(root, "profiles", "traffic.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps({k: v for k, v in out.items() if not k.startswith("_")}))


def source(rep, kernel, top=40):
    txt = _ncu(rep, "source", ("--kernel-name", "regex:" + kernel))
    # the report holds one block per captured launch; keep the first
    blocks = txt.split('"Kernel Name"')
    rows = list(csv.reader(io.StringIO('"Kernel Name"' + blocks[1])))
    hdr = rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    body = [r for r in rows[2:] if len(r) == len(hdr)]
    f = lambda r, k: float(r[idx[k]] or 0)
    tot_inst = sum(f(r, "Instructions Executed") for r in body)
    tot_samp = sum(f(r, "# Samples") for r in body)
    print(f"# {rows[0][1]}: {len(body)} SASS instructions, {tot_inst:.0f} warp instructions executed, {tot_samp:.0f} samples")
    ops = collections.Counter()
    for r in body:
        ops[r[idx["Source"]].split()[0 if not r[idx["Source"]].strip().startswith("@") else 1].split(".")[0]] += f(r, "Instructions Executed")
    print("# warp instructions by opcode: " + ", ".join(f"{k} {v / tot_inst * 100:.1f}%" for k, v in ops.most_common(18)))
    print(f"# top {top} SASS instructions by executed count: share, avg active threads, samples share, SASS")
    for r in sorted(body, key=lambda r: -f(r, "Instructions Executed"))[:top]:
        print(f"{f(r, 'Instructions Executed') / tot_inst * 100:5.2f}%  thr={f(r, 'Avg. Predicated-On Threads Executed'):5.1f}  "
              f"samp={f(r, '# Samples') / max(tot_samp, 1) * 100:5.2f}%  {r[idx['Source']].strip()}")


if __name__ == "__main__":
    mode = sys.argv[1]
    if mode == "launches":
        launches(sys.argv[2])
    elif mode == "raw":
        raw(sys.argv[2])
    elif mode == "traffic":
        traffic(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else None)
    else:
        source(sys.argv[2], sys.argv[3], int(sys.argv[4]) if len(sys.argv) > 4 else 40)
