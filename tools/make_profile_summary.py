"""Turn one GPU-box pass (gpurun_out/<tag>, written by tools/gpu_round.sh) into the tracked summaries under profiles/.

  python tools/make_profile_summary.py r1c

Writes profiles/<tag>_bench.json (the bench line as printed), <tag>_launches.txt (per-kernel share of the step from the
`ncu --metrics gpu__time_duration.sum` launch lists) and <tag>_<kernel>_ncu.txt (key metrics + stall ranking of the
`ncu --set full` captures, exported here with `ncu -i ... --page raw/source --csv`)."""
import collections
import csv
import io
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
src = os.path.join(ROOT, "gpurun_out", tag)
dst = os.path.join(ROOT, "profiles")

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sectors.sum",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sectors_srcunit_tex_op_read.sum",
        "lts__t_sectors_srcunit_tex_op_atom.sum", "lts__t_sectors_srcunit_tex_op_write.sum",
        "lts__t_sectors_srcunit_ltcfabric.sum", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts.sum", "l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_atom.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct"]


def launches(path, out):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    if not rows:
        return
    h = rows[0]
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        k = r[ki].split("(")[0]
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(v[1] for v in agg.values())
    out.write(f"{os.path.basename(path)}: {len(rows) - 1} launches, total {tot / 1e6:.3f} ms ({rows[1][ui]})\n")
    for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        out.write(f"  {k:44s} n={n:5d}  total={t / 1e3:12.1f} us  share={100 * t / tot:5.1f}%  avg={t / n / 1e3:9.1f} us\n")
    out.write("\n")


def export(rep, page):
    return subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], capture_output=True, text=True).stdout


def full(rep, out):
    raw = list(csv.reader(io.StringIO(export(rep, "raw"))))
    hdr, units, vals = raw[0], raw[1], raw[2]
    out.write(f"{os.path.basename(rep)}: {vals[hdr.index('Kernel Name')]}\n")
    for k in KEYS:
        if k in hdr:
            i = hdr.index(k)
            out.write(f"  {k} = {vals[i]} {units[i]}\n")
    rows = list(csv.reader(io.StringIO(export(rep, "source"))))
    h = rows[1]
    idx = {x: i for i, x in enumerate(h)}
    data = [r for r in rows[2:] if len(r) >= len(h)]
    stalls = [x for x in h if x.startswith("stall_") and "Not Issued" not in x]
    ns = sum(int(r[idx["# Samples"]] or 0) for r in data)
    tot = {x: sum(int(r[idx[x]] or 0) for r in data) for x in stalls}
    out.write("  stall samples: " + ", ".join(f"{x[6:]} {100 * v / max(ns, 1):.1f}%" for x, v in
                                              sorted(tot.items(), key=lambda x: -x[1])[:8]) + "\n")
    out.write("  hottest SASS lines (samples, share, executions, instruction, top stalls):\n")
    for r in sorted(data, key=lambda r: -int(r[idx["# Samples"]] or 0))[:12]:
        st = sorted(((x[6:], int(r[idx[x]] or 0)) for x in stalls), key=lambda x: -x[1])[:2]
        out.write(f"  {int(r[idx['# Samples']]):8d} {100 * int(r[idx['# Samples']]) / max(ns, 1):5.1f}%  x{r[idx['Instructions Executed']]:>10}  "
                  f"{r[idx['Source']][:64]:64s} {st}\n")
    out.write("\n")


with open(os.path.join(dst, f"{tag}_launches.txt"), "w") as out:
    out.write("Per-kernel share of the step, from `ncu --metrics gpu__time_duration.sum --clock-control none` (tools/gpu_round.sh).\n"
              "Times under ncu are serialised and cold-cache: read the shares, not the absolutes.\n\n")
    for f in ("sssp_launches.csv", "tick_launches.csv", "tick_c2_launches.csv", "tick_c3_launches.csv"):
        if os.path.exists(os.path.join(src, f)):
            launches(os.path.join(src, f), out)
for f in sorted(os.listdir(src)):
    if f.endswith(".ncu-rep"):
        with open(os.path.join(dst, f"{tag}_{f[:-8]}_ncu.txt"), "w") as out:
            out.write("`ncu --set full --clock-control none --import-source on`, one launch (tools/gpu_round.sh); exported with\n"
                      "`ncu -i <rep> --page raw --csv` / `--page source --csv` (tools/make_profile_summary.py).\n\n")
            full(os.path.join(src, f), out)
# roofline.traffic of bench.py: DRAM bytes of the captured launches, per tree / per tick
import json
traffic = {}
for f, kind, unit, per in (("sssp_warp_full.ncu-rep", "sssp", "bytes_per_tree", int(os.environ.get("BATCH", 2368))),
                           ("tick_node_full.ncu-rep", "tick", "bytes_per_tick", 1)):
    rep = os.path.join(src, f)
    if not os.path.exists(rep):
        continue
    raw = list(csv.reader(io.StringIO(export(rep, "raw"))))
    hdr, units, vals = raw[0], raw[1], raw[2]

    def val(name):
        i = hdr.index(name)
        x = float(vals[i].replace(",", ""))
        return x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(units[i], 1)
    tot = val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
    traffic[kind] = {unit: tot / per, "source": f"profiles/{tag}_{f[:-8]}_ncu.txt (dram__bytes_read.sum + dram__bytes_write.sum of "
                     f"one launch{' of %d trees' % per if per > 1 else ': k_node_phase of one tick'})"}
if traffic:
    json.dump(traffic, open(os.path.join(dst, "traffic.json"), "w"), indent=1)
if os.path.exists(os.path.join(src, "bench.json")):
    open(os.path.join(dst, f"{tag}_bench.json"), "w").write(open(os.path.join(src, "bench.json")).read())
print("wrote", [f for f in os.listdir(dst) if f.startswith(tag)])
