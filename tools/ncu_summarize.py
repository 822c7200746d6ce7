"""Turn ncu outputs into the small tracked summaries under profiles/.

  launches:  python tools/ncu_summarize.py launches gpurun_out/launches_r01.csv  > profiles/r01_launches.md
  full:      python tools/ncu_summarize.py full /tmp/raw.csv                     > profiles/r01_ncu_full.md
             (raw.csv = `ncu -i gpurun_out/prof_r01.ncu-rep --page raw --csv`)
  traffic:   python tools/ncu_summarize.py traffic /tmp/raw.csv                  > profiles/r01_traffic.json
"""
import csv, re, sys
from collections import OrderedDict


def short(name):
    name = re.sub(r"^(void )?(<?unnamed>::|\(anonymous namespace\)::|spanon::)*", "", name)
    m = re.match(r"([A-Za-z0-9_]+)(<[^>(]*>)?", name)
    return (m.group(1) + (m.group(2) or "")) if m else name[:40]


def launches(path):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr, rows = rows[0], rows[1:]
    ix = {n: i for i, n in enumerate(hdr)}
    seq = [(short(r[ix["Kernel Name"]]), r[ix["Grid Size"]], r[ix["Block Size"]], float(r[ix["Metric Value"]]) / 1e3) for r in rows]
    # one step = from one fbank400_kernel launch to the next
    starts = [i for i, s in enumerate(seq) if s[0].startswith("fbank400")]
    print("# ncu launch list (gpu__time_duration.sum, --clock-control none): %d launches captured, %d steps" % (len(seq), len(starts)))
    print("\nPer-launch times under ncu are cold-cache and serialised; compare SHARES with bench.py's CUDA-event profile.\n")
    if len(starts) >= 3:
        a, b = starts[-2], starts[-1]   # a steady-state step (the last complete one)
        step = seq[a:b]
        tot = sum(s[3] for s in step)
        print("## launches of one steady-state step (%d launches, %.1f us total)\n" % (len(step), tot))
        print("| # | kernel | grid | block | us | share |\n|---|---|---|---|---|---|")
        for i, s in enumerate(step):
            print("| %d | %s | %s | %s | %.1f | %.1f%% |" % (i, s[0], s[1], s[2], s[3], 100 * s[3] / tot))
        agg = OrderedDict()
        for s in step:
            k = agg.setdefault(s[0], [0, 0.0]); k[0] += 1; k[1] += s[3]
        print("\n## aggregated by kernel\n\n| kernel | launches | us | share |\n|---|---|---|---|")
        for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            print("| %s | %d | %.1f | %.1f%% |" % (k, n, t, 100 * t / tot))


KEYS = [
    ("gpu__time_duration.sum", "time"), ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"),
    ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu%"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
    ("lts__t_sector_hit_rate.pct", "l2hit%"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_conf"),
    ("smsp__inst_executed.sum", "warp_inst"),
]


def full(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = {}
    for key, label in KEYS:
        c = [i for i, n in enumerate(hdr) if n == key or n.endswith("." + key)]
        if c: ix[label] = c[0]
    kn = hdr.index("Kernel Name")
    print("# ncu --set full --clock-control none, one steady-state step\n")
    print("units: " + ", ".join("%s [%s]" % (l, units[i]) for l, i in ix.items()) + "\n")
    labels = list(ix)
    print("| kernel | " + " | ".join(labels) + " |\n|---|" + "---|" * len(labels))
    for r in data:
        print("| %s | " % short(r[kn]) + " | ".join(r[ix[l]] for l in labels) + " |")


def traffic(path):
    """Per-launch DRAM traffic (dram__bytes_read.sum + dram__bytes_write.sum) by kernel -> JSON for bench.py's
    roofline.traffic (profiles/rNN_traffic.json)."""
    import json
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = lambda key: [i for i, n in enumerate(hdr) if n == key or n.endswith("." + key)][0]   # noqa: E731
    kn, rd, wr, tm = hdr.index("Kernel Name"), col("dram__bytes_read.sum"), col("dram__bytes_write.sum"), col("gpu__time_duration.sum")
    scale = {"Mbyte": 1.0, "Gbyte": 1e3, "Kbyte": 1e-3, "byte": 1e-6}
    agg = OrderedDict()
    for r in data:
        k = agg.setdefault(short(r[kn]), [0.0, 0.0, 0])
        k[0] += float(r[rd]) * scale[units[rd]] + float(r[wr]) * scale[units[wr]]
        k[1] += float(r[tm]) * (1e-3 if units[tm] in ("ns", "nsecond") else 1.0)
        k[2] += 1
    out = {"source": "ncu --set full --clock-control none (tools/ncu_round.sh), batch 60 x 10 s",
           "unit": "MB per launch (dram__bytes_read.sum + dram__bytes_write.sum)",
           "kernels": {k: {"traffic_mb": v[0] / v[2], "time_us": v[1] / v[2], "n": v[2]} for k, v in agg.items()}}
    conv = [(k, v) for k, v in agg.items() if k.startswith("tapgemm_tc") or k.startswith("first_layer")]
    n = sum(v[2] for _, v in conv)
    out["conv_stack"] = {"traffic_mb_per_launch": sum(v[0] for _, v in conv) / n, "launches_captured": n,
                         "traffic_mb_per_step": sum(v[0] for _, v in conv)}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    {"launches": launches, "full": full, "traffic": traffic}[sys.argv[1]](sys.argv[2])
