"""Turn `ncu -i <report> --page raw --csv` (one `--set full` capture of bench.py) into the two summaries kept here.

usage: ncu -i gpurun_out/<name>.ncu-rep --page raw --csv > raw.csv; python profiles/make_summaries.py raw.csv r1
The first launch of every kernel name in the capture is used (the capture starts after the warm-up steps).
"""
import csv, json, re, sys

STAGES = {
    "quantile_select": ["qsel_pivot_kernel", "qsel_sample_kernel", "qsel_partition_kernel", "qsel_cluster_kernel"],
    "normalise+tile_gather": ["tile_gather_tma_kernel", "tile_gather_kernel"],
    "blend_stitch+sigmoid+binarize": ["stitch_blend_mask2d_kernel", "stitch_blend_rows_kernel", "stitch_blend_kernel"],
    "ccl": ["ccl_tile_kernel", "ccl_edge_kernel", "ccl_number_kernel", "ccl_resolve_kernel"],
    "map(label_overlap+match_counts)": ["overlap_init_kernel", "overlap_kernel", "match_kernel"],
}
COLS = {
    "us": ("gpu__time_duration.sum", 1e-3),  # ns -> us (unit checked below)
    "rd": ("dram__bytes_read.sum", 1.0),
    "wr": ("dram__bytes_write.sum", 1.0),
    "dram_pct": ("dram__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    "issue_pct": ("smsp__issue_active.avg.pct_of_peak_sustained_active", 1.0),
    "occ_pct": ("sm__warps_active.avg.pct_of_peak_sustained_active", 1.0),
    "regs": ("launch__registers_per_thread", 1.0),
    "inst": ("smsp__inst_executed.sum", 1.0),
    "xu_pct": ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", 1.0),
    "alu_pct": ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", 1.0),
    "fp64_pct": ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", 1.0),
}
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0,
        "nsecond": 1e-3, "msecond": 1e3}


def short(name):
    m = re.search(r"(\w+)(<[^>]*>)?\(", name)
    return (m.group(1) + (m.group(2) or "")) if m else name


def main(raw, tag):
    rows = list(csv.reader(open(raw)))
    head, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(head)}
    seen, out = set(), []
    for r in rows[2:]:
        if len(r) != len(head):
            continue
        name = short(r[idx["Kernel Name"]])
        if name in seen:
            continue
        seen.add(name)
        rec = {"name": name}
        for key, (col, scale) in COLS.items():
            if col not in idx:
                continue
            try:
                v = float(r[idx[col]].replace(",", ""))
            except ValueError:
                continue
            u = units[idx[col]]
            rec[key] = v * (UNIT[u] if key in ("us", "rd", "wr") and u in UNIT else scale if key != "us" else 1.0)
        out.append(rec)
    json.dump(out, open(f"profiles/{tag}_ncu_full_summary.json", "w"), indent=1)
    stages = {}
    for st, kernels in STAGES.items():
        ks = [o for o in out if any(o["name"].startswith(k) for k in kernels)]
        stages[st] = {"dram_bytes_per_launch": sum(o.get("rd", 0) + o.get("wr", 0) for o in ks),
                      "kernels": [o["name"] for o in ks], "ncu_us": round(sum(o.get("us", 0) for o in ks), 1)}
    json.dump({"source": "ncu --set full --clock-control none, one timed step of `python bench.py --steps 3 --warmup 3` "
                         f"(profiles/{tag}_ncu_full_summary.json)", "stages": stages},
              open(f"profiles/{tag}_traffic.json", "w"), indent=1)
    for o in out:
        print({k: (round(v, 1) if isinstance(v, float) else v) for k, v in o.items()})


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "r1")
