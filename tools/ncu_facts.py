"""profiles/ncu_facts.json from one `ncu --set full` capture of a front-end step: per pipeline stage the DRAM bytes per
frame (dram__bytes_read.sum + dram__bytes_write.sum of the stage's kernels / frames in the captured batch), the duration
and the busiest execution pipe.  bench.py reports the dominant stage's figure as roofline.traffic.

usage: ncu -i gpurun_out/X.ncu-rep --page raw --csv > /tmp/raw.csv
       python tools/ncu_facts.py /tmp/raw.csv <frames in the captured batch> <source label> > profiles/ncu_facts.json"""
import csv
import json
import sys

STAGE_OF = {"k_border0": "pyramid", "k_resize": "pyramid", "k_resize2": "pyramid", "k_ring": "pyramid", "k_fast_nms": "detect", "k_cell_emit": "detect",
            "k_detect": "detect", "k_quota": "select", "k_cell_retain": "select", "k_level_retain": "select", "k_orient": "orient", "k_orient2": "orient",
            "k_blur2": "blur", "k_blur3": "blur", "k_describe": "describe", "k_describe2": "describe", "k_grid_build": "match", "k_candidates": "match", "k_resolve": "match",
            "k_build_info": "info", "k_select": "gf_select", "k_select_ahead": "gf_select"}


def fnum(s):
    try:
        return float(str(s).replace(",", ""))
    except ValueError:
        return float("nan")


def main(path, frames, label):
    with open(path) as f:
        rows = list(csv.reader(l for l in f if not l.startswith("==")))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    sc = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    tsc = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3}
    pipes = [h for h in hdr if h.startswith("sm__inst_executed_pipe_") and h.endswith(".avg.pct_of_peak_sustained_active")]
    stages = {}
    for r in data:
        name = r[col["Kernel Name"]].split("(")[0].split("::")[-1].split("<")[0].strip()
        if name.startswith("void "):
            name = name[5:]
        st = STAGE_OF.get(name)
        if st is None:
            continue
        rb = fnum(r[col["dram__bytes_read.sum"]]) * sc.get(units[col["dram__bytes_read.sum"]], 1)
        wb = fnum(r[col["dram__bytes_write.sum"]]) * sc.get(units[col["dram__bytes_write.sum"]], 1)
        us = fnum(r[col["gpu__time_duration.sum"]]) * tsc.get(units[col["gpu__time_duration.sum"]], 1)
        d = stages.setdefault(st, {"dram_bytes": 0.0, "us": 0.0, "kernels": {}, "top": (0.0, "", "")})
        d["dram_bytes"] += rb + wb
        d["us"] += us
        d["kernels"][name] = d["kernels"].get(name, 0) + 1
        pp = sorted(((fnum(r[col[p]]), p.split("pipe_")[1].split(".")[0]) for p in pipes if fnum(r[col[p]]) == fnum(r[col[p]])), reverse=True)
        if pp and us > d["top"][0]:
            d["top"] = (us, name, "%s: %s pipe %.0f%% of peak, dram %.1f%%" % (
                name, pp[0][1], pp[0][0], fnum(r[col["gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"]])))
    out = {"source": label, "frames": frames, "stages": {}}
    for st, d in stages.items():
        out["stages"][st] = {"dram_bytes_per_frame": round(d["dram_bytes"] / frames, 1), "us_per_batch_under_ncu": round(d["us"], 1),
                             "kernels": d["kernels"], "ncu": d["top"][2]}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]), sys.argv[3] if len(sys.argv) > 3 else "")
