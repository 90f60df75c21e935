"""Summarise `ncu --set full` reports into tracked files under profiles/ (run in the build container):
   python tools/ncu_summarise.py gpurun_out/r02_chunk.ncu-rep gpurun_out/r02_chunk_c3.ncu-rep ...
   python tools/ncu_summarise.py --append gpurun_out/r02_strict.ncu-rep      (adds to the committed files)
writes profiles/r02_ncu_full_summary.csv (one row per profiled launch: duration, DRAM bytes, tensor pipe,
issue activity, registers, grid) and profiles/r02_ncu_dominant.json (per kernel name, from the FIRST report:
what bench.py's `roofline.traffic` / `roofline.ncu` are read from) plus the stall table of k_eig_jacobi."""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COLS = {
    "gpu__time_duration.sum": "duration",
    "dram__bytes_read.sum": "dram_read",
    "dram__bytes_write.sum": "dram_write",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pipe_pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
    "smsp__inst_executed.sum": "warp_instructions",
    "launch__registers_per_thread": "registers",
    "launch__grid_size": "grid",
    "launch__cluster_size": "cluster",
    "launch__block_size": "block",
    "lts__t_bytes.sum": "l2_bytes",
}
UNIT = {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def raw(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    return rows[0], rows[1], rows[2:]


def main():
    summary, dominant = [], {}
    args = sys.argv[1:]
    append = "--append" in args          # keep what profiles/ already holds (earlier captures), add these reports
    args = [a for a in args if a != "--append"]
    csv_path = os.path.join(ROOT, "profiles", "r02_ncu_full_summary.csv")
    old_rows = []
    if append:
        with open(csv_path) as f:
            old_rows = [l.rstrip("\n") for l in f if not l.startswith("#")][1:]
        with open(os.path.join(ROOT, "profiles", "r02_ncu_dominant.json")) as f:
            dominant = json.load(f)
    for ri, path in enumerate(args):
        hdr, units, rows = raw(path)
        ix = {h: i for i, h in enumerate(hdr)}
        stall = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
        for r in rows:
            name = re.sub(r"\(.*", "", r[ix["Kernel Name"]]).replace("void ", "").replace("mused::", "").strip()
            rec = {"report": os.path.basename(path), "kernel": name}
            for col, key in COLS.items():
                if col in ix:
                    v = r[ix[col]].replace(",", "")
                    try:
                        rec[key] = float(v) * UNIT.get(units[ix[col]], 1.0)
                    except ValueError:
                        rec[key] = None
            summary.append(rec)
            base = re.sub(r"<.*", "", name)
            if re.match(r"k_assign_tc<\d+, \d+, 1>", name):
                base = "k_assign_tc_screen"
            if (ri == 0 or append) and base not in dominant:
                d = {"dram_bytes_per_launch": (rec.get("dram_read") or 0) + (rec.get("dram_write") or 0),
                     "duration_ms_under_ncu": 1e3 * rec["duration"], "issue_active_pct": rec.get("issue_active_pct"),
                     "tensor_pipe_pct": rec.get("tensor_pipe_pct"), "warp_instructions": rec.get("warp_instructions"),
                     "registers": rec.get("registers"),
                     "source": "profiles/r02_ncu_full_summary.csv (%s, --cache-control none)" % os.path.basename(path)}
                if base in ("k_eig_jacobi", "k_swfd_rows", "k_assign_tc_screen"):
                    st = {h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]:
                          float(r[ix[h]].replace(",", "")) for h in stall}
                    d["stall_cycles_per_issued_instruction"] = {k: round(v, 3) for k, v in sorted(st.items(), key=lambda kv: -kv[1]) if v >= 0.01}
                    d["cycles_per_issued_instruction"] = round(sum(st.values()), 3)
                dominant[base] = d
    keys = ["report", "kernel"] + list(COLS.values())
    with open(os.path.join(ROOT, "profiles", "r02_ncu_full_summary.csv"), "w") as f:
        f.write("# ncu --set full --clock-control none --cache-control none (tools/ncu_capture.sh); seconds / bytes / percent\n")
        f.write(",".join(keys) + "\n")
        for l in old_rows:
            f.write(l + "\n")
        for rec in summary:
            f.write(",".join("" if rec.get(k) is None else (("%.6g" % rec[k]) if isinstance(rec[k], float) else str(rec[k]))
                             for k in keys) + "\n")
    with open(os.path.join(ROOT, "profiles", "r02_ncu_dominant.json"), "w") as f:
        json.dump(dominant, f, indent=1)
    print(open(os.path.join(ROOT, "profiles", "r02_ncu_full_summary.csv")).read())
    print(json.dumps(dominant.get("k_eig_jacobi"), indent=1))


if __name__ == "__main__":
    main()
