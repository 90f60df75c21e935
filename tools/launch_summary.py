"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list (profiles/r01_launches_*_summary.csv)."""
import csv, sys, collections, re
src, dst = sys.argv[1], sys.argv[2]
rows = [r for r in csv.reader(l for l in open(src) if not l.startswith("==")) if r]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot, cnt, mx = collections.Counter(), collections.Counter(), collections.Counter()
for r in rows[1:]:
    if len(r) <= vi or r[hdr.index("Metric Name")] != "gpu__time_duration.sum":
        continue
    v = float(r[vi].replace(",", ""))
    v *= {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(r[ui], 1)
    name = re.sub(r"\(.*", "", r[ki]).replace("mused::", "")
    tot[name] += v; cnt[name] += 1; mx[name] = max(mx[name], v)
allns = sum(tot.values())
ref = sum(v for k, v in tot.items() if "k_swfd_rows" in k or "k_stage_rows" in k or "k_gather_stack" in k)
with open(dst, "w") as o:
    o.write("# ncu --metrics gpu__time_duration.sum --clock-control none, the bench command of tools/ncu_capture.sh\n")
    o.write("# (python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e --streams= --no-sharded --no-refsched): 4 windows of the\n")
    o.write("# batched device leg + the same-window cov-error check on the fixture stream (one strict run: k_swfd_rows);\n")
    o.write("# share_batched = share of the batched-schedule kernels only\n")
    o.write("kernel,launches,total_ns,share_all,share_batched,max_ns\n")
    for k, v in tot.most_common():
        isref = "k_swfd_rows" in k or "k_stage_rows" in k or "k_gather_stack" in k
        o.write("%s,%d,%d,%.4f,%s,%d\n" % (k, cnt[k], v, v / allns, "" if isref else "%.4f" % (v / (allns - ref)), mx[k]))
print(open(dst).read())
