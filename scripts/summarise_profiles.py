"""Turn the files gpu_profile.sh brought back (gpurun_out/) into the tracked summaries under profiles/."""
import collections, csv, json, os, shutil, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src, dst = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
for a, b in (("bench_4096.json", f"{tag}_bench_4096.json"), ("bench_65536.json", f"{tag}_bench_65536.json"),
             ("bench_snw_8192.json", f"{tag}_bench_semantic_nerfw_8192.json"), ("bench_n2.json", f"{tag}_bench_4096_n2.json"),
             ("bench_n2_64k.json", f"{tag}_bench_65536_n2.json"), ("launches.csv", f"{tag}_launches.csv")):
    if os.path.exists(os.path.join(src, a)):
        shutil.copy(os.path.join(src, a), os.path.join(dst, b))
# launch list -> last complete step, per-kernel shares
rows = list(csv.reader(open(os.path.join(src, "launches.csv"))))
for i, r in enumerate(rows):
    if "Kernel Name" in r:
        h, start = r, i
        break
ik, iv = h.index("Kernel Name"), h.index("Metric Value")
seq = [(r[ik], float(r[iv].replace(",", ""))) for r in rows[start + 2:] if len(r) > iv]
adam = [i for i, (k, _) in enumerate(seq) if "adam" in k]
b = adam[-1] + 1
a = [i for i in adam if i < adam[-1] - 3][-1] + 1
agg = collections.OrderedDict()
for k, v in seq[a:b]:
    k = k.split("(")[0][:80]
    d = agg.setdefault(k, [0.0, 0])
    d[0] += v / 1000
    d[1] += 1
tot = sum(v for v, _ in agg.values())
summary = {"source": "ncu --metrics gpu__time_duration.sum --clock-control none (eager step, 4096 rays; last complete step of the capture)",
           "launches_in_step": b - a, "total_us": round(tot, 1),
           "kernels": [{"kernel": k, "launches": n, "us": round(v, 1), "share": round(v / tot, 4)}
                       for k, (v, n) in sorted(agg.items(), key=lambda kv: -kv[1][0])]}
json.dump(summary, open(os.path.join(dst, f"{tag}_launch_list_summary.json"), "w"), indent=1)
# full capture -> one row per captured kernel with the counters the design argues from
raw = list(csv.reader(open(os.path.join(src, "top_kernels_raw.csv"))))
h = raw[0]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "smsp__inst_executed.sum", "l1tex__t_requests_pipe_lsu_mem_local_op_ld.sum",
        "l1tex__t_requests_pipe_lsu_mem_local_op_st.sum"]
cols = [h.index(w) for w in want if w in h]
with open(os.path.join(dst, f"{tag}_top_kernels_ncu_full.csv"), "w", newline="") as f:
    w = csv.writer(f)
    w.writerow([h[c] for c in cols])
    w.writerow([raw[1][c] for c in cols])
    for r in raw[2:]:
        if len(r) == len(h):
            w.writerow([r[c] for c in cols])
print(json.dumps(summary["kernels"][:12], indent=0))
