"""Turns the ncu outputs of a round into the committed summaries under profiles/ (development aid).
usage: profile_summary.py <tag> <launches.csv> <full.ncu-rep> <kernel key in r02_traffic.json> <rows> <d> <k> [command line]"""
import collections, csv, io, json, subprocess, sys

tag, launches, rep, key, rows_n, d, k = sys.argv[1:8]
cmdline = sys.argv[8] if len(sys.argv) > 8 else "python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu"
rows = list(csv.reader(open(launches)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
hdr = rows[hi]; ix = {h: i for i, h in enumerate(hdr)}
agg = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) < len(hdr) or r[ix["Metric Name"]] != "gpu__time_duration.sum":
        continue
    name = r[ix["Kernel Name"]].split("(")[0]
    v = float(r[ix["Metric Value"]].replace(",", "")); unit = r[ix["Metric Unit"]]
    ms = v / 1e6 if unit in ("ns", "nsecond") else (v / 1e3 if unit in ("us", "usecond") else v)
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += ms
tot = sum(a[1] for a in agg.values())
open(f"profiles/{tag}_launches.csv", "w").write(open(launches).read())
with open(f"profiles/{tag}_launches_summary.md", "w") as f:
    f.write(f"# {tag} - ncu launch list (`ncu --metrics gpu__time_duration.sum --clock-control none {cmdline}`)\n\n")
    f.write(f"Per-launch times are cold-cache and serialised: compare SHARES. The raw list is `{tag}_launches.csv`.\n\n| kernel | launches | total ms | share |\n|---|---|---|---|\n")
    for name, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"| `{name[:150]}` | {n} | {ms:.3f} | {100 * ms / tot:.1f}% |\n")
    lib = [(n, v) for n, v in agg.items() if "b200km" in n] or list(agg.items())  # a -k filtered list is all library
    lt = sum(v[1] for _, v in lib)
    f.write("\nLibrary kernels only (everything `at::` is bench.py's synthetic data generation, outside the timed region):\n\n| kernel | launches | total ms | share of library time |\n|---|---|---|---|\n")
    for name, (n, ms) in sorted(lib, key=lambda kv: -kv[1][1]):
        f.write(f"| `{name}` | {n} | {ms:.3f} | {100 * ms / lt:.2f}% |\n")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(io.StringIO(raw)))
dd = {h: (u, v) for h, u, v in zip(rr[0], rr[1], rr[2])}
want = [l.split(",")[0] for l in open("profiles/r01_kmeans_fused_tc_full.csv").read().split("\n")[1:] if l]
want += ["sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "launch__cluster_size", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
         "l1tex__m_xbar2l1tex_read_bytes.sum", "sm__cycles_elapsed.avg", "sm__icc_request_hit_rate.pct", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
         "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
         "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static"]
kname = key.replace("_kernel", "")
with open(f"profiles/{tag}_{kname}_full.csv", "w") as f:
    f.write("metric,unit,value\n")
    seen = set()
    for m in want:
        if m in dd and m not in seen:
            seen.add(m); f.write(f"{m},{dd[m][0]},{dd[m][1]}\n")
def tobytes(m):
    u, v = dd[m]; return float(v.replace(",", "")) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "Tbyte": 1e12}[u]
import os
t = json.load(open("profiles/r02_traffic.json")) if os.path.exists("profiles/r02_traffic.json") else {}
t[key] = {"rows": int(rows_n), "d": int(d), "k": int(k), "dram_bytes_read": int(tobytes("dram__bytes_read.sum")),
          "dram_bytes_write": int(tobytes("dram__bytes_write.sum"))}
json.dump(t, open("profiles/r02_traffic.json", "w"), indent=1)
print(open(f"profiles/{tag}_{kname}_full.csv").read())
print(open(f"profiles/{tag}_launches_summary.md").read()[-700:])
