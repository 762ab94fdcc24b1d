"""Turn gpurun_out/ ncu artefacts into the committed summaries under profiles/ (round tag = argv[1])."""
import csv, io, json, os, subprocess, sys
tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
go, pr = os.path.join(root, "gpurun_out"), os.path.join(root, "profiles")
os.makedirs(pr, exist_ok=True)
KEEP = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed"]
SCALE = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}
for f in sorted(os.listdir(go)):
    if f.startswith("prof_") and f.endswith(".ncu-rep"):
        raw = subprocess.run(["ncu", "-i", os.path.join(go, f), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        if len(rows) < 3:
            continue
        hdr, units = rows[0], rows[1]
        idx = [i for i, h in enumerate(hdr) if h in KEEP]
        name = f[5:-8]
        with open(os.path.join(pr, f"{tag}_ncu_full_{name}.txt"), "w") as o:
            o.write(f"# ncu --set full --clock-control none, capture '{name}' (see tools/gpu_profile.sh), launches of bench.py\n")
            for r in rows[2:]:
                for i in idx:
                    o.write(f"{hdr[i]:70s} {r[i][:90]} {units[i]}\n")
                o.write("\n")
        if name.startswith("paged_attn_decode"):
            ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
            vals = [float(r[ir]) * SCALE.get(units[ir], 1) + float(r[iw]) * SCALE.get(units[iw], 1) for r in rows[2:]]
            meta = {}
            try:
                meta = json.loads([l for l in open(os.path.join(go, f"plain_{name}.log") if os.path.exists(os.path.join(go, f"plain_{name}.log")) else os.path.join(go, "plain.log")) if l.startswith("{")][-1])["roofline"]
            except Exception:
                pass
            json.dump({"dram_bytes_per_launch": sum(vals) / len(vals),
                       "algorithmic_bytes_same_launch": meta.get("algorithmic_bytes_per_launch"),
                       "us_per_launch": sum(float(r[hdr.index("gpu__time_duration.sum")]) for r in rows[2:]) / len(rows[2:]) /
                                        (1e3 if units[hdr.index("gpu__time_duration.sum")].startswith("ns") else 1),
                       "context": meta.get("context"), "source": f"profiles/{tag}_ncu_full_{name}.txt"},
                      open(os.path.join(pr, "attn_traffic.json"), "w"))
for f, out in (("launches.csv", f"{tag}_launches_B128.txt"), ("launches_B1.csv", f"{tag}_launches_B1.txt"),
               ("launches_B1_chain.csv", f"{tag}_launches_B1_five_kernel_chain.txt"),
               ("launches_B1_two_launch.csv", f"{tag}_launches_B1_two_launch_layer.txt"),
               ("launches_fp8.csv", f"{tag}_launches_B128_fp8w.txt")):
    p = os.path.join(go, f)
    if os.path.exists(p):
        txt = subprocess.run([sys.executable, os.path.join(root, "tools", "launch_summary.py"), p], capture_output=True, text=True).stdout
        open(os.path.join(pr, out), "w").write("# ncu --metrics gpu__time_duration.sum --clock-control none: kernels of the LAST decode step "
                                               "(cold-cache, serialised: compare shares)\n" + txt)
print(sorted(os.listdir(pr)))
