"""Markdown summary of an ncu capture (no GPU needed): first launch of every kernel of this library, with the metrics
the roofline arithmetic uses.
    python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/ncu_summary.md      # needs ncu on PATH
    python tools/ncu_summary.py gpurun_out/ncu_raw.csv  > profiles/ncu_summary.md      # `ncu -i … --page raw --csv`
"""
import csv, io, subprocess, sys

src = sys.argv[1]
raw = open(src).read() if src.endswith(".csv") else \
    subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
cols = [("gpu__time_duration.sum", "time µs"), ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM thr %"), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM thr %"),
        ("dram__bytes_read.sum", "DRAM read MB"), ("dram__bytes_write.sum", "DRAM write MB"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("launch__shared_mem_per_block_dynamic", "dyn smem KB")]
cols = [(k, n) for k, n in cols if k in idx]
scale = {"gpu__time_duration.sum": {"ms": 1e3, "us": 1.0, "ns": 1e-3, "s": 1e6},
         "dram__bytes_read.sum": {"Gbyte": 1e3, "Mbyte": 1.0, "Kbyte": 1e-3, "byte": 1e-6},
         "dram__bytes_write.sum": {"Gbyte": 1e3, "Mbyte": 1.0, "Kbyte": 1e-3, "byte": 1e-6}}
print(f"# ncu --set full --clock-control none: first launch of each kernel (`{src.split('/')[-1]}`)\n")
print("Cold caches, serialised launches, unlocked clocks: use for shares, pipe activity and traffic, not as benchmark times.\n")
print("| kernel | " + " | ".join(n for _, n in cols) + " |")
print("|---|" + "---|" * len(cols))
seen = set()
for r in rows[2:]:
    full = r[idx["Kernel Name"]]
    if "at::" in full.split("(")[0] or full.startswith("void <unnamed>"):
        continue  # torch helper kernels of the target program (randn, fills, copies)
    name = full.replace("b200::", "").split("(")[0][:80]
    if name in seen:
        continue
    seen.add(name)
    vals = []
    for k, _ in cols:
        v, u = r[idx[k]], units[idx[k]]
        try:
            f = float(v) * scale.get(k, {}).get(u, 1.0)
            v = f"{f:.1f}" if (k in scale or "." in v) else v
        except ValueError:
            pass
        vals.append(v)
    print(f"| `{name}` | " + " | ".join(vals) + " |")
