"""Markdown summary of an .ncu-rep (run where ncu is installed; no GPU needed):
    python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/ncu_summary.md"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
cols = [("gpu__time_duration.sum", "time"), ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe % (active)"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM thr %"), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM thr %"),
        ("dram__bytes_read.sum", "DRAM read"), ("dram__bytes_write.sum", "DRAM write"),
        ("l1tex__m_xbar2l1tex_read_bytes.sum.per_second", "L2->SM"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("launch__shared_mem_per_block_dynamic", "dyn smem")]
cols = [(k, n) for k, n in cols if k in idx]
print(f"# ncu --set full summary of `{rep}`\n")
print("| kernel | " + " | ".join(n for _, n in cols) + " |")
print("|---|" + "---|" * len(cols))
seen = {}
for r in rows[2:]:
    name = r[idx["Kernel Name"]].replace("b200::", "").split("(CUtensorMap")[0].split("(const")[0][:70]
    seen[name] = seen.get(name, 0) + 1
    if seen[name] > 1:
        continue  # first launch of each kernel variant
    vals = []
    for k, _ in cols:
        v, u = r[idx[k]], units[idx[k]]
        try:
            v = f"{float(v):.1f}" if "." in v else v
        except ValueError:
            pass
        vals.append(f"{v} {u}".strip())
    print(f"| `{name}` | " + " | ".join(vals) + " |")
