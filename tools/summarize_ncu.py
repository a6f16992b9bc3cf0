"""Summarise an .ncu-rep (raw + source pages) into a small markdown file under profiles/.

    python tools/summarize_ncu.py gpurun_out/prof_tc.ncu-rep profiles/tc_local_round.md "title" [notes...]
"""
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__cluster_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "lts__t_bytes.sum", "sm__cycles_elapsed.max", "smsp__inst_executed.sum", "sm__inst_executed_pipe_uniform.sum"]


def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep, dst, title = sys.argv[1], sys.argv[2], sys.argv[3]
    notes = sys.argv[4:]
    raw = page(rep, "raw")
    hdr, units, rows = raw[0], raw[1], raw[2:]
    lines = [f"# {title}", "", f"Source: `{rep}` (ncu --set full --clock-control none --import-source on; one B200, cold caches per replay).", ""]
    for n in notes:
        lines.append(n)
    lines += ["", "| metric | unit | " + " | ".join(f"launch {i}" for i in range(len(rows))) + " |", "|---|---|" + "---|" * len(rows)]
    ki = hdr.index("Kernel Name")
    lines.insert(3, f"Kernel: `{rows[0][ki]}`")
    for k in KEYS:
        for i, h in enumerate(hdr):
            if h == k:
                lines.append(f"| {h} | {units[i]} | " + " | ".join(r[i] for r in rows) + " |")
    src = page(rep, "source")
    h2 = src[1] if len(src) > 1 else []
    if "# Samples" in h2:
        ni, si = h2.index("# Samples"), h2.index("Source")
        data = [(int(r[ni]), r[si]) for r in src[2:] if len(r) > ni and r[ni].strip().isdigit()]
        tot = sum(d[0] for d in data) or 1
        agg = {}
        for n, s in data:
            op = s.split()[0] if not s.startswith("@") else s.split()[1]
            agg[op] = agg.get(op, 0) + n
        lines += ["", f"Warp-stall samples by SASS opcode (total {tot}):", "", "| opcode | samples | share |", "|---|---|---|"]
        for op, n in sorted(agg.items(), key=lambda kv: -kv[1])[:14]:
            lines.append(f"| {op} | {n} | {100.0 * n / tot:.1f}% |")
    open(dst, "w").write("\n".join(lines) + "\n")
    print("wrote", dst)


if __name__ == "__main__":
    main()
