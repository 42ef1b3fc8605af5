#!/usr/bin/env python
"""Turn an .ncu-rep (ncu --set full) into the small text summary kept under profiles/.
usage: python profiles/summarize.py gpurun_out/x.ncu-rep profiles/x.txt"""
import csv
import io
import subprocess
import sys
from collections import Counter

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "l1tex__m_xbar2l1tex_read_bytes.sum",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]


def main(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, vals = rows[0], rows[1], rows[-1]
    lines = [f"# {rep}", f"kernel: {vals[hdr.index('Kernel Name')]}"]
    got = {}
    for h, u, v in zip(hdr, units, vals):
        if h in KEYS or h in ("l1tex__data_pipe_lsu_wavefronts.sum", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
                             "l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum",
                             "l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_atom.sum",
                             "l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_red.sum"):
            lines.append(f"{h:70s} {v} {u}")
            got[h] = (v, u)
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    h2 = rows[1]
    ix = {h: i for i, h in enumerate(h2)}

    def g(r, k):
        try:
            return float(r[ix[k]])
        except Exception:
            return 0.0
    data = rows[2:]
    tot = sum(g(r, "# Samples") for r in data) or 1.0
    st = Counter()
    for r in data:
        for h in h2:
            if h.startswith("stall_") and "Not Issued" not in h:
                st[h] += g(r, h)
    lines.append("stall reasons (% of warp samples): " + ", ".join(f"{k[6:]} {100 * v / tot:.1f}" for k, v in st.most_common(7)))
    lines.append("hottest SASS lines (% samples, main stall, executions, instruction):")
    for r in sorted(data, key=lambda r: -g(r, "# Samples"))[:12]:
        sts = {h: g(r, h) for h in h2 if h.startswith("stall_") and "Not Issued" not in h}
        lines.append(f"  {100 * g(r, '# Samples') / tot:5.1f}%  {max(sts, key=sts.get)[6:]:18s} {g(r, 'Instructions Executed'):12.0f}  {r[ix['Source']].strip()[:80]}")
    open(out, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines[:24]))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
