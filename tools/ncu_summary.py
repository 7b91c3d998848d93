"""Write the handful of ncu metrics DESIGN.md quotes from a .ncu-rep into a text file under profiles/.
usage: python tools/ncu_summary.py <in.ncu-rep> <out.txt> "<one-line description>"
(reads the report with `ncu -i ... --page raw --csv`; one block per captured launch)"""
import csv, subprocess, sys

WANT = [
    "Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.sum", "sm__inst_executed_pipe_fp64.sum", "sm__inst_executed_pipe_alu.sum",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__cluster_size",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
    "sm__cycles_elapsed.avg", "smsp__cycles_active.avg", "sm__cycles_elapsed.avg.per_second",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__m_xbar2l1tex_read_bytes.sum",
]
SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}


def main():
    rep, out, desc = sys.argv[1], sys.argv[2], sys.argv[3]
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    names, units = rows[0], rows[1]
    lines = ["# " + desc]
    for vals in rows[2:]:
        d = {n: (vals[i], units[i]) for i, n in enumerate(names)}
        for w in WANT:
            if w in d and d[w][0] != "":
                lines.append(f"{w:75s} {d[w][0]:>22s} {d[w][1]}")
        try:
            tr = sum(float(d[k][0].replace(",", "")) * SCALE[d[k][1]] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
            lines.append(f"dram traffic bytes per launch {tr}")
        except (KeyError, ValueError):
            pass
        lines.append("")
    open(out, "w").write("\n".join(lines))
    print("\n".join(lines))


if __name__ == "__main__":
    main()
