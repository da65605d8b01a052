"""Summarise an .ncu-rep (one kernel launch) into the text/JSON files committed under profiles/.

    python tools/ncu_summary.py gpurun_out/s1_prof.ncu-rep profiles/r1b_icp2d

writes <out>_metrics.txt (selected raw metrics), <out>_sass_mix.txt (instruction mix and
stall-sample shares by opcode from the source page) and <out>_traffic.json (DRAM bytes per
launch, read by bench.py for roofline.traffic)."""
import collections
import csv
import io
import json
import re
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__cycles_active.avg", "sm__cycles_elapsed.avg", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct"]


def ncu(rep, *args):
    return subprocess.run(["ncu", "-i", rep, *args, "--csv"], capture_output=True, text=True).stdout


def main():
    rep, out = sys.argv[1], sys.argv[2]
    rows = list(csv.reader(io.StringIO(ncu(rep, "--page", "raw"))))
    head, unit = rows[0], rows[1]
    with open(out + "_metrics.txt", "w") as f:
        f.write(f"source: ncu --set full --clock-control none --import-source on; report {rep}\n")
        traffic = None
        for d in rows[2:]:
            f.write(f"kernel: {d[head.index('Kernel Name')]}\n")
            for k in KEYS:
                if k in head:
                    f.write(f"  {k} [{unit[head.index(k)]}]: {d[head.index(k)]}\n")
            scale = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}
            rd = float(d[head.index("dram__bytes_read.sum")]) * scale[unit[head.index("dram__bytes_read.sum")]]
            wr = float(d[head.index("dram__bytes_write.sum")]) * scale[unit[head.index("dram__bytes_write.sum")]]
            traffic = {"kernel": d[head.index("Kernel Name")], "dram_bytes_read": rd, "dram_bytes_write": wr,
                       "dram_bytes_per_launch": rd + wr, "report": rep}
    json.dump(traffic, open(out + "_traffic.json", "w"), indent=1)
    rows = list(csv.reader(io.StringIO(ncu(rep, "--page", "source"))))
    hdr = rows[1]
    i_s, i_i, i_src = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Source")
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    ops, samp, st = collections.Counter(), collections.Counter(), collections.Counter()
    for r in rows[2:]:
        try:
            n, s = int(r[i_i]), int(r[i_s])
        except (ValueError, IndexError):
            continue
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[i_src])
        op = m.group(2).split(".")[0] if m else "?"
        ops[op] += n
        samp[op] += s
        for h in stalls:
            try:
                st[h] += int(r[hdr.index(h)])
            except ValueError:
                pass
    tot, tots = sum(ops.values()), sum(samp.values())
    with open(out + "_sass_mix.txt", "w") as f:
        f.write(f"warp-instructions executed {tot}, stall samples {tots}\n")
        f.write("opcode      inst%  samples%\n")
        for op, n in ops.most_common(24):
            f.write(f"{op:10s} {100 * n / tot:6.2f} {100 * samp[op] / tots:8.2f}\n")
        f.write("stall reasons (share of samples): " +
                ", ".join(f"{k[6:]} {100 * v / tots:.1f}%" for k, v in st.most_common(10)) + "\n")


if __name__ == "__main__":
    main()
