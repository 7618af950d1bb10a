"""Turn an .ncu-rep into the small per-kernel summary kept under profiles/ (run here, where ncu reads reports without a GPU):
    python tests/ncu_summary.py gpurun_out/x.ncu-rep profiles/prof_x_summary.csv "# header comment"
Launch lists (the --metrics gpu__time_duration.sum --csv log):
    python tests/ncu_summary.py --launches gpurun_out/launches.csv profiles/launches_x.csv "# header comment" """
import csv
import io
import subprocess
import sys

KEEP = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.max", "sm__cycles_active.avg", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor_op_hmma.sum", "sm__pipe_tensor_op_hmma_cycles_active.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__cluster_size", "launch__occupancy_limit_registers",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct", "l1tex__m_xbar2l1tex_read_bytes.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
]


def report(rep, out, header):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    names, units, launches = rows[0], rows[1], rows[2:]
    col = {n: i for i, n in enumerate(names)}
    with open(out, "w") as f:
        f.write(header.rstrip() + "\n")
        f.write("metric,unit," + ",".join(f"launch{i}" for i in range(len(launches))) + "\n")
        for fixed in ("Kernel Name", "Grid Size", "Block Size"):
            if fixed in col:
                f.write(f"{fixed},," + ",".join(l[col[fixed]].replace(",", " ") for l in launches) + "\n")
        for m in KEEP:
            if m in col:
                f.write(f"{m},{units[col[m]]}," + ",".join(l[col[m]].replace(",", "") for l in launches) + "\n")
    print("wrote", out, len(launches), "launch(es)")


def launches(log, out, header):
    lines = [l for l in open(log) if not l.startswith("==")]
    rows = list(csv.reader(io.StringIO("".join(lines))))
    names = rows[0]
    col = {n: i for i, n in enumerate(names)}
    with open(out, "w") as f:
        f.write(header.rstrip() + "\n")
        f.write("idx,kernel,grid,block,duration_ns\n")
        i = 0
        for r in rows[1:]:
            if len(r) <= col["Metric Value"] or r[col["Metric Name"]] != "gpu__time_duration.sum":
                continue
            k = r[col["Kernel Name"]].split("(")[0].replace("void tcq4::", "").replace("tcq4::", "")
            val = float(r[col["Metric Value"]].replace(",", ""))
            unit = r[col["Metric Unit"]]
            ns = val * {"ns": 1, "us": 1e3, "ms": 1e6, "nsecond": 1, "usecond": 1e3, "msecond": 1e6}.get(unit, 1)
            f.write(f'{i},"{k}","{r[col["Grid Size"]]}","{r[col["Block Size"]]}",{int(round(ns))}\n')
            i += 1
    print("wrote", out, i, "launches")


def traffic(prefill_rep, decode_rep, out):
    """profiles/traffic_r02.json for bench.py: DRAM bytes (read + write) per launch of the dominant kernels, averaged over
    the launches of the capture (one layer's GEMM launches / decode GEMV launches of the bench's own graphs)."""
    import json

    def avg(rep, pat):
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        names, units, ls = rows[0], rows[1], rows[2:]
        col = {n: i for i, n in enumerate(names)}
        tot, per = [], []
        for l in ls:
            if pat not in l[col["Kernel Name"]]:
                continue
            b = 0.0
            for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                v = float(l[col[m]].replace(",", ""))
                b += v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[units[col[m]]]
            tot.append(b)
            per.append((l[col["Kernel Name"]].split("(")[0][-40:], l[col["Grid Size"]], int(b)))
        return (sum(tot) / len(tot) if tot else None), per

    pg, pl = avg(prefill_rep, "gemm")
    dg, dl = avg(decode_rep, "gemv")
    d = {"prefill_gemm_bytes_per_launch": pg, "decode_gemv_bytes_per_launch": dg,
         "prefill_note": f"dram__bytes_read.sum + dram__bytes_write.sum, mean over {len(pl)} GEMM launches (the first two layers of the "
                         f"bench's timed step: fan-out q/k/v, attn_output, fan-out gate/up, ffn_down) of one ncu --set full capture "
                         f"(profiles/prof_prefill_r02_summary.csv); weights come from DRAM once, codes and most of y stay in the 126 MB L2",
         "decode_note": f"same, mean over {len(dl)} decode launches (segmented q/k/v GEMV, attn_output, segmented gate/up, ffn_down; "
                        f"profiles/prof_decode_r02_summary.csv)",
         "prefill_launches": pl, "decode_launches": dl}
    with open(out, "w") as f:
        json.dump(d, f, indent=1)
    print("wrote", out, pg, dg)


if __name__ == "__main__":
    if sys.argv[1] == "--traffic":
        traffic(sys.argv[2], sys.argv[3], sys.argv[4])
    elif sys.argv[1] == "--launches":
        launches(sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else "# launch list")
    else:
        report(sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "# ncu summary")
