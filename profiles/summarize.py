"""Turns an .ncu-rep (brought back in gpurun_out/) into the text summaries committed here.

    python profiles/summarize.py gpurun_out/prof.ncu-rep profiles/r01_k1_v3 [launch_index]
"""
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.sum", "sm__cycles_elapsed.max",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]


def ncu(*args):
    return subprocess.run(["ncu", *args], capture_output=True, text=True).stdout


def main():
    rep, out = sys.argv[1], sys.argv[2]
    rows = list(csv.reader(io.StringIO(ncu("-i", rep, "--page", "raw", "--csv"))))
    hdr, units, body = rows[0], rows[1], rows[2:]
    with open(out + "_raw.txt", "w") as f:
        f.write(f"# ncu --set full --clock-control none, source: {rep}\n")
        kn = hdr.index("Kernel Name")
        for li, r in enumerate(body):
            f.write(f"\n## launch {li}: {r[kn][:100]}\n")
            for k in KEYS + [h for h in hdr if h.startswith("smsp__pcsamp_warps_issue_stalled") and not h.endswith("not_issued")]:
                if k in hdr:
                    i = hdr.index(k)
                    f.write(f"{k:75s} {r[i]:>16s} {units[i]}\n")
    launch = sys.argv[3] if len(sys.argv) > 3 else "0"
    txt = ncu("-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-skip", launch,
              "--launch-count", "1")
    lines, h, fname = [], None, None
    for r in csv.reader(io.StringIO(txt)):
        if len(r) == 2 and r[0] == "File Path":
            fname = r[1].split("/")[-1]
        elif r and r[0] == "Line No":
            h = r
        elif h and len(r) == len(h) and r[0] != "":
            g = lambda n: int(r[h.index(n)]) if r[h.index(n)].isdigit() else 0
            lines.append((fname, int(r[0]), r[1].strip()[:90], g("# Samples"), g("Instructions Executed")))
    ti, ts = sum(x[4] for x in lines) or 1, sum(x[3] for x in lines) or 1
    with open(out + "_source.txt", "w") as f:
        f.write(f"# per source line, launch {launch}: % of warp instructions executed, % of stall samples\n")
        for x in sorted(lines, key=lambda x: -(x[3] + x[4] * ts / ti))[:60]:
            f.write(f"{x[4] / ti * 100:5.1f}% inst {x[3] / ts * 100:5.1f}% smp  {x[0]}:{x[1]}  {x[2]}\n")
    print("wrote", out + "_raw.txt", out + "_source.txt")


if __name__ == "__main__":
    main()
