"""Turn an .ncu-rep (ncu --set full) into the markdown summary kept under profiles/.
usage: python tools/ncu_summary.py report.ncu-rep "title / command" > profiles/x_ncu_summary.md"""
import csv
import io
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.max", "launch__grid_size", "launch__block_size",
    "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__inst_executed.sum",
]
STALL = "smsp__pcsamp_warps_issue_stalled_"

rep, title = sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else ""
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
print("# ncu --set full: %s\n" % title)
for vals in rows[2:]:
    d = dict(zip(hdr, zip(units, vals)))
    print("## %s\n" % d.get("Kernel Name", ("", "?"))[1])
    print("| metric | value | unit |\n|---|---|---|")
    for k in WANT:
        if k in d and d[k][1] not in ("", "n/a"):
            print("| %s | %s | %s |" % (k, d[k][1], d[k][0]))
    stalls = sorted(((float(v[1].replace(",", "")), k[len(STALL):]) for k, v in d.items()
                     if k.startswith(STALL) and "not_issued" not in k and v[1] not in ("", "0")), reverse=True)
    tot = sum(s for s, _ in stalls) or 1.0
    print("\nWarp-state samples (pc sampling): " +
          ", ".join("%s %.0f%%" % (n, 100 * s / tot) for s, n in stalls[:8]) + "\n")
