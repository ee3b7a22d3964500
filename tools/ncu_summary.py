"""Summarise an ncu report (raw page) into a small text table for profiles/.
usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/xxx.txt"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
want = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.avg.per_second", "lts__cycles_elapsed.avg.per_second",
    "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum.per_second", "dram__bytes_write.sum.per_second",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
    "l1tex__m_xbar2l1tex_read_bytes.sum", "l1tex__m_xbar2l1tex_read_bytes.sum.per_second",
    "l1tex__m_l1tex2xbar_write_bytes.sum", "smsp__inst_executed.sum",
]
print(f"# ncu --set full summary of {rep} (per launch; cold-cache, serialised replays: compare shares, not absolutes)")
for r in rows[2:]:
    print(f"\n== {r[idx['Kernel Name']]}  (launch id {r[idx['ID']]})")
    for w in want:
        if w in idx:
            print(f"  {w:95s} {r[idx[w]]:>16s} {units[idx[w]]}")
    if len(sys.argv) > 2:       # extra columns by regular expression, e.g. 'pipe_.*pct|issue_active|stall'
        import re
        for h in hdr:
            if h not in want and re.search(sys.argv[2], h):
                print(f"  {h:95s} {r[idx[h]]:>16s} {units[idx[h]]}")
