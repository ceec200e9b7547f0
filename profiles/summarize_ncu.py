"""Text summary of an `ncu --set full` report: headline metrics, stall sampling, tensor / TMA SASS
mnemonics.  Usage: python profiles/summarize_ncu.py report.ncu-rep > profiles/rNN_ncu_<kernel>.txt"""
import csv
import io
import re
import subprocess
import sys
from collections import Counter

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "lts__t_bytes.sum"]
for r in rows[2:]:
    d = dict(zip(hdr, r))
    print("== %s" % d.get("Kernel Name", "?"))
    for k in KEEP:
        if k in d:
            print("   %-75s %s %s" % (k, d[k], units[hdr.index(k)]))
    st = {k.replace("smsp__pcsamp_warps_issue_stalled_", ""): float(v or 0) for k, v in d.items()
          if k.startswith("smsp__pcsamp_warps_issue_stalled_") and "not_issued" not in k}
    tot = sum(st.values()) or 1.0
    print("   warp stall sampling (share of samples):")
    for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:10]:
        print("      %-28s %5.1f %%" % (k, 100 * v / tot))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
ops = Counter()
for line in src.splitlines():
    m = re.search(r'"\s*(?:@!?U?P\d+\s+)?(UTCHMMA|UTCBAR|LDTM|UTCATOMSWS|UBLKCP|LDGSTS|SYNCS|FFMA2|DFMA|DMUL|RED|ATOM)[.A-Z0-9_]*', line)
    if m:
        ops[m.group(1)] += 1
print("   SASS mnemonics of interest (static count): " + ", ".join("%s x%d" % kv for kv in sorted(ops.items())))
