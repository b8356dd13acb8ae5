#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU): key launch metrics, stall reasons, dynamic opcode mix and the hottest
source lines.  usage: python scripts/ncu_summary.py gpurun_out/prof.ncu-rep [out.txt]"""
import collections
import csv
import io
import re
import subprocess
import sys

KEEP = ['gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__warps_active.avg.per_cycle_active', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'sass__inst_executed_local_loads', 'sass__inst_executed_local_stores',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed']


def ncu(rep, *args):
    return subprocess.run(['ncu', '-i', rep] + list(args), capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    out = open(sys.argv[2], 'w') if len(sys.argv) > 2 else sys.stdout
    rows = list(csv.reader(io.StringIO(ncu(rep, '--page', 'raw', '--csv'))))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        out.write("launch %s  %s  grid %s block %s\n" % (r[0], r[hdr.index('Kernel Name')][:70], r[hdr.index('Grid Size')],
                                                         r[hdr.index('Block Size')]))
        for k in KEEP + [h for h in hdr if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('per_issue_active.ratio')]:
            if k in hdr:
                i = hdr.index(k)
                out.write("  %-82s %-14s %s\n" % (k, units[i], r[i]))
    for view, title in (('sass', 'dynamic opcode mix (warp instructions executed)'),):
        rows = list(csv.reader(io.StringIO(ncu(rep, '--page', 'source', '--csv', '--print-source', view))))
        h = rows[1]
        ia, isrc, isamp = h.index("Instructions Executed"), h.index("Source"), h.index("# Samples")
        op, samp, tot = collections.Counter(), collections.Counter(), 0
        for r in rows[2:]:
            if r and r[0] == "Kernel Name":
                break
            if len(r) <= ia or not r[ia].isdigit():
                continue
            m = re.match(r'\s*(@!?U?P\w+\s+)?([A-Z0-9_]+)', r[isrc])
            o = m.group(2) if m else r[isrc][:10]
            n = int(r[ia])
            op[o] += n
            samp[o] += int(r[isamp])
            tot += n
        out.write("\n%s: total %d\n" % (title, tot))
        for o, n in op.most_common(24):
            out.write("  %-12s %12d %6.2f%%   stall samples %d\n" % (o, n, 100.0 * n / tot, samp[o]))
    rows = list(csv.reader(io.StringIO(ncu(rep, '--page', 'source', '--csv', '--print-source', 'cuda'))))
    if len(rows) > 2:
        h = rows[1]
        try:
            ia, isrc, isamp = h.index("Instructions Executed"), h.index("Source"), h.index("# Samples")
            il = h.index("Line") if "Line" in h else 0
            lines = []
            for r in rows[2:]:
                if r and r[0] == "Kernel Name":
                    break
                if len(r) > ia and r[ia].isdigit() and int(r[ia]) > 0:
                    lines.append((int(r[isamp]), int(r[ia]), r[il], r[isrc].strip()[:110]))
            tot_s = sum(x[0] for x in lines) or 1
            out.write("\nhottest source lines by stall samples (samples, %%, warp instrs, line, text)\n")
            for s_, n, ln, txt in sorted(lines, reverse=True)[:40]:
                out.write("  %6d %5.1f%% %11d  %s: %s\n" % (s_, 100.0 * s_ / tot_s, n, ln, txt))
        except ValueError:
            pass


if __name__ == '__main__':
    main()
