#!/usr/bin/env python
"""Summarise an .ncu-rep: headline metrics per launch and (with the matching .so) warp-instruction
counts per CUDA source line.
Usage: python profiles/ncu_summary.py REPORT.ncu-rep [LIB.so kernel-substring]"""
import collections
import csv
import glob
import io
import os
import re
import subprocess
import sys
import tempfile

KEYS = ['Kernel Name', 'Grid Size', 'Block Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'launch__shared_mem_per_block_dynamic', 'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio']


def raw(rep):
    out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print('-' * 60)
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print('%-80s %-12s %s' % (k, units[i], r[i][:90]))


def per_line(rep, so, kernel_sub, top=45):
    tmp = tempfile.mkdtemp()
    subprocess.run(['cuobjdump', '-xelf', 'all', os.path.abspath(so)], cwd=tmp, capture_output=True)
    seq = None
    for cub in sorted(glob.glob(os.path.join(tmp, '*.cubin'))):
        txt = subprocess.run(['nvdisasm', '-g', '-c', cub], capture_output=True, text=True).stdout.split('\n')
        st = [i for i, l in enumerate(txt) if l.startswith('.text.') and kernel_sub in l]
        if not st:
            continue
        cur, seq = None, []
        for l in txt[st[0] + 1:]:
            if l.startswith('.text.') or l.startswith('.section'):
                break
            m = re.search(r'//## File "([^"]+)", line (\d+)', l)
            if m:
                cur = (m.group(1).split('/')[-1], int(m.group(2)))
                continue
            if re.match(r'\s+/\*[0-9a-f]{4,}\*/', l):
                seq.append(cur)
        break
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--launch-skip', '0', '--launch-count', '1'],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, data = rows[1], rows[2:]
    ie, isamp = hdr.index('Instructions Executed'), hdr.index('# Samples')
    agg, samp = collections.Counter(), collections.Counter()
    for k in range(min(len(seq), len(data))):
        agg[seq[k]] += int(data[k][ie])
        samp[seq[k]] += int(data[k][isamp])
    tot, ts = sum(agg.values()), sum(samp.values())
    print('SASS instructions listed %d / profiled rows %d; warp instructions executed %d; stall samples %d'
          % (len(seq), len(data), tot, ts))
    for key, v in agg.most_common(top):
        print('%-28s %12d  %5.1f%% inst  %5.1f%% samples' % (key, v, 100.0 * v / tot, 100.0 * samp[key] / max(ts, 1)))


if __name__ == '__main__':
    raw(sys.argv[1])
    if len(sys.argv) > 3:
        per_line(sys.argv[1], sys.argv[2], sys.argv[3])
