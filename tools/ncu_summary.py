#!/usr/bin/env python
"""`ncu -i REPORT --page raw --csv` -> the per-launch metric table kept under profiles/ (one line per metric,
one column per captured launch).  usage: tools/ncu_summary.py report.ncu-rep > profiles/<name>_summary.txt"""
import csv
import io
import subprocess
import sys

WANT = ['Kernel Name', 'gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__data_pipe_tex_wavefronts.avg.pct_of_peak_sustained_elapsed',
        'l1tex__t_output_wavefronts_pipe_tex_mem_texture.sum.pct_of_peak_sustained_elapsed',
        'l1tex__texin_sm2tex_req_cycles_active.avg.pct_of_peak_sustained_elapsed', 'l1tex__t_requests_pipe_tex_mem_texture.sum',
        'l1tex__t_output_wavefronts_pipe_tex_mem_texture.sum', 'l1tex__t_sectors_pipe_tex_mem_texture.sum',
        'l1tex__t_sector_pipe_tex_mem_texture_op_tex_hit_rate.pct', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'sm__cycles_elapsed.avg', 'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_tex_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio']


def fmt(x):
    try:
        return '%.4g' % float(x)
    except ValueError:
        for junk in ('void ', '(SweepArgs)', '(InitArgs)', '(unsigned long long, int, int, RngPlanes)'):
            x = x.replace(junk, '')
        return x


raw = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
for w in WANT:
    if w in hdr:
        i = hdr.index(w)
        print('%s [%s]: ' % (w, units[i]) + ' | '.join(fmt(r[i]) for r in data))
