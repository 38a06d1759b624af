#!/usr/bin/env python
"""Condense an .ncu-rep (read with `ncu -i ... --page raw --csv`) into a small JSON for profiles/."""
import csv
import io
import json
import subprocess
import sys

PAT = ('gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct',
       'sm__warps_active.avg.pct', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
       'launch__occupancy_limit', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct',
       'sm__inst_executed_pipe_alu.avg.pct', 'sm__inst_executed_pipe_fma.avg.pct', 'sm__inst_executed_pipe_lsu.avg.pct',
       'sm__inst_executed_pipe_xu.avg.pct', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
       'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts.avg.pct',
       'lts__t_sector_hit_rate.pct', 'sm__cycles_elapsed.avg', 'lts__t_bytes.sum', 'lts__t_sectors.sum', 'sm__throughput.avg.pct',
       'smsp__average_warps_issue_stalled', 'dram__cycles_active', 'launch__shared_mem_per_block', 'smsp__warps_eligible.avg',
       'lts__throughput.avg.pct', 'l1tex__throughput.avg.pct', 'lts__t_sectors_srcunit_tex_op_read.sum')


def main(rep, out):
    raw = subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"], text=True)
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    keys = [k for k in hdr if any(p in k for p in PAT)]
    best = {}
    for r in rows[2:]:
        name = r[idx['Kernel Name']].split('(')[0].replace('nx::<unnamed>::', '')
        d = {'kernel': name}
        for k in keys:
            v = r[idx[k]]
            if 'nan' in v or v == '':
                continue
            d[k] = v + ' ' + units[idx[k]]
        tag = name + '|' + d.get('launch__grid_size', '')
        if tag not in best or len(d) > len(best[tag]):
            best[tag] = d
    json.dump(list(best.values()), open(out, 'w'), indent=1)
    for d in best.values():
        print('==', d['kernel'], d.get('launch__grid_size'))
        for k in ('gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_sector_hit_rate.pct',
                  'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
                  'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
                  'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
                  'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'launch__registers_per_thread',
                  'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
                  'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
                  'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
                  'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio',
                  'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio',
                  'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
                  'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio'):
            if k in d:
                print('   ', k, '=', d[k])


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
