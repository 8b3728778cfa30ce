#!/usr/bin/env python
"""Summarise `ncu --set full` reports (no GPU needed): one markdown section per report with the metrics the
profiling recipe asks for -- duration, registers, tensor-pipe activity, DRAM / L2 bytes against the algorithmic
bytes, achieved occupancy and the top warp-stall reasons.

    python tools/ncu_summary.py gpurun_out/ncu_*.ncu-rep > profiles/ncu_kernels.md
"""
import csv
import io
import os
import subprocess
import sys

WANT = [
    ('gpu__time_duration.sum', 'duration'),
    ('launch__grid_size', 'grid'),
    ('launch__block_size', 'block'),
    ('launch__registers_per_thread', 'registers / thread'),
    ('launch__shared_mem_per_block_dynamic', 'dynamic smem / block'),
    ('launch__occupancy_limit_registers', 'occupancy limit (registers), CTAs/SM'),
    ('launch__occupancy_limit_shared_mem', 'occupancy limit (shared memory), CTAs/SM'),
    ('sm__warps_active.avg.pct_of_peak_sustained_active', 'achieved occupancy'),
    ('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'tensor pipe active'),
    ('sm__throughput.avg.pct_of_peak_sustained_elapsed', 'SM throughput'),
    ('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'DRAM throughput'),
    ('dram__bytes_read.sum', 'DRAM bytes read'),
    ('dram__bytes_write.sum', 'DRAM bytes written'),
    ('lts__t_bytes.sum', 'L2 bytes'),
    ('lts__t_sector_hit_rate.pct', 'L2 hit rate'),
    ('l1tex__t_sector_hit_rate.pct', 'L1 hit rate'),
    ('smsp__inst_executed.sum', 'warp instructions'),
    ('smsp__issue_active.avg.pct_of_peak_sustained_active', 'issue slots busy'),
]
STALL_PREFIX = 'smsp__average_warps_issue_stalled_'
STALL_SUFFIX = '_per_issue_active.ratio'


def raw_page(rep):
    out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError(out.stderr[-400:])
    rows = list(csv.reader(io.StringIO(out.stdout)))
    # header row = the one that contains "Kernel Name"; the row after it holds the units
    h = next(i for i, r in enumerate(rows) if 'Kernel Name' in r)
    header, units, data = rows[h], rows[h + 1], rows[h + 2:]
    return header, units, [r for r in data if len(r) == len(header)]


def fmt(v, unit):
    try:
        x = float(v.replace(',', ''))
    except ValueError:
        return v
    if unit in ('byte', 'Kbyte', 'Mbyte', 'Gbyte'):
        x *= {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[unit]
        return '%.2f MB' % (x / 1e6)
    if unit in ('nsecond', 'usecond', 'msecond', 'second'):
        x *= {'nsecond': 1e-3, 'usecond': 1.0, 'msecond': 1e3, 'second': 1e6}[unit]
        return '%.1f us' % x
    if unit == '%':
        return '%.1f %%' % x
    if unit in ('us', 'ns', 'ms', 's', 'KB', 'MB', 'GB', 'B', 'Kbyte/block', 'byte/block', 'Kbyte'):
        return '%.4g %s' % (x, unit)
    return ('%d' % x) if x == int(x) else '%.3g' % x


def short_name(full):
    name = full.replace('(anonymous namespace)::', '').replace('hs::', '').replace('void ', '')
    i = name.find('(')
    return name[:i] if i > 0 else name


def section(rep):
    header, units, data = raw_page(rep)
    col = {name: i for i, name in enumerate(header)}
    groups = {}
    for r in data:                       # launches of the same kernel instance are alike: tabulate the first,
        groups.setdefault(short_name(r[col['Kernel Name']]), []).append(r)       # list the durations of all
    out = []
    tcol = col['gpu__time_duration.sum']
    for name, rows in groups.items():
        r = rows[0]
        out += ['## `%s`' % name[:150], '', 'from `%s`, %d launches captured' % (os.path.basename(rep), len(rows)), '',
                '| metric | value |', '| --- | --- |']
        for key, label in WANT:
            if key in col:
                out.append('| %s (`%s`) | %s |' % (label, key, fmt(r[col[key]], units[col[key]])))
        stalls = []
        for cname, i in col.items():
            if cname.startswith(STALL_PREFIX) and cname.endswith(STALL_SUFFIX) and 'not_issued' not in cname:
                try:
                    stalls.append((float(r[i].replace(',', '')), cname[len(STALL_PREFIX):-len(STALL_SUFFIX)]))
                except ValueError:
                    pass
        stalls.sort(reverse=True)
        if stalls:
            out += ['', 'top warp-stall reasons (warps stalled per issue-active cycle): ' +
                    ', '.join('%s %.2f' % (n, v) for v, n in stalls[:5])]
        if len(rows) > 1:
            out += ['', 'durations of the captured launches: ' + ', '.join(fmt(x[tcol], units[tcol]) for x in rows[:12]) +
                    (' ...' if len(rows) > 12 else '')]
        out.append('')
    return '\n'.join(out)


def main():
    reps = sorted(sys.argv[1:])
    print('# ncu --set full, one capture per kernel family (BERT-large, B=32, S=128, fp16, one eager update)\n')
    print('Captured with `tools/ncu_all.sh` on a B200 (`--clock-control none`, after the same command had exited 0')
    print('without ncu); summarised by `tools/ncu_summary.py`. Durations under ncu are cold-cache, serialised numbers:')
    print('they rank kernels, they are not bench values.\n')
    for rep in reps:
        try:
            print(section(rep))
        except Exception as e:      # noqa: BLE001
            print('## `%s`\n\ncould not be read: %s\n' % (os.path.basename(rep), e))


if __name__ == '__main__':
    main()
