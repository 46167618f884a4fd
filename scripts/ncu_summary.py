"""Summarises an ncu report (gpurun_out/prof_*.ncu-rep) into a small text file for profiles/.

    python scripts/ncu_summary.py gpurun_out/prof_r01a.ncu-rep profiles/ncu_r01a_summary.txt

Per captured launch: duration, DRAM bytes read/written (the `traffic` figure of bench.py's roofline),
DRAM / L2 / SM throughput as % of peak, occupancy, registers, grid/block, and the top warp-stall
reasons; plus, per kernel, the ten SASS instructions with the most stall samples (source page).
Runs here (no GPU): it only reads the report with `ncu -i`.
"""
import csv
import io
import subprocess
import sys

RAW = [
    ('gpu__time_duration.sum', 'duration'),
    ('dram__bytes_read.sum', 'dram_read'),
    ('dram__bytes_write.sum', 'dram_write'),
    ('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram_%'),
    ('lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l2_%'),
    ('sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm_%'),
    ('sm__warps_active.avg.pct_of_peak_sustained_active', 'occupancy_%'),
    ('launch__registers_per_thread', 'regs'),
    ('launch__grid_size', 'grid'),
    ('launch__block_size', 'block'),
    ('launch__waves_per_multiprocessor', 'waves'),
]


def ncu(*args):
    return subprocess.run(['ncu', *args], capture_output=True, text=True).stdout


def main(rep, out):
    rows = list(csv.reader(io.StringIO(ncu('-i', rep, '--page', 'raw', '--csv'))))
    hdr, units, body = rows[0], rows[1], rows[2:]
    lines = [f'# ncu --set full summary of {rep}', '']
    names = []
    traffic = {}

    def to_bytes(v, unit):
        return float(v) * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}.get(unit, 1)

    for r in body:
        name = r[hdr.index('Kernel Name')].split('(')[0].replace('<unnamed>::', '').replace('void ', '')
        names.append(name)
        lines.append(name)
        try:
            ir, iw, it = hdr.index('dram__bytes_read.sum'), hdr.index('dram__bytes_write.sum'), hdr.index('gpu__time_duration.sum')
            t = traffic.setdefault(name, {'launches': 0, 'dram_read_bytes': 0.0, 'dram_write_bytes': 0.0, 'duration_us': 0.0})
            t['launches'] += 1
            t['dram_read_bytes'] += to_bytes(r[ir], units[ir])
            t['dram_write_bytes'] += to_bytes(r[iw], units[iw])
            t['duration_us'] += float(r[it]) * {'ns': 1e-3, 'us': 1, 'ms': 1e3}.get(units[it], 1)
        except ValueError:
            pass
        for key, label in RAW:
            if key in hdr:
                i = hdr.index(key)
                lines.append(f'    {label:12s} {r[i]} {units[i]}')
        stalls = []
        for i, h in enumerate(hdr):
            if 'pcsamp_warps_issue_stalled_' in h and not h.endswith('_not_issued'):
                try:
                    stalls.append((float(r[i]), h.split('stalled_')[1]))
                except ValueError:
                    pass
        tot = sum(s for s, _ in stalls) or 1.0
        lines.append('    stalls       ' + ', '.join(f'{n} {100 * s / tot:.0f}%' for s, n in sorted(stalls, reverse=True)[:6]))
        lines.append('')
    for name in sorted(set(names)):
        short = name.split('<')[0]
        src = list(csv.reader(io.StringIO(ncu('-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + short,
                                              '--launch-count', '1'))))
        hi = [i for i, r in enumerate(src) if r and r[0] == 'Address']
        if not hi:
            continue
        h = src[hi[0]]
        end = [i for i, r in enumerate(src) if i > hi[0] and r and r[0] == 'Kernel Name']
        end = end[0] if end else len(src)
        i_s, i_src, i_n = h.index('# Samples'), h.index('Source'), h.index('Instructions Executed')
        data = [(int(r[i_s]), r[i_src].strip(), r[i_n]) for r in src[hi[0] + 1:end] if len(r) > i_s and r[i_s].isdigit()]
        tot = sum(d[0] for d in data) or 1
        lines.append(f'## {short}: hottest SASS by stall samples (of {tot})')
        for s, text, n in sorted(data, reverse=True)[:10]:
            lines.append(f'    {100 * s / tot:5.1f}%  x{n:>8s}  {text[:90]}')
        lines.append('')
    open(out, 'w').write('\n'.join(lines) + '\n')
    print('wrote', out)
    if len(sys.argv) > 3:  # per-launch averages, read by bench.py for roofline.traffic
        import json
        for t in traffic.values():
            for k in ('dram_read_bytes', 'dram_write_bytes', 'duration_us'):
                t[k] /= t['launches']
        json.dump({'source': rep, 'kernels': traffic}, open(sys.argv[3], 'w'), indent=1)
        print('wrote', sys.argv[3])


if __name__ == '__main__':
    main(sys.argv[1], sys.argv[2])
