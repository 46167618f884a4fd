"""Program-order histogram of warp-stall samples of one kernel in an ncu report (source page).
    python scripts/ncu_phases.py gpurun_out/prof.ncu-rep nms_image [bucket]"""
import csv, io, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
bucket = int(sys.argv[3]) if len(sys.argv) > 3 else 30
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + kern, '--launch-count', '1'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = [i for i, r in enumerate(rows) if r and r[0] == 'Address'][0]
hdr = rows[hi]
end = [i for i, r in enumerate(rows) if i > hi and r and r[0] == 'Kernel Name']
end = end[0] if end else len(rows)
iS, isrc, ix = hdr.index('# Samples'), hdr.index('Source'), hdr.index('Instructions Executed')
data = [(int(r[iS]), r[isrc].strip(), int(r[ix])) for r in rows[hi + 1:end] if len(r) > iS and r[iS].isdigit()]
tot = sum(d[0] for d in data)
print('samples', tot, 'sass', len(data), 'warp-instr', sum(d[2] for d in data))
for i in range(0, len(data), bucket):
    ch = data[i:i + bucket]
    s = sum(c[0] for c in ch)
    if s * 200 > tot:
        top = max(ch, key=lambda c: c[0])
        print(f'{i:5d} {100 * s / tot:5.1f}%  maxexec {max(c[2] for c in ch):7d}  hot {100 * top[0] / tot:4.1f}% {top[1][:70]}')
