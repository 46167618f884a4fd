"""Per-kernel SASS instruction counts of libodk.so (cuobjdump -sass): the mnemonics that show which data-movement
path a kernel uses — UBLKCP (1-D bulk TMA, cp.async.bulk), SYNCS (mbarrier), LDGSTS (cp.async), STG.E.128 / LDG.E.128
(vector global access), ATOMS / ATOMG / REDG (atomics), BAR (block barriers) — plus registers per thread from the ELF.
Run here (no GPU needed):   python scripts/sass_table.py > profiles/sass_r02.txt"""
import collections
import pathlib
import re
import subprocess
import sys

LIB = pathlib.Path(__file__).resolve().parents[1] / 'object_detector_b200' / 'libodk.so'
PATTERNS = ['UBLKCP', 'UTMALDG', 'UTMASTG', 'SYNCS', 'LDGSTS', 'STG.E.128', 'LDG.E.128', 'LDS.128', 'ATOMS', 'ATOMG', 'REDG', 'BAR.', 'SHFL', 'VOTE', 'MUFU', 'DADD|DMUL|DFMA', 'UTC', 'HMMA']


def main():
    sass = subprocess.run(['cuobjdump', '-sass', str(LIB)], capture_output=True, text=True, check=True).stdout
    names = subprocess.run(['cu++filt'], input='\n'.join(re.findall(r'Function : (\S+)', sass)), capture_output=True, text=True).stdout.split('\n')
    res = subprocess.run(['cuobjdump', '-res-usage', str(LIB)], capture_output=True, text=True).stdout
    regs = dict(re.findall(r'Function (\S+):\s*\n\s*REG:(\d+)', res))
    blocks = re.split(r'\n\s*Function : ', sass)[1:]
    print(f'# cuobjdump -sass {LIB.name}: instruction counts per kernel (sm_100a cubin)')
    print('# ' + ' | '.join(['kernel', 'regs', 'instr'] + PATTERNS))
    for blk, name in zip(blocks, names):
        mangled = blk.split('\n', 1)[0].strip()
        lines = [ln for ln in blk.split('\n') if re.search(r'/\*[0-9a-f]{4}\*/', ln)]
        ops = collections.Counter()
        for ln in lines:
            m = re.search(r'/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)', ln)
            if m:
                ops[m.group(1)] += 1
        row = []
        for pat in PATTERNS:
            rx = re.compile(pat.replace('.', r'\.') if '|' not in pat else pat)
            row.append(sum(c for op, c in ops.items() if rx.match(op)))
        short = re.sub(r'\((?:int|bool|unsigned int)\)', '', name).replace('<unnamed>::', '').replace('void ', '')
        short = re.sub(r'\(.*', '', short)
        print(' | '.join([short[:60], regs.get(mangled, '?'), str(sum(ops.values()))] + [str(x) for x in row]))


if __name__ == '__main__':
    sys.exit(main())
