"""Per-kernel counts of the SASS mnemonics that prove what a kernel is made of (B200_PROFILING.md): tcgen05.mma (UTC*MMA),
TMEM loads / stores (LDTM / STTM), tcgen05.commit (UTCBAR), bulk async copies (UBLKCP), TMA tensor loads (UTMALDG), LDGSTS
(cp.async), warp reductions (REDUX / CREDUX), packed fp32 (FFMA2 / FADD2 / FMUL2), from `cuobjdump -sass` of the in-tree
library.  No GPU needed.   python tools/sass_summary.py > profiles/<tag>_sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, 'deepclr_b200', 'libdeepclr_b200.so')
PAT = [('UTC*MMA', r'\bUTC\w*MMA'), ('LDTM', r'\bLDTM'), ('STTM', r'\bSTTM'), ('UTCBAR', r'\bUTCBAR'), ('UBLKCP', r'\bUBLKCP'),
       ('UTMALDG', r'\bUTMALDG'), ('LDGSTS', r'\bLDGSTS'), ('REDUX', r'\bC?REDUX'), ('F*2', r'\bF(FMA|ADD|MUL)2\b'),
       ('BAR', r'\bBAR\.'), ('SYNCS', r'\bSYNCS')]
out = subprocess.run(['cuobjdump', '-sass', lib], capture_output=True, text=True, check=True).stdout
counts, size, name = collections.OrderedDict(), {}, None
for line in out.split('\n'):
    m = re.match(r'\s*Function : (\S+)', line)
    if m:
        name = subprocess.run(['c++filt', '-p', m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
        counts[name] = collections.Counter()
        size[name] = 0
        continue
    if name and re.match(r'\s*/\*[0-9a-f]{4,}\*/', line):
        size[name] += 1
        for key, pat in PAT:
            if re.search(pat, line):
                counts[name][key] += 1
print(f'# {os.path.relpath(lib, ROOT)}: SASS instruction counts per kernel (cuobjdump -sass, sm_100a)')
print('%-60s %7s ' % ('kernel', 'instrs') + ' '.join('%7s' % k for k, _ in PAT))
for n in sorted(counts, key=lambda k: -size[k]):
    print('%-60s %7d ' % (n[:60], size[n]) + ' '.join('%7d' % counts[n][k] for k, _ in PAT))
