"""Writes profiles/r02_sass_evidence.txt: per kernel of libarcb200.so that uses tensor cores, bulk copies or multimem, the counts of
the Blackwell-specific SASS opcodes and the first line of each (see /opt/skills/guides/B200_PROFILING.md, "What proves a
Blackwell-native kernel").  Needs cuobjdump and c++filt; no GPU."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sass = subprocess.run(['cuobjdump', '-sass', os.path.join(ROOT, 'arcnerf_b200', 'libarcb200.so')], capture_output=True, text=True).stdout
pat = re.compile(r'\b(UTCHMMA|UTCBAR|LDTM|STTM|UBLKCP|LDGMC|REDG|UTCATOMSWS|SYNCS|LDGSTS|ST\S*\.MC\S*)')
out = ['SASS evidence for the round-2 kernels: cuobjdump -sass arcnerf_b200/libarcb200.so (sm_100a), opcode counts per kernel and the',
       'first line of each Blackwell-specific opcode.  UTCHMMA = tcgen05.mma (kind::f16), LDTM = tcgen05.ld, UTCBAR = tcgen05.commit,',
       'UTCATOMSWS = tcgen05.alloc/dealloc, UBLKCP = cp.async.bulk (TMA bulk copy), SYNCS = mbarrier ops, LDGSTS = cp.async,',
       'LDGMC = multimem.ld_reduce (NVLS).  Regenerate: python scripts/sass_evidence.py', '']
cur, counts, first = None, None, None


def flush():
    if cur and counts and any(k in counts for k in ('UTCHMMA', 'UBLKCP', 'LDGMC')):
        name = subprocess.run(['c++filt', cur], capture_output=True, text=True).stdout.strip()
        out.append(name[:160])
        out.append('    counts: ' + ', '.join('%s x%d' % kv for kv in sorted(counts.items())))
        out.extend('    ' + line for line in first.values())
        out.append('')


for line in sass.split('\n'):
    m = re.search(r'Function : (\S+)', line)
    if m:
        flush()
        cur, counts, first = m.group(1), collections.Counter(), collections.OrderedDict()
        continue
    if cur:
        m = pat.search(line)
        if m:
            op = m.group(1)
            counts[op] += 1
            if op not in first and op not in ('SYNCS', 'REDG', 'LDGSTS'):
                first[op] = re.sub(r'\s+/\*.*$', '', line.strip())[:150]
flush()
open(os.path.join(ROOT, 'profiles', 'r02_sass_evidence.txt'), 'w').write('\n'.join(out) + '\n')
print(len(out), 'lines')
