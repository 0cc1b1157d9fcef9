#!/usr/bin/env python
"""Extract the SASS of the communication / optimizer kernels from the built library into profiles/.

    python scripts/dump_sass.py

For every kernel of interest the first matching instantiation is written to profiles/sass_<name>.txt, and one
summary (profiles/sass_comm_mnemonics.txt) lists, per kernel, the instructions that show where its data goes:
multimem (LDGMC / ST*.MC / RED on multicast addresses), 16-byte peer loads / stores (LDG.E.*.128 / STG.E.*.128),
system-scope release / acquire flags."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, 'kungfu_b200', 'lib', 'libkungfu_cuda.so')
OUT = os.path.join(ROOT, 'profiles')
WANT = {
    'ssgd_fused_nvls_bf16': r'k_ssgd_fused_nvlsI13__nv_bfloat16S1_Lb0E',
    'ssgd_fused_bf16_np8': r'k_ssgd_fusedI13__nv_bfloat16S1_Li8ELb0E',
    'sma_fused_nvls_bf16': r'k_sma_fusedI13__nv_bfloat16S1_Lb1ELi1E',
    'sma_fused_p2p_bf16': r'k_sma_fusedI13__nv_bfloat16S1_Lb0ELi0E',
    'pair_avg_bf16': r'k_pair_avgI13__nv_bfloat16S1_Lb1E',
    'pair_pull': r'k_pair_pull',
    'broadcast': r'k_broadcast',
    'allreduce_nvls_bf16': r'k_allreduce_nvlsI13__nv_bfloat16',
    'allreduce_twoshot_bf16_np8': r'k_allreduce_twoshotI13__nv_bfloat16Li0ELi8E',
    'allreduce_oneshot_bf16_np8': r'k_allreduce_oneshotI13__nv_bfloat16Li0ELi8E',
    'graph_allreduce_f32': r'k_graph_allreduceIfLi0E',
    'grad_variance_bf16': r'k_grad_varianceI13__nv_bfloat16',
    'peer_signal': r'k_peer_signal',
    'consensus': r'k_consensus',
}
KEY = re.compile(r'\b(LDGMC|MULTIMEM|REDG?|ST[G]?\.E\.[A-Z0-9.]*MC|STG\.E\.[A-Z0-9.]*128|LDG\.E\.[A-Z0-9.]*128|'
                 r'[A-Z.]*STRONG\.SYS|MEMBAR\.[A-Z.]*SYS|ERRBAR|CCTL\.IVALL|NANOSLEEP)')


def main():
    sass = subprocess.run(['cuobjdump', '-sass', LIB], stdout=subprocess.PIPE, text=True, check=True).stdout
    funcs = {}
    cur = None
    for line in sass.splitlines():
        m = re.match(r'\s*Function : (\S+)', line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
        elif cur is not None:
            funcs[cur].append(line)
    summary = []
    for name, pat in WANT.items():
        hit = next((f for f in funcs if re.search(pat, f)), None)
        if hit is None:
            summary.append('%s: NOT FOUND (%s)' % (name, pat))
            continue
        body = funcs[hit]
        with open(os.path.join(OUT, 'sass_%s.txt' % name), 'w') as f:
            f.write('// %s\n// cuobjdump -sass kungfu_b200/lib/libkungfu_cuda.so (sm_100a)\n' % hit)
            f.write('\n'.join(l for l in body if '/*' in l and not re.match(r'\s*/\* 0x', l)) + '\n')
        counts = {}
        for l in body:
            m = re.search(r'\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)', l)
            if m and KEY.search(m.group(1) + ' '):
                counts[m.group(1)] = counts.get(m.group(1), 0) + 1
        summary.append('%s  (%s)\n    %s' % (name, hit[:90], '  '.join('%s x%d' % kv for kv in sorted(counts.items()))))
    with open(os.path.join(OUT, 'sass_comm_mnemonics.txt'), 'w') as f:
        f.write('\n'.join(summary) + '\n')
    print('\n'.join(summary))


if __name__ == '__main__':
    sys.exit(main())
