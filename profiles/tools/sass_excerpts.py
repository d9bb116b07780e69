"""Excerpts of the SASS of the hot kernels (cuobjdump -sass on the built objects), kept under
profiles/sass/ so that the instruction-level claims of DESIGN.md can be checked from the tracked
tree: per kernel the count of every interesting opcode and the first few lines of each with
their offsets.   python profiles/tools/sass_excerpts.py   (needs only the CUDA toolkit, no GPU)"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
LIB = os.path.join(ROOT, 'kiosk-autoscaler_b200', 'lib')
OUT = os.path.join(ROOT, 'profiles', 'sass')

KERNELS = [
    ('kas_match.o', r'match_count_kernelILi2ELb1ELb1E', 'k1_match_count_2_wide_joint'),
    ('kas_replay.o', r'replay_identity_kernelIiLi4ELi1ELi0ELi2E', 'k2_identity_int32_vec4_k1'),
    ('kas_replay.o', r'replay_packed16_kernelILb1E', 'k2_packed16_k1'),
    ('kas_replay.o', r'replay_sweep_kernelIiLb0ELi1ELi2E', 'k2_sweep_int32_packed'),
    ('kas_replay.o', r'replay_sweep_kernelItLb0ELi1ELi2E', 'k2_sweep_uint16_packed'),
    ('kas_replay.o', r'replay_grouped_sorted_kernelIiLi1ELi0ELb1ELi512E', 'k2_multi_queue_lean_int32_k1_512'),
    ('kas_ingest.o', r'pack_heads_kernel', 'pack_heads'),
]
INTEREST = ['LDG.E', 'UBLKCP', 'SYNCS', 'VIMNMX3', 'VIMNMX', 'VIADDMNMX', 'VIADD.16x2', 'IMAD.HI', 'LDS', 'ATOMS',
            'RED.', 'ATOMG', 'SHFL', 'VOTE', 'REDUX', 'PRMT', 'LOP3', 'BAR.SYNC', 'STL', 'LDL']


def functions(obj):
    out = subprocess.run(['cuobjdump', '-sass', os.path.join(LIB, obj)], capture_output=True, text=True).stdout
    cur, body = None, collections.OrderedDict()
    for line in out.splitlines():
        m = re.match(r'\s*Function : (\S+)', line)
        if m:
            cur = m.group(1)
            body[cur] = []
        elif cur and re.match(r'\s+/\*[0-9a-f]{4,}\*/', line):
            body[cur].append(re.sub(r'/\* 0x[0-9a-f]+ \*/\s*$', '', line).rstrip())
    return body


def main():
    os.makedirs(OUT, exist_ok=True)
    cache = {}
    for obj, pat, name in KERNELS:
        if obj not in cache:
            cache[obj] = functions(obj)
        hits = [f for f in cache[obj] if re.search(pat, f)]
        if not hits:
            print('no function matches', pat)
            continue
        fn = hits[0]
        lines = cache[obj][fn]
        ops = collections.Counter()
        first = collections.defaultdict(list)
        for ln in lines:
            m = re.search(r'\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.x]+)', ln)
            if not m:
                continue
            op = m.group(1)
            for key in INTEREST:
                if op.startswith(key):
                    full = op
                    ops[full] += 1
                    if len(first[full]) < 3:
                        first[full].append(ln.strip())
                    break
        with open(os.path.join(OUT, name + '.txt'), 'w') as f:
            f.write('%s\n%s: %d SASS instructions (sm_100a, cuobjdump -sass %s)\n\n' % (fn, name, len(lines), obj))
            f.write('opcode counts of interest:\n')
            for op, n in sorted(ops.items(), key=lambda kv: (-kv[1], kv[0])):
                f.write('  %-28s %5d\n' % (op, n))
            f.write('\nfirst occurrences (offset, instruction):\n')
            for op in sorted(first):
                for ln in first[op]:
                    f.write('  ' + ln + '\n')
        print(name, len(lines), 'instructions;', ', '.join('%s x%d' % kv for kv in ops.most_common(6)))


if __name__ == '__main__':
    main()
