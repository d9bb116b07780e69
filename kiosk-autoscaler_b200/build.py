"""Build lib/libkas.so (sm_100a only) from csrc/*.cu with nvcc.

    python kiosk-autoscaler_b200/build.py [--force] [--verbose]

The .so is git-ignored but travels to the GPU box with the tree.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIBDIR = os.path.join(HERE, 'lib')
OUT = os.path.join(LIBDIR, 'libkas.so')
SOURCES = ['kas_match.cu', 'kas_replay.cu', 'kas_events.cu', 'kas_glob.cu', 'kas_ingest.cu', 'kas_comm.cu', 'kas_rdb.cu']
HEADERS = ['kas_common.cuh', 'kas_table.h', 'kas_decide.cuh', 'kas_replay_scan.cuh', os.path.join('..', '..', 'include', 'kas.h')]
NVCC = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')
FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
         '-Xcompiler', '-fPIC,-fvisibility=hidden,-Wall', '-Xptxas', '-v',
         '--expt-relaxed-constexpr']


def _stale():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not _stale():
        return OUT
    os.makedirs(LIBDIR, exist_ok=True)
    objs = []
    log = []
    for s in SOURCES:
        o = os.path.join(LIBDIR, s.replace('.cu', '.o'))
        cmd = [NVCC] + FLAGS + ['-c', os.path.join(CSRC, s), '-o', o]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log.append(r.stderr)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError('nvcc failed on %s' % s)
        objs.append(o)
    cmd = [NVCC, '-shared', '-gencode', 'arch=compute_100a,code=sm_100a', '-o', OUT] + objs + ['-ldl']
    subprocess.check_call(cmd)
    with open(os.path.join(LIBDIR, 'ptxas.log'), 'w') as f:
        f.write('\n'.join(log))
    if verbose:
        sys.stderr.write('\n'.join(log))
    return OUT


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='--verbose' in sys.argv))
