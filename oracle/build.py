"""Build oracle/libkas_oracle.so from oracle/kas_oracle.c (gcc, OpenMP).

TEST INFRASTRUCTURE.  The reference is pure Python, so there is no C reference
to compile into oracle/_ref/; the C file here is this repository's own
restatement (see its header).  Run: ``python oracle/build.py``.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, 'kas_oracle.c')
OUT = os.path.join(HERE, 'libkas_oracle.so')


def build(force=False):
    if (not force and os.path.exists(OUT)
            and os.path.getmtime(OUT) >= os.path.getmtime(SRC)):
        return OUT
    cmd = ['gcc', '-O3', '-march=x86-64-v2', '-fopenmp', '-shared', '-fPIC',
           '-fvisibility=hidden', '-std=c11', '-Wall', '-Wextra', '-o', OUT, SRC]
    subprocess.check_call(cmd)
    return OUT


if __name__ == '__main__':
    print(build(force='--force' in sys.argv))
