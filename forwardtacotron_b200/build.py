"""Builds libftb200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m forwardtacotron_b200.build [--force]
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIB = os.path.join(HERE, 'libftb200.so')
SOURCES = ['ftb_api.cu', 'ftb_gemm.cu', 'ftb_rnn.cu', 'ftb_lr.cu', 'ftb_misc.cu', 'ftb_fft.cu', 'ftb_tail.cu', 'ftb_attn.cu', 'ftb_attn5.cu', 'ftb_gl.cu']
HEADERS = ['ftb_internal.h', 'ftb_ptx.cuh', os.path.join('..', '..', 'include', 'ftb200.h')]
# fast-math is deliberately NOT enabled: the gate math needs accurate expf/tanhf.
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
              '-Xcompiler', '-fPIC'] + os.environ.get('FTB_NVCC_EXTRA', '').split()   # e.g. -DFTB_RNN_PROFILE


def _nvcc():
    for cand in (os.environ.get('NVCC'), '/usr/local/cuda/bin/nvcc', 'nvcc'):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return 'nvcc'


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force=False, verbose=False):
    """Compile every .cu into one shared library next to this file.  Returns the path."""
    if not force and not needs_build():
        return LIB
    objs = []
    procs = []
    os.makedirs(os.path.join(HERE, 'build'), exist_ok=True)
    for s in SOURCES:
        o = os.path.join(HERE, 'build', s.replace('.cu', '.o'))
        objs.append(o)
        cmd = [_nvcc()] + NVCC_FLAGS + (['-Xptxas', '-v'] if verbose else []) + ['-c', os.path.join(CSRC, s), '-o', o]
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for s, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode:
            sys.stderr.write(out)
        if p.returncode:
            failed = True
            sys.stderr.write(f'nvcc failed on {s}\n')
    if failed:
        raise RuntimeError('libftb200.so: compilation failed')
    link = [_nvcc(), '-shared', '-gencode', 'arch=compute_100a,code=sm_100a', '-o', LIB] + objs
    subprocess.check_call(link)
    return LIB


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv))
