"""Build libnes_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIB = os.path.join(HERE, 'libnes_b200.so')
SOURCES = ['noise.cu', 'master.cu', 'decode.cu', 'cider.cu', 'mnist.cu', 'capi.cu']
HEADERS = ['common.cuh', 'tc_ptx.cuh', 'member_gemm.cuh', 'dense.cuh', 'hybrid.cuh', 'lstm_pair.cuh', os.path.join('..', '..', 'include', 'nes_b200.h')]
FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17',
         '-Xcompiler', '-fPIC', '--expt-relaxed-constexpr', '-Xptxas', '-v']


def _digest():
    h = hashlib.sha256()
    for f in SOURCES + HEADERS:
        with open(os.path.join(CSRC, f), 'rb') as fh:
            h.update(fh.read())
    h.update(' '.join(FLAGS).encode())
    return h.hexdigest()


def build(force=False, verbose=False):
    stamp = LIB + '.stamp'
    dig = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(stamp) and open(stamp).read() == dig:
        return LIB
    nvcc = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')
    objdir = os.path.join(HERE, 'build')
    os.makedirs(objdir, exist_ok=True)
    procs = []
    for s in SOURCES:
        obj = os.path.join(objdir, s.replace('.cu', '.o'))
        cmd = [nvcc] + FLAGS + ['-c', os.path.join(CSRC, s), '-o', obj]
        procs.append((s, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    objs, log = [], []
    for s, obj, p in procs:
        out, _ = p.communicate()
        log.append('==== %s\n%s' % (s, out))
        if p.returncode != 0:
            sys.stderr.write('\n'.join(log))
            raise RuntimeError('nvcc failed on %s' % s)
        objs.append(obj)
    cmd = [nvcc, '-shared', '-o', LIB] + objs + ['-gencode', 'arch=compute_100a,code=sm_100a']
    subprocess.check_call(cmd)
    with open(os.path.join(objdir, 'ptxas.log'), 'w') as fh:
        fh.write('\n'.join(log))
    with open(stamp, 'w') as fh:
        fh.write(dig)
    if verbose:
        print('\n'.join(log))
    return LIB


if __name__ == '__main__':
    build(force='--force' in sys.argv, verbose=True)
    print('built', LIB)
