"""Build libpifpaf_b200.so (sm_100a only) in-tree with nvcc.

The library is a plain C-ABI shared object (include/pifpaf_b200.h); it does not
link against torch.  ``-fmad=false`` keeps every float expression a sequence of
single IEEE operations, which the bit-exact parity of CifHr / seeds / CafScored
depends on.
"""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIB = os.path.join(HERE, 'libpifpaf_b200.so')
SOURCES = ['api.cu', 'cifhr.cu', 'seeds.cu', 'caf_scored.cu', 'assemble.cu', 'nms.cu', 'shim.cu', 'head.cu', 'inverse.cu', 'det.cu', 'pack.cu', 'hrmem.cu']
HEADERS = ['common.cuh', 'stages.cuh', 'nms.cuh', 'occupancy.cuh', 'caf_prescan.cuh', '../../include/pifpaf_b200.h',
           '../../include/pifpaf_detmath.h']
NVCC_FLAGS = [
    '-gencode', 'arch=compute_100a,code=sm_100a',
    '-O3', '-lineinfo', '-std=c++17', '-fmad=false',
    '-Xcompiler', '-fPIC', '-Xcompiler', '-O2',
    '--expt-relaxed-constexpr',
]


def nvcc_path():
    for cand in (os.environ.get('NVCC'), shutil.which('nvcc'), '/usr/local/cuda/bin/nvcc'):
        if cand and os.path.exists(cand):
            return cand
    raise FileNotFoundError('nvcc not found; libpifpaf_b200.so cannot be built')


def is_stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """Compile every .cu into one shared library. Returns the library path."""
    if not force and not is_stale():
        return LIB
    nvcc = nvcc_path()
    objdir = os.path.join(HERE, 'csrc', '_obj')
    os.makedirs(objdir, exist_ok=True)
    procs = []
    for src in SOURCES:
        obj = os.path.join(objdir, src.replace('.cu', '.o'))
        cmd = [nvcc] + NVCC_FLAGS + os.environ.get('PP_EXTRA_NVCC', '').split() + \
            (['-Xptxas', '-v'] if verbose else []) + \
            ['-c', os.path.join(CSRC, src), '-o', obj]
        procs.append((src, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE,
                                                 stderr=subprocess.STDOUT, text=True)))
    objs = []
    for src, obj, proc in procs:
        out, _ = proc.communicate()
        if proc.returncode != 0:
            raise RuntimeError('nvcc failed on %s:\n%s' % (src, out))
        if verbose and out:
            print(out)
        objs.append(obj)
    tmp = LIB + '.tmp'
    subprocess.run([nvcc, '-shared', '-o', tmp] + objs + ['-gencode', 'arch=compute_100a,code=sm_100a',
                                                          '-lcudart_static', '-lrt', '-lpthread', '-ldl'],
                   check=True)
    os.replace(tmp, LIB)
    return LIB


if __name__ == '__main__':
    import sys
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv))
