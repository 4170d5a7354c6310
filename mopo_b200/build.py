"""Build libmopo_b200.so in-tree with nvcc for sm_100a (no torch headers: the boundary is a plain C ABI).

The library is rebuilt when the SHA-256 of its inputs (every file under csrc/, include/mopo_b200*.h, the flags) differs
from the hash recorded beside it at build time (`<lib>.hash`); the same hash is compiled into the library
(`mopo_build_hash()`), so a loaded binary can be checked against the tree (tests/test_abi.py).
`build(probes=True)` builds libmopo_b200_dbg.so: the same sources plus the developer probes (-DMOPO_DEBUG_PROBES:
umma_probe.cu, microbench.cu, the forward kernel's timeline variant; declared in include/mopo_b200_debug.h)."""
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
INCLUDE = os.path.join(HERE, '..', 'include')
LIB = os.path.join(HERE, 'libmopo_b200.so')
LIB_DBG = os.path.join(HERE, 'libmopo_b200_dbg.so')
SOURCES = ['api.cu', 'ens_fp32.cu', 'ens_tc.cu', 'ens_tc3.cu', 'tail.cu', 'policy.cu', 'policy_tc.cu', 'train.cu', 'sac.cu']
PROBE_SOURCES = ['umma_probe.cu', 'microbench.cu']
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
              '-Xcompiler', '-fPIC', '-Xcompiler', '-fvisibility=hidden']


def _nvcc():
    for c in (os.environ.get('NVCC'), '/usr/local/cuda/bin/nvcc', 'nvcc'):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    raise RuntimeError('nvcc not found')


def _flags(probes):
    return NVCC_FLAGS + (['-DMOPO_DEBUG_PROBES'] if probes else []) + os.environ.get('MOPO_NVCC_EXTRA', '').split()


def tree_hash(probes=False):
    """SHA-256 over the build inputs, in a fixed order."""
    h = hashlib.sha256()
    files = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC))]
    files += [os.path.join(INCLUDE, f) for f in sorted(os.listdir(INCLUDE)) if f.endswith('.h')]
    for f in files:
        h.update(os.path.basename(f).encode() + b'\0')
        with open(f, 'rb') as fh:
            h.update(fh.read())
    h.update(' '.join(_flags(probes)).encode())
    return h.hexdigest()[:32]


def needs_build(probes=False):
    lib = LIB_DBG if probes else LIB
    if not os.path.exists(lib) or not os.path.exists(lib + '.hash'):
        return True
    with open(lib + '.hash') as fh:
        return fh.read().strip() != tree_hash(probes)


def build(force=False, verbose=False, probes=False):
    lib = LIB_DBG if probes else LIB
    if not force and not needs_build(probes):
        return lib
    digest = tree_hash(probes)
    flags = _flags(probes) + ['-DMOPO_BUILD_HASH="%s"' % digest]
    objs = []
    procs = []
    odir = os.path.join(HERE, 'build_dbg' if probes else 'build')
    os.makedirs(odir, exist_ok=True)
    for s in SOURCES + (PROBE_SOURCES if probes else []):
        o = os.path.join(odir, s.replace('.cu', '.o'))
        cmd = [_nvcc()] + flags + (['-Xptxas', '-v'] if verbose else []) + ['-c', os.path.join(CSRC, s), '-o', o]
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(o)
    for s, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode:
            sys.stderr.write(out)
        if p.returncode:
            raise RuntimeError('nvcc failed on %s' % s)
    cmd = [_nvcc(), '-shared', '-Xlinker', '-Bsymbolic', '-o', lib] + objs + ['-lcudart', '-lcuda']   # -Bsymbolic: the probe build can sit beside the product library in one process
    subprocess.check_call(cmd)
    with open(lib + '.hash', 'w') as fh:
        fh.write(digest + '\n')
    return lib


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv, probes='--probes' in sys.argv))
