"""Builds libp2c_b200.so in-tree with nvcc for sm_100a (no JIT cache: the .so travels with the repo)."""
import os
import shutil
import subprocess

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, 'csrc')
INCLUDE = os.path.join(PKG_DIR, '..', 'include')
LIB_PATH = os.path.join(PKG_DIR, 'libp2c_b200.so')

NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
              '--use_fast_math=false', '-Xcompiler', '-fPIC', '-I', INCLUDE, '-I', CSRC]


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith('.cu'))


def _stale():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = sources() + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith('.cuh')]
    deps.append(os.path.join(INCLUDE, 'p2c_b200.h'))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, defines=(), out=None, overrides=None):
    """Compile every .cu under csrc/ into one shared library. Returns the library path.
    `defines` / `out`: experiment builds (tools/variants.py) -- extra -D flags into another .so beside the product's."""
    lib_path = out or LIB_PATH
    if not force and not out and not _stale():
        return LIB_PATH
    nvcc = shutil.which('nvcc') or '/usr/local/cuda/bin/nvcc'
    flags = [f for f in NVCC_FLAGS if f != '--use_fast_math=false']
    if os.environ.get('P2C_BUILD_CLOCKS') == '1':   # debug build: phase clocks of the tile kernel (tools/tile_clocks.py)
        flags.append('-DP2C_CLOCKS_BUILD')
    flags += ['-D' + d for d in defines]
    objs = []
    build_dir = os.path.join(PKG_DIR, 'build' if not out else 'build_' + os.path.basename(out).replace('.', '_'))
    os.makedirs(build_dir, exist_ok=True)
    procs = []
    for src in sources():
        obj = os.path.join(build_dir, os.path.basename(src)[:-3] + '.o')
        objs.append(obj)
        src = (overrides or {}).get(os.path.basename(src), src)   # experiment builds: another version of one file
        cmd = [nvcc] + flags + (['-Xptxas', '-v'] if verbose else []) + ['-c', src, '-o', obj]
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for cmd, p in procs:
        out, _ = p.communicate()
        if verbose and out:
            print(out)
        if p.returncode != 0:
            raise RuntimeError('nvcc failed: %s\n%s' % (' '.join(cmd), out))
    # the SHARED CUDA runtime (the one PyTorch has already loaded; /usr/local/cuda as the fallback search path): the
    # artefact then carries none of the runtime's own entry points
    cmd = [nvcc, '-shared', '--cudart', 'shared', '-o', lib_path + '.tmp'] + objs + [
        '-gencode', 'arch=compute_100a,code=sm_100a', '-Xlinker', '-rpath,/usr/local/cuda/lib64']
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError('link failed: %s\n%s' % (' '.join(cmd), r.stdout))
    os.replace(lib_path + '.tmp', lib_path)
    return lib_path


if __name__ == '__main__':
    print(build(force=True, verbose=True))
