"""Build `libtransformer_engine.so` (the C-ABI library) in-tree with nvcc for sm_100a.

Usage: python -m transformerengine_b200.build_ext [--force]
Objects go to transformerengine_b200/_build/, the library next to this file so
that it travels to the GPU box with the repo snapshot.
"""
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
BUILD = os.path.join(HERE, "_build")
LIB = os.path.join(HERE, "libtransformer_engine.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-std=c++17", "-O3", "-lineinfo", "--expt-relaxed-constexpr",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-Xcompiler", "-fPIC,-Wall,-Wno-unknown-pragmas",
    "-Xptxas", "-v",
]
EXPORT_MAP = os.path.join(CSRC, "exports.map")


def _sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith((".cu", ".cpp")))


def _stamp(path):
    h = hashlib.sha256()
    for f in sorted(os.listdir(CSRC)) + [os.path.join("..", "..", "include", "transformerengine_b200.h")]:
        p = os.path.join(CSRC, f)
        if os.path.isfile(p) and (f.endswith((".h", ".cuh")) or os.path.abspath(p) == os.path.abspath(path)):
            h.update(open(p, "rb").read())
    h.update(" ".join(FLAGS).encode())
    return h.hexdigest()


def _compile(src):
    obj = os.path.join(BUILD, src + ".o")
    stamp_file = obj + ".stamp"
    stamp = _stamp(os.path.join(CSRC, src))
    if os.path.exists(obj) and os.path.exists(stamp_file) and open(stamp_file).read() == stamp:
        return obj, False, ""
    cmd = [NVCC, *FLAGS, "-x", "cu", "-c", os.path.join(CSRC, src), "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    open(stamp_file, "w").write(stamp)
    return obj, True, r.stderr


def build(force=False, verbose=False):
    os.makedirs(BUILD, exist_ok=True)
    if force:
        for f in os.listdir(BUILD):
            os.remove(os.path.join(BUILD, f))
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        results = list(ex.map(_compile, _sources()))
    objs = [r[0] for r in results]
    if verbose:
        for _, rebuilt, log in results:
            if rebuilt:
                sys.stderr.write(log)
    if any(r[1] for r in results) or not os.path.exists(LIB):
        cmd = [NVCC, "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a",
               "-Xlinker", f"--version-script={EXPORT_MAP}", "-ldl"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
