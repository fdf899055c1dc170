"""Build libvq_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python vqvae_b200/build.py [--force] [--verbose]
    python vqvae_b200/build.py --variant NAME DEF1=V DEF2 ...     # tuning builds: libvq_b200_NAME.so, -DVQ_TUNING

Every translation unit is compiled to an object under ``vqvae_b200/build/`` (git-ignored) in parallel, then linked.
The .so is git-ignored but travels to the GPU box with the gpurun snapshot.  The ptxas resource report of the last
build lands in ``vqvae_b200/build/ptxas_v.log`` (not tracked; profiles/ holds the copies that are meant to be read).
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")
LIB = os.path.join(HERE, "libvq_b200.so")
OBJDIR = os.path.join(HERE, "build")
STAMP = os.path.join(OBJDIR, ".build_stamp")
SOURCES = ["vq_abi.cu", "vq_assign_tc.cu", "vq_assign_tc3.cu", "vq_assign_simt.cu", "vq_pass.cu", "vq_ema.cu",
           "vq_layout.cu", "vq_knn.cu", "vq_noalloc.cu"]
HEADERS = ["vq_common.cuh", "vq_assign_tc.cuh"]
CC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
            "-Xcompiler", "-fPIC", "-Xptxas", "-v"]
LD_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-Xcompiler", "-fPIC", "-cudart", "static"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def _sources():
    return [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def _digest(extra: str = "") -> str:
    h = hashlib.sha256()
    files = [os.path.join(CSRC, s) for s in _sources()] + \
            [os.path.join(CSRC, x) for x in HEADERS if os.path.exists(os.path.join(CSRC, x))] + \
            [os.path.join(INCLUDE, "vq_b200.h"), __file__]
    for f in files:
        with open(f, "rb") as fh:
            h.update(fh.read())
    h.update((" ".join(CC_FLAGS + LD_FLAGS) + extra).encode())
    return h.hexdigest()


def _compile(out: str, defines, objdir: str, verbose: bool) -> str:
    """Compile every source to an object (in parallel), link `out`; returns the concatenated ptxas report."""
    os.makedirs(objdir, exist_ok=True)
    nvcc = _nvcc()
    dflags = [f"-D{d}" for d in defines]

    def one(src: str):
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        cmd = [nvcc] + CC_FLAGS + dflags + ["-I", INCLUDE, "-c", "-o", obj, os.path.join(CSRC, src)]
        proc = subprocess.run(cmd, capture_output=True, text=True)
        return src, obj, cmd, proc

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        results = list(ex.map(one, _sources()))
    report = ""
    for src, obj, cmd, proc in results:
        if verbose or proc.returncode != 0:
            sys.stderr.write(" ".join(cmd) + "\n" + proc.stdout + proc.stderr)
        if proc.returncode != 0:
            raise RuntimeError(f"nvcc failed compiling {src}")
        report += f"==== {src}\n{proc.stderr}"
    cmd = [nvcc] + LD_FLAGS + ["-o", out] + [r[1] for r in results]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or proc.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + proc.stdout + proc.stderr)
    if proc.returncode != 0:
        raise RuntimeError("nvcc failed linking " + out)
    return report


def build_variant(name: str, defines) -> str:
    """Tuning aid: compile libvq_b200_<name>.so with -DVQ_TUNING (environment knobs, timing-experiment switches whose
    results are WRONG) plus extra -D flags; selected at run time with VQ_B200_LIB.  Never the product build."""
    out = os.path.join(HERE, f"libvq_b200_{name}.so")
    _compile(out, ["VQ_TUNING"] + list(defines), os.path.join(OBJDIR, "variant_" + name), verbose=False)
    return out


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile if sources changed; returns the path of the shared library."""
    digest = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(STAMP):
        with open(STAMP) as fh:
            if fh.read().strip() == digest:
                return LIB
    report = _compile(LIB, [], OBJDIR, verbose)
    with open(os.path.join(OBJDIR, "ptxas_v.log"), "w") as fh:
        fh.write(report)
    with open(STAMP, "w") as fh:
        fh.write(digest)
    return LIB


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--variant":      # build.py --variant NAME DEF1=V DEF2=V ...
        print(build_variant(sys.argv[2], sys.argv[3:]))
        sys.exit(0)
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
