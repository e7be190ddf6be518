"""Builds wheat_3dgs_b200/libwheatsplat.so in-tree with nvcc for sm_100a (cross-compiles
without a GPU).  `python -m wheat_3dgs_b200.build [--force] [--verbose]`."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libwheatsplat.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
SOURCES = ["ws_api.cu", "ws_preprocess.cu", "ws_radix_sort.cu", "ws_binning.cu", "ws_render_fwd.cu",
           "ws_render_bwd.cu", "ws_preprocess_bwd.cu", "ws_knn_labels.cu", "ws_loss.cu", "ws_optim.cu", "ws_objeval.cu", "ws_collective.cu",
           "ws_densify.cu"]
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC", "-ccbin", "/usr/bin/g++"]


def _deps():
    d = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    d.append(os.path.join(HERE, "..", "include", "wheatsplat.h"))
    d.append(os.path.abspath(__file__))
    return d


def build_variant(name, defines):
    """A/B builds: wheat_3dgs_b200/variants/lib<name>.so compiled with -D<defines> (git-ignored, loaded through
    WHEATSPLAT_LIB by tools/ab.py).  Not part of the product."""
    vdir = os.path.join(HERE, "variants")
    os.makedirs(os.path.join(vdir, name), exist_ok=True)
    out = os.path.join(vdir, "lib%s.so" % name)
    def cc(src):
        obj = os.path.join(vdir, name, src.replace(".cu", ".o"))
        cmd = [NVCC] + FLAGS + ["-D" + d for d in defines] + ["-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(r.stderr[-4000:])
        return obj

    with ThreadPoolExecutor(max_workers=8) as ex:
        objs = list(ex.map(cc, SOURCES))
    r = subprocess.run([NVCC, "-shared", "-o", out] + objs + ["-ccbin", "/usr/bin/g++"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(r.stderr[-4000:])
    return out


def build(force=False, verbose=False, extra_flags=()):
    if not force and os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(p) for p in _deps()):
        return LIB
    if not os.path.exists(NVCC):
        if os.path.exists(LIB):
            return LIB  # GPU box without a toolchain problem: use the shipped build
        raise RuntimeError("nvcc not found at %s and no prebuilt %s" % (NVCC, LIB))
    objdir = os.path.join(HERE, "csrc", "_obj")
    os.makedirs(objdir, exist_ok=True)

    def cc(src):
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        cmd = [NVCC] + FLAGS + list(extra_flags) + ["-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed: %s\n%s" % (" ".join(cmd), r.stderr[-6000:]))
        if verbose and r.stderr.strip():
            print(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=8) as ex:
        objs = list(ex.map(cc, SOURCES))
    cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-ccbin", "/usr/bin/g++"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed: %s\n%s" % (" ".join(cmd), r.stderr[-4000:]))
    return LIB


if __name__ == "__main__":
    flags = ["-Xptxas", "-v"] if "--verbose" in sys.argv else []
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv, extra_flags=flags))
