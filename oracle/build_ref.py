"""Builds the UNMODIFIED reference CUDA rasterizers / simple-knn for sm_100 into oracle/_ref/.

The reference sources are compiled from where they lie under /root/reference (nothing is
copied into this repository); only our C-ABI shims (oracle/ref_shim_*.cu) are added.  The
outputs (oracle/_ref/*.so) are git-ignored but travel to the GPU box with the snapshot.
The GPU box has no /root/reference: there the prebuilt files are used as they are.

  libref_dgr.so  diff-gaussian-rasterization  cuda_rasterizer/{rasterizer_impl,forward,backward}.cu
  libref_fsr.so  flashsplat-rasterization     same three files (+ -include cstdint: its
                                              rasterizer_impl.h lacks <cstdint>; -DREF_FLASH)
  libref_knn.so  simple-knn                   simple_knn.cu

Flags follow the reference's setup.py (DGR/setup.py:29: only -I third_party/glm; default
-O3, -fmad=true, no fast-math); arch = sm_100 (what TORCH_CUDA_ARCH_LIST=10.0 would give).
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/submodules"
OUT = os.path.join(HERE, "_ref")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
COMMON = ["-std=c++17", "-O3", "-gencode", "arch=compute_100,code=sm_100", "-Xcompiler", "-fPIC",
          "-ccbin", "/usr/bin/g++", "-w"]


def _obj(src, obj, extra):
    cmd = [NVCC] + COMMON + extra + ["-c", src, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed: %s\n%s" % (" ".join(cmd), r.stderr[-4000:]))
    return obj


def _raster(name, sub, extra):
    d = os.path.join(REF, sub)
    inc = ["-I", os.path.join(d, "third_party/glm"), "-I", os.path.join(d, "cuda_rasterizer")] + extra
    jobs = [(os.path.join(d, "cuda_rasterizer", f + ".cu"), os.path.join(OUT, "%s_%s.o" % (name, f)), inc)
            for f in ("rasterizer_impl", "forward", "backward")]
    jobs.append((os.path.join(HERE, "ref_shim_raster.cu"), os.path.join(OUT, "%s_shim.o" % name), inc))
    return name, jobs


def build(force=False):
    libs = {n: os.path.join(OUT, "libref_%s.so" % n) for n in ("dgr", "fsr", "knn")}
    if not os.path.isdir(REF):
        return {n: p for n, p in libs.items() if os.path.exists(p)}  # GPU box: prebuilt only
    if not force and all(os.path.exists(p) for p in libs.values()):
        newest_shim = max(os.path.getmtime(os.path.join(HERE, f))
                          for f in ("ref_shim_raster.cu", "ref_shim_knn.cu", "build_ref.py"))
        if all(os.path.getmtime(p) >= newest_shim for p in libs.values()):
            return libs
    os.makedirs(OUT, exist_ok=True)
    groups = [_raster("dgr", "diff-gaussian-rasterization", []),
              _raster("fsr", "flashsplat-rasterization", ["-include", "cstdint", "-DREF_FLASH"])]
    kd = os.path.join(REF, "simple-knn")
    groups.append(("knn", [(os.path.join(kd, "simple_knn.cu"), os.path.join(OUT, "knn_simple_knn.o"), ["-I", kd]),
                           (os.path.join(HERE, "ref_shim_knn.cu"), os.path.join(OUT, "knn_shim.o"), ["-I", kd])]))
    with ThreadPoolExecutor(max_workers=8) as ex:
        futs = [ex.submit(_obj, *j) for _, jobs in groups for j in jobs]
        for f in futs:
            f.result()
    for name, jobs in groups:
        cmd = [NVCC, "-shared", "-o", libs[name]] + [j[1] for j in jobs] + ["-ccbin", "/usr/bin/g++"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed: %s\n%s" % (" ".join(cmd), r.stderr[-4000:]))
        for j in jobs:
            os.remove(j[1])
    return libs


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
