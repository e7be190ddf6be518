"""Builds the CPU oracle shared library (test infrastructure only).

-ffp-contract=off: every FMA in the restatement is an explicit fmaf() mirroring the
reference's SASS; gcc must add none of its own.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_build", "libsplat_oracle.so")
SRC = os.path.join(HERE, "splat_oracle.c")


def build(force=False, verbose=False):
    if (not force and os.path.exists(OUT) and os.path.getmtime(OUT) >= os.path.getmtime(SRC)):
        return OUT
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    base = ["-O2", "-std=c11", "-fPIC", "-shared", "-ffp-contract=off", "-fno-fast-math",
            "-Wall", "-Wextra", "-Wno-unknown-pragmas"]
    last = None
    # /opt/gcc (first on PATH in this image) ships without libgomp.spec; /usr/bin/gcc has it.
    for cc in ("/usr/bin/gcc", "gcc", "cc"):
        for omp in (["-fopenmp"], []):
            cmd = [cc] + base + omp + ["-o", OUT, SRC, "-lm"]
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode == 0:
                if verbose:
                    print(" ".join(cmd))
                return OUT
            last = r.stderr
    raise RuntimeError("oracle build failed:\n" + str(last))


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
