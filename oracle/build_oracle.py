"""Compile oracle/raster_oracle.c -> oracle/liboracle.so (TEST INFRASTRUCTURE ONLY)."""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "raster_oracle.c")
OUT = os.path.join(HERE, "liboracle.so")
BASE = ["-O2", "-fPIC", "-shared", "-std=c11", "-ffp-contract=off", "-fno-fast-math", "-mfma"]


def build(force=False):
    if not force and os.path.exists(OUT) and os.path.getmtime(OUT) >= os.path.getmtime(SRC):
        return OUT
    compilers = [c for c in ("/usr/bin/gcc", shutil.which("gcc"), shutil.which("cc")) if c]
    last = None
    for cc in compilers:
        for extra in (["-fopenmp"], []):
            r = subprocess.run([cc, *BASE, *extra, "-o", OUT, SRC, "-lm"], capture_output=True, text=True)
            if r.returncode == 0:
                return OUT
            last = r.stderr
    raise RuntimeError("could not build the CPU oracle:\n" + str(last))


if __name__ == "__main__":
    print(build(force=True))
