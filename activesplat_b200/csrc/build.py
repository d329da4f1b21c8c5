"""Build libactivesplat_b200.so in-tree with nvcc for sm_100a (no torch headers involved)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "libactivesplat_b200.so")
SOURCES = ["preprocess.cu", "scan_sort.cu", "blend.cu", "fused.cu", "loss.cu", "scoring.cu", "densify.cu", "capi.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "--use_fast_math=false",
         "-Xcompiler", "-fPIC", "-Xptxas", "-v"]


def build_variant(out, extra_flags, only=None):
    """Developer tool: a second copy of the library with extra nvcc flags (e.g. -DASB_BW_STAGES=1), objects kept in
    a directory of their own; ASB_LIB_PATH points activesplat_b200.capi at it for an A/B measurement."""
    objdir = out + ".obj"
    os.makedirs(objdir, exist_ok=True)
    procs, objs = [], []
    for s in SOURCES:
        o = os.path.join(objdir, s[:-3] + ".o")
        objs.append(o)
        flags = [f for f in FLAGS if f != "--use_fast_math=false"] + (list(extra_flags) if (only is None or s in only) else [])
        procs.append((s, subprocess.Popen([NVCC, *flags, "-c", os.path.join(HERE, s), "-o", o], stdout=subprocess.PIPE,
                                          stderr=subprocess.STDOUT, text=True)))
    for s, p in procs:
        out_txt, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError("nvcc failed on %s:\n%s" % (s, out_txt))
    r = subprocess.run([NVCC, "-shared", "-o", out, *objs, "-lcudart"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    return out


def build(force=False, verbose=False):
    srcs = [os.path.join(HERE, s) for s in SOURCES if os.path.exists(os.path.join(HERE, s))]
    deps = srcs + [os.path.join(HERE, "common.cuh"), os.path.join(HERE, "raster_math.cuh"),
                   os.path.join(HERE, "..", "..", "include", "activesplat_b200.h")]
    if not force and os.path.exists(OUT) and all(os.path.getmtime(OUT) >= os.path.getmtime(d) for d in deps):
        return OUT
    objs = []
    procs = []
    for s in srcs:
        o = s[:-3] + ".o"
        objs.append(o)
        extra = os.environ.get("ASB_NVCC_EXTRA", "").split()  # experiments only (e.g. -DASB_FW_UNROLL=4)
        cmd = [NVCC, *[f for f in FLAGS if f != "--use_fast_math=false"], *extra, "-c", s, "-o", o]
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    log = []
    for s, p in procs:
        out, _ = p.communicate()
        log.append(out)
        if p.returncode != 0:
            raise RuntimeError("nvcc failed on %s:\n%s" % (s, out))
    r = subprocess.run([NVCC, "-shared", "-o", OUT, *objs, "-lcudart"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    if verbose:
        print("".join(log))
    with open(os.path.join(HERE, "ptxas.log"), "w") as f:
        f.write("".join(log))
    return OUT


if __name__ == "__main__":
    print(build(force=True, verbose="-v" in sys.argv))
