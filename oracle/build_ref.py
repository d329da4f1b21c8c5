"""Build the UNMODIFIED reference rasterizer (CUDA, torch extension) into oracle/_ref/.

TEST INFRASTRUCTURE ONLY. The sources are compiled from where they lie under
/root/reference/submodules/diff-gaussian-rasterization (nothing is copied into this repo);
the only output is oracle/_ref/ref_dgr_C.so (git-ignored, shipped to the GPU box by gpurun).
It exports the reference's three pybind functions (ext.cpp:15-18):
rasterize_gaussians, rasterize_gaussians_backward, mark_visible.

The one deviation from the reference's own setup.py is `-include cstdint`
(rasterizer_impl.h uses std::uintptr_t without including it; gcc 13 rejects that).
"""
import os
import sys

REF = "/root/reference/submodules/diff-gaussian-rasterization"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def build(verbose=False):
    if not os.path.isdir(REF):
        return None
    os.makedirs(OUT, exist_ok=True)
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    os.environ.setdefault("MAX_JOBS", "8")
    from torch.utils.cpp_extension import load
    srcs = [os.path.join(REF, p) for p in (
        "cuda_rasterizer/rasterizer_impl.cu", "cuda_rasterizer/forward.cu",
        "cuda_rasterizer/backward.cu", "rasterize_points.cu", "ext.cpp")]
    mod = load(name="ref_dgr_C", sources=srcs, build_directory=OUT,
               extra_include_paths=[os.path.join(REF, "third_party/glm")],
               extra_cflags=["-O3", "-include", "cstdint"],
               extra_cuda_cflags=["-O3", "-include", "cstdint", "-lineinfo",
                                  "-gencode", "arch=compute_100a,code=sm_100a"],
               verbose=verbose)
    return mod


if __name__ == "__main__":
    m = build(verbose="-v" in sys.argv)
    print("built" if m is not None else "reference tree not present; nothing built", OUT)
