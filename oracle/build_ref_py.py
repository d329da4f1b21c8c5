"""Byte-compile the UNMODIFIED reference Python of the SplaTAM layer into oracle/_ref/refpy/.

TEST INFRASTRUCTURE ONLY.  Like oracle/build_ref.py for the CUDA sources: the files are compiled from
where they lie under /root/reference (nothing is copied into this repository); the only outputs are
byte-code files (`*.bytecode`, the .pyc format) under oracle/_ref/refpy/ (git-ignored, shipped to the GPU box by gpurun, where
/root/reference does not exist).  `load()` imports them there with the reference's own module names:

  mapper.splatam.splatam                 src/mapper/splatam/splatam.py      (get_loss, add_new_gaussians,
                                         get_pointcloud, initialize_new_params, render, get_rendervars)
  mapper.splatam.utils.slam_helpers      transform_to_frame, quat_mult, rendervar builders
  mapper.splatam.utils.slam_external     build_rotation, calc_ssim
  mapper.splatam.utils.recon_helpers     setup_camera
  diff_gaussian_rasterization            dgr/diff_gaussian_rasterization/__init__.py, with `_C` bound to
                                         whichever extension module the caller names: the reference CUDA
                                         build (oracle/_ref/ref_dgr_C.so) to produce goldens, or
                                         activesplat_b200._C to show the wrapper runs unmodified on ours.
"""
import importlib.machinery
import importlib.util
import os
import py_compile
import sys
import types

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref", "refpy")

# module name -> source path below /root/reference
MODULES = {
    "mapper.splatam.splatam": "src/mapper/splatam/splatam.py",
    "mapper.splatam.utils.slam_helpers": "src/mapper/splatam/utils/slam_helpers.py",
    "mapper.splatam.utils.slam_external": "src/mapper/splatam/utils/slam_external.py",
    "mapper.splatam.utils.recon_helpers": "src/mapper/splatam/utils/recon_helpers.py",
    "diff_gaussian_rasterization": "submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py",
}


def _pyc(name):
    return os.path.join(OUT, name + ".bytecode")   # not "*.pyc": gpurun's snapshot leaves those behind


def build():
    if not os.path.isdir(REF):
        return None
    os.makedirs(OUT, exist_ok=True)
    for name, rel in MODULES.items():
        py_compile.compile(os.path.join(REF, rel), cfile=_pyc(name), dfile=rel, doraise=True,
                           invalidation_mode=py_compile.PycInvalidationMode.UNCHECKED_HASH)
    return OUT


def available():
    return all(os.path.exists(_pyc(n)) for n in MODULES)


def _exec(name, is_pkg=False, preset=None):
    loader = importlib.machinery.SourcelessFileLoader(name, _pyc(name))
    spec = importlib.util.spec_from_loader(name, loader, is_package=is_pkg)
    mod = importlib.util.module_from_spec(spec)
    if is_pkg:
        mod.__path__ = []
    for k, v in (preset or {}).items():
        setattr(mod, k, v)
    sys.modules[name] = mod
    loader.exec_module(mod)
    return mod


def load(c_module):
    """Import the reference modules with diff_gaussian_rasterization._C = c_module.  Returns a namespace with
    .splatam, .slam_helpers, .slam_external, .recon_helpers, .dgr.  Every call re-imports (so the two bindings
    can be used one after the other in one process); modules of an earlier call keep working."""
    if not available():
        raise FileNotFoundError("oracle/_ref/refpy is empty: run oracle/build_ref_py.py where /root/reference exists")
    # plotting back ends the reference imports at module level and only touches in its debug-visualisation branch
    for stub in ("matplotlib", "matplotlib.pyplot", "cv2"):
        try:
            importlib.import_module(stub)
        except Exception:
            sys.modules[stub] = types.ModuleType(stub)
    if not hasattr(sys.modules["matplotlib"], "pyplot"):
        sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    for pkg in ("mapper", "mapper.splatam", "mapper.splatam.utils"):
        m = types.ModuleType(pkg)
        m.__path__ = []
        sys.modules[pkg] = m
    shadowed = {k: sys.modules.get(k) for k in ("diff_gaussian_rasterization", "diff_gaussian_rasterization._C")}
    sys.modules["diff_gaussian_rasterization._C"] = c_module
    ns = types.SimpleNamespace()
    try:
        ns.dgr = _exec("diff_gaussian_rasterization", is_pkg=True, preset={"_C": c_module})
        ns.slam_external = _exec("mapper.splatam.utils.slam_external")
        ns.slam_helpers = _exec("mapper.splatam.utils.slam_helpers")
        ns.recon_helpers = _exec("mapper.splatam.utils.recon_helpers")
        ns.splatam = _exec("mapper.splatam.splatam")
    finally:
        # the reference modules hold their imports by reference; give the import table back to whoever had it
        # (this repository's own diff_gaussian_rasterization alias package)
        for k, v in shadowed.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return ns


def load_ref_extension():
    """oracle/_ref/ref_dgr_C.so (the reference CUDA build made by oracle/build_ref.py)."""
    import torch  # noqa: F401  (the extension links against libtorch)
    path = os.path.join(HERE, "_ref", "ref_dgr_C.so")
    spec = importlib.util.spec_from_file_location("ref_dgr_C", path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


if __name__ == "__main__":
    out = build()
    print("compiled" if out else "reference tree not present; nothing built", OUT)
