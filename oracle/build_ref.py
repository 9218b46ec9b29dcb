"""Build recipe for the UNMODIFIED reference rasterizer extensions (test infrastructure only).

This is oracle/test infrastructure: nothing in the product path imports it.  It compiles
the two reference CUDA extensions straight from their sources where they lie under
/root/reference (never copied into this repo) for sm_100a, and writes ONLY into
oracle/_ref/ (git-ignored, but it travels to the GPU box with the gpurun snapshot):

  oracle/_ref/ref_hx_C.so  <- hessian-diff-gaussian-rasterization-w-depth   (near plane 0.001, fused
                              "Hessian" backward, HX/setup.py:18-29)
  oracle/_ref/ref_dg_C.so  <- diff-gaussian-rasterization-w-depth.git       (near plane 0.2, two-stage
                              backward, DG/setup.py:18-29)

The reference's own build system (setup.py/CMake) is not run; this is the short explicit
nvcc/g++ recipe.  The only deviation from the reference flags is `-include cstdint`
(rasterizer_impl.h uses uint32_t without the header) and the sm_100a
-gencode.  No --use_fast_math, as in the reference (HX/setup.py:29).

The GPU tests (tests/test_vs_reference_gpu.py) and bench.py --impl reference load these
modules through oracle/ref_ext.py.
"""
import os
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF_ROOT = os.environ.get("SPLAT_REFERENCE_ROOT", "/root/reference")
VARIANTS = {
    "ref_hx_C": "hessian-diff-gaussian-rasterization-w-depth",
    "ref_dg_C": "diff-gaussian-rasterization-w-depth.git",
}
CU_SOURCES = ["cuda_rasterizer/rasterizer_impl.cu", "cuda_rasterizer/forward.cu",
              "cuda_rasterizer/backward.cu", "rasterize_points.cu"]
CPP_SOURCES = ["ext.cpp"]


def _run(cmd):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("command failed: %s\n%s" % (" ".join(cmd), r.stdout[-4000:]))
    return r.stdout


def available():
    return os.path.isdir(REF_ROOT)


def build(force=False, verbose=True):
    """Compile both reference variants. Returns dict name -> .so path (only those built)."""
    if not available():
        return {n: os.path.join(OUT, n + ".so") for n in VARIANTS
                if os.path.exists(os.path.join(OUT, n + ".so"))}
    import torch  # headers + libs only
    from torch.utils.cpp_extension import include_paths, library_paths
    os.makedirs(OUT, exist_ok=True)
    tinc = include_paths()
    tlib = library_paths()
    pyinc = sysconfig.get_paths()["include"]
    cxx11 = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    built = {}
    jobs = []
    for name, sub in VARIANTS.items():
        src_root = os.path.join(REF_ROOT, sub)
        so = os.path.join(OUT, name + ".so")
        built[name] = so
        srcs = [os.path.join(src_root, s) for s in CU_SOURCES + CPP_SOURCES]
        if (not force and os.path.exists(so)
                and os.path.getmtime(so) > max(os.path.getmtime(s) for s in srcs)):
            continue
        objdir = os.path.join(OUT, "obj_" + name)
        os.makedirs(objdir, exist_ok=True)
        common = ["-DTORCH_EXTENSION_NAME=" + name, "-DTORCH_API_INCLUDE_EXTENSION_H",
                  "-D_GLIBCXX_USE_CXX11_ABI=%d" % cxx11, "-std=c++17", "-O3",
                  "-I" + os.path.join(src_root, "third_party/glm"), "-I" + pyinc]
        common += ["-I" + p for p in tinc]
        for s in CU_SOURCES:
            o = os.path.join(objdir, os.path.basename(s) + ".o")
            jobs.append(["nvcc", "-c", os.path.join(src_root, s), "-o", o,
                         "-gencode", "arch=compute_100a,code=sm_100a", "-include", "cstdint",
                         "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC", "-w"] + common)
        for s in CPP_SOURCES:
            o = os.path.join(objdir, os.path.basename(s) + ".o")
            jobs.append(["g++", "-c", os.path.join(src_root, s), "-o", o, "-fPIC", "-w",
                         "-I/usr/local/cuda/include"] + common)
    if jobs:
        if verbose:
            print("[build_ref] compiling %d reference translation units ..." % len(jobs), flush=True)
        with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
            list(ex.map(_run, jobs))
        for name in VARIANTS:
            objdir = os.path.join(OUT, "obj_" + name)
            if not os.path.isdir(objdir):
                continue
            objs = [os.path.join(objdir, os.path.basename(s) + ".o") for s in CU_SOURCES + CPP_SOURCES]
            if not all(os.path.exists(o) for o in objs):
                continue
            link = ["g++", "-shared", "-o", built[name]] + objs
            for p in tlib:
                link += ["-L" + p, "-Wl,-rpath," + p]
            link += ["-L/usr/local/cuda/lib64", "-Wl,-rpath,/usr/local/cuda/lib64",
                     "-lc10", "-ltorch", "-ltorch_cpu", "-ltorch_python", "-lc10_cuda",
                     "-ltorch_cuda", "-lcudart"]
            _run(link)
            if verbose:
                print("[build_ref] built", built[name], flush=True)
    return {n: p for n, p in built.items() if os.path.exists(p)}


# ---------------------------------------------------------------------------------------------------------------
# The reference's own PYTHON, staged next to the compiled extensions (oracle/_ref is git-ignored: nothing of this
# enters the repository) so that the GPU box -- which has no /root/reference -- can run it UNCHANGED on top of either
# rasterizer (tests/test_gpu_reference_python.py):
#   _ref/refpy/utils/{slam_helpers,slam_external,recon_helpers,keyframe_selection}.py   verbatim copies
#   _ref/refpy/splatam_fns.py    the functions of scripts/splatam.py listed in SPLATAM_FUNCS, cut out of the file with
#                                `ast` (the module itself is not importable: cv2, wandb, datasets ...), under the
#                                reference's own `from utils...` / `from diff_gaussian_rasterization ...` import lines
#   _ref/refpy_dg/diff_gaussian_rasterization/__init__.py   verbatim copy of DG's Python package; its `from . import _C`
#                                resolves to a generated two-line _C.py that loads _ref/ref_dg_C.so
PY_UTILS = ["slam_helpers.py", "slam_external.py", "recon_helpers.py", "keyframe_selection.py"]
SPLATAM_FUNCS = ["get_pointcloud", "initialize_params", "initialize_optimizer", "get_loss", "initialize_new_params",
                 "add_new_gaussians", "initialize_camera_pose"]


def stage_python(verbose=True):
    """Returns the staging directory (or None when /root/reference is absent and nothing was staged before)."""
    import ast
    import shutil
    dst = os.path.join(OUT, "refpy")
    if not available():
        return dst if os.path.isdir(dst) else None
    os.makedirs(os.path.join(dst, "utils"), exist_ok=True)
    open(os.path.join(dst, "utils", "__init__.py"), "w").close()
    for f in PY_UTILS:
        shutil.copyfile(os.path.join(REF_ROOT, "utils", f), os.path.join(dst, "utils", f))
    src = open(os.path.join(REF_ROOT, "scripts", "splatam.py")).read()
    tree = ast.parse(src)
    keep_imports, funcs = [], {}
    for node in tree.body:
        if isinstance(node, ast.ImportFrom) and node.module and (
                node.module in ("utils.keyframe_selection", "utils.recon_helpers", "utils.slam_helpers",
                                "utils.slam_external", "diff_gaussian_rasterization")):
            keep_imports.append(ast.get_source_segment(src, node))
        elif isinstance(node, ast.FunctionDef) and node.name in SPLATAM_FUNCS:
            funcs[node.name] = ast.get_source_segment(src, node)
    missing = [f for f in SPLATAM_FUNCS if f not in funcs]
    if missing:
        raise RuntimeError("scripts/splatam.py no longer defines %s" % missing)
    with open(os.path.join(dst, "splatam_fns.py"), "w") as f:
        f.write("# GENERATED by oracle/build_ref.py::stage_python from scripts/splatam.py of the reference (ast source\n"
                "# segments, unchanged).  Third-party modules the cut functions only touch on their visualisation\n"
                "# branches (cv2, matplotlib) are not imported.\n"
                "import os\nimport numpy as np\nimport torch\nimport torch.nn.functional as F\n\n")
        f.write("\n".join(keep_imports) + "\n\n\n")
        f.write("\n\n\n".join(funcs[n] for n in SPLATAM_FUNCS) + "\n")
    pkg = os.path.join(OUT, "refpy_dg", "diff_gaussian_rasterization")
    os.makedirs(pkg, exist_ok=True)
    shutil.copyfile(os.path.join(REF_ROOT, VARIANTS["ref_dg_C"], "diff_gaussian_rasterization", "__init__.py"),
                    os.path.join(pkg, "__init__.py"))
    with open(os.path.join(pkg, "_C.py"), "w") as f:
        f.write("# GENERATED by oracle/build_ref.py: the reference package's `from . import _C`, bound to the compiled\n"
                "# reference extension oracle/_ref/ref_dg_C.so\n"
                "from oracle import ref_ext as _r\n_m = _r.load('dg')\n"
                "rasterize_gaussians = _m.rasterize_gaussians\n"
                "rasterize_gaussians_backward = _m.rasterize_gaussians_backward\nmark_visible = _m.mark_visible\n")
    if verbose:
        print("[build_ref] staged the reference's Python into", dst, flush=True)
    return dst


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
    print(stage_python())
