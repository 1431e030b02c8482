"""TEST INFRASTRUCTURE ONLY -- builds oracle/liboracle.so from oracle/oracle.c (gcc), and, when
/root/reference is present (build container only), oracle/_ref/interval_op*.so from the
reference's own csrc/interval_op/interval_op.cpp where it lies (pybind11; one file, no
external deps).  The reference's two .cu files need torch's headers and >5 min of nvcc time
each; they are built only on request (`python oracle/build_oracle.py --ref-cuda`) as the
"reference GPU kernel" timing baseline of BASELINE.md section 6 -- never as the oracle."""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("AREAL_REFERENCE_ROOT", "/root/reference")


def _stale(out, srcs):
    return (not os.path.exists(out)) or any(os.path.getmtime(s) > os.path.getmtime(out) for s in srcs)


def build_c_oracle(verbose=False):
    src = os.path.join(HERE, "oracle.c")
    out = os.path.join(HERE, "liboracle.so")
    if _stale(out, [src]):
        cmd = ["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-std=c11", "-o", out, src]
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
    return out


def build_ref_merge_intervals(verbose=False):
    """csrc/interval_op/interval_op.cpp -> oracle/_ref/interval_op<ext>.so (host-only pybind11)."""
    src = os.path.join(REF_ROOT, "csrc/interval_op/interval_op.cpp")
    if not os.path.isfile(src):
        return None
    import pybind11
    outdir = os.path.join(HERE, "_ref")
    os.makedirs(outdir, exist_ok=True)
    out = os.path.join(outdir, "interval_op" + sysconfig.get_config_var("EXT_SUFFIX"))
    if _stale(out, [src]):
        cmd = ["g++", "-O2", "-fPIC", "-shared", "-std=c++17", "-I", pybind11.get_include(),
               "-I", sysconfig.get_paths()["include"], "-o", out, src]
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
    return out


def build_ref_cuda(verbose=True):
    """Reference csrc/cugae/gae.cu and csrc/interval_op/interval_op.cu as torch extensions for
    sm_100a (timing baseline on the GPU box; optional)."""
    import torch
    from torch.utils import cpp_extension as ce
    outdir = os.path.join(HERE, "_ref")
    os.makedirs(outdir, exist_ok=True)
    ext = sysconfig.get_config_var("EXT_SUFFIX")
    inc = []
    for p in ce.include_paths("cuda"):
        inc += ["-I", p]
    inc += ["-I", sysconfig.get_paths()["include"]]
    for name, rel in (("ref_cugae", "csrc/cugae/gae.cu"), ("ref_interval_op_cuda", "csrc/interval_op/interval_op.cu")):
        src = os.path.join(REF_ROOT, rel)
        out = os.path.join(outdir, name + ext)
        if not os.path.isfile(src) or not _stale(out, [src]):
            continue
        libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
        cmd = ["nvcc", "-O3", "-std=c++17", "-shared", "-Xcompiler", "-fPIC",
               "-gencode", "arch=compute_100a,code=sm_100a",
               f"-DTORCH_EXTENSION_NAME={name}", "-DTORCH_API_INCLUDE_EXTENSION_H",
               "-D_GLIBCXX_USE_CXX11_ABI=" + str(int(torch._C._GLIBCXX_USE_CXX11_ABI)),
               *inc, "-L", libdir, "-lc10", "-ltorch", "-ltorch_cpu", "-ltorch_python", "-lc10_cuda", "-ltorch_cuda",
               "-o", out, src]
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)


if __name__ == "__main__":
    print(build_c_oracle(verbose=True))
    print(build_ref_merge_intervals(verbose=True))
    if "--ref-cuda" in sys.argv:
        build_ref_cuda()
