"""Builds areal_b200/libareal_b200.so in-tree with nvcc for sm_100a (no torch headers: the kernels
sit behind a plain C ABI, so a full rebuild takes seconds).  `python -m areal_b200.build [-v] [-f]`."""
import concurrent.futures as cf
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libareal_b200.so")
OBJ = os.path.join(HERE, "csrc", "_obj")
SOURCES = ["c_api.cu", "logits_fused.cu", "logits_rowres.cu", "gae_scan.cu", "interval_copy.cu", "ppo_loss.cu", "masked_norm.cu",
           "masked_stats.cu", "clamped_entropy.cu", "lm_head_tc.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
         "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--expt-relaxed-constexpr"]
if os.environ.get("AREAL_B200_NVTX") == "1":      # NVTX ranges around the C-ABI calls (c_api.cu); link needs libdl only
    FLAGS += ["-DAREAL_B200_NVTX"]


def _deps():
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    hdrs.append(os.path.join(os.path.dirname(HERE), "include", "areal_b200.h"))
    return hdrs


def _compile(src, verbose):
    obj = os.path.join(OBJ, src.replace(".cu", ".o"))
    cmd = [NVCC, *FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0 or verbose:
        sys.stderr.write(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed on {src}")
    return obj


def build(verbose=False, force=False):
    os.makedirs(OBJ, exist_ok=True)
    srcs = [os.path.join(CSRC, s) for s in SOURCES]
    newest = max(os.path.getmtime(p) for p in srcs + _deps() + [os.path.abspath(__file__)])
    if not force and os.path.exists(OUT) and os.path.getmtime(OUT) >= newest:
        return OUT
    with cf.ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(lambda s: _compile(s, verbose), SOURCES))
    cmd = [NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "shared", "-Xlinker",
           "-rpath,/usr/local/cuda/lib64", "-o", OUT, *objs, "-ldl"]
    if verbose:
        print(" ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
