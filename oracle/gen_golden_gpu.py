"""TEST INFRASTRUCTURE ONLY.  Golden vectors that can only come from a GPU: the reference's 2D GAE kernels
(csrc/cugae/gae.cu:62-206: `gae_2d_olp`, `gae_2d_nolp`) have no Python caller, no CPU path and no test upstream, so
they are pinned by RUNNING them -- the reference's own `gae.cu`, compiled for sm_100a from where it lies
(`python oracle/build_oracle.py --ref-cuda` -> oracle/_ref/ref_cugae*.so, which travels to the GPU box) -- on seeded
inputs.  Run on a B200 (`python oracle/gen_golden_gpu.py`); writes gpurun_out/gae2d.npz, committed as
tests/golden/gae2d.npz."""
import glob
import importlib.util
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load_ref_cugae():
    so = glob.glob(os.path.join(ROOT, "oracle", "_ref", "ref_cugae*.so"))
    assert so, "oracle/_ref/ref_cugae*.so missing: run `python oracle/build_oracle.py --ref-cuda` where /root/reference exists"
    spec = importlib.util.spec_from_file_location("ref_cugae", so[0])
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def case_inputs(k):
    B, Hh, p_done = ((3, 40, 0.05), (5, 1030, 0.01), (16, 513, 0.01), (4, 96, 0.0))[k]
    rng = np.random.RandomState(40 + k)
    r = rng.randn(B, Hh).astype(np.float32)
    v = rng.randn(B, Hh + 1).astype(np.float32)
    done = rng.rand(B, Hh + 1) < p_done
    done[:, 0] = False
    idx, cu, trunc = [], [0], []
    for b in range(B):
        ys = np.nonzero(done[b])[0]
        idx.extend(ys.tolist())
        cu.append(len(idx))
        trunc.extend((rng.rand(len(ys)) < 0.3).tolist())
    return (r, v, done, np.array(idx, np.int32), np.array(cu, np.int32), np.array(trunc, bool),
            (0.99, 0.95) if k != 2 else (1.0, 1.0))


def main():
    ref = load_ref_cugae()
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    z = {"n_cases": np.array(4)}
    for k in range(4):
        r, v, done, idx, cu, trunc, (gamma, lam) = case_inputs(k)
        mx = int(np.diff(cu).max(initial=0))
        a, ret = ref.gae_2d_olp(t(r), t(v), t(done), t(idx), t(cu), mx, t(trunc), gamma, lam)
        a2, ret2 = ref.gae_2d_nolp(t(r), t(v), t(done), t(idx), t(cu), mx, t(trunc), gamma, lam)
        torch.cuda.synchronize()
        p = f"c{k}_"
        z.update({p + "rewards": r, p + "values": v, p + "done": done, p + "idx": idx, p + "cu": cu, p + "trunc": trunc,
                  p + "gamma": np.array(gamma), p + "lam": np.array(lam),
                  p + "adv_olp": a.cpu().numpy(), p + "ret_olp": ret.cpu().numpy(),
                  p + "adv_nolp": a2.cpu().numpy(), p + "ret_nolp": ret2.cpu().numpy()})
    out = os.path.join(ROOT, "gpurun_out", "gae2d.npz")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    np.savez_compressed(out, **z)
    print("gae2d golden cases: 4 ->", out)


if __name__ == "__main__":
    main()
