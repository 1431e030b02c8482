"""Shared by the CPU and GPU tests of the AEnt clamped entropy: golden cases and the tie-aware gradient check."""
import numpy as np

from tests import helpers as H


def load():
    z = H.load("clamped_entropy.npz")
    out = []
    for k in range(int(z["n_cases"])):
        clamp, temp = [float(v) for v in z[f"c{k}_cfg"]]
        out.append(dict(bits=z[f"c{k}_logits_bits"], labels=z[f"c{k}_labels"], clamp=clamp, temp=temp,
                        w_lp=z[f"c{k}_w_lp"], w_h=z[f"c{k}_w_h"], logp=z[f"c{k}_logp"], ent=z[f"c{k}_ent"],
                        grad=z[f"c{k}_grad"]))
    return out


def check_grad(got, ref, logits_f32, clamp, tol_rel, what):
    """The reference's top-k removes SOME `r` of the logits equal to the threshold value; ours spreads the
    kept weight over all of them.  So: every element that is not tied at the threshold must match, and the
    gradient summed over the tie class must match."""
    T, V = ref.shape
    k = int(V * clamp)
    for t in range(T):
        row = logits_f32[t]
        scale = max(float(np.abs(ref[t]).max()), 1e-12)
        if k > 0:
            tau = np.partition(row, k - 1)[k - 1]
            tie = row == tau
        else:
            tie = np.zeros(V, bool)
        err = np.abs(got[t][~tie].astype(np.float64) - ref[t][~tie])
        assert float(err.max(initial=0.0)) <= tol_rel * scale, (what, t, float(err.max()), scale)
        if tie.any():
            a, b = float(got[t][tie].astype(np.float64).sum()), float(ref[t][tie].astype(np.float64).sum())
            assert abs(a - b) <= tol_rel * scale * max(1.0, np.sqrt(tie.sum())), (what, t, a, b, int(tie.sum()))
