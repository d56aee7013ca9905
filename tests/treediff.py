"""Where does a tree grown by the engine first differ from the oracle's, and how close was the oracle's decision there?

north_star allows a different tree "wherever a split-gain tie falls within fp32 epsilon".  The engine also quantises
the gradients of a round to 2^-19 of their largest magnitude, which can flip a decision whose two best candidates are
closer than that noise.  This module finds the ROOT-CAUSE nodes (first tree of a fold that differs, and in it the
nodes whose ancestors all agree) and grades each by the oracle's own margin:

  * split vs split with another (feature, bin): margin = (best - runner_up) / best   (oracle dumps the runner-up in `value`)
  * split vs leaf: margin = |best - threshold| / max(best, threshold), threshold = max(gamma, 1e-6)
"""
import numpy as np

FP32_TIE = 4 * 2.0 ** -23      # a few ulps of the gain
QUANT_FLIP = 2e-3              # anything above this is not explainable by 2^-19 gradient quantisation


def root_causes(mine, ref, gamma):
    """mine, ref: [rounds][MAX_NODES] dumps of one fold.  Returns (n_compared_nodes, [(round, node, kind, margin)]):
    nodes compared are those of the trees up to and including the first differing one."""
    compared, out = 0, []
    for r in range(len(ref)):
        a, b = mine[r], ref[r]
        exists = b["feat"] != -2
        same = (a["feat"] == b["feat"]) & (a["bin"] == b["bin"])
        if np.all(same[exists]) and np.all((a["feat"] != -2) == exists):
            compared += int(exists.sum())
            continue
        # first differing tree: walk from the root, stop descending at the first mismatch
        stack = [0]
        while stack:
            nid = stack.pop()
            if nid >= len(b) or b["feat"][nid] == -2 and a["feat"][nid] == -2:
                continue
            compared += 1
            if a["feat"][nid] == b["feat"][nid] and a["bin"][nid] == b["bin"][nid]:
                if b["feat"][nid] >= 0:
                    stack += [2 * nid + 1, 2 * nid + 2]
                continue
            best = float(b["gain"][nid])
            if a["feat"][nid] >= 0 and b["feat"][nid] >= 0:
                margin = (best - float(b["value"][nid])) / max(abs(best), 1e-30)
                kind = "other split"
            else:
                thr = max(float(gamma), 1e-6)
                margin = abs(best - thr) / max(abs(best), thr)
                kind = "split vs leaf"
            out.append((r, nid, kind, margin))
        break       # later trees start from different gradients: not comparable
    return compared, out


def grade(causes):
    ties = sum(1 for c in causes if c[3] <= FP32_TIE)
    flips = sum(1 for c in causes if FP32_TIE < c[3] <= QUANT_FLIP)
    wrong = [c for c in causes if c[3] > QUANT_FLIP]
    return ties, flips, wrong
