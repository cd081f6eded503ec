"""Synthetic DBoW2 vocabulary files in the text format of the reference's loadFromTextFile
(R/Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1362-1436; ORBvoc.txt itself is not shipped) -- TEST INFRASTRUCTURE.

Header "k L scoring weighting"; one line per node in breadth-first order: parent id, leaf flag, 32 descriptor bytes,
weight. Children differ from their parent by a few flipped bits so that the descent is meaningful; some siblings are
exact duplicates (first-minimum tie-break), some leaves have weight 0 (features that are dropped), branching varies
and some leaves sit one level above L. No trailing newline: the reference's `while(!f.eof())` loop would turn it into
a phantom child of the root."""
import numpy as np


def write(path, k=10, L=4, scoring=0, weighting=0, seed=1, min_leaf_level=3):
    rng = np.random.Generator(np.random.PCG64(seed))
    lines = [f"{k} {L} {scoring} {weighting}"]
    level_nodes = [(0, rng.integers(0, 256, 32, dtype=np.uint8))]   # (id, descriptor) of the current level
    next_id = 1
    n_words = 0
    for level in range(1, L + 1):
        new_nodes = []
        for pid, pdesc in level_nodes:
            nchild = k if level == 1 else int(rng.integers(2, k + 1))
            prev = None
            for c in range(nchild):
                d = pdesc.copy()
                bits = rng.choice(256, int(rng.integers(4, 60 // level + 8)), replace=False)
                np.bitwise_xor.at(d, bits >> 3, (1 << (bits & 7)).astype(np.uint8))
                if prev is not None and rng.random() < 0.08:
                    d = prev.copy()                                   # duplicate sibling
                prev = d
                leaf = level == L or (level >= min_leaf_level and rng.random() < 0.15)
                w = 0.0 if (leaf and rng.random() < 0.03) else float(rng.uniform(0.05, 9.0))
                lines.append(f"{pid} {int(leaf)} " + " ".join(str(int(b)) for b in d) + f" {w!r}")
                if leaf:
                    n_words += 1
                else:
                    new_nodes.append((next_id, d))
                next_id += 1
        level_nodes = new_nodes
    with open(path, "w") as f:
        f.write("\n".join(lines))
    return next_id, n_words


def features(n, voc_path, seed=2):
    """Descriptors near the vocabulary's leaves (plus some uniform noise rows)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    rows = [ln.split() for ln in open(voc_path).read().split("\n")[1:]]
    leaves = np.array([[int(x) for x in r[2:34]] for r in rows if r[1] == "1"], np.uint8)
    out = leaves[rng.integers(0, len(leaves), n)].copy()
    for r in range(n):
        bits = rng.choice(256, int(rng.integers(0, 50)), replace=False)
        np.bitwise_xor.at(out[r], bits >> 3, (1 << (bits & 7)).astype(np.uint8))
    out[::11] = rng.integers(0, 256, (len(out[::11]), 32), dtype=np.uint8)
    return out
