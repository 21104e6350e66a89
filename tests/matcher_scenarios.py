"""Synthetic inputs for the matcher parity tests: two consecutive frames of one agent's stream with their oracle-extracted
keypoints / descriptors, pseudo-BoW feature vectors (the ORB vocabulary file is absent from the reference checkout, so node
ids are a hash of descriptor bits — SURVEY.md 8d config 3), and per-element flags standing for the MapPoint predicates."""
import functools
import numpy as np

from mam3slam_b200 import synth

KP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"),
               ("class_id", "<i4")])


def to_kp(arr6):
    k = np.zeros(len(arr6), KP)
    k["x"], k["y"], k["size"], k["angle"], k["response"] = arr6[:, 0], arr6[:, 1], arr6[:, 2], arr6[:, 3], arr6[:, 4]
    k["octave"] = arr6[:, 5].astype(np.int32)
    k["class_id"] = -1
    return k


@functools.lru_cache(maxsize=8)
def frame_pair(w=752, h=480, nfeatures=1000, seed=0):
    import oracle
    frames = synth.stream(w, h, seed, 2)
    ex = oracle.Extractor(nfeatures, 1.2, 8, 20, 7)
    out = []
    for f in frames:
        _, k, d = ex(f)
        out.append((to_kp(k), d))
    tabs = ex.tables()
    return out[0], out[1], (0.0, 0.0, float(w), float(h)), tabs


def feat_vec(desc, bits=6):
    """pseudo DBoW2::FeatureVector as CSR (node asc, off, idx in keypoint order)."""
    node_of = (desc[:, 0].astype(np.int32) ^ (desc[:, 7].astype(np.int32) << 1)) & ((1 << bits) - 1)
    nodes = np.unique(node_of)
    off = [0]
    idx = []
    for nd in nodes:
        m = np.nonzero(node_of == nd)[0]
        idx.extend(m.tolist())
        off.append(len(idx))
    return nodes.astype(np.int32) * 3 + 5, np.array(off, np.int32), np.array(idx, np.int32)


def knn_data(ndb, nq, seed=1234, planted=True):
    rng = np.random.Generator(np.random.PCG64(seed))
    db = rng.integers(0, 256, (ndb, 32), dtype=np.uint8)
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    if planted:
        # exact duplicates of some rows later in the db: equal distances must resolve to the lowest row
        dup = rng.integers(0, ndb // 2, max(ndb // 50, 1))
        db[ndb // 2 + np.arange(len(dup))] = db[dup]
        src = rng.integers(0, ndb, nq)
        q = db[src].copy()
        for i in range(nq):
            nflip = int(rng.integers(0, 41))
            bits = rng.choice(256, nflip, replace=False)
            for b in bits:
                q[i, b >> 3] ^= 1 << (b & 7)
    return q, db


def make_vocab(k=10, L=4, seed=3, stop_frac=0.05, ragged=True, weighting=0, norm=1):
    """Synthetic DBoW2 vocabulary tree in the array form of include/orbm.h (the real ORBvoc.txt is absent from the reference
    checkout).  Node ids are a random permutation so that siblings are NOT contiguous, some inner nodes have fewer than k
    children and some leaves sit above level L (k-means yields both), sibling descriptors contain exact duplicates (ties go to
    the first child) and a fraction of the words is stopped (weight 0)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    parents = [(-1, 0)]                        # (parent position, depth) in creation order; position 0 = root
    desc = [np.zeros(32, np.uint8)]
    kids = [[]]
    frontier = [0]
    for depth in range(1, L + 1):
        nxt = []
        for p in frontier:
            if ragged and depth > 1 and rng.random() < 0.03:
                continue                       # p stays a leaf above level L
            nk = k if not ragged or depth == 1 or rng.random() < 0.8 else int(rng.integers(1, k + 1))
            base = desc[p] if depth > 1 else rng.integers(0, 256, 32, dtype=np.uint8)
            for c in range(nk):
                d = base.copy() if depth > 1 else rng.integers(0, 256, 32, dtype=np.uint8)
                flips = rng.choice(256, int(rng.integers(8, 60)), replace=False)
                np.bitwise_xor.at(d, flips // 8, (1 << (flips % 8)).astype(np.uint8))
                if c > 0 and rng.random() < 0.05:
                    d = desc[kids[p][int(rng.integers(0, c))]].copy()      # duplicate sibling: tie
                desc.append(d)
                kids.append([])
                kids[p].append(len(desc) - 1)
                nxt.append(len(desc) - 1)
        frontier = nxt
    nn = len(desc)
    perm = np.concatenate([[0], 1 + rng.permutation(nn - 1)]).astype(np.int64)     # position -> node id (root stays 0)
    inv = np.empty(nn, np.int64)
    inv[perm] = np.arange(nn)
    child_off = np.zeros(nn + 1, np.int32)
    children = []
    node_desc = np.zeros((nn, 32), np.uint8)
    word_id = np.full(nn, -1, np.int32)
    weight = np.zeros(nn, np.float64)
    nw = 0
    for nid in range(nn):                      # DBoW2 numbers the words in node-id order (createWords)
        pos = int(inv[nid])
        node_desc[nid] = desc[pos]
        children.extend(int(perm[c]) for c in kids[pos])
        child_off[nid + 1] = len(children)
        if not kids[pos]:
            word_id[nid] = nw
            nw += 1
            weight[nid] = 0.0 if rng.random() < stop_frac else float(rng.uniform(0.01, 9.0))
    return dict(L=L, child_off=child_off, children=np.array(children, np.int32), desc=node_desc, word_id=word_id, weight=weight,
                weighting=weighting, norm=norm, nwords=nw)


def vocab_features(voc, n, seed=11):
    """descriptors near random leaves of the tree (so that the descent is not a coin toss at every level), plus pure noise"""
    rng = np.random.Generator(np.random.PCG64(seed))
    leaves = np.nonzero(voc["word_id"] >= 0)[0]
    src = voc["desc"][rng.choice(leaves, n)].copy()
    for i in range(n):
        flips = rng.choice(256, int(rng.integers(0, 50)), replace=False)
        np.bitwise_xor.at(src[i], flips // 8, (1 << (flips % 8)).astype(np.uint8))
    src[::17] = rng.integers(0, 256, (len(src[::17]), 32), dtype=np.uint8)
    if n > 40:
        src[5] = src[23]                       # identical features land in the same word
        src[31] = src[23]
    return src
