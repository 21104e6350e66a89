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
