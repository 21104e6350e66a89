"""ORACLE — TEST INFRASTRUCTURE ONLY.

Second, independent restatement of DistributeOctTree (src/ORBextractor.cc:555-779) in plain Python, used only to
cross-check oracle/cpp's std::list version (tests/test_oracle_tree.py).  The list is a Python list with index 0 =
front.  The one thing Python cannot supply is libstdc++'s introsort element movement for compareNodes ties, so the
sort step is delegated to the real std::sort through oracle.sort_nodes (orc_sort_nodes).
"""
import math
import numpy as np


class _Node:
    __slots__ = ("x0", "y0", "x1", "y1", "keys", "no_more")

    def __init__(self, x0, y0, x1, y1):
        self.x0, self.y0, self.x1, self.y1 = x0, y0, x1, y1
        self.keys = []
        self.no_more = False

    def divide(self):                                   # ExtractorNode::DivideNode (:480-536)
        hx = int(math.ceil(np.float32(self.x1 - self.x0) / np.float32(2)))
        hy = int(math.ceil(np.float32(self.y1 - self.y0) / np.float32(2)))
        mx, my = self.x0 + hx, self.y0 + hy
        c = [_Node(self.x0, self.y0, mx, my), _Node(mx, self.y0, self.x1, my),
             _Node(self.x0, my, mx, self.y1), _Node(mx, my, self.x1, self.y1)]
        for k in self.keys:
            if k[0] < mx:
                c[0 if k[1] < my else 2].keys.append(k)
            else:
                c[1 if k[1] < my else 3].keys.append(k)
        for n in c:
            if len(n.keys) == 1:
                n.no_more = True
        return c


def distribute(keys, min_x, max_x, min_y, max_y, N, sort_nodes):
    """keys: iterable of (x, y, response) ints relative to the min border.  Returns the kept keys in list order."""
    keys = [tuple(int(v) for v in k) for k in keys]
    n_ini = int(np.round(np.float32(max_x - min_x) / np.float32(max_y - min_y)))      # :559 (round half away: values never tie)
    hx = np.float32(max_x - min_x) / np.float32(n_ini)
    roots = [_Node(int(hx * np.float32(i)), 0, int(hx * np.float32(i + 1)), max_y - min_y) for i in range(n_ini)]
    for k in keys:
        roots[int(np.float32(k[0]) / hx)].keys.append(k)
    nodes = []
    for r in roots:
        if len(r.keys) == 1:
            r.no_more = True
        if r.keys:
            nodes.append(r)

    def adopt(children, pending):
        for ch in children:
            if ch.keys:
                nodes.insert(0, ch)
                if len(ch.keys) > 1:
                    pending.append(ch)

    finish = False
    while not finish:
        prev = len(nodes)
        pending = []
        # walk the list front to back; children go to the front and are not visited in this pass
        snapshot = list(nodes)
        for nd in snapshot:
            if nd.no_more:
                continue
            adopt(nd.divide(), pending)
            nodes.remove(nd)
        if len(nodes) >= N or len(nodes) == prev:
            finish = True
        elif len(nodes) + 3 * len(pending) > N:
            while not finish:
                prev = len(nodes)
                old = pending
                pending = []
                perm = sort_nodes([len(p.keys) for p in old], [p.x0 for p in old])
                ordered = [old[i] for i in perm]
                for nd in reversed(ordered):
                    adopt(nd.divide(), pending)
                    nodes.remove(nd)
                    if len(nodes) >= N:
                        break
                if len(nodes) >= N or len(nodes) == prev:
                    finish = True
    out = []
    for nd in nodes:
        best = nd.keys[0]
        for k in nd.keys[1:]:
            if k[2] > best[2]:
                best = k
        out.append(best)
    return np.array(out, np.int32).reshape(-1, 3)
