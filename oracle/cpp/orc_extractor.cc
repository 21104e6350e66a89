// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_extractor.h).
#include "orc_extractor.h"

namespace orc {

static const int kPattern[1024] = {
#include "../../include/orb_pattern_31.inc"
};

// ---- constructor tables: src/ORBextractor.cc:409-469 ---------------------------------------------------------
Extractor::Extractor(int nf, float sf, int nl, int ini, int mn)
    : nfeatures(nf), nlevels(nl), iniTh(ini), minTh(mn), scaleFactor(sf) {
    scale.assign(nl, 1.f); sigma2.assign(nl, 1.f); invScale.resize(nl); invSigma2.resize(nl);
    for (int i = 1; i < nl; ++i) {
        scale[i] = (float)(scale[i - 1] * scaleFactor);      // float * double member -> float   (:421)
        sigma2[i] = scale[i] * scale[i];
    }
    for (int i = 0; i < nl; ++i) { invScale[i] = 1.0f / scale[i]; invSigma2[i] = 1.0f / sigma2[i]; }

    featuresPerLevel.resize(nl);
    float factor = (float)(1.0f / scaleFactor);                                               // (:433)
    float want = nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nl)); // (:434)
    int sum = 0;
    for (int l = 0; l < nl - 1; ++l) {
        featuresPerLevel[l] = cv_round(want);
        sum += featuresPerLevel[l];
        want *= factor;
    }
    featuresPerLevel[nl - 1] = std::max(nfeatures - sum, 0);

    pattern.assign(kPattern, kPattern + 1024);

    // circular patch row half-widths (:453-468)
    umax.assign(kHalfPatch + 1, 0);
    int vmax = cv_floor(kHalfPatch * std::sqrt(2.f) / 2 + 1);
    int vmin = cv_ceil(kHalfPatch * std::sqrt(2.f) / 2);
    const double hp2 = kHalfPatch * kHalfPatch;
    for (int v = 0; v <= vmax; ++v) umax[v] = cv_round(std::sqrt(hp2 - v * v));
    for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
        while (umax[v0] == umax[v0 + 1]) ++v0;
        umax[v] = v0;
        ++v0;
    }
    levels.resize(nl);
}

// ---- ComputePyramid: src/ORBextractor.cc:1170-1195 -----------------------------------------------------------
void Extractor::computePyramid(const uint8_t* img, int w, int h, int stride) {
    for (int l = 0; l < nlevels; ++l) {
        Level& L = levels[l];
        float s = invScale[l];
        L.w = cv_round((float)w * s);
        L.h = cv_round((float)h * s);
        L.pstride = L.w + 2 * kEdge;
        L.padded.assign((size_t)L.pstride * (L.h + 2 * kEdge), 0);
        uint8_t* roi = L.padded.data() + (size_t)kEdge * L.pstride + kEdge;
        if (l == 0) {
            make_border_101(img, w, h, stride, L.padded.data(), L.pstride, kEdge);
        } else {
            const Level& P = levels[l - 1];
            resize_linear_8u(P.roi(), P.w, P.h, P.pstride, roi, L.w, L.h, L.pstride);
            make_border_101(roi, L.w, L.h, L.pstride, L.padded.data(), L.pstride, kEdge);
        }
    }
}

// ---- quadtree culling: src/ORBextractor.cc:480-779 -----------------------------------------------------------
namespace {
struct Node {
    int x0, y0, x1, y1;                    // UL=(x0,y0) UR=(x1,y0) BL=(x0,y1) BR=(x1,y1)
    std::vector<KeyPoint> keys;
    std::list<Node>::iterator self;
    bool noMore = false;
    void split(Node c[4]) const {          // DivideNode (:480-536)
        const int hx = (int)std::ceil(static_cast<float>(x1 - x0) / 2);
        const int hy = (int)std::ceil(static_cast<float>(y1 - y0) / 2);
        const int mx = x0 + hx, my = y0 + hy;
        c[0] = Node{x0, y0, mx, my, {}, {}, false};
        c[1] = Node{mx, y0, x1, my, {}, {}, false};
        c[2] = Node{x0, my, mx, y1, {}, {}, false};
        c[3] = Node{mx, my, x1, y1, {}, {}, false};
        for (const KeyPoint& k : keys) {
            int q = (k.x < mx) ? ((k.y < my) ? 0 : 2) : ((k.y < my) ? 1 : 3);
            c[q].keys.push_back(k);
        }
        for (int q = 0; q < 4; ++q) if (c[q].keys.size() == 1) c[q].noMore = true;
    }
};
typedef std::pair<int, Node*> SizedNode;
// compareNodes (:538-553): by key count, then by UL.x.  Takes non-const refs like the reference.
bool smallerNode(SizedNode& a, SizedNode& b) {
    if (a.first != b.first) return a.first < b.first;
    return a.second->x0 < b.second->x0;
}
}  // namespace

std::vector<KeyPoint> Extractor::distributeOctTree(const std::vector<KeyPoint>& keys, int minX, int maxX,
                                                   int minY, int maxY, int N) {
    const int nIni = (int)std::round(static_cast<float>(maxX - minX) / (maxY - minY));
    const float hX = static_cast<float>(maxX - minX) / nIni;

    std::list<Node> nodes;
    std::vector<Node*> roots(nIni);
    for (int i = 0; i < nIni; ++i) {
        Node n{(int)(hX * static_cast<float>(i)), 0, (int)(hX * static_cast<float>(i + 1)), maxY - minY, {}, {}, false};
        nodes.push_back(n);
        roots[i] = &nodes.back();
    }
    for (const KeyPoint& k : keys) roots[(int)(k.x / hX)]->keys.push_back(k);

    for (auto it = nodes.begin(); it != nodes.end();) {
        if (it->keys.size() == 1) { it->noMore = true; ++it; }
        else if (it->keys.empty()) it = nodes.erase(it);
        else ++it;
    }

    std::vector<SizedNode> pending;      // vSizeAndPointerToNode
    // push the non-empty children to the FRONT of the list in n1..n4 order; remember the splittable ones
    auto adopt = [&](Node c[4], int* nToExpand) {
        for (int q = 0; q < 4; ++q) {
            if (c[q].keys.empty()) continue;
            nodes.push_front(c[q]);
            if (c[q].keys.size() > 1) {
                if (nToExpand) ++*nToExpand;
                pending.push_back(std::make_pair((int)c[q].keys.size(), &nodes.front()));
                nodes.front().self = nodes.begin();
            }
        }
    };

    bool finish = false;
    while (!finish) {
        int prevSize = (int)nodes.size();
        int nToExpand = 0;
        pending.clear();
        for (auto it = nodes.begin(); it != nodes.end();) {
            if (it->noMore) { ++it; continue; }
            Node c[4];
            it->split(c);
            adopt(c, &nToExpand);
            it = nodes.erase(it);
        }
        if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) {
            finish = true;
        } else if ((int)nodes.size() + nToExpand * 3 > N) {
            while (!finish) {
                prevSize = (int)nodes.size();
                std::vector<SizedNode> prev = pending;
                pending.clear();
                std::sort(prev.begin(), prev.end(), smallerNode);
                for (size_t t = 1; t < prev.size(); ++t)
                    if (prev[t].first == prev[t - 1].first && prev[t].second->x0 == prev[t - 1].second->x0) ++sortTieCount;
                for (int j = (int)prev.size() - 1; j >= 0; --j) {
                    Node c[4];
                    prev[j].second->split(c);
                    adopt(c, nullptr);
                    nodes.erase(prev[j].second->self);
                    if ((int)nodes.size() >= N) break;
                }
                if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) finish = true;
            }
        }
    }

    std::vector<KeyPoint> out;
    out.reserve(nodes.size());
    for (const Node& n : nodes) {
        const KeyPoint* best = &n.keys[0];
        for (size_t k = 1; k < n.keys.size(); ++k)
            if (n.keys[k].response > best->response) best = &n.keys[k];
        out.push_back(*best);
    }
    return out;
}

// ---- IC_Angle: src/ORBextractor.cc:76-103 --------------------------------------------------------------------
float Extractor::icAngle(const Level& L, float px, float py) const {
    return icAngleAt(L.roi() + (size_t)cv_round(py) * L.pstride + cv_round(px), L.pstride);
}
float Extractor::icAngleAt(const uint8_t* c, int step) const {
    int m01 = 0, m10 = 0;
    for (int u = -kHalfPatch; u <= kHalfPatch; ++u) m10 += u * c[u];
    for (int v = 1; v <= kHalfPatch; ++v) {
        int vs = 0, d = umax[v];
        for (int u = -d; u <= d; ++u) {
            int lo = c[u + v * step], hi = c[u - v * step];
            vs += lo - hi;
            m10 += u * (lo + hi);
        }
        m01 += v * vs;
    }
    return fast_atan2_deg((float)m01, (float)m10);
}

// ---- computeOrbDescriptor: src/ORBextractor.cc:106-146 ---------------------------------------------------------
// Convention: products and sum rounded separately in float (no FMA contraction; this
// file is compiled with -ffp-contract=off), cos/sin evaluated in double and narrowed.
void Extractor::orbDescriptor(const KeyPoint& kp, const uint8_t* img, int step, uint8_t* out) const {
    const float factorPI = (float)(M_PI / 180.f);
    float angle = (float)kp.angle * factorPI;
    float a = (float)std::cos((double)angle), b = (float)std::sin((double)angle);
    const uint8_t* c = img + (size_t)cv_round(kp.y) * step + cv_round(kp.x);
    const int* p = pattern.data();
    auto sample = [&](int i) -> int {
        float fx = (float)p[2 * i], fy = (float)p[2 * i + 1];
        int row = cv_round(fx * b + fy * a);
        int col = cv_round(fx * a - fy * b);
        return c[row * step + col];
    };
    for (int byte = 0; byte < 32; ++byte, p += 32) {
        int v = 0;
        for (int bit = 0; bit < 8; ++bit) v |= (sample(2 * bit) < sample(2 * bit + 1)) << bit;
        out[byte] = (uint8_t)v;
    }
}

// ---- ComputeKeyPointsOctTree: src/ORBextractor.cc:781-896 ----------------------------------------------------
void Extractor::computeKeyPointsOctTree() {
    const float W = 35;
    std::vector<FastPt> cell;
    for (int l = 0; l < nlevels; ++l) {
        Level& L = levels[l];
        const int minBX = kEdge - 3, minBY = minBX;
        const int maxBX = L.w - kEdge + 3, maxBY = L.h - kEdge + 3;
        L.cand.clear();
        const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
        const int nCols = (int)(width / W), nRows = (int)(height / W);
        const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
        for (int i = 0; i < nRows; ++i) {
            const float iniY = (float)(minBY + i * hCell);
            float maxY = iniY + hCell + 6;
            if (iniY >= maxBY - 3) continue;
            if (maxY > maxBY) maxY = (float)maxBY;
            for (int j = 0; j < nCols; ++j) {
                const float iniX = (float)(minBX + j * wCell);
                float maxX = iniX + wCell + 6;
                if (iniX >= maxBX - 6) continue;
                if (maxX > maxBX) maxX = (float)maxBX;
                const int cx = (int)iniX, cy = (int)iniY, cw = (int)maxX - cx, ch = (int)maxY - cy;
                const uint8_t* sub = L.roi() + (size_t)cy * L.pstride + cx;
                fast9_nms(sub, cw, ch, L.pstride, iniTh, cell);
                if (cell.empty()) fast9_nms(sub, cw, ch, L.pstride, minTh, cell);
                for (const FastPt& p : cell)
                    L.cand.push_back(KeyPoint{(float)(p.x + j * wCell), (float)(p.y + i * hCell), 7.f, -1.f, (float)p.score, 0});
            }
        }
        L.kept = distributeOctTree(L.cand, minBX, maxBX, minBY, maxBY, featuresPerLevel[l]);
        const int scaledPatch = (int)(kPatch * scale[l]);
        for (KeyPoint& k : L.kept) {
            k.x += minBX; k.y += minBY; k.octave = l; k.size = (float)scaledPatch;
        }
    }
    for (int l = 0; l < nlevels; ++l)
        for (KeyPoint& k : levels[l].kept) k.angle = icAngle(levels[l], k.x, k.y);
}

// ---- operator(): src/ORBextractor.cc:1086-1168 -----------------------------------------------------------------
int Extractor::extract(const uint8_t* img, int w, int h, int stride, int lap0, int lap1,
                       std::vector<KeyPoint>& kps, std::vector<uint8_t>& desc) {
    if (!img || w <= 0 || h <= 0) return -1;
    computePyramid(img, w, h, stride);
    computeKeyPointsOctTree();
    int total = 0;
    for (const Level& L : levels) total += (int)L.kept.size();
    kps.assign(total, KeyPoint{0, 0, 0, -1, 0, 0});
    desc.assign((size_t)total * 32, 0);
    int mono = 0, stereo = total - 1;
    std::vector<uint8_t> d(32);
    for (int l = 0; l < nlevels; ++l) {
        Level& L = levels[l];
        if (L.kept.empty()) { L.blurred.clear(); continue; }
        L.blurred.resize((size_t)L.w * L.h);
        gaussian7_8u(L.roi(), L.w, L.h, L.pstride, L.blurred.data(), L.w);   // clone() drops the padding (:1132)
        const float s = scale[l];
        for (const KeyPoint& k0 : L.kept) {
            orbDescriptor(k0, L.blurred.data(), L.w, d.data());
            KeyPoint k = k0;
            if (l != 0) { k.x *= s; k.y *= s; }
            int dst = (k.x >= lap0 && k.x <= lap1) ? stereo-- : mono++;
            kps[dst] = k;
            std::memcpy(&desc[(size_t)dst * 32], d.data(), 32);
        }
    }
    return mono;
}

}  // namespace orc
