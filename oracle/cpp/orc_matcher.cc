// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_matcher.h).
#include "orc_matcher.h"
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>
#include <map>
#include <vector>
#include <thread>

namespace orc {

// ---- DescriptorDistance: src/ORBmatcher.cc:2058-2074 -------------------------------------------------------------
int descriptor_distance(const uint8_t* a, const uint8_t* b) {
    int32_t pa[8], pb[8];
    std::memcpy(pa, a, 32);
    std::memcpy(pb, b, 32);
    int dist = 0;
    for (int i = 0; i < 8; i++) {
        unsigned int v = pa[i] ^ pb[i];
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}

// ---- Frame grid: src/Frame.cc:251-252 (cell size), :385-416 (AssignFeaturesToGrid), :725-735 (PosInGrid) ---------
void FrameView::build(int n_, const KP* k, const uint8_t* d, float minX_, float minY_, float maxX_, float maxY_) {
    n = n_; kps = k; desc = d;
    minX = minX_; minY = minY_; maxX = maxX_; maxY = maxY_;
    invW = static_cast<float>(GRID_COLS) / static_cast<float>(maxX - minX);
    invH = static_cast<float>(GRID_ROWS) / static_cast<float>(maxY - minY);
    grid.assign(GRID_COLS * GRID_ROWS, std::vector<int>());
    for (int i = 0; i < n; i++) {
        const KP& kp = kps[i];
        int posX = (int)std::round((kp.x - minX) * invW);
        int posY = (int)std::round((kp.y - minY) * invH);
        if (posX < 0 || posX >= GRID_COLS || posY < 0 || posY >= GRID_ROWS) continue;
        grid[posX * GRID_ROWS + posY].push_back(i);
    }
}

// ---- GetFeaturesInArea: src/Frame.cc:657-723 ------------------------------------------------------------------------
void FrameView::features_in_area(float x, float y, float r, int minLevel, int maxLevel, std::vector<int>& out) const {
    out.clear();
    const float factorX = r, factorY = r;
    const int nMinCellX = std::max(0, (int)std::floor((x - minX - factorX) * invW));
    if (nMinCellX >= GRID_COLS) return;
    const int nMaxCellX = std::min((int)GRID_COLS - 1, (int)std::ceil((x - minX + factorX) * invW));
    if (nMaxCellX < 0) return;
    const int nMinCellY = std::max(0, (int)std::floor((y - minY - factorY) * invH));
    if (nMinCellY >= GRID_ROWS) return;
    const int nMaxCellY = std::min((int)GRID_ROWS - 1, (int)std::ceil((y - minY + factorY) * invH));
    if (nMaxCellY < 0) return;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++) {
        for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
            const std::vector<int>& vCell = grid[ix * GRID_ROWS + iy];
            for (size_t j = 0; j < vCell.size(); j++) {
                const KP& kpUn = kps[vCell[j]];
                if (bCheckLevels) {
                    if (kpUn.octave < minLevel) continue;
                    if (maxLevel >= 0)
                        if (kpUn.octave > maxLevel) continue;
                }
                const float distx = kpUn.x - x;
                const float disty = kpUn.y - y;
                if (std::fabs(distx) < factorX && std::fabs(disty) < factorY) out.push_back(vCell[j]);
            }
        }
    }
}

// ---- ComputeThreeMaxima: src/ORBmatcher.cc:2012-2053 ---------------------------------------------------------------
void three_maxima(const int* sizes, int L, int& ind1, int& ind2, int& ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; i++) {
        const int s = sizes[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

int rot_bin(float a1, float a2) {
    const float factor = 1.0f / HISTO_LENGTH;
    float rot = a1 - a2;
    if (rot < 0.0) rot += 360.0f;
    int bin = (int)std::round(rot * factor);
    if (bin == HISTO_LENGTH) bin = 0;
    return bin;
}

namespace {
struct RotHist {
    std::vector<int> h[HISTO_LENGTH];
    void add(float a1, float a2, int v) { h[rot_bin(a1, a2)].push_back(v); }
    // calls drop(v) for every entry outside the three dominant bins
    template <class Fn> void filter(Fn drop) {
        int sizes[HISTO_LENGTH];
        for (int i = 0; i < HISTO_LENGTH; i++) sizes[i] = (int)h[i].size();
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(sizes, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int v : h[i]) drop(v);
        }
    }
};
}  // namespace

// ---- M1: src/ORBmatcher.cc:43-213 (mono branch), RadiusByViewingCos :215-221 -------------------------------------
int search_local_map(const FrameView& F, uint8_t* occupied, const float* scaleFactors, int nq, const uint8_t* valid,
                     const float* projX, const float* projY, const int* level, const float* viewCos,
                     const uint8_t* qdesc, float th, float nnratio, int* kpMatch, const uint8_t* qocc) {
    int nmatches = 0;
    const bool bFactor = th != 1.0;
    std::vector<int> vIndices;
    for (int iMP = 0; iMP < nq; iMP++) {
        if (!valid[iMP]) continue;
        const int nPredictedLevel = level[iMP];
        float r = (viewCos[iMP] > 0.998) ? 2.5f : 4.0f;
        if (bFactor) r *= th;
        F.features_in_area(projX[iMP], projY[iMP], r * scaleFactors[nPredictedLevel], nPredictedLevel - 1, nPredictedLevel, vIndices);
        if (vIndices.empty()) continue;
        const uint8_t* MPdescriptor = qdesc + 32 * (size_t)iMP;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int idx : vIndices) {
            if (occupied[idx]) continue;
            const int dist = descriptor_distance(MPdescriptor, F.desc + 32 * (size_t)idx);
            if (dist < bestDist) {
                bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = F.kps[idx].octave; bestIdx = idx;
            } else if (dist < bestDist2) {
                bestLevel2 = F.kps[idx].octave; bestDist2 = dist;
            }
        }
        if (bestDist <= TH_HIGH) {
            if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;
            if (bestLevel != bestLevel2 || bestDist <= nnratio * bestDist2) {
                kpMatch[bestIdx] = iMP;
                // F.mvpMapPoints[bestIdx] = pMP (:188): later queries skip the slot iff THAT point has Observations() > 0 (:92-94)
                occupied[bestIdx] = qocc ? qocc[iMP] : 1;
                nmatches++;
            }
        }
    }
    return nmatches;
}

// ---- M2 / M3: src/ORBmatcher.cc:1676-1887 (mono: neither bForward nor bBackward), :1889-2010 ---------------------
int search_frame(const FrameView& F, uint8_t* occupied, const float* scaleFactors, int nq, const uint8_t* valid,
                 const float* u, const float* v, const int* octave, const float* qangle, const uint8_t* qdesc,
                 float th, int thDist, bool checkOri, int* kpMatch, const uint8_t* qocc) {
    int nmatches = 0;
    RotHist rot;
    std::vector<int> vIndices2;
    for (int i = 0; i < nq; i++) {
        if (!valid[i]) continue;
        const int nLastOctave = octave[i];
        const float radius = th * scaleFactors[nLastOctave];
        F.features_in_area(u[i], v[i], radius, nLastOctave - 1, nLastOctave + 1, vIndices2);
        if (vIndices2.empty()) continue;
        const uint8_t* dMP = qdesc + 32 * (size_t)i;
        int bestDist = 256, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            if (occupied[i2]) continue;
            const int dist = descriptor_distance(dMP, F.desc + 32 * (size_t)i2);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= thDist) {
            kpMatch[bestIdx2] = i;
            // CurrentFrame.mvpMapPoints[bestIdx2] = pMP (:1770): it blocks later queries iff pMP->Observations() > 0 (:1745-1747);
            // qocc == nullptr: any assigned point blocks (the KeyFrame variant, :1955)
            occupied[bestIdx2] = qocc ? qocc[i] : 1;
            nmatches++;
            if (checkOri) rot.add(qangle[i], F.kps[bestIdx2].angle, bestIdx2);
        }
    }
    if (checkOri) rot.filter([&](int kp) { kpMatch[kp] = -1; nmatches--; });
    return nmatches;
}

// ---- M7: src/ORBmatcher.cc:427-532 / :534-646 ---------------------------------------------------------------------
int search_sim3(const FrameView& KF, uint8_t* matched, const float* scaleFactors, int nq, const uint8_t* valid,
                const float* u, const float* v, const int* level, const uint8_t* qdesc, int th, float ratioHamming,
                int* kpMatch) {
    int nmatches = 0;
    std::vector<int> vIndices;
    for (int iMP = 0; iMP < nq; iMP++) {
        if (!valid[iMP]) continue;
        const int nPredictedLevel = level[iMP];
        const float radius = th * scaleFactors[nPredictedLevel];
        KF.features_in_area(u[iMP], v[iMP], radius, -1, -1, vIndices);
        if (vIndices.empty()) continue;
        const uint8_t* dMP = qdesc + 32 * (size_t)iMP;
        int bestDist = 256, bestIdx = -1;
        for (int idx : vIndices) {
            if (matched[idx]) continue;
            const int kpLevel = KF.kps[idx].octave;
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            const int dist = descriptor_distance(dMP, KF.desc + 32 * (size_t)idx);
            if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
        }
        if (bestDist <= TH_LOW * ratioHamming) {
            matched[bestIdx] = 1;
            kpMatch[bestIdx] = iMP;
            nmatches++;
        }
    }
    return nmatches;
}

// ---- M5: src/ORBmatcher.cc:648-763 -----------------------------------------------------------------------------------
int search_initialization(const FrameView& F1, const FrameView& F2, float* prevMatched, int windowSize, float nnratio,
                          bool checkOri, int* vnMatches12) {
    int nmatches = 0;
    for (int i = 0; i < F1.n; i++) vnMatches12[i] = -1;
    RotHist rot;
    std::vector<int> vMatchedDistance(F2.n, INT_MAX), vnMatches21(F2.n, -1), vIndices2;
    for (int i1 = 0; i1 < F1.n; i1++) {
        const int level1 = F1.kps[i1].octave;
        if (level1 > 0) continue;
        F2.features_in_area(prevMatched[2 * i1], prevMatched[2 * i1 + 1], (float)windowSize, level1, level1, vIndices2);
        if (vIndices2.empty()) continue;
        const uint8_t* d1 = F1.desc + 32 * (size_t)i1;
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            const int dist = descriptor_distance(d1, F2.desc + 32 * (size_t)i2);
            if (vMatchedDistance[i2] <= dist) continue;
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
            else if (dist < bestDist2) { bestDist2 = dist; }
        }
        if (bestDist <= TH_LOW) {
            if (bestDist < (float)bestDist2 * nnratio) {
                if (vnMatches21[bestIdx2] >= 0) { vnMatches12[vnMatches21[bestIdx2]] = -1; nmatches--; }
                vnMatches12[i1] = bestIdx2;
                vnMatches21[bestIdx2] = i1;
                vMatchedDistance[bestIdx2] = bestDist;
                nmatches++;
                if (checkOri) rot.add(F1.kps[i1].angle, F2.kps[bestIdx2].angle, i1);
            }
        }
    }
    if (checkOri)
        rot.filter([&](int idx1) { if (vnMatches12[idx1] >= 0) { vnMatches12[idx1] = -1; nmatches--; } });
    for (int i1 = 0; i1 < F1.n; i1++)
        if (vnMatches12[i1] >= 0) {
            prevMatched[2 * i1] = F2.kps[vnMatches12[i1]].x;
            prevMatched[2 * i1 + 1] = F2.kps[vnMatches12[i1]].y;
        }
    return nmatches;
}

// merge-join of two DBoW2::FeatureVector maps (std::map ordered by node id; lower_bound skips), e.g. :241-392
template <class Fn> static void join_nodes(const FeatVec& a, const FeatVec& b, Fn fn) {
    int ia = 0, ib = 0;
    while (ia < a.nn && ib < b.nn) {
        if (a.node[ia] == b.node[ib]) { fn(ia, ib); ia++; ib++; }
        else if (a.node[ia] < b.node[ib]) ia = (int)(std::lower_bound(a.node, a.node + a.nn, b.node[ib]) - a.node);
        else ib = (int)(std::lower_bound(b.node, b.node + b.nn, a.node[ia]) - b.node);
    }
}

// ---- M4a: src/ORBmatcher.cc:223-425 (F.Nleft == -1) ----------------------------------------------------------------
int search_bow_kf_frame(const FrameView& KF, const FeatVec& fvKF, const uint8_t* elig1, const FrameView& F,
                        const FeatVec& fvF, float nnratio, bool checkOri, int* matchF) {
    for (int i = 0; i < F.n; i++) matchF[i] = -1;
    int nmatches = 0;
    RotHist rot;
    join_nodes(fvKF, fvF, [&](int a, int b) {
        for (int iKF = fvKF.off[a]; iKF < fvKF.off[a + 1]; iKF++) {
            const int realIdxKF = fvKF.idx[iKF];
            if (!elig1[realIdxKF]) continue;
            const uint8_t* dKF = KF.desc + 32 * (size_t)realIdxKF;
            int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
            for (int iF = fvF.off[b]; iF < fvF.off[b + 1]; iF++) {
                const int realIdxF = fvF.idx[iF];
                if (matchF[realIdxF] >= 0) continue;
                const int dist = descriptor_distance(dKF, F.desc + 32 * (size_t)realIdxF);
                if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = realIdxF; }
                else if (dist < bestDist2) { bestDist2 = dist; }
            }
            if (bestDist1 <= TH_LOW) {
                if (static_cast<float>(bestDist1) < nnratio * static_cast<float>(bestDist2)) {
                    matchF[bestIdxF] = realIdxKF;
                    if (checkOri) rot.add(KF.kps[realIdxKF].angle, F.kps[bestIdxF].angle, bestIdxF);
                    nmatches++;
                }
            }
        }
    });
    if (checkOri) rot.filter([&](int idxF) { matchF[idxF] = -1; nmatches--; });
    return nmatches;
}

// ---- M4b: src/ORBmatcher.cc:765-905 ----------------------------------------------------------------------------------
int search_bow_kf_kf(const FrameView& K1, const FeatVec& fv1, const uint8_t* elig1, const FrameView& K2,
                     const FeatVec& fv2, const uint8_t* elig2, float nnratio, bool checkOri, int* matches12) {
    for (int i = 0; i < K1.n; i++) matches12[i] = -1;
    std::vector<uint8_t> vbMatched2(K2.n, 0);
    int nmatches = 0;
    RotHist rot;
    join_nodes(fv1, fv2, [&](int a, int b) {
        for (int p1 = fv1.off[a]; p1 < fv1.off[a + 1]; p1++) {
            const int idx1 = fv1.idx[p1];
            if (!elig1[idx1]) continue;
            const uint8_t* d1 = K1.desc + 32 * (size_t)idx1;
            int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
            for (int p2 = fv2.off[b]; p2 < fv2.off[b + 1]; p2++) {
                const int idx2 = fv2.idx[p2];
                if (vbMatched2[idx2] || !elig2[idx2]) continue;
                const int dist = descriptor_distance(d1, K2.desc + 32 * (size_t)idx2);
                if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = idx2; }
                else if (dist < bestDist2) { bestDist2 = dist; }
            }
            if (bestDist1 < TH_LOW) {
                if (static_cast<float>(bestDist1) < nnratio * static_cast<float>(bestDist2)) {
                    matches12[idx1] = bestIdx2;
                    vbMatched2[bestIdx2] = 1;
                    if (checkOri) rot.add(K1.kps[idx1].angle, K2.kps[bestIdx2].angle, idx1);
                    nmatches++;
                }
            }
        }
    });
    if (checkOri) rot.filter([&](int idx1) { matches12[idx1] = -1; nmatches--; });
    return nmatches;
}

// ---- M6: src/ORBmatcher.cc:907-1146, mono; epipolar test = src/CameraModels/Pinhole.cpp:107-129 -------------------
int search_triangulation(const FrameView& K1, const FeatVec& fv1, const uint8_t* free1, const FrameView& K2,
                         const FeatVec& fv2, const uint8_t* free2, const float* ep, const float* scaleFactors2,
                         const float* F12, const float* sigma2, bool coarse, bool checkOri, int* vMatches12) {
    for (int i = 0; i < K1.n; i++) vMatches12[i] = -1;
    int nmatches = 0;
    RotHist rot;
    join_nodes(fv1, fv2, [&](int a, int b) {
        for (int p1 = fv1.off[a]; p1 < fv1.off[a + 1]; p1++) {
            const int idx1 = fv1.idx[p1];
            if (!free1[idx1]) continue;
            const KP& kp1 = K1.kps[idx1];
            const uint8_t* d1 = K1.desc + 32 * (size_t)idx1;
            int bestDist = TH_LOW, bestIdx2 = -1;
            for (int p2 = fv2.off[b]; p2 < fv2.off[b + 1]; p2++) {
                const int idx2 = fv2.idx[p2];
                if (!free2[idx2]) continue;                      // vbMatched2 is never set (:949,1004)
                const int dist = descriptor_distance(d1, K2.desc + 32 * (size_t)idx2);
                if (dist > TH_LOW || dist > bestDist) continue;
                const KP& kp2 = K2.kps[idx2];
                const float distex = ep[0] - kp2.x;
                const float distey = ep[1] - kp2.y;
                if (distex * distex + distey * distey < 100 * scaleFactors2[kp2.octave]) continue;
                bool ok = coarse;
                if (!ok) {
                    const float la = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
                    const float lb = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
                    const float lc = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
                    const float num = la * kp2.x + lb * kp2.y + lc;
                    const float den = la * la + lb * lb;
                    if (den == 0) ok = false;
                    else {
                        const float dsqr = num * num / den;
                        ok = dsqr < 3.84 * sigma2[kp2.octave];
                    }
                }
                if (ok) { bestIdx2 = idx2; bestDist = dist; }
            }
            if (bestIdx2 >= 0) {
                vMatches12[idx1] = bestIdx2;
                nmatches++;
                if (checkOri) rot.add(kp1.angle, K2.kps[bestIdx2].angle, idx1);
            }
        }
    });
    if (checkOri) rot.filter([&](int idx1) { vMatches12[idx1] = -1; nmatches--; });
    return nmatches;
}

// ---- M9 (candidate part): src/ORBmatcher.cc:1233-1291 (mono), :1400-1437 (Sim3 variant, no chi2 gate) -------------
void fuse_candidates(const FrameView& KF, const float* scaleFactors, const float* invSigma2, int nq,
                     const uint8_t* valid, const float* u, const float* v, const int* level, const uint8_t* qdesc,
                     float th, bool chi2Gate, int* bestIdxOut, int* bestDistOut) {
    std::vector<int> vIndices;
    for (int i = 0; i < nq; i++) {
        bestIdxOut[i] = -1; bestDistOut[i] = 256;
        if (!valid[i]) continue;
        const int nPredictedLevel = level[i];
        const float radius = th * scaleFactors[nPredictedLevel];
        KF.features_in_area(u[i], v[i], radius, -1, -1, vIndices);
        if (vIndices.empty()) continue;
        const uint8_t* dMP = qdesc + 32 * (size_t)i;
        int bestDist = 256, bestIdx = -1;
        for (int idx : vIndices) {
            const KP& kp = KF.kps[idx];
            const int kpLevel = kp.octave;
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            if (chi2Gate) {
                const float ex = u[i] - kp.x;
                const float ey = v[i] - kp.y;
                const float e2 = ex * ex + ey * ey;
                if (e2 * invSigma2[kpLevel] > 5.99) continue;
            }
            const int dist = descriptor_distance(dMP, KF.desc + 32 * (size_t)idx);
            if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
        }
        bestIdxOut[i] = bestIdx; bestDistOut[i] = bestDist;
    }
}

// ---- M8: src/ORBmatcher.cc:1457-1674 (no caller in the tree; kept for the class surface) ---------------------------
// The per-point pre-checks (map point present / not yet matched / not bad, positive depth, IsInImage, distance
// invariance, PredictScale) are the caller's: valid/u/v/level per keypoint slot of each keyframe.
static void sim3_one_way(const FrameView& KF, const float* scaleFactors, int nq, const uint8_t* valid, const float* u,
                         const float* v, const int* level, const uint8_t* qdesc, float th, std::vector<int>& vnMatch) {
    std::vector<int> vIndices;
    vnMatch.assign(nq, -1);
    for (int i = 0; i < nq; i++) {
        if (!valid[i]) continue;
        const int nPredictedLevel = level[i];
        const float radius = th * scaleFactors[nPredictedLevel];
        KF.features_in_area(u[i], v[i], radius, -1, -1, vIndices);
        if (vIndices.empty()) continue;
        const uint8_t* dMP = qdesc + 32 * (size_t)i;
        int bestDist = INT_MAX, bestIdx = -1;
        for (int idx : vIndices) {
            const KP& kp = KF.kps[idx];
            if (kp.octave < nPredictedLevel - 1 || kp.octave > nPredictedLevel) continue;
            const int dist = descriptor_distance(dMP, KF.desc + 32 * (size_t)idx);
            if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
        }
        if (bestDist <= TH_HIGH) vnMatch[i] = bestIdx;
    }
}
int search_by_sim3(const FrameView& K1, const FrameView& K2, const float* scaleFactors1, const float* scaleFactors2,
                   const uint8_t* valid1, const float* u12, const float* v12, const int* level12, const uint8_t* qdesc1,
                   const uint8_t* valid2, const float* u21, const float* v21, const int* level21, const uint8_t* qdesc2, float th,
                   int* matches12) {
    std::vector<int> vnMatch1, vnMatch2;
    sim3_one_way(K2, scaleFactors2, K1.n, valid1, u12, v12, level12, qdesc1, th, vnMatch1);   // KF1 points into KF2
    sim3_one_way(K1, scaleFactors1, K2.n, valid2, u21, v21, level21, qdesc2, th, vnMatch2);   // KF2 points into KF1
    int nFound = 0;
    for (int i1 = 0; i1 < K1.n; i1++) {
        matches12[i1] = -1;
        const int idx2 = vnMatch1[i1];
        if (idx2 >= 0 && vnMatch2[idx2] == i1) { matches12[i1] = idx2; nFound++; }
    }
    return nFound;
}

// ---- 8(f) rank 4: Frame::ComputeStereoMatches (src/Frame.cc:811-981) ------------------------------------------------
// mvImagePyramid[l] of each extractor is given as a plain plane (the level ROI; every access below stays inside it for
// keypoints the extractor can produce: >= 16 level-pixels from the border).  cv::norm(IL, IR, NORM_L1) on 8-bit windows is
// an exact integer sum.  The row table vRowIndices is not materialised: its row `vL` lists, in ascending iR, exactly the
// right keypoints with floor(y - r) <= (int)vL <= ceil(y + r), r = 2 * scale[octave] (:825-838).
int stereo_matches(int N, const KP* keysL, const uint8_t* descL, int Nr, const KP* keysR, const uint8_t* descR, int nlevels,
                   const uint8_t* const* pyrL, const int* strideL, const uint8_t* const* pyrR, const int* strideR,
                   const int* colsR, const float* scaleFactors, const float* invScaleFactors, float mb, float mbf,
                   float* mvuRight, float* mvDepth) {
    (void)nlevels;
    for (int i = 0; i < N; i++) { mvuRight[i] = -1.0f; mvDepth[i] = -1.0f; }
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    const float minZ = mb;
    const float minD = 0;
    const float maxD = mbf / minZ;
    std::vector<std::pair<int, int>> vDistIdx;
    vDistIdx.reserve(N);
    for (int iL = 0; iL < N; iL++) {
        const KP& kpL = keysL[iL];
        const int levelL = kpL.octave;
        const float vL = kpL.y;
        const float uL = kpL.x;
        const int row = (int)vL;                                    // vRowIndices[vL]
        const float minU = uL - maxD;
        const float maxU = uL - minD;
        if (maxU < 0) continue;
        int bestDist = TH_HIGH;
        size_t bestIdxR = 0;
        const uint8_t* dL = descL + 32 * (size_t)iL;
        for (int iR = 0; iR < Nr; iR++) {
            const KP& kpR = keysR[iR];
            const float r = 2.0f * scaleFactors[kpR.octave];
            const int maxr = (int)std::ceil(kpR.y + r);
            const int minr = (int)std::floor(kpR.y - r);
            if (row < minr || row > maxr) continue;                 // iR is not in vRowIndices[row]
            if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
            const float uR = kpR.x;
            if (uR >= minU && uR <= maxU) {
                const int dist = descriptor_distance(dL, descR + 32 * (size_t)iR);
                if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
            }
        }
        if (bestDist < thOrbDist) {
            const float uR0 = keysR[bestIdxR].x;
            const float scaleFactor = invScaleFactors[kpL.octave];
            const float scaleduL = std::round(kpL.x * scaleFactor);
            const float scaledvL = std::round(kpL.y * scaleFactor);
            const float scaleduR0 = std::round(uR0 * scaleFactor);
            const int w = 5;
            const uint8_t* IL = pyrL[kpL.octave] + (size_t)((int)(scaledvL - w)) * strideL[kpL.octave] + (int)(scaleduL - w);
            int bestDistS = INT_MAX;
            int bestincR = 0;
            const int L = 5;
            float vDists[2 * 5 + 1];
            const float iniu = scaleduR0 + L - w;
            const float endu = scaleduR0 + L + w + 1;
            if (iniu < 0 || endu >= colsR[kpL.octave]) continue;
            for (int incR = -L; incR <= +L; incR++) {
                const uint8_t* IR = pyrR[kpL.octave] + (size_t)((int)(scaledvL - w)) * strideR[kpL.octave] +
                                    (int)(scaleduR0 + incR - w);
                long sum = 0;
                for (int y = 0; y < 2 * w + 1; y++)
                    for (int x = 0; x < 2 * w + 1; x++)
                        sum += std::abs((int)IL[y * strideL[kpL.octave] + x] - (int)IR[y * strideR[kpL.octave] + x]);
                const float dist = (float)(double)sum;
                if (dist < bestDistS) { bestDistS = dist; bestincR = incR; }
                vDists[L + incR] = dist;
            }
            if (bestincR == -L || bestincR == L) continue;
            const float dist1 = vDists[L + bestincR - 1];
            const float dist2 = vDists[L + bestincR];
            const float dist3 = vDists[L + bestincR + 1];
            const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
            if (deltaR < -1 || deltaR > 1) continue;
            float bestuR = scaleFactors[kpL.octave] * ((float)scaleduR0 + (float)bestincR + deltaR);
            float disparity = (uL - bestuR);
            if (disparity >= minD && disparity < maxD) {
                if (disparity <= 0) { disparity = 0.01; bestuR = uL - 0.01; }
                mvDepth[iL] = mbf / disparity;
                mvuRight[iL] = bestuR;
                vDistIdx.push_back(std::pair<int, int>(bestDistS, iL));
            }
        }
    }
    if (vDistIdx.empty()) return 0;                                  // the reference would index an empty vector here
    std::sort(vDistIdx.begin(), vDistIdx.end());
    const float median = vDistIdx[vDistIdx.size() / 2].first;
    const float thDist = 1.5f * 1.4f * median;
    int kept = (int)vDistIdx.size();
    for (int i = (int)vDistIdx.size() - 1; i >= 0; i--) {
        if (vDistIdx[i].first < thDist) break;
        mvuRight[vDistIdx[i].second] = -1;
        mvDepth[vDistIdx[i].second] = -1;
        kept--;
    }
    return kept;
}

// ---- M11: brute-force k-NN, strict '<' scan order => ties go to the lowest row --------------------------------------
static inline int popdist(const uint64_t* a, const uint64_t* b) {
    return __builtin_popcountll(a[0] ^ b[0]) + __builtin_popcountll(a[1] ^ b[1]) +
           __builtin_popcountll(a[2] ^ b[2]) + __builtin_popcountll(a[3] ^ b[3]);
}
void knn(const uint8_t* q, int nq, const uint8_t* db, long ndb, int k, uint16_t* dist, int32_t* idx, int threads) {
    auto work = [&](int q0, int q1) {
        std::vector<int> bd(k), bi(k);
        for (int i = q0; i < q1; i++) {
            uint64_t qa[4];
            std::memcpy(qa, q + 32 * (size_t)i, 32);
            std::fill(bd.begin(), bd.end(), 65535);
            std::fill(bi.begin(), bi.end(), -1);
            for (long j = 0; j < ndb; j++) {
                uint64_t da[4];
                std::memcpy(da, db + 32 * (size_t)j, 32);
                const int d = popdist(qa, da);
                if (d < bd[k - 1]) {
                    int p = k - 1;
                    while (p > 0 && bd[p - 1] > d) { bd[p] = bd[p - 1]; bi[p] = bi[p - 1]; p--; }
                    bd[p] = d; bi[p] = (int)j;
                }
            }
            for (int t = 0; t < k; t++) { dist[(size_t)i * k + t] = (uint16_t)bd[t]; idx[(size_t)i * k + t] = bi[t]; }
        }
    };
    threads = std::max(1, std::min(threads, nq));
    if (threads == 1) { work(0, nq); return; }
    std::vector<std::thread> th;
    for (int t = 0; t < threads; t++) th.emplace_back(work, (int)((long)nq * t / threads), (int)((long)nq * (t + 1) / threads));
    for (auto& x : th) x.join();
}

// ---- MapPoint::ComputeDistinctiveDescriptors: src/MapPoint.cc:366-396 -------------------------------------------------
int distinctive_descriptor(const uint8_t* desc, int N) {
    std::vector<float> D((size_t)N * N);
    for (int i = 0; i < N; i++) {
        D[(size_t)i * N + i] = 0;
        for (int j = i + 1; j < N; j++) {
            int d = descriptor_distance(desc + 32 * (size_t)i, desc + 32 * (size_t)j);
            D[(size_t)i * N + j] = d; D[(size_t)j * N + i] = d;
        }
    }
    int BestMedian = INT_MAX, BestIdx = 0;
    for (int i = 0; i < N; i++) {
        std::vector<int> vDists(D.begin() + (size_t)i * N, D.begin() + (size_t)(i + 1) * N);
        std::sort(vDists.begin(), vDists.end());
        int median = vDists[0.5 * (N - 1)];
        if (median < BestMedian) { BestMedian = median; BestIdx = i; }
    }
    return BestIdx;
}

// ---- Frame::UndistortKeyPoints (src/Frame.cc:747-780): cv::undistortPoints restated ---------------------------------------
void undistort_points(const float* xy, int n, const float* K, const float* dist, int ndist, float* out) {
    double k[12] = {0};
    for (int i = 0; i < ndist && i < 12; i++) k[i] = dist[i];
    const double fx = K[0], fy = K[1], cx = K[2], cy = K[3];
    const double ifx = 1. / fx, ify = 1. / fy;
    for (int i = 0; i < n; i++) {
        const double u = xy[2 * i], v = xy[2 * i + 1];
        double x = (u - cx) * ifx, y = (v - cy) * ify;
        const double x0 = x, y0 = y;
        for (int j = 0; j < 5; j++) {
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
            if (icdist < 0) { x = (u - cx) * ifx; y = (v - cy) * ify; break; }
            const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
            const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
            x = (x0 - deltaX) * icdist;
            y = (y0 - deltaY) * icdist;
        }
        // RR = P * R = K: third row (0, 0, 1) -> ww = 1
        const double xx = fx * x + 0. * y + cx, yy = 0. * x + fy * y + cy, ww = 1. / (0. * x + 0. * y + 1.);
        out[2 * i] = (float)(xx * ww);
        out[2 * i + 1] = (float)(yy * ww);
    }
}

// ---- DBoW2 transform (TemplatedVocabulary.h:1218-1260) -------------------------------------------------------------------------
void bow_descend(const VocabView& V, const uint8_t* feature, int levelsup, int* wordId, double* weight, int* nodeId) {
    const int nid_level = V.L - levelsup;
    if (nid_level <= 0) *nodeId = 0;  // root
    int final_id = 0, current_level = 0;
    do {
        ++current_level;
        const int b = V.childOff[final_id], e = V.childOff[final_id + 1];
        final_id = V.children[b];
        double best_d = descriptor_distance(feature, V.desc + 32 * (size_t)final_id);
        for (int c = b + 1; c < e; c++) {
            const int id = V.children[c];
            const double d = descriptor_distance(feature, V.desc + 32 * (size_t)id);
            if (d < best_d) { best_d = d; final_id = id; }
        }
        if (current_level == nid_level) *nodeId = final_id;
    } while (V.childOff[final_id + 1] != V.childOff[final_id]);
    *wordId = V.wordId[final_id];
    *weight = V.weight[final_id];
}

void bow_transform(const VocabView& V, const uint8_t* desc, int n, int levelsup, int* featWord, int* featNode,
                   int* nBow, int* bowWord, double* bowValue, int* nFv, int* fvNode, int* fvOff, int* fvIdx) {
    std::map<int, double> v;                         // DBoW2::BowVector
    std::map<int, std::vector<unsigned>> fv;         // DBoW2::FeatureVector
    const bool tf = V.weighting == 0 || V.weighting == 1;
    for (int i = 0; i < n; i++) {
        int id = 0, nid = 0; double w = 0;
        bow_descend(V, desc + 32 * (size_t)i, levelsup, &id, &w, &nid);
        featWord[i] = id; featNode[i] = nid;
        if (w > 0) {  // not stopped
            if (tf) v[id] += w;                      // addWeight
            else v.insert({id, w});                  // addIfNotExist
            fv[nid].push_back((unsigned)i);
        } else {
            featWord[i] = -1 - id;                   // stopped word: reported as -(id+1)
        }
    }
    if (tf && !v.empty() && V.norm == 0) {
        const double nd = (double)v.size();
        for (auto& kv : v) kv.second /= nd;
    }
    if (V.norm != 0) {
        double norm = 0.0;
        if (V.norm == 1) { for (auto& kv : v) norm += std::fabs(kv.second); }
        else { for (auto& kv : v) norm += kv.second * kv.second; norm = std::sqrt(norm); }
        if (norm > 0.0) for (auto& kv : v) kv.second /= norm;
    }
    int a = 0;
    for (auto& kv : v) { bowWord[a] = kv.first; bowValue[a] = kv.second; a++; }
    *nBow = a;
    int b = 0, o = 0;
    for (auto& kv : fv) {
        fvNode[b] = kv.first; fvOff[b] = o;
        for (unsigned f : kv.second) fvIdx[o++] = (int)f;
        b++;
    }
    fvOff[b] = o;
    *nFv = b;
}

}  // namespace orc
