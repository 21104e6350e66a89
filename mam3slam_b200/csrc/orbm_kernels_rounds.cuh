// Grid-window searches against an occupancy state under the rules "best unblocked candidate, accept if distance <= th"
// (SearchByProjection(Cur, Last), src/ORBmatcher.cc:1676-1887; SearchByProjection(Frame, KeyFrame, found), :1889-2010;
// the Sim3 projections, :427-646) and "best / second best with the NN ratio" (SearchByProjection(Frame, local map points),
// :43-213) resolved in CLAIM ROUNDS instead of one warp passing over the queries in order.
//
// The reference loop is sequential: query q takes the best keypoint of its window that no earlier query has taken, if the rule
// accepts it.  Two facts make that parallel:
//   (1) only candidates below a distance bound that are free at the start can ever matter to q (rounds_keepable), so q's whole
//       future is a short list: its first four such candidates in (distance, scan order), plus how many there were;
//   (2) blocks only grow.  So q's current pick K (first unblocked entry of its list; under the ratio rule the first two) is
//       FINAL as soon as no earlier unresolved occupying query has K anywhere in its list: nobody who comes before q can take K
//       any more, and everything better than K on q's list is already blocked for good.
// Phase A (all CTAs): a warp per query flattens the window's column runs over its lanes (walk_candidates, the reference's order)
//   and pops the four best by min-reductions.  Phase B (one CTA, a thread per query): rounds over all unresolved queries in
//   parallel - every occupying query registers min(q) on the keypoints of its unblocked entries (round-tagged atomicMax, no
//   clearing), a query whose picks carry no smaller registrant commits; the lowest unresolved query always does, so the rounds
//   end, and their number is the longest chain of queries that actually compete (14-26 on tracking data) instead of nq / 32
//   batches in sequence.
//   - A query with more than four candidates on file may have to look beyond its list.  It can only run short if all (but
//     one, under the ratio rule) of its unblocked entries are also on the list of an earlier unresolved occupying query; only
//     while that holds does it keep every LATER query from committing (step 1b), and when its entries are used up a warp
//     rebuilds its list against the state as it is (step 1r).
//   - A query that does not occupy what it matches (a map point without observations) holds nobody back, so later queries may
//     commit before it is resolved: keypoints carry the index of the query that took them and block only later queries.
//   - A run of rounds that settle almost nothing (hundreds of queries piled on a few dozen keypoints) hands the open queries,
//     compacted in order, to one warp that finishes them in the reference's own order.
// Results are identical to the sequential loop (tests/test_matcher_gpu.py: contention and pile-up cases against the oracle;
// tests/test_env_switches_gpu.py: byte-identical to the in-order kernel).
#pragma once
#include "orbm_kernels.cuh"

namespace orbm {

constexpr int kRoundThreads = 1024, kRoundQueries = kRoundThreads / 32;     // a warp per query; the resolving CTA wants a thread per query
constexpr uint8_t kRfResolved = 1, kRfIncomplete = 2, kRfOccupies = 4, kRfAccepted = 8;

// dynamic shared memory of the resolving CTA: owners u16 [n] | registrations u32 [n] | lists uint4 [nq] | best and second picks u32 [2 nq] | flags [nq] | re-walk queue u16 [nq]
__host__ __device__ inline size_t rounds_smem_bytes(int n, int nq) {
    const size_t a = ((size_t)n * 2 + 15) & ~(size_t)15;
    return a + (((size_t)n * 4 + 15) & ~(size_t)15) + (size_t)nq * 16 + (size_t)nq * 8 + (((size_t)nq + 15) & ~(size_t)15) + (size_t)nq * 2;
}
constexpr size_t kRoundsSmemMax = 200 * 1024;
constexpr int kRoundsMaxQueries = 65000;       // query index + 1 and 0xffff - query index live in 16 bits

// owner of a keypoint during the rounds: 0 = free, 0xffff = occupied before the search, else 1 + the occupying query that took it.
// A NON-occupying query does not hold later queries back, so later ones may commit before it is resolved: what it sees as
// blocked must be the state at ITS turn, i.e. only keypoints taken by earlier queries.
constexpr uint16_t kOwnerInitial = 0xffff;
__device__ __forceinline__ bool owner_blocks(const uint16_t* owner, int idx, int q) {
    const unsigned o = owner[idx];
    return o != 0 && (o == kOwnerInitial || (int)o - 1 < q);
}

// Which rules the rounds cover, and which candidates can matter under them.  A candidate matters if it could be accepted as the
// best (its distance passes the rule's threshold) or, under the best / second-best rule of SearchByProjection(Frame, local map
// points) (src/ORBmatcher.cc:107-139), if it could reject a best as the runner-up: bestDist > nnratio * d2 with bestDist <= TH_HIGH
// needs nnratio * d2 < TH_HIGH.  Both tests are monotone in the distance, so the kept candidates are exactly the closest ones
// and the first two unblocked entries of a list are the best and the second best the reference would see.
__host__ __device__ inline bool rounds_rule(int rule) { return rule == RULE_BEST_LE || rule == RULE_LOCAL_MAP; }
__device__ __forceinline__ bool rounds_keepable(const SearchParams& P, int d) {
    if (P.rule == RULE_BEST_LE) return (float)d <= P.thF;
    return d <= ORBM_TH_HIGH || __fmul_rn(P.nnratio, (float)d) < (float)ORBM_TH_HIGH;
}

// The list of one query, by a whole warp: walk_candidates flattens the window's column runs over the lanes (reference
// order), every lane keeps the candidates it saw that pass the tests, can matter under the rule and are not blocked, and the four
// smallest (distance, scan position) of the warp are popped by min-reductions.  Every lane returns the same list.
template <class Blocked>
__device__ __forceinline__ void warp_list(const SearchParams& P, const FrameDev& T, int q, Blocked&& isBlocked, Top4& out,
                                          unsigned long long& nEval) {
    const int lane = threadIdx.x & 31;
    uint32_t qd[8];
    {
        const uint4 a = reinterpret_cast<const uint4*>(P.qdesc)[2 * q], b = reinterpret_cast<const uint4*>(P.qdesc)[2 * q + 1];
        qd[0] = a.x; qd[1] = a.y; qd[2] = a.z; qd[3] = a.w; qd[4] = b.x; qd[5] = b.y; qd[6] = b.z; qd[7] = b.w;
    }
    // lane-private four best as (key = dist << 20 | scan position, entry); positions grow, so '<' keeps the earlier one first
    uint32_t k0 = 0xffffffffu, k1 = 0xffffffffu, k2 = 0xffffffffu, k3 = 0xffffffffu;
    uint32_t e0 = kNoEntry, e1 = kNoEntry, e2 = kNoEntry, e3 = kNoEntry;
    int count = 0, pos0 = 0;
    walk_candidates(P, T, q, qd, nEval, [&](bool ok, uint32_t entry) {
        const bool keep = ok && rounds_keepable(P, (int)(entry >> 23)) && !isBlocked((int)(entry & 0x7ffff));
        count += __popc(__ballot_sync(0xffffffffu, keep));
        if (keep) {
            const uint32_t key = ((entry >> 23) << 20) | (uint32_t)((pos0 + lane) & 0xfffff);
            if (key < k3) {
                if (key < k2) {
                    k3 = k2; e3 = e2;
                    if (key < k1) {
                        k2 = k1; e2 = e1;
                        if (key < k0) { k1 = k0; e1 = e0; k0 = key; e0 = entry; } else { k1 = key; e1 = entry; }
                    } else { k2 = key; e2 = entry; }
                } else { k3 = key; e3 = entry; }
            }
        }
        pos0 += 32;
    });
    uint32_t r[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t m = __reduce_min_sync(0xffffffffu, k0);
        r[i] = kNoEntry;
        if (m != 0xffffffffu) {                          // warp-uniform
            const unsigned w = __ballot_sync(0xffffffffu, k0 == m);         // keys are distinct: one lane
            r[i] = __shfl_sync(0xffffffffu, e0, __ffs(w) - 1);
            if (k0 == m) { k0 = k1; e0 = e1; k1 = k2; e1 = e2; k2 = k3; e2 = e3; k3 = 0xffffffffu; e3 = kNoEntry; }
        }
    }
    out.t0 = r[0]; out.t1 = r[1]; out.t2 = r[2]; out.t3 = r[3]; out.count = count;
}

// Phase A for the queries [q0, q0 + warps of the CTA) of one search, a warp each; lists / counts in global memory.
__device__ __forceinline__ void rounds_walk_cta(const SearchParams& P, int q0, uint4* __restrict__ lists, int* __restrict__ counts,
                                                unsigned long long& nEval) {
    const int q = q0 + (threadIdx.x >> 5);
    if (q >= P.nq) return;                               // whole warp
    Top4 m;
    warp_list(P, P.tgt, q, [&](int idx) { return P.occ[idx] != 0; }, m, nEval);
    if ((threadIdx.x & 31) == 0) { lists[q] = make_uint4(m.t0, m.t1, m.t2, m.t3); counts[q] = m.count; }
}

// Phase B, one CTA: resolves all queries of the search; leaves results in qResult / qDist / qBin and the final occupancy in
// P.occ.  `dyn` = rounds_smem_bytes(n, nq) bytes.
__device__ __forceinline__ void rounds_resolve_cta(const SearchParams& P, uint8_t* dyn, const uint4* lists, const int* counts) {
    __shared__ unsigned int sBarrier;
    __shared__ int sQueued;
    __shared__ int sSettled[2];          // queries settled in a round, by round parity
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int n = P.tgt.n, nq = P.nq;
    uint16_t* owner = reinterpret_cast<uint16_t*>(dyn);
    uint32_t* reg = reinterpret_cast<uint32_t*>(dyn + (((size_t)n * 2 + 15) & ~(size_t)15));
    uint4* L = reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(reg) + (((size_t)n * 4 + 15) & ~(size_t)15));
    uint32_t* pick = reinterpret_cast<uint32_t*>(L + nq);
    uint32_t* pick2 = pick + nq;
    uint8_t* flag = reinterpret_cast<uint8_t*>(pick2 + nq);
    const int needed = rule_needs_second(P.rule) ? 2 : 1;
    uint16_t* queue = reinterpret_cast<uint16_t*>(flag + (((size_t)nq + 15) & ~(size_t)15));
    for (int i = tid; i < n; i += nthr) { owner[i] = P.occ[i] ? kOwnerInitial : (uint16_t)0; reg[i] = 0; }
    for (int q = tid; q < nq; q += nthr) {
        L[q] = __ldcg(lists + q);
        const int c = __ldcg(counts + q);
        flag[q] = (uint8_t)((c > 4 ? kRfIncomplete : 0) | ((!P.qocc || P.qocc[q]) ? kRfOccupies : 0));
    }
    if (tid == 0) { sBarrier = 0; sQueued = 0; sSettled[0] = sSettled[1] = 0; }
    bool serial = false;
    unsigned roundsRun = 0;
    int slowRounds = 0;
    __syncthreads();
#ifdef ORBM_STATS
    const long long rT0 = clock64();
    long long st1 = 0, st1r = 0, st1b = 0, st2 = 0, c1 = 0, c1b = 0, c2 = 0;
#endif
    for (unsigned round = 1;; ++round) {
        const unsigned tag = round << 16;
        roundsRun = round;
#ifdef ORBM_STATS
        const long long rc0 = clock64();
#endif
        // step 1: every unresolved query picks its first unblocked entry and registers on all its unblocked entries
        auto settle = [&](int q, uint8_t f, const uint4 l, bool b0, bool b1, bool b2, bool b3) {
            if (!(b0 || b1 || b2 || b3)) { flag[q] = f | kRfResolved; return; }      // nothing it could accept is left: no match, for good
            // best and second best = the first two unblocked entries
            const unsigned m = (unsigned)b0 | ((unsigned)b1 << 1) | ((unsigned)b2 << 2) | ((unsigned)b3 << 3);
            const int ia = __ffs(m) - 1, ib = __ffs(m & (m - 1)) - 1;
            pick[q] = ia == 0 ? l.x : ia == 1 ? l.y : ia == 2 ? l.z : l.w;
            pick2[q] = ib == 1 ? l.y : ib == 2 ? l.z : ib == 3 ? l.w : kNoEntry;
            if (f & kRfOccupies) {
                const unsigned v = tag | (0xffffu - (unsigned)q);
                if (b0) atomicMax(&reg[l.x & 0x7ffff], v);
                if (b1) atomicMax(&reg[l.y & 0x7ffff], v);
                if (b2) atomicMax(&reg[l.z & 0x7ffff], v);
                if (b3) atomicMax(&reg[l.w & 0x7ffff], v);
            }
        };
        for (int q = tid; q < nq; q += nthr) {
            const uint8_t f = flag[q];
            if (f & kRfResolved) continue;
            const uint4 l = L[q];
            const bool b0 = l.x != kNoEntry && !owner_blocks(owner, l.x & 0x7ffff, q), b1 = l.y != kNoEntry && !owner_blocks(owner, l.y & 0x7ffff, q);
            const bool b2 = l.z != kNoEntry && !owner_blocks(owner, l.z & 0x7ffff, q), b3 = l.w != kNoEntry && !owner_blocks(owner, l.w & 0x7ffff, q);
            if ((int)b0 + (int)b1 + (int)b2 + (int)b3 < needed && (f & kRfIncomplete)) { queue[atomicAdd(&sQueued, 1)] = (uint16_t)q; continue; }
            settle(q, f, l, b0, b1, b2, b3);
        }
        __syncthreads();
        if (tid == 0) sSettled[(round + 1) & 1] = 0;        // last read before this barrier (end of the previous round), next written in round + 1
#ifdef ORBM_STATS
        c1 = clock64(); st1 += c1 - rc0;
#endif
        // step 1r: a list that is used up although the window held more is rebuilt against the state as it is now, a warp each
        const int queued = sQueued;
        if (queued) {                                       // CTA-uniform
#ifdef ORBM_STATS
            const long long q0c = clock64();
#endif
            for (int i = tid >> 5; i < queued; i += nthr >> 5) {
                const int q = queue[i];
                Top4 t; unsigned long long dummy = 0;
                warp_list(P, P.tgt, q, [&](int idx) { return owner_blocks(owner, idx, q); }, t, dummy);
                if ((tid & 31) == 0) {
                    uint8_t f = flag[q];
                    const uint4 l = make_uint4(t.t0, t.t1, t.t2, t.t3);
                    L[q] = l;
                    if (t.count <= 4) { f &= (uint8_t)~kRfIncomplete; flag[q] = f; }
                    settle(q, f, l, l.x != kNoEntry, l.y != kNoEntry, l.z != kNoEntry, l.w != kNoEntry);
                }
            }
            __syncthreads();
            if (tid == 0) sQueued = 0;
#ifdef ORBM_STATS
            if (tid == 0) { P.evals[5] += queued; P.evals[6] += 1; P.evals[7] += (unsigned long long)(clock64() - q0c); }
#endif
        }
#ifdef ORBM_STATS
        c1b = clock64(); st1r += c1b - c1;
#endif
        // step 1b: an occupying query with more candidates than its list holds can only run short of entries (one, or two under
        // a best / second-best rule) if all but that many of its unblocked entries also sit on the list of an earlier unresolved
        // occupying query; only then must the later queries wait
        for (int q = tid; q < nq; q += nthr) {
            const uint8_t f = flag[q];
            if ((f & (kRfResolved | kRfIncomplete | kRfOccupies)) != (kRfIncomplete | kRfOccupies)) continue;
            const uint4 l = L[q];
            const unsigned mine = tag | (0xffffu - (unsigned)q);
            int safe = 0;                             // unblocked entries nobody before q can take
            if (l.x != kNoEntry && !owner_blocks(owner, l.x & 0x7ffff, q) && reg[l.x & 0x7ffff] == mine) ++safe;
            if (l.y != kNoEntry && !owner_blocks(owner, l.y & 0x7ffff, q) && reg[l.y & 0x7ffff] == mine) ++safe;
            if (l.z != kNoEntry && !owner_blocks(owner, l.z & 0x7ffff, q) && reg[l.z & 0x7ffff] == mine) ++safe;
            if (l.w != kNoEntry && !owner_blocks(owner, l.w & 0x7ffff, q) && reg[l.w & 0x7ffff] == mine) ++safe;
            if (safe < needed) atomicMax(&sBarrier, mine);
        }
        __syncthreads();
#ifdef ORBM_STATS
        c2 = clock64(); st1b += c2 - c1b;
#endif
        // step 2: a pick without an earlier registrant is final
        const unsigned bar = sBarrier;
        const int barQ = (bar >> 16) == round ? (int)(0xffffu - (bar & 0xffffu)) : INT_MAX;
        int open = 0, settled = 0;
        for (int q = tid; q < nq; q += nthr) {
            const uint8_t f = flag[q];
            if (f & kRfResolved) continue;
            const uint32_t e = pick[q], e2 = needed == 2 ? pick2[q] : kNoEntry;
            const int K = (int)(e & 0x7ffff);
            const unsigned r = reg[K];
            int minQ = (r >> 16) == round ? (int)(0xffffu - (r & 0xffffu)) : INT_MAX;
            if (e2 != kNoEntry) {                     // the runner-up must be as settled as the best
                const unsigned r2 = reg[e2 & 0x7ffff];
                if ((r2 >> 16) == round) minQ = min(minQ, (int)(0xffffu - (r2 & 0xffffu)));
            }
            if (minQ >= q && q <= barQ) {
                const bool acc = decide_query(P, e, e2);
                if (acc && (f & kRfOccupies)) owner[K] = (uint16_t)(q + 1);
                flag[q] = f | kRfResolved | (acc ? kRfAccepted : 0);
                ++settled;
            } else ++open;
        }
        if (settled) atomicAdd(&sSettled[round & 1], settled);
        const int stillOpen = __syncthreads_count(open);
#ifdef ORBM_STATS
        if (tid == 0) { st2 += clock64() - c2; }
#endif
        if (!stillOpen) {
#ifdef ORBM_STATS
            if (tid == 0) { P.evals[3] = (unsigned long long)(clock64() - rT0); P.evals[8] = st1; P.evals[9] = st1r; P.evals[10] = st1b; P.evals[11] = st2; }
#endif
            break;
        }
        // Long chains of queries that all compete for the same few keypoints settle a query or two per round: from there on the
        // reference's own order is the faster way (below).  The count is complete after the barrier and read by everyone before
        // the next one, so the decision is uniform.  A single slow round is usually one query holding the later ones back
        // (step 1b) with hundreds ready to go in the next round, so only a run of slow rounds counts.
        slowRounds = (sSettled[round & 1] < 4) ? slowRounds + 1 : 0;
        if (round >= 8 && slowRounds >= 4) { serial = true; break; }
    }
    if (serial) {
        // in-order finish by one warp (all lanes do the same scalar work; the lanes only share the window walks): every query
        // before q is settled, so q's first unblocked entries are what the reference sees
        // the open queries in ascending order (ballot ranks inside a warp, warp counts summed across the CTA)
        __shared__ int sWarpOpen[32], sOpenBase;
        if (tid == 0) sOpenBase = 0;
        __syncthreads();
        for (int q0 = 0; q0 < nq; q0 += nthr) {
            const int q = q0 + tid;
            const bool isOpen = q < nq && !(flag[q] & kRfResolved);
            const unsigned bal = __ballot_sync(0xffffffffu, isOpen);
            if ((tid & 31) == 0) sWarpOpen[tid >> 5] = __popc(bal);
            __syncthreads();
            int before = sOpenBase;
            for (int w = 0; w < (tid >> 5); ++w) before += sWarpOpen[w];
            if (isOpen) queue[before + __popc(bal & ((1u << (tid & 31)) - 1))] = (uint16_t)q;
            __syncthreads();
            if (tid == nthr - 1) sOpenBase = before + __popc(bal);
            __syncthreads();
        }
        const int nOpen = sOpenBase;
        if (tid < 32) {
            for (int i = 0; i < nOpen; ++i) {
                const int q = queue[i];
                const uint8_t f = flag[q];
                uint4 l = L[q];
                bool b0 = l.x != kNoEntry && !owner_blocks(owner, l.x & 0x7ffff, q), b1 = l.y != kNoEntry && !owner_blocks(owner, l.y & 0x7ffff, q);
                bool b2 = l.z != kNoEntry && !owner_blocks(owner, l.z & 0x7ffff, q), b3 = l.w != kNoEntry && !owner_blocks(owner, l.w & 0x7ffff, q);
                if ((int)b0 + (int)b1 + (int)b2 + (int)b3 < needed && (f & kRfIncomplete)) {
                    Top4 t; unsigned long long dummy = 0;
                    warp_list(P, P.tgt, q, [&](int idx) { return owner_blocks(owner, idx, q); }, t, dummy);
                    l = make_uint4(t.t0, t.t1, t.t2, t.t3);
                    b0 = l.x != kNoEntry; b1 = l.y != kNoEntry; b2 = l.z != kNoEntry; b3 = l.w != kNoEntry;
                }
                uint8_t nf = f | kRfResolved;
                if (b0 || b1 || b2 || b3) {
                    const unsigned m = (unsigned)b0 | ((unsigned)b1 << 1) | ((unsigned)b2 << 2) | ((unsigned)b3 << 3);
                    const int ia = __ffs(m) - 1, ib = __ffs(m & (m - 1)) - 1;
                    const uint32_t e = ia == 0 ? l.x : ia == 1 ? l.y : ia == 2 ? l.z : l.w;
                    const uint32_t e2 = needed == 2 ? (ib == 1 ? l.y : ib == 2 ? l.z : ib == 3 ? l.w : kNoEntry) : kNoEntry;
                    if (decide_query(P, e, e2)) {
                        if (f & kRfOccupies) owner[e & 0x7ffff] = (uint16_t)(q + 1);
                        pick[q] = e;
                        nf |= kRfAccepted;
                    }
                }
                flag[q] = nf;
                __syncwarp();
            }
        }
        __syncthreads();
    }
    if (tid == 0) { P.evals[1] = roundsRun; P.evals[2] = serial ? 1ull : 0ull; }       // reporting (orbm_debug_last_rounds)
#ifdef ORBM_STATS
    { int inc = 0; for (int q = tid; q < nq; q += nthr) inc += __ldcg(counts + q) > 4; if (inc) atomicAdd(P.evals + 12, (unsigned long long)inc); }
#endif
    // results in one pass (the rotation bin needs the accepted keypoint's angle from global memory: one round trip, not one per round)
    for (int q = tid; q < nq; q += nthr) {
        const bool acc = (flag[q] & kRfAccepted) != 0;
        write_result(P, P.tgt, q, acc ? pick[q] : kNoEntry, acc, (acc && P.checkOri) ? P.qangle[q] : 0.f);
    }
    for (int i = tid; i < n; i += nthr) P.occ[i] = owner[i] != 0;
    __syncthreads();
}

// One search (orbm_search_frame / orbm_search_sim3): grid = ceil(nq / kRoundQueries) CTAs; the last one to finish resolves and
// runs the rotation-consistency epilogue.
// mirrorDst / mirrorSrc / mirrorVec: the block (occupancy, counters, results) is written to the caller's page-locked host
// mirror by this CTA at the end - no separate device-to-host copy on the latency path.
__global__ void __launch_bounds__(kRoundThreads) k_search_rounds(const SearchParams P, uint4* lists, int* counts, uint4* mirrorDst,
                                                                 const uint4* mirrorSrc, int mirrorVec) {
    extern __shared__ __align__(16) uint8_t dyn[];
    __shared__ int isLast;
    unsigned long long nEval = 0;
#ifdef ORBM_STATS
    const long long wT0 = clock64();
#endif
    rounds_walk_cta(P, blockIdx.x * kRoundQueries, lists, counts, nEval);
#ifdef ORBM_STATS
    __syncthreads();
    if (threadIdx.x == 0) atomicMax(P.evals + 4, (unsigned long long)(clock64() - wT0));
#endif
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) nEval += __shfl_xor_sync(0xffffffffu, nEval, o);
    if ((threadIdx.x & 31) == 0 && nEval) atomicAdd(P.evals, nEval);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) isLast = atomicAdd(P.ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!isLast) return;
    __threadfence();
    rounds_resolve_cta(P, dyn, lists, counts);
    if (P.checkOri) rot_filter_cta(P.nq, P.qBin, P.qResult, true);
    if (mirrorDst) {
        __syncthreads();
        for (int i = threadIdx.x; i < mirrorVec; i += blockDim.x) mirrorDst[i] = mirrorSrc[i];
    }
}

// The tracking step of a batch (orbm_track_batch_device) in two launches: the walk over small CTAs (a warp per query of every
// pair: grid = (ceil(stride / 8), pairs) x 256), then one resolving CTA per pair that also filters by rotation and scatters into
// the reference's output (what k_track_finish does for the in-order path).
__global__ void __launch_bounds__(256) k_track_walk(const TrackBatch T) {
    const int pair = blockIdx.y, cur = pair + 1;
    SearchParams P;
    track_pair_params(T, pair, P);
    unsigned long long nEval = 0;
    rounds_walk_cta(P, blockIdx.x * (blockDim.x >> 5), T.lists + (size_t)cur * T.stride, T.listCounts + (size_t)cur * T.stride, nEval);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) nEval += __shfl_xor_sync(0xffffffffu, nEval, o);
    if ((threadIdx.x & 31) == 0 && nEval) atomicAdd(P.evals, nEval);
}

__global__ void __launch_bounds__(kRoundThreads) k_track_resolve(const TrackBatch T) {
    extern __shared__ __align__(16) uint8_t dyn[];
    __shared__ int cnt;
    const int pair = blockIdx.x, cur = pair + 1;
    SearchParams P;
    track_pair_params(T, pair, P);
    if (threadIdx.x == 0) cnt = 0;
    if (P.nq > 0) rounds_resolve_cta(P, dyn, T.lists + (size_t)cur * T.stride, T.listCounts + (size_t)cur * T.stride);
    if (T.checkOri) rot_filter_cta(P.nq, P.qBin, P.qResult);
    __syncthreads();
    int mine = 0;
    for (int q = threadIdx.x; q < P.nq; q += blockDim.x) {
        const int r = P.qResult[q];
        if (r >= 0) { T.kpMatch[(size_t)cur * T.stride + r] = q; ++mine; }
    }
    if (mine) atomicAdd(&cnt, mine);
    __syncthreads();
    if (threadIdx.x == 0) { T.nMatches[cur] = cnt; if (cur == 1) T.nMatches[0] = 0; }
}

}  // namespace orbm
