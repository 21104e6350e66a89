// Exact emulation of libstdc++'s std::sort (introsort: median-of-3 quicksort to a 16-element threshold with a
// 2*floor(log2 n) depth limit and heap-sort fallback, then insertion sort) on an array of (key, id) pairs ordered by
// `key` alone.  The reference's DistributeOctTree sorts (size, node*) pairs with compareNodes
// (src/ORBextractor.cc:538-553,712): that predicate is a strict order on (size, UL.x) only, std::sort is unstable,
// so WHICH of two equal nodes is split first — and therefore which keypoints survive — is decided by the sort's
// internal element movement.  Bit-exact keypoint parity needs the same movement; this header restates the published
// algorithm (GCC libstdc++ bits/stl_algo.h, bits/stl_heap.h) over index arithmetic so that it runs in one GPU thread
// on shared memory.  tests/test_sort_emulation.py checks it against the real std::sort / std::partial_sort.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define ISE_HD __host__ __device__ __forceinline__
#else
#define ISE_HD inline
#endif

namespace ise {

struct Elem {
    uint64_t key;   // (size << 32) | UL.x  — compareNodes' lexicographic order
    int id;
    int pad;
};

ISE_HD bool less(const Elem& a, const Elem& b) { return a.key < b.key; }
ISE_HD void swap_elem(Elem* a, int i, int j) { Elem t = a[i]; a[i] = a[j]; a[j] = t; }

// std::__adjust_heap followed by its inlined std::__push_heap
ISE_HD void adjust_heap(Elem* a, int first, int hole, int len, Elem value) {
    const int top = hole;
    int child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (less(a[first + child], a[first + child - 1])) --child;
        a[first + hole] = a[first + child];
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
        child = 2 * (child + 1);
        a[first + hole] = a[first + child - 1];
        hole = child - 1;
    }
    int parent = (hole - 1) / 2;
    while (hole > top && less(a[first + parent], value)) {
        a[first + hole] = a[first + parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    a[first + hole] = value;
}

// std::__partial_sort(first, last, last): make_heap + sort_heap
ISE_HD void heap_sort(Elem* a, int first, int last) {
    const int len = last - first;
    if (len >= 2) {
        int parent = (len - 2) / 2;
        while (true) {
            Elem v = a[first + parent];
            adjust_heap(a, first, parent, len, v);
            if (parent == 0) break;
            --parent;
        }
    }
    int end = last;
    while (end - first > 1) {
        --end;
        Elem v = a[end];
        a[end] = a[first];
        adjust_heap(a, first, 0, end - first, v);
    }
}

// std::__move_median_to_first(result, a, b, c)
ISE_HD void median_to_first(Elem* e, int result, int a, int b, int c) {
    if (less(e[a], e[b])) {
        if (less(e[b], e[c])) swap_elem(e, result, b);
        else if (less(e[a], e[c])) swap_elem(e, result, c);
        else swap_elem(e, result, a);
    } else if (less(e[a], e[c])) swap_elem(e, result, a);
    else if (less(e[b], e[c])) swap_elem(e, result, c);
    else swap_elem(e, result, b);
}

// std::__unguarded_partition(first, last, pivot)
ISE_HD int unguarded_partition(Elem* e, int first, int last, int pivot) {
    while (true) {
        while (less(e[first], e[pivot])) ++first;
        --last;
        while (less(e[pivot], e[last])) --last;
        if (!(first < last)) return first;
        swap_elem(e, first, last);
        ++first;
    }
}

ISE_HD void unguarded_linear_insert(Elem* e, int last) {
    Elem v = e[last];
    int next = last - 1;
    while (less(v, e[next])) {
        e[last] = e[next];
        last = next;
        --next;
    }
    e[last] = v;
}

ISE_HD void insertion_sort(Elem* e, int first, int last) {
    if (first == last) return;
    for (int i = first + 1; i != last; ++i) {
        if (less(e[i], e[first])) {
            Elem v = e[i];
            for (int j = i; j > first; --j) e[j] = e[j - 1];
            e[first] = v;
        } else {
            unguarded_linear_insert(e, i);
        }
    }
}

// std::sort(e, e+n).  The recursion of __introsort_loop (right part recursed, left part looped) is replaced by an
// explicit stack: sub-ranges are disjoint, so the order in which they are processed does not change the result.
ISE_HD void sort(Elem* e, int n) {
    if (n <= 1) return;
    int lg = 0;
    for (int t = n; t > 1; t >>= 1) ++lg;
    int stF[64], stL[64], stD[64];
    int sp = 0;
    stF[0] = 0; stL[0] = n; stD[0] = 2 * lg; sp = 1;
    while (sp > 0) {
        --sp;
        int first = stF[sp], last = stL[sp], depth = stD[sp];
        while (last - first > 16) {
            if (depth == 0) { heap_sort(e, first, last); break; }
            --depth;
            int mid = first + (last - first) / 2;
            median_to_first(e, first, first + 1, mid, last - 1);
            int cut = unguarded_partition(e, first + 1, last, first);
            stF[sp] = cut; stL[sp] = last; stD[sp] = depth; ++sp;
            last = cut;
        }
    }
    if (n > 16) {
        insertion_sort(e, 0, 16);
        for (int i = 16; i < n; ++i) unguarded_linear_insert(e, i);
    } else {
        insertion_sort(e, 0, n);
    }
}

// ---- the same sort as independent work items (what the GPU runs on the lanes of one warp) ------------------------------
// The sub-ranges produced by the introsort loop are disjoint, so they can be partitioned by different workers in any order.
// The final insertion pass never moves an element across a range boundary either: a partition leaves  left <= pivot <=
// right, and insertion stops at the first element that is not greater (strict '<'), so it is the same as a stable
// insertion sort of every leaf range (<= 16 elements, or heap-sorted) on its own.  The decomposition below is checked
// against the real std::sort in tests/test_abi_and_host.py (sort_rounds) and on the device (orbx_debug_sort_nodes mode 2).

// one introsort step on [first, last) (more than 16 elements, depth > 0): returns the cut
ISE_HD int split_step(Elem* e, int first, int last) {
    const int mid = first + (last - first) / 2;
    median_to_first(e, first, first + 1, mid, last - 1);
    return unguarded_partition(e, first + 1, last, first);
}

#ifndef __CUDACC__
// host form of the round-based schedule (ranges of one round in arbitrary order)
inline void sort_rounds(Elem* e, int n, bool reverse_order) {
    if (n <= 1) return;
    int lg = 0;
    for (int t = n; t > 1; t >>= 1) ++lg;
    struct R { int first, last, depth; };
    R* cur = new R[n + 1]; R* nxt = new R[n + 1]; R* leaf = new R[n + 1];
    int nc = 0, nl = 0;
    if (n > 16) cur[nc++] = R{0, n, 2 * lg}; else leaf[nl++] = R{0, n, 0};
    while (nc > 0) {
        int nn = 0;
        for (int k = 0; k < nc; ++k) {
            const R r = cur[reverse_order ? nc - 1 - k : k];
            if (r.depth == 0) { heap_sort(e, r.first, r.last); continue; }
            const int cut = split_step(e, r.first, r.last);
            const R ch[2] = {R{r.first, cut, r.depth - 1}, R{cut, r.last, r.depth - 1}};
            for (const R& c : ch) {
                if (c.last - c.first > 16) nxt[nn++] = c;
                else if (c.last - c.first > 1) leaf[nl++] = c;
            }
        }
        R* t = cur; cur = nxt; nxt = t; nc = nn;
    }
    for (int k = 0; k < nl; ++k) { const R r = leaf[reverse_order ? nl - 1 - k : k]; insertion_sort(e, r.first, r.last); }
    delete[] cur; delete[] nxt; delete[] leaf;
}
#endif

}  // namespace ise
