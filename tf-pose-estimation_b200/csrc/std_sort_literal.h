// std_sort_literal.h -- libstdc++'s std::sort restated for the assembly kernel's tie path; compiled for the device
// (k3_assemble.cu) and for the host (tests/native/std_sort_harness.cpp runs it against the real std::sort).
#pragma once
#if defined(__CUDACC__)
#define SS_HD __host__ __device__
#else
#define SS_HD
#endif

namespace tfpd {

// ------------------------------------------------------------------------------------------
// std::sort, literally (libstdc++: bits/stl_algo.h __sort / __introsort_loop / __unguarded_partition_pivot /
// __move_median_to_first / __final_insertion_sort, bits/stl_heap.h for the depth-limit fallback).
// pafprocess.cpp:97 sorts a limb's candidates with std::sort and comp_candidate (a.score > b.score, :239-241).
// std::sort is not stable, so where two candidates have bit-equal scores their order -- and with it which of them the
// greedy step accepts -- is whatever this algorithm makes of the order the candidates were pushed in (peak_a outer,
// peak_b inner, :57-60).  Run by one lane on a limb segment that holds such a tie, it reproduces the reference's own
// order; segments without a tie have only one possible order and never come here.
// The key stores ~score in bits 24..55 (scores are > 0, so float order == bit order): a.score > b.score is
// (a >> 24) < (b >> 24) for two keys of one limb.
// ------------------------------------------------------------------------------------------
typedef unsigned long long CandKey;
SS_HD inline bool ss_comp(CandKey a, CandKey b) { return (a >> 24) < (b >> 24); }

SS_HD inline void ss_push_heap(CandKey *first, int hole, int top, CandKey value) {
    int parent = (hole - 1) / 2;
    while (hole > top && ss_comp(first[parent], value)) {
        first[hole] = first[parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    first[hole] = value;
}
SS_HD inline void ss_adjust_heap(CandKey *first, int hole, int len, CandKey value) {
    const int top = hole;
    int child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (ss_comp(first[child], first[child - 1])) --child;
        first[hole] = first[child];
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
        child = 2 * (child + 1);
        first[hole] = first[child - 1];
        hole = child - 1;
    }
    ss_push_heap(first, hole, top, value);
}
// __partial_sort(first, last, last): __heap_select (= __make_heap, its loop over [middle, last) is empty) + __sort_heap
#ifdef SS_TEST_HOOK
static int ss_heap_sort_calls = 0;          // host test harness only: how often the depth-limit fallback ran
#endif
SS_HD inline void ss_heap_sort(CandKey *first, int len) {
#ifdef SS_TEST_HOOK
    ++ss_heap_sort_calls;
#endif
    if (len >= 2) {
        int parent = (len - 2) / 2;
        for (;;) {
            const CandKey value = first[parent];
            ss_adjust_heap(first, parent, len, value);
            if (parent == 0) break;
            --parent;
        }
    }
    int last = len;
    while (last > 1) {
        --last;
        const CandKey value = first[last];          // __pop_heap(first, last, last)
        first[last] = first[0];
        ss_adjust_heap(first, 0, last, value);
    }
}
SS_HD inline void ss_swap(CandKey *a, int i, int j) {
    const CandKey t = a[i];
    a[i] = a[j];
    a[j] = t;
}
SS_HD inline void ss_move_median_to_first(CandKey *a, int result, int ia, int ib, int ic) {
    if (ss_comp(a[ia], a[ib])) {
        if (ss_comp(a[ib], a[ic])) ss_swap(a, result, ib);
        else if (ss_comp(a[ia], a[ic])) ss_swap(a, result, ic);
        else ss_swap(a, result, ia);
    } else if (ss_comp(a[ia], a[ic])) ss_swap(a, result, ia);
    else if (ss_comp(a[ib], a[ic])) ss_swap(a, result, ic);
    else ss_swap(a, result, ib);
}
SS_HD inline int ss_unguarded_partition(CandKey *a, int first, int last, int pivot) {
    for (;;) {
        while (ss_comp(a[first], a[pivot])) ++first;
        --last;
        while (ss_comp(a[pivot], a[last])) --last;
        if (!(first < last)) return first;
        ss_swap(a, first, last);
        ++first;
    }
}
SS_HD inline void ss_unguarded_linear_insert(CandKey *a, int last) {
    const CandKey val = a[last];
    int next = last - 1;
    while (ss_comp(val, a[next])) {
        a[last] = a[next];
        last = next;
        --next;
    }
    a[last] = val;
}
SS_HD inline void ss_insertion_sort(CandKey *a, int first, int last) {
    if (first == last) return;
    for (int i = first + 1; i != last; ++i) {
        if (ss_comp(a[i], a[first])) {
            const CandKey val = a[i];
            for (int k = i; k > first; --k) a[k] = a[k - 1];      // move_backward(first, i, i + 1)
            a[first] = val;
        } else {
            ss_unguarded_linear_insert(a, i);
        }
    }
}
SS_HD inline void std_sort_literal(CandKey *a, int n) {
    if (n <= 0) return;
    constexpr int kThreshold = 16;                  // _S_threshold
    int st_first[64], st_last[64], st_depth[64];    // __introsort_loop recurses on the right part, loops on the left:
    int sp = 0;                                     // the sub-ranges are disjoint, so the order they are taken in is free
    st_first[0] = 0;
    st_last[0] = n;
    int lg = 0;
    while ((n >> (lg + 1)) != 0) ++lg;
    st_depth[0] = 2 * lg;                           // std::__lg(last - first) * 2
    sp = 1;
    while (sp > 0) {
        --sp;
        int first = st_first[sp], last = st_last[sp], depth = st_depth[sp];
        while (last - first > kThreshold) {
            if (depth == 0) {
                ss_heap_sort(a + first, last - first);
                break;
            }
            --depth;
            const int mid = first + (last - first) / 2;             // __unguarded_partition_pivot
            ss_move_median_to_first(a, first, first + 1, mid, last - 1);
            const int cut = ss_unguarded_partition(a, first + 1, last, first);
            if (sp < 64) {
                st_first[sp] = cut;
                st_last[sp] = last;
                st_depth[sp] = depth;
                ++sp;
            }
            last = cut;
        }
    }
    if (n > kThreshold) {                           // __final_insertion_sort
        ss_insertion_sort(a, 0, kThreshold);
        for (int i = kThreshold; i != n; ++i) ss_unguarded_linear_insert(a, i);
    } else {
        ss_insertion_sort(a, 0, n);
    }
}

}  // namespace tfpd
