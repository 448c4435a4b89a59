// retainBest emulation: a faithful port of the selection algorithm behind std::nth_element in libstdc++
// (introselect: median-of-3 to first, unguarded Hoare partition, depth limit 2*floor(log2 n) -> heap-select,
// final insertion sort on <=3 elements), because which tied-score keypoints survive
// cv::KeyPointsFilter::retainBest (called at reference src/ORBextractor.cc:734,750) -- and in which order --
// is defined by that exact permutation.  Keys are packed u32 (score in the top 8 bits); comp(a,b) = score(a) >
// score(b) ("response greater").  Usable from host (unit test against std::nth_element) and device.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define GF_HD __host__ __device__ __forceinline__
#else
#define GF_HD inline
#endif

namespace gforb {

struct ScoreGreater {
    GF_HD bool operator()(uint32_t a, uint32_t b) const { return (a >> 24) > (b >> 24); }
};

template <class T> GF_HD void gf_swap(T& a, T& b) { T t = a; a = b; b = t; }

template <class T, class C>
GF_HD void is_move_median_to_first(T* result, T* a, T* b, T* c, C comp) {
    if (comp(*a, *b)) {
        if (comp(*b, *c)) gf_swap(*result, *b);
        else if (comp(*a, *c)) gf_swap(*result, *c);
        else gf_swap(*result, *a);
    } else if (comp(*a, *c)) gf_swap(*result, *a);
    else if (comp(*b, *c)) gf_swap(*result, *c);
    else gf_swap(*result, *b);
}

template <class T, class C>
GF_HD T* is_unguarded_partition(T* first, T* last, T* pivot, C comp) {
    while (true) {
        while (comp(*first, *pivot)) ++first;
        --last;
        while (comp(*pivot, *last)) --last;
        if (!(first < last)) return first;
        gf_swap(*first, *last);
        ++first;
    }
}

template <class T, class C>
GF_HD void is_push_heap(T* first, int holeIndex, int topIndex, T value, C comp) {
    int parent = (holeIndex - 1) / 2;
    while (holeIndex > topIndex && comp(first[parent], value)) {
        first[holeIndex] = first[parent];
        holeIndex = parent;
        parent = (holeIndex - 1) / 2;
    }
    first[holeIndex] = value;
}

template <class T, class C>
GF_HD void is_adjust_heap(T* first, int holeIndex, int len, T value, C comp) {
    const int topIndex = holeIndex;
    int secondChild = holeIndex;
    while (secondChild < (len - 1) / 2) {
        secondChild = 2 * (secondChild + 1);
        if (comp(first[secondChild], first[secondChild - 1])) secondChild--;
        first[holeIndex] = first[secondChild];
        holeIndex = secondChild;
    }
    if ((len & 1) == 0 && secondChild == (len - 2) / 2) {
        secondChild = 2 * (secondChild + 1);
        first[holeIndex] = first[secondChild - 1];
        holeIndex = secondChild - 1;
    }
    is_push_heap(first, holeIndex, topIndex, value, comp);
}

template <class T, class C>
GF_HD void is_heap_select(T* first, T* middle, T* last, C comp) {
    const int len = (int)(middle - first);
    if (len >= 2) {                                   // __make_heap
        int parent = (len - 2) / 2;
        while (true) {
            T value = first[parent];
            is_adjust_heap(first, parent, len, value, comp);
            if (parent == 0) break;
            parent--;
        }
    }
    for (T* i = middle; i < last; ++i)
        if (comp(*i, *first)) {                       // __pop_heap(first, middle, i)
            T value = *i;
            *i = *first;
            is_adjust_heap(first, 0, len, value, comp);
        }
}

template <class T, class C>
GF_HD void is_insertion_sort(T* first, T* last, C comp) {
    if (first == last) return;
    for (T* i = first + 1; i != last; ++i) {
        if (comp(*i, *first)) {
            T val = *i;
            for (T* p = i; p != first; --p) *p = *(p - 1);
            *first = val;
        } else {
            T val = *i;
            T* lastp = i;
            T* next = i - 1;
            while (comp(val, *next)) { *lastp = *next; lastp = next; --next; }
            *lastp = val;
        }
    }
}

// std::nth_element(first, nth, last, comp)
template <class T, class C>
GF_HD void nth_element_libstdcxx(T* first, T* nth, T* last, C comp) {
    if (first == last || nth == last) return;
    int n = (int)(last - first);
    int lg = 0;
    while ((n >> (lg + 1)) != 0) lg++;                // std::__lg
    int depth_limit = lg * 2;
    while (last - first > 3) {
        if (depth_limit == 0) {
            is_heap_select(first, nth + 1, last, comp);
            gf_swap(*first, *nth);
            return;
        }
        --depth_limit;
        T* mid = first + (last - first) / 2;
        is_move_median_to_first(first, first + 1, mid, last - 1, comp);
        T* cut = is_unguarded_partition(first + 1, last, first, comp);
        if (cut <= nth) first = cut;
        else last = cut;
    }
    is_insertion_sort(first, last, comp);
}

// cv::KeyPointsFilter::retainBest(v, n) followed by the reference's resize(n): the survivors are the first n
// elements of the nth_element permutation (std::partition only touches [n, end)).  variant 0: nth at begin()+n
// (OpenCV 2.4.13 / 3.4.1), variant 1: begin()+n-1 (newer OpenCV).  Returns the new size.
template <class T, class C>
GF_HD int retain_best_first_n(T* v, int size, int n, int variant, C comp) {
    if (n < 0 || size <= n) return size;
    if (n == 0) return 0;
    nth_element_libstdcxx(v, v + (variant ? n - 1 : n), v + size, comp);
    return n;
}

}  // namespace gforb
