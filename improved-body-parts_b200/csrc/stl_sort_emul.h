// Emulation of libstdc++'s std::sort (bits/stl_algo.h: __introsort_loop + __final_insertion_sort, threshold 16,
// heap-sort fallback) for the one call the reference makes with a NON-STRICT comparator:
//     sort(candidates.begin(), candidates.end(), comp_candidate)   with   comp(a,b) = a.overall_score >= b.overall_score
// (/root/reference utils/pafprocess/pafprocess.cpp:109, 333-335).  With duplicate keys the resulting order is whatever
// this particular algorithm does, and the greedy connection step consumes candidates in that order, so a drop-in
// replacement has to walk the same steps.  Elements are 64-bit words: high 32 bits = IEEE-754 bits of the key, low
// 32 bits = payload; only the key takes part in comparisons.
//
// Where the real std::sort would step outside [first,last) because the comparator is not a strict weak ordering
// (a duplicate of the maximum chosen as pivot, or an unguarded insertion running past the front), the emulation stops
// the scan at the boundary and sets *undefined = true: the reference's result is undefined there.
//
// Host+device: the same code is unit-tested on the CPU against the real std::sort (tests/test_sort_emul.py) and runs
// in the limb kernel on one thread for the rare (n > 16 && duplicate keys) limbs.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define IBP_HD __host__ __device__ __forceinline__
#else
#define IBP_HD inline
#endif

namespace ibp_sortemu {

struct Ctx {
    uint64_t *a;   // array base
    int n;         // number of elements
    bool undefined;
};

IBP_HD float key_of(uint64_t v) {
    union { uint32_t u; float f; } c;
    c.u = (uint32_t)(v >> 32);
    return c.f;
}
// comp(x, y) of the reference
IBP_HD bool ge(uint64_t x, uint64_t y) { return key_of(x) >= key_of(y); }

IBP_HD void swp(uint64_t *a, int i, int j) {
    uint64_t t = a[i];
    a[i] = a[j];
    a[j] = t;
}

// __move_median_to_first(result, a, b, c)
IBP_HD void median_to_first(uint64_t *v, int result, int a, int b, int c) {
    if (ge(v[a], v[b])) {
        if (ge(v[b], v[c])) swp(v, result, b);
        else if (ge(v[a], v[c])) swp(v, result, c);
        else swp(v, result, a);
    } else if (ge(v[a], v[c])) swp(v, result, a);
    else if (ge(v[b], v[c])) swp(v, result, c);
    else swp(v, result, b);
}

// __unguarded_partition(first, last, pivot)
IBP_HD int unguarded_partition(Ctx &c, int first, int last, int pivot) {
    uint64_t *v = c.a;
    while (true) {
        while (true) {
            if (first >= c.n) { c.undefined = true; break; }
            if (!ge(v[first], v[pivot])) break;
            ++first;
        }
        --last;
        while (true) {
            if (last < 0) { c.undefined = true; break; }
            if (!ge(v[pivot], v[last])) break;
            --last;
        }
        if (!(first < last)) return first;
        swp(v, first, last);
        ++first;
    }
}

// __adjust_heap(first, holeIndex, len, value) followed by __push_heap, offsets relative to `base`
IBP_HD void adjust_heap(uint64_t *v, int base, int hole, int len, uint64_t value) {
    const int top = hole;
    int child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (ge(v[base + child], v[base + child - 1])) child--;
        v[base + hole] = v[base + child];
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
        child = 2 * (child + 1);
        v[base + hole] = v[base + child - 1];
        hole = child - 1;
    }
    // __push_heap(first, hole, top, value)
    int parent = (hole - 1) / 2;
    while (hole > top && ge(v[base + parent], value)) {
        v[base + hole] = v[base + parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    v[base + hole] = value;
}

// __partial_sort(first, last, last) == __heap_select(first,last,last) + __sort_heap(first,last)
IBP_HD void heap_sort(uint64_t *v, int first, int last) {
    const int len = last - first;
    if (len >= 2) {  // __make_heap
        int parent = (len - 2) / 2;
        while (true) {
            uint64_t value = v[first + parent];
            adjust_heap(v, first, parent, len, value);
            if (parent == 0) break;
            parent--;
        }
    }
    // __heap_select's loop over [middle,last) is empty because middle == last
    int end = last;
    while (end - first > 1) {  // __sort_heap: __pop_heap(first, end-1, end-1)
        --end;
        uint64_t value = v[end];
        v[end] = v[first];
        adjust_heap(v, first, 0, end - first, value);
    }
}

// __unguarded_linear_insert(last)
IBP_HD void unguarded_linear_insert(Ctx &c, int last) {
    uint64_t *v = c.a;
    uint64_t val = v[last];
    int next = last - 1;
    while (true) {
        if (next < 0) { c.undefined = true; break; }
        if (!ge(val, v[next])) break;
        v[last] = v[next];
        last = next;
        --next;
    }
    v[last] = val;
}

// __insertion_sort(first, last)
IBP_HD void insertion_sort(Ctx &c, int first, int last) {
    uint64_t *v = c.a;
    if (first == last) return;
    for (int i = first + 1; i != last; ++i) {
        if (ge(v[i], v[first])) {
            uint64_t val = v[i];
            for (int k = i; k > first; --k) v[k] = v[k - 1];  // move_backward(first, i, i + 1)
            v[first] = val;
        } else {
            unguarded_linear_insert(c, i);
        }
    }
}

// std::sort(a, a + n, comp).  Returns true when the real std::sort would have stayed inside the array.
IBP_HD bool sort_like_libstdcxx(uint64_t *a, int n) {
    Ctx c;
    c.a = a;
    c.n = n;
    c.undefined = false;
    if (n <= 1) return true;
    // __introsort_loop, recursion on the right part turned into an explicit stack (depth <= 2*lg n + 1)
    int depth0 = 0;
    for (int t = n; t > 1; t >>= 1) depth0++;  // std::__lg(n)
    depth0 *= 2;
    int stack_first[64], stack_last[64], stack_depth[64];
    int sp = 0;
    stack_first[0] = 0;
    stack_last[0] = n;
    stack_depth[0] = depth0;
    sp = 1;
    while (sp > 0) {
        --sp;
        int first = stack_first[sp], last = stack_last[sp], depth = stack_depth[sp];
        // body of __introsort_loop(first, last, depth)
        while (last - first > 16) {
            if (depth == 0) {
                heap_sort(a, first, last);
                break;
            }
            --depth;
            int mid = first + (last - first) / 2;
            median_to_first(a, first, first + 1, mid, last - 1);
            int cut = unguarded_partition(c, first + 1, last, first);
            // the recursive call on [cut, last) runs to completion BEFORE the loop continues on [first, cut):
            // push the continuation of the left part below the right part on the stack.
            if (sp + 2 > 64) { c.undefined = true; break; }
            stack_first[sp] = first;
            stack_last[sp] = cut;
            stack_depth[sp] = depth;
            ++sp;
            stack_first[sp] = cut;
            stack_last[sp] = last;
            stack_depth[sp] = depth;
            ++sp;
            first = last = 0;  // this frame is finished; both halves are on the stack
        }
    }
    // __final_insertion_sort
    if (n > 16) {
        insertion_sort(c, 0, 16);
        for (int i = 16; i != n; ++i) unguarded_linear_insert(c, i);
    } else {
        insertion_sort(c, 0, n);
    }
    return !c.undefined;
}

}  // namespace ibp_sortemu
