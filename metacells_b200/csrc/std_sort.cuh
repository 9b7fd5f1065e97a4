// libstdc++ 13's std::nth_element and std::sort run by ONE thread, step for step (bits/stl_algo.h, bits/stl_heap.h): the
// order they leave equal elements in is part of the results of choose_seeds.cpp (:87-124, :330-337) and gaps.cpp (:88-101),
// so it is reproduced rather than replaced by a parallel sort.
#pragma once

namespace mcb {

// ---- libstdc++ 13 selection / sorting, one thread (bits/stl_algo.h, bits/stl_heap.h) ------------------

template <typename T, typename Less>
__device__ void median_to_first(T* a, int result, int x, int y, int z, Less less) {
    int chosen;
    if (less(a[x], a[y]))
        chosen = less(a[y], a[z]) ? y : (less(a[x], a[z]) ? z : x);
    else
        chosen = less(a[x], a[z]) ? x : (less(a[y], a[z]) ? z : y);
    const T tmp = a[result];
    a[result] = a[chosen];
    a[chosen] = tmp;
}

template <typename T, typename Less>
__device__ int hoare_partition(T* a, int first, int last, int pivot, Less less) {
    const T pivot_value = a[pivot];
    for (;;) {
        while (less(a[first], pivot_value)) ++first;
        --last;
        while (less(pivot_value, a[last])) --last;
        if (!(first < last)) return first;
        const T tmp = a[first];
        a[first] = a[last];
        a[last] = tmp;
        ++first;
    }
}

template <typename T, typename Less>
__device__ void sift_up(T* a, int first, int hole, int top, T value, Less less) {
    int parent = (hole - 1) / 2;
    while (hole > top && less(a[first + parent], value)) {
        a[first + hole] = a[first + parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    a[first + hole] = value;
}

template <typename T, typename Less>
__device__ void sift_down(T* a, int first, int hole, int length, T value, Less less) {
    const int top = hole;
    int child = hole;
    while (child < (length - 1) / 2) {
        child = 2 * (child + 1);
        if (less(a[first + child], a[first + child - 1])) child--;
        a[first + hole] = a[first + child];
        hole = child;
    }
    if ((length & 1) == 0 && child == (length - 2) / 2) {
        child = 2 * (child + 1);
        a[first + hole] = a[first + child - 1];
        hole = child - 1;
    }
    sift_up(a, first, hole, top, value, less);
}

template <typename T, typename Less>
__device__ void pop_to(T* a, int first, int last, int result, Less less) {
    const T value = a[result];
    a[result] = a[first];
    sift_down(a, first, 0, last - first, value, less);
}

template <typename T, typename Less>
__device__ void heap_select(T* a, int first, int middle, int last, Less less) {
    const int length = middle - first;
    if (length >= 2)
        for (int parent = (length - 2) / 2;; --parent) {
            sift_down(a, first, parent, length, a[first + parent], less);
            if (parent == 0) break;
        }
    for (int i = middle; i < last; ++i)
        if (less(a[i], a[first])) pop_to(a, first, middle, i, less);
}

template <typename T, typename Less>
__device__ void linear_insert(T* a, int last, Less less) {
    const T value = a[last];
    int next = last - 1;
    while (less(value, a[next])) {
        a[last] = a[next];
        last = next;
        --next;
    }
    a[last] = value;
}

template <typename T, typename Less>
__device__ void insertion_sort(T* a, int first, int last, Less less) {
    if (first == last) return;
    for (int i = first + 1; i != last; ++i) {
        if (less(a[i], a[first])) {
            const T value = a[i];
            for (int j = i; j > first; --j) a[j] = a[j - 1];
            a[first] = value;
        } else {
            linear_insert(a, i, less);
        }
    }
}

__device__ __forceinline__ int floor_lg(int n) { return 31 - __clz(n); }

// __introselect from a given depth limit on
template <typename T, typename Less>
__device__ void serial_select(T* a, int first, int nth, int last, int depth_limit, Less less) {
    while (last - first > 3) {
        if (depth_limit == 0) {
            heap_select(a, first, nth + 1, last, less);
            const T tmp = a[first];
            a[first] = a[nth];
            a[nth] = tmp;
            return;
        }
        --depth_limit;
        median_to_first(a, first, first + 1, first + (last - first) / 2, last - 1, less);
        const int cut = hoare_partition(a, first + 1, last, first, less);
        if (cut <= nth)
            first = cut;
        else
            last = cut;
    }
    insertion_sort(a, first, last, less);
}

// std::sort: __introsort_loop (the recursion is a stack here: the two sides of a cut are independent) and
// __final_insertion_sort
template <typename T, typename Less>
__device__ void serial_sort(T* a, int count, Less less) {
    if (count == 0) return;
    int stack_first[72], stack_last[72], stack_depth[72];
    int top = 0;
    stack_first[0] = 0, stack_last[0] = count, stack_depth[0] = floor_lg(count) * 2;
    ++top;
    while (top > 0) {
        --top;
        int first = stack_first[top], last = stack_last[top], depth_limit = stack_depth[top];
        while (last - first > 16) {
            if (depth_limit == 0) {
                heap_select(a, first, last, last, less);
                while (last - first > 1) {
                    --last;
                    pop_to(a, first, last, last, less);
                }
                break;
            }
            --depth_limit;
            median_to_first(a, first, first + 1, first + (last - first) / 2, last - 1, less);
            const int cut = hoare_partition(a, first + 1, last, first, less);
            stack_first[top] = cut, stack_last[top] = last, stack_depth[top] = depth_limit;
            ++top;
            last = cut;
        }
    }
    if (count > 16) {
        insertion_sort(a, 0, 16, less);
        for (int i = 16; i != count; ++i) linear_insert(a, i, less);
    } else {
        insertion_sort(a, 0, count, less);
    }
}

}  // namespace mcb
