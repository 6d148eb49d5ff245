// tie_sort.h — libstdc++'s std::sort algorithm on one row of (id, distance) pairs, comparator "distance <" (see
// vamana_build.cu: rows whose distances tie must end up where the reference's std::sort leaves them).  Host + device: the
// kernel runs it per row, tests/test_tie_sort_cpu.py compiles the same code with g++ and checks it against std::sort.
#pragma once
#include <stdint.h>
#ifdef __CUDACC__
#define PANN_HD __host__ __device__
#else
#define PANN_HD
#endif
#define PANN_TIE_MAXDEG 128
struct TieRow {
  uint32_t id[PANN_TIE_MAXDEG];
  float d[PANN_TIE_MAXDEG];
  PANN_HD inline bool lt(int a, int b) const { return d[a] < d[b]; }
  PANN_HD inline void swp(int a, int b) { const uint32_t ti = id[a]; id[a] = id[b]; id[b] = ti; const float td = d[a]; d[a] = d[b]; d[b] = td; }
  // the heap-sort fallback, on [f, l): std::__partial_sort(first, last, last) = __make_heap + __sort_heap (bits/stl_heap.h)
  PANN_HD void push_heap(int f, int hole, int top, uint32_t vi, float vd) {
    int parent = (hole - 1) / 2;
    while (hole > top && d[f + parent] < vd) { id[f + hole] = id[f + parent]; d[f + hole] = d[f + parent]; hole = parent; parent = (hole - 1) / 2; }
    id[f + hole] = vi; d[f + hole] = vd;
  }
  PANN_HD void adjust_heap(int f, int hole, int len, uint32_t vi, float vd) {
    const int top = hole;
    int child = hole;
    while (child < (len - 1) / 2) {
      child = 2 * (child + 1);
      if (d[f + child] < d[f + child - 1]) child--;
      id[f + hole] = id[f + child]; d[f + hole] = d[f + child]; hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
      child = 2 * (child + 1);
      id[f + hole] = id[f + child - 1]; d[f + hole] = d[f + child - 1]; hole = child - 1;
    }
    push_heap(f, hole, top, vi, vd);
  }
  PANN_HD void heap_sort(int f, int l) {
    const int len = l - f;
    if (len >= 2) { int parent = (len - 2) / 2; while (true) { adjust_heap(f, parent, len, id[f + parent], d[f + parent]); if (parent == 0) break; parent--; } }
    while (l - f > 1) { --l; const uint32_t vi = id[l]; const float vd = d[l]; id[l] = id[f]; d[l] = d[f]; adjust_heap(f, 0, l - f, vi, vd); }
  }
};
// false: the explicit stack would overflow (cannot happen: the depth limit bounds it) — the caller sorts the row elsewhere
PANN_HD inline bool pann_tie_sort_row(TieRow& r, int n) {
  // __introsort_loop(first, last, 2 * lg(n)); the recursion on the right part becomes a stack (the parts are disjoint)
  int stk_f[16], stk_l[16], stk_d[16], sp = 0;
  int depth0 = 0; for (int m = n; m > 1; m >>= 1) depth0++;
  if (n > 1) { stk_f[0] = 0; stk_l[0] = n; stk_d[0] = 2 * depth0; sp = 1; }
  while (sp > 0) {
    sp--;
    int first = stk_f[sp], last = stk_l[sp], depth = stk_d[sp];
    while (last - first > 16) {
      if (sp >= 15) return false;   // cannot happen (the depth limit bounds the stack at 2 lg n <= 14)
      if (depth == 0) { r.heap_sort(first, last); break; }   // the depth limit ran out (pruned rows arrive nearly sorted: this is not rare)
      --depth;
      const int mid = first + (last - first) / 2;
      {  // __move_median_to_first(first, first + 1, mid, last - 1)
        const int a = first + 1, b = mid, c = last - 1;
        if (r.lt(a, b)) {
          if (r.lt(b, c)) r.swp(first, b);
          else if (r.lt(a, c)) r.swp(first, c);
          else r.swp(first, a);
        } else if (r.lt(a, c)) r.swp(first, a);
        else if (r.lt(b, c)) r.swp(first, c);
        else r.swp(first, b);
      }
      int lo = first + 1, hi = last;   // __unguarded_partition(first + 1, last, pivot = first)
      while (true) {
        while (r.lt(lo, first)) ++lo;
        --hi;
        while (r.lt(first, hi)) --hi;
        if (!(lo < hi)) break;
        r.swp(lo, hi);
        ++lo;
      }
      stk_f[sp] = lo; stk_l[sp] = last; stk_d[sp] = depth; sp++;   // __introsort_loop(cut, last, depth_limit)
      last = lo;
    }
  }
  // __final_insertion_sort: __insertion_sort over the first 16, __unguarded_insertion_sort over the rest
  const int head = n > 16 ? 16 : n;
  for (int i = 1; i < head; i++) {
    const uint32_t vi = r.id[i]; const float vd = r.d[i];
    if (vd < r.d[0]) {
      for (int j = i; j > 0; j--) { r.id[j] = r.id[j - 1]; r.d[j] = r.d[j - 1]; }
      r.id[0] = vi; r.d[0] = vd;
    } else {
      int j = i;
      while (vd < r.d[j - 1]) { r.id[j] = r.id[j - 1]; r.d[j] = r.d[j - 1]; j--; }
      r.id[j] = vi; r.d[j] = vd;
    }
  }
  for (int i = head; i < n; i++) {
    const uint32_t vi = r.id[i]; const float vd = r.d[i];
    int j = i;
    while (vd < r.d[j - 1]) { r.id[j] = r.id[j - 1]; r.d[j] = r.d[j - 1]; j--; }
    r.id[j] = vi; r.d[j] = vd;
  }
  return true;
}
