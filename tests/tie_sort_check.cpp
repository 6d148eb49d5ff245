// test infrastructure: parlayann_b200/csrc/tie_sort.h (the code pann_tie_sort_kernel runs per row) against this machine's
// std::sort / std::partial_sort on random, sorted, nearly sorted, organ-pipe and pair-swapped rows with ties
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>
#include "tie_sort.h"
int main(int argc, char** argv) {
  const int iters = argc > 1 ? atoi(argv[1]) : 300000;
  std::mt19937 g(1);
  long bad = 0, badheap = 0, guard = 0;
  auto cmp = [](const std::pair<uint32_t, float>& a, const std::pair<uint32_t, float>& b) { return a.second < b.second; };
  for (int it = 0; it < iters; it++) {
    const int n = 1 + g() % PANN_TIE_MAXDEG;
    const int range = (it % 3 == 0) ? 8 : (it % 3 == 1 ? 1000 : 100000);
    TieRow r, h;
    std::vector<std::pair<uint32_t, float>> v(n);
    for (int j = 0; j < n; j++) { r.id[j] = j; r.d[j] = (float)(g() % range); }
    switch (it % 5) {
      case 1: std::sort(r.d, r.d + n); break;
      case 2: std::sort(r.d, r.d + n * 2 / 3); break;                                  // a pruned row + appended reverse edges
      case 3: std::sort(r.d, r.d + n); std::reverse(r.d, r.d + n / 2); break;
      case 4: std::sort(r.d, r.d + n); for (int j = 0; j + 1 < n; j += 2) std::swap(r.d[j], r.d[j + 1]); break;
    }
    for (int j = 0; j < n; j++) v[j] = {r.id[j], r.d[j]};
    h = r;
    auto vh = v;
    std::sort(v.begin(), v.end(), cmp);
    if (!pann_tie_sort_row(r, n)) { guard++; continue; }
    for (int j = 0; j < n; j++) if (r.id[j] != v[j].first) { bad++; break; }
    std::partial_sort(vh.begin(), vh.end(), vh.end(), cmp);      // = the heap-sort fallback of std::sort
    h.heap_sort(0, n);
    for (int j = 0; j < n; j++) if (h.id[j] != vh[j].first) { badheap++; break; }
  }
  printf("rows %d differ %ld heap_differ %ld guard %ld\n", iters, bad, badheap, guard);
  return (bad || badheap || guard) ? 1 : 0;
}
