// host_util.cpp — host-side pieces of the Vamana build that decide WHAT the device inserts and WHEN:
// the insertion order and the batch schedule.  Both are part of the result (a different order or
// schedule gives a different graph), so they follow the reference exactly.
#include <math.h>
#include <stdint.h>

#include <algorithm>
#include <limits>
#include <random>
#include <thread>
#include <utility>
#include <vector>

// parlay::hash64 (parlay/utilities.h:132-142)
static inline uint64_t hash64(uint64_t u) {
  uint64_t v = u * 3935559000370003845ull + 2691343689449507681ull;
  v ^= v >> 21; v ^= v << 37; v ^= v >> 4;
  v *= 4768777513237032717ull;
  v ^= v << 20; v ^= v >> 41; v ^= v << 5;
  return v;
}

namespace {
struct Rnd {  // parlay::random (parlay/random.h:62-75)
  uint64_t state;
  uint64_t ith(uint64_t i) const { return hash64(i + state); }
  Rnd fork(uint64_t i) const { return Rnd{hash64(hash64(i + state))}; }
};
void knuth_shuffle(int32_t* a, size_t m, Rnd r) {  // parlay/random.h:80-89
  for (size_t i = m; i-- > 1;) std::swap(a[i], a[r.ith(i) % (i + 1)]);
}
}  // namespace

// parlay::random_permutation<int>(n) with the default generator (parlay/random.h:154-159, :91-137):
// below 8192 elements one sequential Knuth shuffle; otherwise a stable counting sort into 2^bits random
// buckets followed by an independent Knuth shuffle of each bucket.
void pann_host_random_permutation(int64_t n, int32_t* out) {
  Rnd r{0};
  if (n < 8192) {
    for (int64_t i = 0; i < n; i++) out[i] = (int32_t)i;
    if (n > 1) knuth_shuffle(out, (size_t)n, r);
    return;
  }
  size_t lg = 0;
  for (uint64_t b = (uint64_t)n - 1; b > 0; b >>= 1) lg++;
  const size_t bits = ((uint64_t)n < (1ull << 27)) ? (lg - 7) / 2 : (lg - 17);
  const size_t nb = (size_t)1 << bits, mask = nb - 1;
  // the stable counting sort, blocked over threads (per-block counts -> offsets, each block scatters its slice in index order:
  // the same permutation as the sequential loop), then the buckets' shuffles side by side.  10M: ~0.12 s on one core.
  unsigned nt = std::thread::hardware_concurrency();
  nt = std::max(1u, std::min(nt, 32u));
  if ((uint64_t)n < (1ull << 20)) nt = 1;
  const int64_t chunk = (n + nt - 1) / nt;
  std::vector<std::vector<size_t>> cnt(nt, std::vector<size_t>(nb, 0));
  auto run = [&](auto&& fn) {
    if (nt == 1) { fn(0u); return; }
    std::vector<std::thread> ts;
    for (unsigned t = 0; t < nt; t++) ts.emplace_back(fn, t);
    for (auto& t : ts) t.join();
  };
  run([&](unsigned t) {
    const int64_t lo = (int64_t)t * chunk, hi = std::min<int64_t>(n, lo + chunk);
    for (int64_t i = lo; i < hi; i++) cnt[t][r.ith((uint64_t)i) & mask]++;
  });
  std::vector<size_t> start(nb + 1, 0);
  for (size_t b = 0; b < nb; b++) {
    size_t acc = start[b];
    for (unsigned t = 0; t < nt; t++) { const size_t c = cnt[t][b]; cnt[t][b] = acc; acc += c; }   // cnt becomes this block's write cursor
    start[b + 1] = acc;
  }
  run([&](unsigned t) {
    const int64_t lo = (int64_t)t * chunk, hi = std::min<int64_t>(n, lo + chunk);
    for (int64_t i = lo; i < hi; i++) out[cnt[t][r.ith((uint64_t)i) & mask]++] = (int32_t)i;
  });
  run([&](unsigned t) {
    for (size_t b = t; b < nb; b += nt)
      if (start[b + 1] - start[b] > 1) knuth_shuffle(out + start[b], start[b + 1] - start[b], r.fork(b));
  });
}

// batch_insert's schedule (algorithms/vamana/index.h:206-209, :223-234) with base = 2, max_fraction = .02:
// prefix-doubling batches [2^i - 1, 2^(i+1) - 1) while 2^i <= max_batch, then fixed strides of max_batch.
int64_t pann_host_batch_schedule_f(uint64_t n, uint64_t m, double max_fraction, int64_t* lo, int64_t* hi, int64_t cap);
int64_t pann_host_batch_schedule(uint64_t n, uint64_t m, int64_t* lo, int64_t* hi, int64_t cap) {
  return pann_host_batch_schedule_f(n, m, 0.02, lo, hi, cap);
}
// the same with batch_insert's max_fraction argument (index.h:191, :206-207: double max_fraction * float n)
int64_t pann_host_batch_schedule_f(uint64_t n, uint64_t m, double max_fraction, int64_t* lo, int64_t* hi, int64_t cap) {
  size_t max_batch = std::min((size_t)(max_fraction * (double)(float)n), (size_t)1000000);
  if (max_batch == 0) max_batch = n;
  size_t inc = 0, count = 0;
  int64_t nb = 0;
  while (count < m) {
    size_t f, c;
    if (pow(2.0, (double)inc) <= (double)max_batch) {
      f = (size_t)pow(2.0, (double)inc) - 1;
      c = std::min((size_t)pow(2.0, (double)(inc + 1)) - 1, (size_t)m);
      count = c;
    } else {
      f = count;
      c = std::min(count + max_batch, (size_t)m);
      count += max_batch;
    }
    if (nb < cap) { lo[nb] = (int64_t)f; hi[nb] = (int64_t)c; }
    nb++; inc++;
  }
  return nb;
}

// build_index's final step (algorithms/vamana/index.h:180-185, graph.h:111-113): std::sort of every adjacency
// list with a comparator on the distance to the owner ONLY.  The distances come from the device (K6); the
// sort runs here with the same std::sort, so equal-distance neighbours end up where the reference leaves them.
#include <thread>
void pann_host_sort_rows(uint32_t* graph, const float* dists, uint64_t n, uint32_t maxdeg) {
  unsigned nt = std::thread::hardware_concurrency();
  if (nt == 0) nt = 1;
  if (nt > 64) nt = 64;
  auto work = [&](uint64_t lo, uint64_t hi) {
    std::vector<std::pair<uint32_t, float>> tmp(maxdeg);
    for (uint64_t i = lo; i < hi; i++) {
      uint32_t* row = graph + i * (maxdeg + 1);
      const float* d = dists + i * maxdeg;
      uint32_t deg = row[0];
      for (uint32_t j = 0; j < deg; j++) tmp[j] = {row[1 + j], d[j]};
      std::sort(tmp.begin(), tmp.begin() + deg, [](const std::pair<uint32_t, float>& a, const std::pair<uint32_t, float>& b) { return a.second < b.second; });
      for (uint32_t j = 0; j < deg; j++) row[1 + j] = tmp[j].first;
    }
  };
  std::vector<std::thread> ts;
  uint64_t chunk = (n + nt - 1) / nt;
  for (unsigned t = 0; t < nt; t++) {
    uint64_t lo = t * chunk, hi = std::min<uint64_t>(n, lo + chunk);
    if (lo < hi) ts.emplace_back(work, lo, hi);
  }
  for (auto& t : ts) t.join();
}

// beamSearchRandom's per-query start points (algorithms/utils/beamSearch.h:327-332):
//   parlay::random_generator gen;  std::uniform_int_distribution<long> dis(0, n - 1);  indices[i] = dis(gen[i]);
// gen[i] is a generator seeded with hash64((i+1) * 0x7fffffff + state), state = 0 (parlay/random.h:48-61); the draw itself
// is libstdc++'s uniform_int_distribution over that 64-bit engine — used here as is, so the values are the reference's
// wherever it is built with the same standard library.
namespace {
struct ParlayGen {  // parlay::random_generator
  using result_type = size_t;
  size_t state;
  result_type operator()() { return state = (size_t)hash64(state); }
  static constexpr result_type max() { return std::numeric_limits<result_type>::max(); }
  static constexpr result_type min() { return std::numeric_limits<result_type>::lowest(); }
};
}  // namespace
void pann_host_random_starts(uint64_t nq, uint64_t n, uint32_t* out) {
  std::uniform_int_distribution<long> dis(0, (long)n - 1);
  for (uint64_t i = 0; i < nq; i++) {
    ParlayGen r{(size_t)hash64((i + 1) * 0x7fffffffull + 0)};
    out[i] = (uint32_t)dis(r);
  }
}
