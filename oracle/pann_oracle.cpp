// oracle/pann_oracle.cpp — TEST INFRASTRUCTURE ONLY (never linked or imported by the product path).
//
// A plain single-file C++ restatement of the reference's CPU algorithms for the hot path
// (landrumb/ParlayANN @ 7e52da5), written from the behaviour described in SURVEY.md §8 /
// Appendix A+B.  It is the checker the CUDA path is compared with where the real reference
// binary (oracle/_ref/ref_driver, built from /root/reference by oracle/Makefile) cannot travel
// or is inconvenient to call in-process.  PARITY IS PINNED: every entry point below is checked
// against outputs of the unmodified reference (tests/golden/*, made by tests/golden/make_golden.py
// through oracle/_ref/ref_driver) in tests/test_oracle_vs_reference.py.
//
// Each function cites the reference file:line it follows (paths relative to /root/reference).
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <random>
#include <thread>
#include <unordered_map>
#include <unordered_set>
#include <utility>
#include <vector>

#include "../parlayann_b200/csrc/datagen.h"

namespace {

enum { DT_F32 = 0, DT_U8 = 1, DT_I8 = 2 };
enum { M_L2 = 0, M_MIPS = 1 };

struct PointSet {
  int dtype, metric;
  const uint8_t* base;  // n rows
  uint64_t n;
  uint32_t d;
  uint64_t stride;  // bytes between rows
  const void* row(uint64_t i) const { return base + i * stride; }
};

// algorithms/utils/euclidian_point.h:54-62 (u8), :74-81 (i8), :83-90 (f32)
// algorithms/utils/mips_point.h:43-49 (u8), :51-57 (i8), :59-65 (f32)
// 8-bit: int32 accumulate then one cast to float.  f32: strictly sequential float accumulate.
// NOTE the f32 loops are compiled with -ffp-contract=off here so that every square/product is
// rounded before the add, which is what g++ -O3 emits for the reference's main loop when d%8==0
// (vector sub/mul, scalar add chain).
float dist_rows(int dtype, int metric, const void* a, const void* b, uint32_t d) {
  if (dtype == DT_U8) {
    const uint8_t *p = (const uint8_t*)a, *q = (const uint8_t*)b;
    int32_t r = 0;
    if (metric == M_L2) { for (uint32_t i = 0; i < d; i++) { int32_t t = (int32_t)p[i] - (int32_t)q[i]; r += t * t; } return (float)r; }
    for (uint32_t i = 0; i < d; i++) r += (int32_t)p[i] * (int32_t)q[i];
    return -((float)r);
  }
  if (dtype == DT_I8) {
    const int8_t *p = (const int8_t*)a, *q = (const int8_t*)b;
    int32_t r = 0;
    if (metric == M_L2) { for (uint32_t i = 0; i < d; i++) { int32_t t = (int32_t)p[i] - (int32_t)q[i]; r += t * t; } return (float)r; }
    for (uint32_t i = 0; i < d; i++) r += (int32_t)p[i] * (int32_t)q[i];
    return -((float)r);
  }
  const float *p = (const float*)a, *q = (const float*)b;
  float r = 0.0f;
  if (metric == M_L2) { for (uint32_t i = 0; i < d; i++) { float t = q[i] - p[i]; float s = t * t; r = r + s; } return r; }
  for (uint32_t i = 0; i < d; i++) { float s = q[i] * p[i]; r = r + s; }
  return -r;
}

using idd = std::pair<uint32_t, float>;
inline bool less_idd(const idd& a, const idd& b) {  // beamSearch.h:46-48, vamana/index.h:80-82
  return a.second < b.second || (a.second == b.second && a.first < b.first);
}

// parlay/utilities.h:145-150
inline uint64_t hash64_2(uint64_t x) {
  x = (x ^ (x >> 30)) * UINT64_C(0xbf58476d1ce4e5b9);
  x = (x ^ (x >> 27)) * UINT64_C(0x94d049bb133111eb);
  return x ^ (x >> 31);
}
// parlay/utilities.h:132-142
inline uint64_t hash64(uint64_t u) {
  uint64_t v = u * 3935559000370003845ul + 2691343689449507681ul;
  v ^= v >> 21; v ^= v << 37; v ^= v >> 4;
  v *= 4768777513237032717ul;
  v ^= v << 20; v ^= v >> 41; v ^= v << 5;
  return v;
}

struct QP { long k, beam; double cut; long limit, degree_limit; };

struct SearchOut {
  std::vector<idd> frontier, visited;
  size_t cmps = 0;
};

// algorithms/utils/beamSearch.h:22-214 with use_filtering == false, Points == Q_Points.
// graph: n x (maxdeg+1) u32, slot 0 = degree (algorithms/utils/graph.h:134-141).
// self >= 0: the query IS base row `self` (build mode) -> same_as() skips it (euclidian_point.h:178-180).
// rowfn(v, deg) -> pointer to v's neighbour ids (flat Graph rows, or HNSW's per-node per-level lists)
template <class RowFn>
void beam_search_g(const PointSet& P, RowFn rowfn, uint32_t maxdeg, const void* q, long self,
                   const uint32_t* starts, uint32_t nstarts, const QP& qp, SearchOut& out) {
  const int beam = (int)qp.beam;
  auto dist = [&](uint32_t a) { return dist_rows(P.dtype, P.metric, P.row(a), q, P.d); };
  int bits = std::max<int>(10, (int)std::ceil(std::log2((double)(beam * beam))) - 2);  // :52
  std::vector<uint32_t> filter((size_t)1 << bits, 0xffffffffu);
  auto seen = [&](uint32_t a) {  // :54-59
    size_t loc = hash64_2(a) & (((uint64_t)1 << bits) - 1);
    if (filter[loc] == a) return true;
    filter[loc] = a;
    return false;
  };
  std::vector<idd> frontier, unvisited(std::max<size_t>(beam, nstarts)), visited, cand, nf;
  for (uint32_t i = 0; i < nstarts; i++) { frontier.push_back(idd(starts[i], dist(starts[i]))); seen(starts[i]); }  // :64-69
  std::sort(frontier.begin(), frontier.end(), less_idd);
  for (size_t i = 0; i < frontier.size(); i++) unvisited[i] = frontier[i];
  size_t cmps = nstarts;  // :84
  int remain = (int)frontier.size(), nvis = 0, offset = 0;
  nf.resize(2 * std::max<size_t>(beam, nstarts) + maxdeg);
  while (remain > offset && nvis < qp.limit) {  // :107
    idd cur = unvisited[offset];
    visited.insert(std::upper_bound(visited.begin(), visited.end(), cur, less_idd), cur);  // :112-113
    nvis++;
    bool full = (int)frontier.size() == beam;  // :115
    uint32_t deg = 0;
    const uint32_t* row = rowfn(cur.first, deg);
    long ne = std::min<long>(deg, qp.degree_limit);  // :130
    float cutoff = full ? frontier.back().second : (float)std::numeric_limits<int>::max();  // :150-152
    for (long i = 0; i < ne; i++) {
      uint32_t a = row[i];
      if (seen(a) || (long)a == self) continue;  // :133
      float dd = dist(a);
      cmps++;  // :155
      if (dd >= cutoff) continue;
      cand.push_back(idd(a, dd));
    }
    if (cand.empty() || (qp.limit >= 2 * beam && (long)cand.size() < beam / 8 && offset + 1 < remain)) {  // :162-168
      offset++;
      continue;
    }
    offset = 0;
    std::sort(cand.begin(), cand.end(), less_idd);  // :173
    auto ce = std::unique(cand.begin(), cand.end(), [](const idd& a, const idd& b) { return a.first == b.first; });
    size_t nfs = std::set_union(frontier.begin(), frontier.end(), cand.begin(), ce, nf.begin(), less_idd) - nf.begin();
    cand.clear();
    nfs = std::min<size_t>(beam, nfs);  // :185
    if (qp.k > 0 && (long)nfs > qp.k && P.metric == M_L2) {  // :190-195
      idd thr((uint32_t)0, (float)(qp.cut * (double)nf[qp.k].second));
      size_t c = std::upper_bound(nf.begin(), nf.begin() + nfs, thr, less_idd) - nf.begin();
      nfs = std::max<uint32_t>((uint32_t)c, (uint32_t)frontier.size());
    }
    frontier.assign(nf.begin(), nf.begin() + nfs);  // :198-200
    remain = (int)(std::set_difference(frontier.begin(), frontier.begin() + std::min<long>(frontier.size(), qp.beam),
                                       visited.begin(), visited.end(), unvisited.begin(), less_idd) - unvisited.begin());
  }
  out.frontier = frontier;
  out.visited = visited;
  out.cmps = cmps;
}

void beam_search(const PointSet& P, const uint32_t* graph, uint32_t maxdeg, const void* q, long self,
                 const uint32_t* starts, uint32_t nstarts, const QP& qp, SearchOut& out) {
  beam_search_g(P, [&](uint32_t v, uint32_t& deg) { const uint32_t* r = graph + (uint64_t)v * (maxdeg + 1); deg = r[0]; return r + 1; },
                maxdeg, q, self, starts, nstarts, qp, out);
}

// instrumentation for DESIGN.md (not part of the algorithm): log2 histogram of prune candidate counts
std::atomic<long> g_prune_hist[40];
std::atomic<long> g_prune_max(0);
inline void prune_hist_add(size_t c) {
  int b = 0; while (((size_t)1 << b) < c) b++;
  g_prune_hist[b]++;
  long m = g_prune_max.load(); while ((long)c > m && !g_prune_max.compare_exchange_weak(m, (long)c)) {}
}

// algorithms/vamana/index.h:63-120.  cand already carries d(c,p).
long robust_prune(const PointSet& P, const uint32_t* graph, uint32_t maxdeg, uint32_t R, uint32_t p,
                  std::vector<idd> cand, double alpha, bool add, std::vector<uint32_t>& out) {
  long comps = 0;
  const uint32_t* row = graph + (uint64_t)p * (maxdeg + 1);
  if (add)
    for (uint32_t i = 0; i < row[0]; i++) {  // :72-77
      comps++;
      cand.push_back(idd(row[1 + i], dist_rows(P.dtype, P.metric, P.row(row[1 + i]), P.row(p), P.d)));
    }
  prune_hist_add(cand.size());
  std::sort(cand.begin(), cand.end(), less_idd);  // :80-83
  cand.erase(std::unique(cand.begin(), cand.end(), [](const idd& a, const idd& b) { return a.first == b.first; }), cand.end());
  std::vector<int> ids(cand.size());
  for (size_t i = 0; i < cand.size(); i++) ids[i] = (int)cand[i].first;
  out.clear();
  size_t idx = 0;
  while (out.size() < R && idx < cand.size()) {  // :95-116
    int ps = ids[idx++];
    if (ps == (int)p || ps == -1) continue;
    out.push_back((uint32_t)ps);
    for (size_t i = idx; i < cand.size(); i++) {
      int pp = ids[i];
      if (pp == -1) continue;
      comps++;
      float dsp = dist_rows(P.dtype, P.metric, P.row((uint32_t)ps), P.row((uint32_t)pp), P.d);
      if (alpha * dsp <= cand[i].second) ids[i] = -1;  // double * float -> double compare (:111)
    }
  }
  return comps;
}

// parlay/random.h:154-159 -> random_shuffle_ :91-137, seq_random_shuffle_ :80-89, count_sort (stable)
void random_permutation(int64_t n, int32_t* out) {
  struct rnd { uint64_t state; uint64_t ith(uint64_t i) const { return hash64(i + state); }
               rnd fork(uint64_t i) const { return rnd{hash64(hash64(i + state))}; } };
  auto knuth = [](int32_t* a, size_t m, rnd r) {
    if (m < 2) return;
    for (size_t i = m - 1; i > 0; i--) std::swap(a[i], a[r.ith(i) % (i + 1)]);
  };
  rnd r{0};
  if (n < 8192) {  // parlay/internal/counting_sort.h:29
    for (int64_t i = 0; i < n; i++) out[i] = (int32_t)i;
    knuth(out, (size_t)n, r);
    return;
  }
  auto log2_up = [](uint64_t i) { size_t a = 0; uint64_t b = i - 1; while (b > 0) { b >>= 1; a++; } return a; };
  size_t bits = ((uint64_t)n < (1ull << 27)) ? (log2_up(n) - 7) / 2 : (log2_up(n) - 17);
  size_t nb = (size_t)1 << bits, mask = nb - 1;
  std::vector<size_t> off(nb + 1, 0);
  for (int64_t i = 0; i < n; i++) off[(r.ith(i) & mask) + 1]++;
  for (size_t b = 0; b < nb; b++) off[b + 1] += off[b];
  std::vector<size_t> cur(off.begin(), off.end() - 1);
  for (int64_t i = 0; i < n; i++) out[cur[r.ith(i) & mask]++] = (int32_t)i;
  for (size_t b = 0; b < nb; b++) knuth(out + off[b], off[b + 1] - off[b], r.fork(b));
}

template <class F>
void par_for(size_t lo, size_t hi, int nthreads, F f) {
  if (nthreads <= 1 || hi - lo < 2) { for (size_t i = lo; i < hi; i++) f(i); return; }
  std::atomic<size_t> next(lo);
  std::vector<std::thread> ts;
  for (int t = 0; t < nthreads; t++)
    ts.emplace_back([&] { for (;;) { size_t i = next.fetch_add(16); if (i >= hi) return; for (size_t j = i; j < std::min(hi, i + 16); j++) f(j); } });
  for (auto& t : ts) t.join();
}


// ---------------------------------------------------------------------------------------------------------------------
// HNSW (algorithms/HNSW/HNSW.hpp): layered graph, per node per level a neighbour list.
struct Hnsw {
  uint32_t dim = 0, m = 0, efc = 0, n = 0;
  float m_l = 0, alpha = 0;
  std::vector<uint32_t> level;                          // node::level
  std::vector<std::vector<std::vector<uint32_t>>> nbr;  // nbr[v][l], l = 0..level[v]
  std::vector<uint32_t> entrance;
  uint32_t cap(uint32_t l) const { return l == 0 ? 2 * m : m; }   // get_threshold_m, :527-531
};

// search_layer (:1088-1109): beam_search_impl over level l with QP(ef, ef, 1.35, limit, cap(l)); returns the frontier
void hnsw_search_layer(const Hnsw& H, const PointSet& P, const void* q, long self, const std::vector<uint32_t>& eps, uint32_t ef,
                       uint32_t l, long limit, SearchOut& o) {
  QP qp{(long)ef, (long)ef, 1.35, limit, (long)H.cap(l)};
  beam_search_g(P, [&](uint32_t v, uint32_t& deg) { const auto& r = H.nbr[v][l]; deg = (uint32_t)r.size(); return r.data(); },
                H.cap(l), q, self, eps.data(), (uint32_t)eps.size(), qp, o);
}

// prune_heuristic (:430-486) as select_neighbors calls it (:488-503): candidates sorted by distance; c is dropped when some
// kept r has d(c, r) < alpha * d(c, u) (float product, strict <); stop at `size` kept
std::vector<uint32_t> hnsw_prune_heuristic(const Hnsw& H, const PointSet& P, std::vector<idd> cand, uint32_t size) {
  std::stable_sort(cand.begin(), cand.end(), [](const idd& a, const idd& b) { return a.second < b.second; });
  std::vector<uint32_t> res;
  for (const idd& c : cand) {
    const float d_cu = c.second * H.alpha;
    bool pruned = false;
    for (uint32_t r : res)
      if (dist_rows(P.dtype, P.metric, P.row(c.first), P.row(r), P.d) < d_cu) { pruned = true; break; }
    if (!pruned) { res.push_back(c.first); if (res.size() == size) break; }
  }
  return res;
}

// get_level_random (:505-516) on ONE worker: thread_local mt19937{worker_id = 0}, uniform_real_distribution<>(FLT_MIN, 1)
struct HnswLevels {
  std::mt19937 gen{0};
  std::uniform_real_distribution<> dis{std::numeric_limits<float>::min(), 1.0};
  uint32_t next(float m_l) { return (uint32_t)(-log(dis(gen)) * m_l); }
};

// insert (:828-1014), from_blank = true, nodes [b0, b1) — their levels are already in H.level
void hnsw_insert(Hnsw& H, const PointSet& P, uint32_t b0, uint32_t b1, int nthreads) {
  const uint32_t level_ep = H.level[H.entrance[0]];
  const uint32_t bs = b1 - b0;
  std::vector<std::vector<uint32_t>> eps(bs), nbh_new(bs);
  for (uint32_t i = 0; i < bs; i++) H.nbr[b0 + i].assign(H.level[b0 + i] + 1, {});
  par_for(0, bs, nthreads, [&](size_t i) {  // :865-881: ef = 1 descent to the node's own top level
    const uint32_t u = b0 + (uint32_t)i;
    eps[i] = H.entrance;
    for (uint32_t l = level_ep; l > H.level[u]; --l) {
      SearchOut o;
      hnsw_search_layer(H, P, P.row(u), (long)u, eps[i], 1, l, (long)H.n, o);
      eps[i].assign(1, o.frontier[0].first);
    }
  });
  for (int l_c = (int)level_ep; l_c >= 0; --l_c) {  // :885-1000
    par_for(0, bs, nthreads, [&](size_t i) {
      const uint32_t u = b0 + (uint32_t)i;
      if ((uint32_t)l_c > H.level[u]) return;
      SearchOut o;
      hnsw_search_layer(H, P, P.row(u), (long)u, eps[i], H.efc, (uint32_t)l_c, (long)H.n, o);
      nbh_new[i] = hnsw_prune_heuristic(H, P, o.frontier, H.cap((uint32_t)l_c));
      eps[i].clear();
      for (const idd& e : o.frontier) eps[i].push_back(e.first);
    });
    for (uint32_t i = 0; i < bs; i++)  // forward edges (:938-942)
      if ((uint32_t)l_c <= H.level[b0 + i]) H.nbr[b0 + i][l_c] = nbh_new[i];
    // reverse edges (:944-999): grouped by target, values in batch order
    std::vector<uint32_t> keys;
    std::unordered_map<uint32_t, std::vector<uint32_t>> groups;
    for (uint32_t i = 0; i < bs; i++) {
      if ((uint32_t)l_c > H.level[b0 + i]) continue;
      for (uint32_t pv : nbh_new[i]) {
        auto it = groups.find(pv);
        if (it == groups.end()) { keys.push_back(pv); it = groups.emplace(pv, std::vector<uint32_t>()).first; }
        it->second.push_back(b0 + i);
      }
    }
    par_for(0, keys.size(), nthreads, [&](size_t j) {
      const uint32_t pv = keys[j];
      std::vector<uint32_t>& nb = H.nbr[pv][l_c];
      const std::vector<uint32_t>& add = groups.find(pv)->second;
      const uint32_t cap = H.cap((uint32_t)l_c);
      if (nb.size() + add.size() > cap) {  // :963-987: keep the cap closest of old + new (plain sort by distance)
        std::vector<idd> c;
        for (uint32_t x : nb) c.push_back(idd(x, dist_rows(P.dtype, P.metric, P.row(x), P.row(pv), P.d)));
        for (uint32_t x : add) c.push_back(idd(x, dist_rows(P.dtype, P.metric, P.row(x), P.row(pv), P.d)));
        std::stable_sort(c.begin(), c.end(), [](const idd& a, const idd& b) { return a.second < b.second; });
        nb.resize(cap);
        for (uint32_t k = 0; k < cap; k++) nb[k] = c[k].first;
      } else {
        nb.insert(nb.end(), add.begin(), add.end());
      }
    });
  }
  // entrance update (:1002-1013): first node of the batch with the highest level
  uint32_t hi = b0;
  for (uint32_t u = b0; u < b1; u++) if (H.level[u] > H.level[hi]) hi = u;
  if (H.level[hi] > level_ep) H.entrance.assign(1, hi);
  else if (H.level[hi] == level_ep) H.entrance.push_back(hi);
}

}  // namespace

extern "C" {

// out[0..40) = #prunes with candidate count in (2^(b-1), 2^b]; out[40] = max count; resets the counters
void oracle_prune_hist(long* out) { for (int i = 0; i < 40; i++) out[i] = g_prune_hist[i].exchange(0); out[40] = g_prune_max.exchange(0); }

int oracle_num_threads() { unsigned h = std::thread::hardware_concurrency(); return h ? (int)h : 1; }

float oracle_distance(int dtype, int metric, const void* a, const void* b, uint32_t d) { return dist_rows(dtype, metric, a, b, d); }

void oracle_random_permutation(int64_t n, int32_t* out) { random_permutation(n, out); }

// batch driver = qsearchAll (beamSearch.h:537-565) / beam_search (:217-223) over nq queries.
// self_base >= 0: query i is base row self_base+i? no -> selfs[i] (may be NULL) gives the base row a query aliases.
// Outputs (any may be NULL): out_ids/out_dists nq*k (first k of the frontier), visited/cmps counters nq,
// full lists: fr_* nq*fcap, vis_* nq*vcap with counts in nfr/nvis.
// starts_stride == 0: one start list for the batch (searchAll); else query i starts from
// starts[i * starts_stride .. + (nstarts_q ? nstarts_q[i] : nstarts)) (beamSearchRandom, beamSearch.h:327-338; HNSW).
int oracle_search_batch_starts(int dtype, int metric, const void* points, uint64_t n, uint32_t d, uint64_t stride,
                        const uint32_t* graph, uint32_t maxdeg, const void* queries, uint64_t nq, uint64_t qstride,
                        const int64_t* selfs, const uint32_t* starts, uint32_t nstarts, uint32_t starts_stride, const uint32_t* nstarts_q,
                        long k, long beam, double cut, long limit, long degree_limit,
                        uint32_t* out_ids, float* out_dists, uint32_t* out_visited, uint32_t* out_cmps,
                        uint32_t fcap, uint32_t* fr_ids, float* fr_dists, uint32_t* nfr,
                        uint32_t vcap, uint32_t* vis_ids, float* vis_dists, uint32_t* nvis, int nthreads) {
  PointSet P{dtype, metric, (const uint8_t*)points, n, d, stride};
  QP qp{k, beam, cut, limit, degree_limit};
  std::atomic<int> bad(0);
  par_for(0, nq, nthreads, [&](size_t i) {
    SearchOut o;
    beam_search(P, graph, maxdeg, (const uint8_t*)queries + i * qstride, selfs ? selfs[i] : -1, starts + i * (size_t)starts_stride,
                (starts_stride && nstarts_q) ? nstarts_q[i] : nstarts, qp, o);
    if ((long)o.frontier.size() < k) { bad = 1; return; }  // beamSearch.h:416-419 aborts
    for (long j = 0; j < k; j++) { if (out_ids) out_ids[i * k + j] = o.frontier[j].first; if (out_dists) out_dists[i * k + j] = o.frontier[j].second; }
    if (out_visited) out_visited[i] = (uint32_t)o.visited.size();
    if (out_cmps) out_cmps[i] = (uint32_t)o.cmps;
    if (nfr) nfr[i] = (uint32_t)o.frontier.size();
    if (nvis) nvis[i] = (uint32_t)o.visited.size();
    for (size_t j = 0; j < o.frontier.size() && j < fcap; j++) { if (fr_ids) fr_ids[i * fcap + j] = o.frontier[j].first; if (fr_dists) fr_dists[i * fcap + j] = o.frontier[j].second; }
    for (size_t j = 0; j < o.visited.size() && j < vcap; j++) { if (vis_ids) vis_ids[i * vcap + j] = o.visited[j].first; if (vis_dists) vis_dists[i * vcap + j] = o.visited[j].second; }
  });
  return bad;
}

int oracle_search_batch(int dtype, int metric, const void* points, uint64_t n, uint32_t d, uint64_t stride,
                        const uint32_t* graph, uint32_t maxdeg, const void* queries, uint64_t nq, uint64_t qstride,
                        const int64_t* selfs, const uint32_t* starts, uint32_t nstarts,
                        long k, long beam, double cut, long limit, long degree_limit,
                        uint32_t* out_ids, float* out_dists, uint32_t* out_visited, uint32_t* out_cmps,
                        uint32_t fcap, uint32_t* fr_ids, float* fr_dists, uint32_t* nfr,
                        uint32_t vcap, uint32_t* vis_ids, float* vis_dists, uint32_t* nvis, int nthreads) {
  return oracle_search_batch_starts(dtype, metric, points, n, d, stride, graph, maxdeg, queries, nq, qstride, selfs, starts, nstarts, 0, nullptr,
                                    k, beam, cut, limit, degree_limit, out_ids, out_dists, out_visited, out_cmps, fcap, fr_ids, fr_dists, nfr,
                                    vcap, vis_ids, vis_dists, nvis, nthreads);
}

// knn_index::build_index (vamana/index.h:150-186) + batch_insert (:188-316), start_point = 0.
// graph: n x (R+1) u32, zero-filled by the caller for a fresh build.  stats (may be NULL): n u32 each.
// batch_insert (vamana/index.h:188-316): `order` = the vertices in insertion order (already shuffled, :211-218)
static void vamana_batch_insert(const PointSet& P, uint32_t* graph, uint32_t maxdeg, uint32_t R, uint32_t L, double a,
                                const std::vector<uint32_t>& order, double max_fraction, uint32_t* st_visited, uint32_t* st_dists, int nthreads) {
  const uint64_t n = P.n;
  const uint32_t start = 0;
  {
    size_t m = order.size(), inc = 0, count = 0;
    size_t max_batch = std::min((size_t)(max_fraction * (double)(float)n), (size_t)1000000);  // :206-207 (double * float n)
    if (max_batch == 0) max_batch = n;
    while (count < m) {
      size_t lo, hi;
      if (pow(2.0, (double)inc) <= (double)max_batch) {  // :226-229
        lo = (size_t)pow(2.0, (double)inc) - 1;
        hi = std::min((size_t)pow(2.0, (double)(inc + 1)) - 1, m);
        count = hi;
      } else {  // :231-233
        lo = count;
        hi = std::min(count + max_batch, m);
        count += max_batch;
      }
      std::vector<std::vector<uint32_t>> new_out(hi - lo);
      par_for(lo, hi, nthreads, [&](size_t i) {  // phase 1 :247-266
        uint32_t p = order[i];
        QP qp{0, (long)L, 0.0, (long)n, (long)maxdeg};  // :250
        SearchOut o;
        beam_search(P, graph, maxdeg, P.row(p), (long)p, &start, 1, qp, o);
        long c2 = robust_prune(P, graph, maxdeg, R, p, o.visited, a, true, new_out[i - lo]);
        if (st_dists) st_dists[p] += (uint32_t)(o.cmps + c2);
        if (st_visited) st_visited[p] += (uint32_t)o.visited.size();
      });
      for (size_t i = lo; i < hi; i++) {  // phase 2 :268-270
        uint32_t* row = graph + (uint64_t)order[i] * (maxdeg + 1);
        row[0] = (uint32_t)new_out[i - lo].size();
        std::copy(new_out[i - lo].begin(), new_out[i - lo].end(), row + 1);
      }
      // phase 3 :278-300 — group (nbr, p) by nbr, values keep batch order (parlay group_by_key)
      std::vector<uint32_t> keys;
      std::unordered_map<uint32_t, std::vector<uint32_t>> groups;
      for (size_t i = lo; i < hi; i++)
        for (uint32_t nb : new_out[i - lo]) {
          auto it = groups.find(nb);
          if (it == groups.end()) { keys.push_back(nb); it = groups.emplace(nb, std::vector<uint32_t>()).first; }
          it->second.push_back(order[i]);
        }
      par_for(0, keys.size(), nthreads, [&](size_t j) {
        uint32_t t = keys[j];
        std::vector<uint32_t>& c = groups.find(t)->second;
        uint32_t* row = graph + (uint64_t)t * (maxdeg + 1);
        if (c.size() + row[0] <= R) {  // :291-294, :140-146
          std::unordered_set<uint32_t> have(c.begin(), c.end());
          std::vector<uint32_t> merged = c;
          for (uint32_t i = 0; i < row[0]; i++) if (!have.count(row[1 + i])) merged.push_back(row[1 + i]);
          row[0] = (uint32_t)merged.size();
          std::copy(merged.begin(), merged.end(), row + 1);
        } else {  // :296-298, ids-only overload :124-137
          std::vector<idd> cc;
          long c0 = 0;
          for (uint32_t x : c) { c0++; cc.push_back(idd(x, dist_rows(P.dtype, P.metric, P.row(x), P.row(t), P.d))); }
          std::vector<uint32_t> res;
          long c1 = robust_prune(P, graph, maxdeg, R, t, cc, a, true, res);
          row[0] = (uint32_t)res.size();
          std::copy(res.begin(), res.end(), row + 1);
          if (st_dists) st_dists[t] += (uint32_t)(c0 + c1);
        }
      });
      inc++;
    }
  }
}

int oracle_vamana_build(int dtype, int metric, const void* points, uint64_t n, uint32_t d, uint64_t stride,
                        uint32_t R, uint32_t L, double alpha, int num_passes, int sort_neighbors,
                        uint32_t* graph, uint32_t* st_visited, uint32_t* st_dists, int nthreads) {
  PointSet P{dtype, metric, (const uint8_t*)points, n, d, stride};
  const uint32_t maxdeg = R;
  std::vector<int32_t> perm(n);
  random_permutation((int64_t)n, perm.data());  // :211-216 (inserts = identity, so shuffled = perm)
  std::vector<uint32_t> order(perm.begin(), perm.end());
  for (int pass = 0; pass < num_passes; pass++) {
    double a = (pass == num_passes - 1) ? alpha : 1.0;  // :173-178
    vamana_batch_insert(P, graph, maxdeg, R, L, a, order, 0.02, st_visited, st_dists, nthreads);
  }
  if (sort_neighbors)  // :180-185 — dist-only comparator; ties unordered in the reference (std::sort)
    par_for(0, n, nthreads, [&](size_t i) {
      uint32_t* row = graph + (uint64_t)i * (maxdeg + 1);
      std::sort(row + 1, row + 1 + row[0], [&](uint32_t j, uint32_t k) {
        return dist_rows(P.dtype, P.metric, P.row(i), P.row(j), P.d) < dist_rows(P.dtype, P.metric, P.row(i), P.row(k), P.d);
      });
    });
  return 0;
}

// knn_index::batch_insert (vamana/index.h:188-316) of `inserts` into an existing graph (alpha as given, no final sort)
int oracle_vamana_batch_insert(int dtype, int metric, const void* points, uint64_t n, uint32_t d, uint64_t stride, uint32_t R, uint32_t L,
                               double alpha, const uint32_t* inserts, uint64_t m, int random_order, double max_fraction, uint32_t* graph,
                               uint32_t* st_visited, uint32_t* st_dists, int nthreads) {
  PointSet P{dtype, metric, (const uint8_t*)points, n, d, stride};
  std::vector<int32_t> rperm(m);
  if (random_order) random_permutation((int64_t)m, rperm.data());
  else for (uint64_t i = 0; i < m; i++) rperm[i] = (int32_t)i;
  std::vector<uint32_t> order(m);
  for (uint64_t i = 0; i < m; i++) order[i] = inserts[rperm[i]];  // :217-218
  vamana_batch_insert(P, graph, R, R, L, alpha, order, max_fraction, st_visited, st_dists, nthreads);
  return 0;
}

// the batch schedule alone (vamana/index.h:206-209, :223-234), for host-logic tests: returns #batches
int64_t oracle_batch_schedule(uint64_t n, uint64_t m, int64_t* lo_out, int64_t* hi_out, int64_t cap) {
  size_t inc = 0, count = 0;
  size_t max_batch = std::min((size_t)(0.02 * (double)(float)n), (size_t)1000000);
  if (max_batch == 0) max_batch = n;
  int64_t nb = 0;
  while (count < m) {
    size_t lo, hi;
    if (pow(2.0, (double)inc) <= (double)max_batch) { lo = (size_t)pow(2.0, (double)inc) - 1; hi = std::min((size_t)pow(2.0, (double)(inc + 1)) - 1, (size_t)m); count = hi; }
    else { lo = count; hi = std::min(count + max_batch, (size_t)m); count += max_batch; }
    if (nb < cap) { lo_out[nb] = (int64_t)lo; hi_out[nb] = (int64_t)hi; }
    nb++; inc++;
  }
  return nb;
}

// data_tools/compute_groundtruth.cpp:22-60 then write_ibin's per-row sort by dist (:68-73).
// Output rows are sorted by (dist, id) — the reference sorts by dist only with a non-stable sort, so
// the order inside a tie group is unspecified there; tests compare tie groups as sets.
int oracle_groundtruth(int dtype, int metric, const void* base, uint64_t nb, const void* queries, uint64_t nq,
                       uint32_t d, uint64_t bstride, uint64_t qstride, uint32_t k, int32_t* out_ids, float* out_dists, int nthreads) {
  if (nb < k) return 1;
  const float dmin = metric == M_L2 ? 0.0f : -std::numeric_limits<float>::max();  // euclidian_point.h:111, mips_point.h:81
  par_for(0, nq, nthreads, [&](size_t i) {
    const void* q = (const uint8_t*)queries + i * qstride;
    std::vector<std::pair<int, float>> top;
    float topdist = dmin;
    int toppos = 0;
    for (uint64_t j = 0; j < nb; j++) {
      float dd = dist_rows(dtype, metric, q, (const uint8_t*)base + j * bstride, d);  // Q[i].distance(B[j]) :34
      if (top.size() < k) {  // :35-41
        if (dd > topdist) { topdist = dd; toppos = (int)top.size(); }
        top.push_back({(int)j, dd});
      } else if (dd < topdist) {  // :42-54
        float nt = dmin; int np = 0;
        top[toppos] = {(int)j, dd};
        for (size_t l = 0; l < top.size(); l++) if (top[l].second > nt) { nt = top[l].second; np = (int)l; }
        topdist = nt; toppos = np;
      }
    }
    std::sort(top.begin(), top.end(), [](auto& a, auto& b) { return a.second < b.second || (a.second == b.second && a.first < b.first); });
    for (uint32_t j = 0; j < k; j++) { out_ids[i * k + j] = top[j].first; out_dists[i * k + j] = top[j].second; }
  });
  return 0;
}

// Synthetic data — CPU evaluation of parlayann_b200/csrc/datagen.h; the CUDA library
// evaluates the same header on the device, bit-identically.
void oracle_gen_points(int kind, uint64_t model_seed, uint64_t stream, uint64_t row0, uint64_t nrows, uint32_t d,
                       uint32_t r, uint32_t nc, void* out, uint64_t stride, int nthreads) {
  std::vector<int32_t> A((size_t)r * d), C((size_t)nc * r);
  for (uint32_t k = 0; k < r; k++) for (uint32_t j = 0; j < d; j++) A[(size_t)k * d + j] = gen_A(model_seed, k, j);
  for (uint32_t c = 0; c < nc; c++) for (uint32_t k = 0; k < r; k++) C[(size_t)c * r + k] = gen_C(model_seed, c, k);
  par_for(0, nrows, nthreads, [&](size_t i) {
    uint64_t row = row0 + i;
    int32_t z[PANN_GEN_MAX_R];
    gen_latent(model_seed, stream, row, r, nc, nc ? C.data() : nullptr, z);
    uint8_t* o = (uint8_t*)out + i * stride;
    std::vector<int32_t> x(d);
    for (uint32_t j = 0; j < d; j++) x[j] = gen_coord(model_seed, stream, row, j, r, z, A.data() + j, d);
    if (kind == PANN_GEN_U8) for (uint32_t j = 0; j < d; j++) o[j] = gen_to_u8(x[j]);
    else if (kind == PANN_GEN_I8) for (uint32_t j = 0; j < d; j++) ((int8_t*)o)[j] = gen_to_i8(x[j]);
    else if (kind == PANN_GEN_F32) for (uint32_t j = 0; j < d; j++) ((float*)o)[j] = gen_to_f32(x[j]);
    else gen_normalize_row(x.data(), d, (float*)o);
  });
}

// ---- 1-byte scalar quantisation of float rows + exact re-rank (the -quantize_mode 1 path, vamana/neighbors.h:117-127)
// Euclidian_Point<uint8_t>::generate_parameters (euclidian_point.h:211-237) and parameters(min,max,dims) (:104-108):
// per row min/max START AT 0 (so min_val <= 0 <= max_val), all-integer non-negative data maps to [0,255] unscaled.
void oracle_quant_params_u8(const float* pts, uint64_t n, uint32_t d, uint64_t stride, float* slope, int32_t* offset) {
  float min_val = 0.0f, max_val = 0.0f;
  bool all_ints = true;
  for (uint64_t i = 0; i < n; i++) {
    const float* r = (const float*)((const uint8_t*)pts + i * stride);
    for (uint32_t j = 0; j < d; j++) {
      all_ints = all_ints && (r[j] >= 0) && (r[j] - (long)r[j]) == 0;
      min_val = std::min<float>(min_val, r[j]);
      max_val = std::max<float>(max_val, r[j]);
    }
  }
  if (all_ints) { if (max_val < 256) max_val = 255; min_val = 0; }
  const long range = 255;
  *slope = range / (max_val - min_val);
  *offset = (int32_t)round(min_val * *slope);
}
// Euclidian_Point<uint8_t>::translate_point (euclidian_point.h:183-209)
void oracle_quantize_u8(const float* pts, uint64_t n, uint32_t d, uint64_t stride, float slope, int32_t offset, uint8_t* out) {
  const long range = 255;
  for (uint64_t i = 0; i < n; i++) {
    const float* r = (const float*)((const uint8_t*)pts + i * stride);
    for (uint32_t j = 0; j < d; j++) {
      if (slope == 1.0 && offset == 0) { out[i * d + j] = (uint8_t)r[j]; continue; }
      float x = r[j];
      int64_t v = (int64_t)(std::round(x * slope)) - offset;
      if (v < 0) v = 0;
      if (v > range) v = range;
      out[i * d + j] = (uint8_t)v;
    }
  }
}
// ---- Quantized_Mips_Point<8,trim> over normalised float rows (python/graph_index.cpp:88-95, python/builder.cpp:45-54, :73-86)
// Mips_Point<float>::normalize (mips_point.h:115-124): float products summed into a double in element order, sqrt,
// 1/norm rounded to float, float products written back in place.
void oracle_normalize_f32(float* pts, uint64_t n, uint32_t d, uint64_t stride) {
  for (uint64_t i = 0; i < n; i++) {
    float* r = (float*)((uint8_t*)pts + i * stride);
    double norm = 0.0;
    for (uint32_t j = 0; j < d; j++) { float sq = r[j] * r[j]; norm += sq; }
    norm = std::sqrt(norm);
    if (norm == 0) norm = 1.0;
    float inv_norm = 1.0 / norm;
    for (uint32_t j = 0; j < d; j++) r[j] = r[j] * inv_norm;
  }
}
// Quantized_Mips_Point::generate_parameters (mips_point.h:437-462): all values sorted; trim drops 1e-4 of each tail
// (lower index = (long)(float cutoff * float len), upper = (long)((1.0 - cutoff) * (len - 1)) in double); bound = max(max, -min).
float oracle_quant_params_mips8(const float* pts, uint64_t n, uint32_t d, uint64_t stride, int trim) {
  const long len = (long)(n * d);
  if (len == 0) return 1.0f;
  std::vector<float> vals((size_t)len);
  for (uint64_t i = 0; i < n; i++) {
    const float* r = (const float*)((const uint8_t*)pts + i * stride);
    for (uint32_t j = 0; j < d; j++) vals[i * d + j] = r[j];
  }
  std::sort(vals.begin(), vals.end());
  float min_val, max_val;
  if (trim) {
    float cutoff = .0001;
    min_val = vals[(long)(cutoff * len)];
    max_val = vals[(long)((1.0 - cutoff) * (len - 1))];
  } else {
    min_val = vals[0];
    max_val = vals[len - 1];
  }
  return std::max(max_val, -min_val);
}
// Quantized_Mips_Point<8>::translate_point (mips_point.h:421-434), range = 255: scale = 127 / max_val in float,
// values beyond +-max_val cap at +-127, the rest std::round(x * scale) (halves away from zero).
void oracle_quantize_mips8(const float* pts, uint64_t n, uint32_t d, uint64_t stride, float max_val, int8_t* out) {
  const int range = 255;
  for (uint64_t i = 0; i < n; i++) {
    const float* r = (const float*)((const uint8_t*)pts + i * stride);
    for (uint32_t j = 0; j < d; j++) {
      float mv = max_val;
      float scale = (range / 2) / mv;
      float pj = r[j];
      if (pj < -mv) out[i * d + j] = (int8_t)(-range / 2);
      else if (pj > mv) out[i * d + j] = (int8_t)(range / 2);
      else { int32_t v = std::round(pj * scale); out[i * d + j] = (int8_t)v; }
    }
  }
}
// beam_search_rerank's re-rank step (beamSearch.h:426-444): exact distances of the first min(k*rerank_factor, |frontier|)
// frontier entries, std::sort by (dist, id), first k.
int oracle_rerank(int dtype, int metric, const void* base, uint64_t n, uint32_t d, uint64_t stride, const void* queries,
                  uint64_t nq, uint64_t qstride, const uint32_t* fr_ids, const uint32_t* nfr, uint32_t fcap, long k,
                  long rerank_factor, uint32_t* out_ids, float* out_dists, int nthreads) {
  PointSet P{dtype, metric, (const uint8_t*)base, n, d, stride};
  std::atomic<int> bad(0);
  par_for(0, nq, nthreads, [&](size_t i) {
    if ((long)nfr[i] < k) { bad = 1; return; }
    const int num_check = (int)std::min<long>(k * rerank_factor, (long)nfr[i]);
    std::vector<idd> pts;
    for (int t = 0; t < num_check; t++) {
      const uint32_t j = fr_ids[i * fcap + t];
      pts.push_back(idd(j, dist_rows(dtype, metric, P.row(j), (const uint8_t*)queries + i * qstride, d)));
    }
    std::sort(pts.begin(), pts.end(), less_idd);
    for (long t = 0; t < k; t++) { out_ids[i * k + t] = pts[t].first; if (out_dists) out_dists[i * k + t] = pts[t].second; }
  });
  return bad;
}

// ---- HNSW entry points.  The model travels as flat arrays: level[n]; adjacency in the model file's own order (for each node,
// for l = 0..level: a count, then ids), as `counts` (one per (node, level)) and `edges`; the entrance list.
// oracle_hnsw_build: HNSW(begin, end, dim, m_l, m, efc, alpha) (HNSW.hpp:746-824) with the levels of a single worker.
// Returns the number of edges; call with edges == NULL first to size the buffers (counts needs sum(level)+n entries).
static Hnsw g_hnsw;   // the last built / loaded model (one at a time: test infrastructure)
static void hnsw_export(const Hnsw& H, uint32_t* level, uint64_t* counts, uint32_t* edges, uint32_t* entrance, uint64_t* sizes) {
  uint64_t nc = 0, ne = 0;
  for (uint32_t v = 0; v < H.n; v++) {
    if (level) level[v] = H.level[v];
    for (uint32_t l = 0; l <= H.level[v]; l++) {
      if (counts) counts[nc] = H.nbr[v][l].size();
      nc++;
      for (uint32_t x : H.nbr[v][l]) { if (edges) edges[ne] = x; ne++; }
    }
  }
  if (entrance) for (size_t i = 0; i < H.entrance.size(); i++) entrance[i] = H.entrance[i];
  sizes[0] = nc; sizes[1] = ne; sizes[2] = H.entrance.size();
}
int oracle_hnsw_build(int dtype, int metric, const void* points, uint64_t n, uint32_t d, uint64_t stride, float m_l, uint32_t m,
                      uint32_t efc, float alpha, const uint32_t* levels_in, uint64_t* sizes, int nthreads) {
  PointSet P{dtype, metric, (const uint8_t*)points, n, d, stride};
  Hnsw& H = g_hnsw;
  H = Hnsw();
  H.dim = d; H.m = m; H.efc = efc; H.n = (uint32_t)n; H.m_l = m_l; H.alpha = alpha;
  H.level.resize(n); H.nbr.resize(n);
  HnswLevels L;
  for (uint64_t i = 0; i < n; i++) H.level[i] = levels_in ? levels_in[i] : L.next(m_l);   // node order = draw order on one worker
  H.nbr[0].assign(H.level[0] + 1, {});
  H.entrance.assign(1, 0);
  uint32_t batch_begin = 0, batch_end = 1, size_limit = (uint32_t)(n * 0.02);  // :771
  while (batch_end < n) {  // :773-799, batch_base = 2
    batch_begin = batch_end;
    batch_end = std::min({(uint32_t)n, (uint32_t)std::ceil(batch_begin * 2.0f) + 1, batch_begin + size_limit});
    hnsw_insert(H, P, batch_begin, batch_end, nthreads);
  }
  hnsw_export(H, nullptr, nullptr, nullptr, nullptr, sizes);
  return 0;
}
void oracle_hnsw_export(uint32_t* level, uint64_t* counts, uint32_t* edges, uint32_t* entrance) {
  uint64_t sizes[3];
  hnsw_export(g_hnsw, level, counts, edges, entrance, sizes);
}
void oracle_hnsw_import(uint32_t n, uint32_t dim, float m_l, uint32_t m, uint32_t efc, float alpha, const uint32_t* level,
                        const uint64_t* counts, const uint32_t* edges, const uint32_t* entrance, uint64_t nent) {
  Hnsw& H = g_hnsw;
  H = Hnsw();
  H.dim = dim; H.m = m; H.efc = efc; H.n = n; H.m_l = m_l; H.alpha = alpha;
  H.level.assign(level, level + n); H.nbr.resize(n);
  uint64_t nc = 0, ne = 0;
  for (uint32_t v = 0; v < n; v++) {
    H.nbr[v].resize(level[v] + 1);
    for (uint32_t l = 0; l <= level[v]; l++) { H.nbr[v][l].assign(edges + ne, edges + ne + counts[nc]); ne += counts[nc]; nc++; }
  }
  H.entrance.assign(entrance, entrance + nent);
}
// HNSW::search (:1570-1613) for every query: ef = 1 descent from the entrances to level 1 (search_layer_to :1545-1568), then
// search_layer(ef) on level 0; the result is cut to k plus the entries tied with the k-th distance.  out_* hold up to `ocap`
// entries per query, out_cnt the number the reference returns, out_cmps the summed distance comparisons of all layers.
int oracle_hnsw_search(int dtype, int metric, const void* points, uint64_t stride, const void* queries, uint64_t nq, uint64_t qstride,
                       uint32_t k, uint32_t ef, long limit, uint32_t ocap, uint32_t* out_ids, float* out_dists, uint32_t* out_cnt,
                       uint32_t* out_cmps, int nthreads) {
  const Hnsw& H = g_hnsw;
  PointSet P{dtype, metric, (const uint8_t*)points, H.n, H.dim, stride};
  par_for(0, nq, nthreads, [&](size_t i) {
    const void* q = (const uint8_t*)queries + i * qstride;
    std::vector<uint32_t> eps = H.entrance;
    size_t cmps = 0;
    for (uint32_t l = H.level[H.entrance[0]]; l > 0; --l) {
      SearchOut o;
      hnsw_search_layer(H, P, q, -1, eps, 1, l, (long)H.n, o);   // limit_eval is not passed down (:1553-1556)
      cmps += o.cmps;
      eps.assign(1, o.frontier[0].first);
    }
    SearchOut o;
    hnsw_search_layer(H, P, q, -1, eps, ef, 0, limit > 0 ? limit : (long)H.n, o);
    cmps += o.cmps;
    std::vector<idd> R = o.frontier;
    uint32_t kk = k;
    if (R.size() > k) {  // :1594-1600: sorted by distance; everything tied with the k-th stays
      std::stable_sort(R.begin(), R.end(), [](const idd& a, const idd& b) { return a.second < b.second; });
      if (k > 0) { kk = k; while (kk < R.size() && !(R[k - 1].second < R[kk].second)) kk++; }
      R.resize(kk);
    }
    out_cnt[i] = (uint32_t)R.size();
    out_cmps[i] = (uint32_t)cmps;
    for (size_t j = 0; j < R.size() && j < ocap; j++) { out_ids[i * ocap + j] = R[j].first; out_dists[i * ocap + j] = R[j].second; }
  });
  return 0;
}
// the level sequence of one worker (get_level_random, HNSW.hpp:505-516)
void oracle_hnsw_levels(uint64_t n, float m_l, uint32_t* out) { HnswLevels L; for (uint64_t i = 0; i < n; i++) out[i] = L.next(m_l); }

}  // extern "C"
