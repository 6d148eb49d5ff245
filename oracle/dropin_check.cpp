// oracle/dropin_check.cpp — TEST INFRASTRUCTURE.  Compiles the drop-in header (include/parlayann_b200/dropin.h) inside
// the UNMODIFIED reference (headers included by path) and checks, in ONE process, that switching a call site from
// parlayANN::X to parlayANN::b200::X gives the same bytes: the built graph and build statistics, qsearchAll's ids and
// per-query counters, and the ground truth.  Built by oracle/Makefile into oracle/_ref/ (needs /root/reference);
// run by tests/test_dropin_gpu.py on the GPU box.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>

#include "utils/beamSearch.h"
#include "utils/euclidian_point.h"
#include "utils/graph.h"
#include "utils/mips_point.h"
#include "utils/point_range.h"
#include "utils/stats.h"
#include "utils/types.h"
#include "vamana/index.h"

#define main ref_compute_groundtruth_main
#define pid gt_pid
#include "../data_tools/compute_groundtruth.cpp"
#undef pid
#undef main

#include "parlayann_b200/dropin.h"

using namespace parlayANN;
using uint = unsigned int;

static const char* arg(int argc, char** argv, const char* name, const char* dflt) {
  for (int i = 1; i + 1 < argc; i++) if (!strcmp(argv[i], name)) return argv[i + 1];
  return dflt;
}

// float MIPS data the way python/graph_index.cpp:88-95 and python/builder.cpp:45-54, :73-86 treat it: rows normalised,
// Quantized_Mips_Point<8,true> copy; the graph is built over the copy, the search runs on the copy and is re-ranked with
// the exact rows.  Reference on the CPU and drop-in on the GPU, same process.
template <typename PR>
static int run_qmips(PR& Points, PR& Q, long R, long L, long k, int passes) {
  int bad = 0;
  for (size_t i = 0; i < Points.size(); i++) Points[i].normalize();
  for (size_t i = 0; i < Q.size(); i++) Q[i].normalize();
  using QPR = PointRange<Quantized_Mips_Point<8, true>>;
  QPR Q_Points(Points);
  QPR Q_Q(Q, Q_Points.params);
  double alpha = Points.dimension() < 100 ? (float)1.0 : (float).98;
  BuildParams BP(R, L, alpha, passes, false);
  Graph<uint> Gc(BP.max_degree(), Points.size()), Gg(BP.max_degree(), Points.size());
  stats<uint> Sc(Points.size()), Sg(Points.size());
  { knn_index<QPR, QPR, uint> I(BP); I.build_index(Gc, Q_Points, Q_Points, Sc, true); }
  { b200::knn_index<QPR, QPR, uint> I(BP); I.build_index(Gg, Q_Points, Q_Points, Sg, true); }
  size_t diff_rows = 0;
  for (size_t i = 0; i < Points.size(); i++) {
    bool same = Gc[i].size() == Gg[i].size();
    for (size_t j = 0; same && j < Gc[i].size(); j++) same = Gc[i][j] == Gg[i][j];
    diff_rows += !same;
  }
  bool st_same = std::equal(Sc.visited.begin(), Sc.visited.end(), Sg.visited.begin()) && std::equal(Sc.distances.begin(), Sc.distances.end(), Sg.distances.begin());
  printf("DROPIN build over Quantized_Mips_Point<8,true>: %zu of %zu rows differ, stats %s\n", diff_rows, Points.size(), st_same ? "identical" : "DIFFER");
  bad += diff_rows != 0 || !st_same;
  for (long beam : {10L, 20L, 40L, 100L}) {
    if (beam < k) continue;
    QueryParams QP(k, beam, 1.35, (long)Gc.size(), (long)Gc.max_degree());
    stats<uint> qc(Q.size()), qg(Q.size());
    auto rc = qsearchAll<PR, QPR, QPR, uint>(Q, Q_Q, Q_Q, Gc, Points, Q_Points, Q_Points, qc, 0u, QP);
    auto rg = b200::qsearchAll<PR, QPR, QPR, uint>(Q, Q_Q, Q_Q, Gc, Points, Q_Points, Q_Points, qg, 0u, QP);
    size_t dq = 0;
    for (size_t i = 0; i < Q.size(); i++) dq += !(rc[i] == rg[i]);
    bool s2 = std::equal(qc.visited.begin(), qc.visited.end(), qg.visited.begin()) && std::equal(qc.distances.begin(), qc.distances.end(), qg.distances.begin());
    printf("DROPIN qsearchAll Quantized_Mips_Point<8,true> + re-rank, beam %ld: %zu of %zu queries differ, stats %s\n", beam, dq, Q.size(), s2 ? "identical" : "DIFFER");
    bad += dq != 0 || !s2;
  }
  printf(bad ? "DROPIN_FAIL\n" : "DROPIN_OK\n");
  return bad ? 1 : 0;
}

template <typename PR>
static int run(int argc, char** argv) {
  PR Points((char*)arg(argc, argv, "-base", ""));
  PR Q((char*)arg(argc, argv, "-query", ""));
  long R = atol(arg(argc, argv, "-R", "32")), L = atol(arg(argc, argv, "-L", "64")), k = atol(arg(argc, argv, "-k", "10"));
  int passes = atoi(arg(argc, argv, "-passes", "1"));
  int bad = 0;
  if (!strcmp(arg(argc, argv, "-mode", ""), "qmips")) {
    if constexpr (std::is_same<PR, PointRange<Mips_Point<float>>>::value) return run_qmips(Points, Q, R, L, k, passes);
    else { fprintf(stderr, "dropin_check: -mode qmips needs -data_type float -dist_func mips\n"); return 2; }
  }
  BuildParams BP(R, L, 1.2, passes, false);
  // --- build: reference on the CPU, drop-in on the GPU
  Graph<uint> Gc(BP.max_degree(), Points.size()), Gg(BP.max_degree(), Points.size());
  stats<uint> Sc(Points.size()), Sg(Points.size());
  { knn_index<PR, PR, uint> I(BP); I.build_index(Gc, Points, Points, Sc, true); }
  { b200::knn_index<PR, PR, uint> I(BP); I.build_index(Gg, Points, Points, Sg, true); }
  size_t diff_rows = 0;
  for (size_t i = 0; i < Points.size(); i++) {
    bool same = Gc[i].size() == Gg[i].size();
    for (size_t j = 0; same && j < Gc[i].size(); j++) same = Gc[i][j] == Gg[i][j];
    diff_rows += !same;
  }
  bool st_same = std::equal(Sc.visited.begin(), Sc.visited.end(), Sg.visited.begin()) && std::equal(Sc.distances.begin(), Sc.distances.end(), Sg.distances.begin());
  printf("DROPIN build: %zu of %zu rows differ, stats %s\n", diff_rows, Points.size(), st_same ? "identical" : "DIFFER");
  bad += diff_rows != 0 || !st_same;
  // --- qsearchAll over the reference's sweep of beams, same graph on both sides
  for (long beam : {10L, 13L, 20L, 40L, 100L, 300L}) {
    if (beam < k) continue;
    QueryParams QP(k, beam, 1.35, (long)Gc.size(), (long)Gc.max_degree());
    stats<uint> qc(Q.size()), qg(Q.size());
    auto rc = qsearchAll<PR, PR, PR, uint>(Q, Q, Q, Gc, Points, Points, Points, qc, 0u, QP);
    auto rg = b200::qsearchAll<PR, PR, PR, uint>(Q, Q, Q, Gc, Points, Points, Points, qg, 0u, QP);
    size_t dq = 0;
    for (size_t i = 0; i < Q.size(); i++) dq += !(rc[i] == rg[i]);
    bool s2 = std::equal(qc.visited.begin(), qc.visited.end(), qg.visited.begin()) && std::equal(qc.distances.begin(), qc.distances.end(), qg.distances.begin());
    printf("DROPIN qsearchAll beam %ld: %zu of %zu queries differ, stats %s\n", beam, dq, Q.size(), s2 ? "identical" : "DIFFER");
    bad += dq != 0 || !s2;
  }
  // --- 1-byte quantised copy + exact re-rank (-quantize_mode 1, vamana/neighbors.h:117-127): float Euclidean data only
  if constexpr (std::is_same<PR, PointRange<Euclidian_Point<float>>>::value) {
    using QPR = PointRange<Euclidian_Point<uint8_t>>;
    QPR Q_Points(Points);
    QPR Q_Q(Q, Q_Points.params);
    for (long beam : {10L, 20L, 40L, 100L}) {
      if (beam < k) continue;
      QueryParams QP(k, beam, 1.35, (long)Gc.size(), (long)Gc.max_degree());
      stats<uint> qc(Q.size()), qg(Q.size());
      auto rc = qsearchAll<PR, QPR, QPR, uint>(Q, Q_Q, Q_Q, Gc, Points, Q_Points, Q_Points, qc, 0u, QP);
      auto rg = b200::qsearchAll<PR, QPR, QPR, uint>(Q, Q_Q, Q_Q, Gc, Points, Q_Points, Q_Points, qg, 0u, QP);
      size_t dq = 0;
      for (size_t i = 0; i < Q.size(); i++) dq += !(rc[i] == rg[i]);
      bool s2 = std::equal(qc.visited.begin(), qc.visited.end(), qg.visited.begin()) && std::equal(qc.distances.begin(), qc.distances.end(), qg.distances.begin());
      printf("DROPIN qsearchAll quantised + re-rank, beam %ld: %zu of %zu queries differ, stats %s\n", beam, dq, Q.size(), s2 ? "identical" : "DIFFER");
      bad += dq != 0 || !s2;
    }
  }
  // --- batch_insert used directly (vamana/index.h:188-192): first half in the given order, second half in random order, a
  //     scattered subset again with another max_fraction — graphs and statistics after every call
  {
    BuildParams BP1(R, L, 1.2, 1, false);
    Graph<uint> Hc(BP1.max_degree(), Points.size()), Hg(BP1.max_degree(), Points.size());
    stats<uint> Tc(Points.size()), Tg(Points.size());
    knn_index<PR, PR, uint> Ic(BP1); Ic.start_point = 0;
    b200::knn_index<PR, PR, uint> Ig(BP1);
    const uint n = (uint)Points.size();
    parlay::sequence<uint> a = parlay::tabulate(n / 2, [&](size_t i) { return (uint)i; });
    parlay::sequence<uint> b = parlay::tabulate(n - n / 2, [&](size_t i) { return (uint)(n / 2 + i); });
    parlay::sequence<uint> c = parlay::tabulate(n / 7, [&](size_t i) { return (uint)((i * 7 + 3) % n); });
    struct Step { parlay::sequence<uint>* ins; double alpha; bool rnd; double frac; } steps[3] = {{&a, 1.2, false, .02}, {&b, 1.2, true, .02}, {&c, 1.0, false, .05}};
    for (int si = 0; si < 3; si++) {
      Ic.batch_insert(*steps[si].ins, Hc, Points, Points, Tc, steps[si].alpha, steps[si].rnd, 2, steps[si].frac, false);
      Ig.batch_insert(*steps[si].ins, Hg, Points, Points, Tg, steps[si].alpha, steps[si].rnd, 2, steps[si].frac, false);
      size_t dr = 0;
      for (size_t i = 0; i < Points.size(); i++) {
        bool same = Hc[i].size() == Hg[i].size();
        for (size_t j = 0; same && j < Hc[i].size(); j++) same = Hc[i][j] == Hg[i][j];
        dr += !same;
      }
      bool s3 = std::equal(Tc.visited.begin(), Tc.visited.end(), Tg.visited.begin()) && std::equal(Tc.distances.begin(), Tc.distances.end(), Tg.distances.begin());
      printf("DROPIN batch_insert step %d: %zu of %zu rows differ, stats %s\n", si, dr, Points.size(), s3 ? "identical" : "DIFFER");
      bad += dr != 0 || !s3;
    }
    // --- the resident-index cache must notice a graph that changed in place behind the same addresses: search Hc, rewrite
    //     its rows with Gc's (same buffer, same max degree), search again — the second answer must be the one for Gc
    QueryParams QP(k, 20, 1.35, (long)Gc.size(), (long)Gc.max_degree());
    stats<uint> q1(Q.size()), q2(Q.size()), q3(Q.size());
    auto r1 = b200::qsearchAll<PR, PR, PR, uint>(Q, Q, Q, Hc, Points, Points, Points, q1, 0u, QP);
    for (size_t i = 0; i < Points.size(); i++) {
      parlay::sequence<uint> row = parlay::tabulate(Gc[i].size(), [&](size_t j) { return Gc[i][j]; });
      Hc[i].update_neighbors(row);
    }
    auto r2 = b200::qsearchAll<PR, PR, PR, uint>(Q, Q, Q, Hc, Points, Points, Points, q2, 0u, QP);
    auto r3 = qsearchAll<PR, PR, PR, uint>(Q, Q, Q, Gc, Points, Points, Points, q3, 0u, QP);
    size_t dq = 0, changed = 0;
    for (size_t i = 0; i < Q.size(); i++) { dq += !(r2[i] == r3[i]); changed += !(r1[i] == r2[i]); }
    printf("DROPIN cache after an in-place graph rewrite: %zu of %zu queries differ from the reference on the new graph (%zu answers changed)\n", dq, Q.size(), changed);
    bad += dq != 0;
    b200::clear_cache();
  }
  // --- ground truth: sorted distance rows must be identical
  {
    int gk = 20;
    auto ac = compute_groundtruth<PR>(Points, Q, gk);
    auto ag = b200::compute_groundtruth<PR>(Points, Q, gk);
    size_t dr = 0;
    for (size_t i = 0; i < Q.size(); i++) {
      std::vector<float> a, b;
      for (auto& e : ac[i]) a.push_back(e.second);
      for (auto& e : ag[i]) b.push_back(e.second);
      std::sort(a.begin(), a.end());
      dr += !(a == b);
    }
    printf("DROPIN compute_groundtruth: %zu of %zu rows differ\n", dr, Q.size());
    bad += dr != 0;
  }
  printf(bad ? "DROPIN_FAIL\n" : "DROPIN_OK\n");
  return bad ? 1 : 0;
}

int main(int argc, char** argv) {
  std::string dt = arg(argc, argv, "-data_type", "uint8"), df = arg(argc, argv, "-dist_func", "Euclidian");
  bool l2 = df == "Euclidian";
  if (dt == "uint8") return l2 ? run<PointRange<Euclidian_Point<uint8_t>>>(argc, argv) : run<PointRange<Mips_Point<uint8_t>>>(argc, argv);
  if (dt == "int8") return l2 ? run<PointRange<Euclidian_Point<int8_t>>>(argc, argv) : run<PointRange<Mips_Point<int8_t>>>(argc, argv);
  if (dt == "float") return l2 ? run<PointRange<Euclidian_Point<float>>>(argc, argv) : run<PointRange<Mips_Point<float>>>(argc, argv);
  return 2;
}
