// oracle/ref_driver.cpp — TEST INFRASTRUCTURE, not product code.
//
// A thin command-line driver around the UNMODIFIED reference headers
// (landrumb/ParlayANN, included by path from /root/reference at compile time;
// no reference source is copied into this repo).  It calls the reference's own
// entry points and dumps raw results so that the CUDA path and the C++
// restatement in oracle/pann_oracle.cpp can be compared with them bit for bit:
//
//   perm    -> parlay::random_permutation<int>(n)                    (parlay/random.h:154-159)
//   build   -> knn_index::build_index                                (algorithms/vamana/index.h:150-186)
//   insert  -> knn_index::batch_insert called directly, three times        (algorithms/vamana/index.h:188-316)
//   search  -> qsearchAll + stats                                    (algorithms/utils/beamSearch.h:537-565)
//   full    -> beam_search (frontier + visited lists per query)      (algorithms/utils/beamSearch.h:217-223)
//   random  -> beamSearchRandom (one random start point per query)     (algorithms/utils/beamSearch.h:309-351)
//   recall  -> checkRecall (timed qsearchAll + tie-aware recall)     (algorithms/utils/check_nn_recall.h:17-125)
//   qsearch -> qsearchAll over a 1-byte quantised copy + exact re-rank (vamana/neighbors.h:117-127, beamSearch.h:409-444;
//              MIPS: Quantized_Mips_Point<8,true> over normalised points, python/graph_index.cpp:88-95, :166-181)
//   qbuild  -> the build python/builder.cpp:36-95 intends for float data: graph over the 1-byte copy
//   gt      -> compute_groundtruth + write_ibin                      (data_tools/compute_groundtruth.cpp:22-100)
//
// Built by oracle/Makefile into oracle/_ref/ (git-ignored, travels to the GPU box).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <type_traits>
#include <vector>

#include "utils/beamSearch.h"
#include "utils/check_nn_recall.h"
#include "utils/euclidian_point.h"
#include "utils/graph.h"
#include "utils/mips_point.h"
#include "utils/point_range.h"
#include "utils/stats.h"
#include "utils/types.h"
#include "vamana/index.h"

// compute_groundtruth.cpp has its own main(); rename it and reuse its two functions.
#define main ref_compute_groundtruth_main
#define pid gt_pid
#include "../data_tools/compute_groundtruth.cpp"
#undef pid
#undef main

using namespace parlayANN;
using uint = unsigned int;

static double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

static void dump(const std::string& path, const void* p, size_t bytes) {
  FILE* f = fopen(path.c_str(), "wb");
  if (!f) { perror(path.c_str()); exit(2); }
  if (bytes && fwrite(p, 1, bytes, f) != bytes) { perror("fwrite"); exit(2); }
  fclose(f);
}

static const char* arg(int argc, char** argv, const char* name, const char* dflt = nullptr) {
  for (int i = 2; i + 1 < argc; i++) if (!strcmp(argv[i], name)) return argv[i + 1];
  return dflt;
}
static const char* need(int argc, char** argv, const char* name) {
  const char* v = arg(argc, argv, name);
  if (!v) { fprintf(stderr, "ref_driver: missing %s\n", name); exit(2); }
  return v;
}

template <typename PR>
static int run(const std::string& cmd, int argc, char** argv) {
  using Point = typename PR::Point;
  using findex = knn_index<PR, PR, uint>;
  PR Points((char*)need(argc, argv, "-base"));
  std::string out = arg(argc, argv, "-out", "/tmp/ref_out");

  if (cmd == "build") {
    long R = atol(need(argc, argv, "-R")), L = atol(need(argc, argv, "-L"));
    double alpha = atof(arg(argc, argv, "-alpha", "1.2"));
    int passes = atoi(arg(argc, argv, "-passes", "1"));
    bool sortn = atoi(arg(argc, argv, "-sort", "1")) != 0;
    BuildParams BP(R, L, alpha, passes, false);
    Graph<uint> G(BP.max_degree(), Points.size());
    findex I(BP);
    stats<uint> BuildStats(G.size());
    double t0 = now_s();
    I.build_index(G, Points, Points, BuildStats, sortn);
    double t1 = now_s();
    printf("REF_BUILD_SECONDS %.6f\n", t1 - t0);
    G.save((char*)(out + ".graph").c_str());
    dump(out + ".bvisited", BuildStats.visited.begin(), 4 * G.size());
    dump(out + ".bdistances", BuildStats.distances.begin(), 4 * G.size());
    return 0;
  }

  if (cmd == "insert") {
    // knn_index::batch_insert (vamana/index.h:188-316) used directly, three calls on one graph: the first half of the points
    // (given order), the second half (random order), then a scattered subset again with another max_fraction; the graph is
    // dumped after every call
    long R = atol(need(argc, argv, "-R")), L = atol(need(argc, argv, "-L"));
    double alpha = atof(arg(argc, argv, "-alpha", "1.2"));
    BuildParams BP(R, L, alpha, 1, false);
    Graph<uint> G(BP.max_degree(), Points.size());
    findex I(BP);
    I.start_point = 0;   // set_start (index.h:148); build_index does it before its first batch_insert
    stats<uint> BuildStats(G.size());
    const uint n = (uint)Points.size();
    auto dumpg = [&](int step) { G.save((char*)(out + ".g" + std::to_string(step)).c_str()); };
    parlay::sequence<uint> a = parlay::tabulate(n / 2, [&](size_t i) { return (uint)i; });
    I.batch_insert(a, G, Points, Points, BuildStats, alpha, false, 2, .02, false);
    dumpg(0);
    parlay::sequence<uint> b = parlay::tabulate(n - n / 2, [&](size_t i) { return (uint)(n / 2 + i); });
    I.batch_insert(b, G, Points, Points, BuildStats, alpha, true, 2, .02, false);
    dumpg(1);
    parlay::sequence<uint> c = parlay::tabulate(n / 7, [&](size_t i) { return (uint)((i * 7 + 3) % n); });
    I.batch_insert(c, G, Points, Points, BuildStats, 1.0, false, 2, .05, false);
    dumpg(2);
    dump(out + ".bvisited", BuildStats.visited.begin(), 4 * G.size());
    dump(out + ".bdistances", BuildStats.distances.begin(), 4 * G.size());
    return 0;
  }

  if (cmd == "gt") {
    PR Q((char*)need(argc, argv, "-query"));
    int k = atoi(need(argc, argv, "-k"));
    long nq_lim = atol(arg(argc, argv, "-nq", "0"));
    (void)nq_lim;
    double t0 = now_s();
    auto answers = compute_groundtruth<PR>(Points, Q, k);
    double t1 = now_s();
    printf("REF_GT_SECONDS %.6f\n", t1 - t0);
    // raw (pre-sort) buffer order too: ids then dists, as the reference left them
    {
      std::vector<int> ids; std::vector<float> ds;
      for (auto& row : answers) for (auto& e : row) { ids.push_back(e.first); ds.push_back(e.second); }
      dump(out + ".gt_raw_ids", ids.data(), ids.size() * 4);
      dump(out + ".gt_raw_dists", ds.data(), ds.size() * 4);
    }
    write_ibin(answers, out + ".ibin", k);
    return 0;
  }

  if (cmd == "qbuild") {
    // what python/builder.cpp:36-95 intends for 4-byte data (it does not compile at this commit): normalise non-metric
    // points and override alpha (:45-54), build over the 1-byte copy (Euclidian_Point<uint8_t> or
    // Quantized_Mips_Point<8,true>, :60-87), save that graph.
    if constexpr (std::is_same<typename Point::T, float>::value) {
      long R = atol(need(argc, argv, "-R")), L = atol(need(argc, argv, "-L"));
      double alpha = (float)atof(arg(argc, argv, "-alpha", "1.2"));   // builder.cpp takes alpha as a float
      int passes = atoi(arg(argc, argv, "-passes", "1"));
      if (!Point::is_metric()) {
        for (size_t i = 0; i < Points.size(); i++) Points[i].normalize();
        if (Points.dimension() <= 200) alpha = Points.dimension() < 100 ? (float)1.0 : (float).98;
      }
      BuildParams BP(R, L, alpha, passes, false);
      stats<uint> BuildStats(Points.size());
      auto go = [&](auto& Q_Points) {
        using QPR = std::remove_reference_t<decltype(Q_Points)>;
        Graph<uint> G(BP.max_degree(), Q_Points.size());
        knn_index<QPR, QPR, uint> I(BP);
        I.build_index(G, Q_Points, Q_Points, BuildStats, true);
        G.save((char*)(out + ".graph").c_str());
        const size_t d = Q_Points.dimension();
        std::vector<uint8_t> qb(Q_Points.size() * d);
        for (size_t i = 0; i < Q_Points.size(); i++) for (size_t j = 0; j < d; j++) qb[i * d + j] = (uint8_t)Q_Points[i][j];
        dump(out + ".qbase", qb.data(), qb.size());
      };
      if constexpr (std::is_same<Point, Euclidian_Point<float>>::value) {
        PointRange<Euclidian_Point<uint8_t>> Q_Points(Points);
        printf("REF_QUANT slope %.9g offset %d\n", (double)Q_Points.params.slope, (int)Q_Points.params.offset);
        go(Q_Points);
      } else {
        PointRange<Quantized_Mips_Point<8, true>> Q_Points(Points);
        printf("REF_QUANT max_val %.9g\n", (double)Q_Points.params.max_val);
        go(Q_Points);
      }
      return 0;
    } else {
      fprintf(stderr, "ref_driver: qbuild needs -data_type float\n");
      return 2;
    }
  }

  // the remaining commands need a graph and queries
  Graph<uint> G((char*)need(argc, argv, "-graph"));
  PR Q((char*)need(argc, argv, "-query"));
  long k = atol(arg(argc, argv, "-k", "10"));
  long beam = atol(arg(argc, argv, "-beam", "10"));
  double cut = atof(arg(argc, argv, "-cut", "1.35"));
  if (arg(argc, argv, "-cutf")) cut = (double)(float)atof(arg(argc, argv, "-cutf"));  // the sweep's float-rounded cut
  long limit = atol(arg(argc, argv, "-limit", "-1"));
  if (limit < 0) limit = (long)G.size();
  long dl = atol(arg(argc, argv, "-degree_limit", "-1"));
  if (dl < 0) dl = G.max_degree();
  uint start = (uint)atol(arg(argc, argv, "-start", "0"));
  QueryParams QP(k, beam, cut, limit, dl);
  size_t nq = Q.size();

  if (cmd == "search") {
    stats<uint> QS(nq);
    int reps = atoi(arg(argc, argv, "-reps", "1"));
    parlay::sequence<parlay::sequence<uint>> res;
    double best = 1e30;
    for (int r = 0; r < reps; r++) {
      QS.clear();
      double t0 = now_s();
      res = qsearchAll<PR, PR, PR, uint>(Q, Q, Q, G, Points, Points, Points, QS, start, QP);
      double t1 = now_s();
      best = std::min(best, t1 - t0);
    }
    printf("REF_SEARCH_SECONDS %.6f\nREF_SEARCH_QPS %.3f\n", best, nq / best);
    std::vector<uint> ids(nq * k);
    for (size_t i = 0; i < nq; i++) for (long j = 0; j < k; j++) ids[i * k + j] = res[i][j];
    dump(out + ".ids", ids.data(), ids.size() * 4);
    dump(out + ".visited", QS.visited.begin(), 4 * nq);
    dump(out + ".distances", QS.distances.begin(), 4 * nq);
    return 0;
  }

  if (cmd == "random") {
    // beamSearchRandom (beamSearch.h:309-351): one random start per query; the start indices are re-derived with the very
    // expressions of :327-332 and dumped beside the results
    stats<uint> QS(nq);
    auto res = beamSearchRandom<PR, uint>(Q, G, Points, QS, QP);
    std::vector<uint> ids(nq * k), st(nq);
    for (size_t i = 0; i < nq; i++) for (long j = 0; j < k; j++) ids[i * k + j] = res[i][j];
    parlay::random_generator gen;
    std::uniform_int_distribution<long> dis(0, G.size() - 1);
    for (size_t i = 0; i < nq; i++) { auto r = gen[i]; st[i] = (uint)dis(r); }
    dump(out + ".ids", ids.data(), ids.size() * 4);
    dump(out + ".starts", st.data(), st.size() * 4);
    dump(out + ".visited", QS.visited.begin(), 4 * nq);
    dump(out + ".distances", QS.distances.begin(), 4 * nq);
    return 0;
  }

  if (cmd == "qsearch") {
    // float Euclidean data only: search the reference's own uint8 scalar-quantised copy (euclidian_point.h:182-235,
    // point_range.h:50-63) and re-rank with the exact points, as ANN() does for -quantize_mode 1
    if constexpr (std::is_same<typename Point::T, float>::value && std::is_same<Point, Euclidian_Point<float>>::value) {
      using QPoint = Euclidian_Point<uint8_t>;
      using QPR = PointRange<QPoint>;
      QPR Q_Points(Points);
      QPR Q_Q(Q, Q_Points.params);
      stats<uint> QS(nq);
      auto res = qsearchAll<PR, QPR, QPR, uint>(Q, Q_Q, Q_Q, G, Points, Q_Points, Q_Points, QS, start, QP);
      std::vector<uint> ids(nq * k);
      for (size_t i = 0; i < nq; i++) for (long j = 0; j < k; j++) ids[i * k + j] = res[i][j];
      dump(out + ".ids", ids.data(), ids.size() * 4);
      dump(out + ".visited", QS.visited.begin(), 4 * nq);
      dump(out + ".distances", QS.distances.begin(), 4 * nq);
      const size_t d = Points.dimension();
      std::vector<uint8_t> qb(Points.size() * d), qq(nq * d);
      for (size_t i = 0; i < Points.size(); i++) for (size_t j = 0; j < d; j++) qb[i * d + j] = Q_Points[i][j];
      for (size_t i = 0; i < nq; i++) for (size_t j = 0; j < d; j++) qq[i * d + j] = Q_Q[i][j];
      dump(out + ".qbase", qb.data(), qb.size());
      dump(out + ".qquery", qq.data(), qq.size());
      printf("REF_QUANT slope %.9g offset %d\n", (double)Q_Points.params.slope, (int)Q_Points.params.offset);
      return 0;
    } else if constexpr (std::is_same<Point, Mips_Point<float>>::value) {
      // python/graph_index.cpp:88-95, :166-181: points and queries normalised, Quantized_Mips_Point<8,true> copies,
      // beam_search_rerank with the exact (normalised) points (dims <= 200: no second-level filter)
      for (size_t i = 0; i < Points.size(); i++) Points[i].normalize();
      for (size_t i = 0; i < nq; i++) Q[i].normalize();
      using QPoint = Quantized_Mips_Point<8, true>;
      using QPR = PointRange<QPoint>;
      QPR Q_Points(Points);
      QPR Q_Q(Q, Q_Points.params);
      stats<uint> QS(nq);
      auto res = qsearchAll<PR, QPR, QPR, uint>(Q, Q_Q, Q_Q, G, Points, Q_Points, Q_Points, QS, start, QP);
      std::vector<uint> ids(nq * k);
      for (size_t i = 0; i < nq; i++) for (long j = 0; j < k; j++) ids[i * k + j] = res[i][j];
      dump(out + ".ids", ids.data(), ids.size() * 4);
      dump(out + ".visited", QS.visited.begin(), 4 * nq);
      dump(out + ".distances", QS.distances.begin(), 4 * nq);
      const size_t d = Points.dimension();
      std::vector<uint8_t> qb(Points.size() * d), qq(nq * d);
      std::vector<float> nb(64 * d), nqv(nq * d);
      for (size_t i = 0; i < Points.size(); i++) for (size_t j = 0; j < d; j++) qb[i * d + j] = (uint8_t)Q_Points[i][j];
      for (size_t i = 0; i < nq; i++) for (size_t j = 0; j < d; j++) { qq[i * d + j] = (uint8_t)Q_Q[i][j]; nqv[i * d + j] = Q[i][j]; }
      for (size_t i = 0; i < 64 && i < Points.size(); i++) for (size_t j = 0; j < d; j++) nb[i * d + j] = Points[i][j];
      dump(out + ".qbase", qb.data(), qb.size());
      dump(out + ".qquery", qq.data(), qq.size());
      dump(out + ".nbase_head", nb.data(), nb.size() * 4);
      dump(out + ".nquery", nqv.data(), nqv.size() * 4);
      printf("REF_QUANT max_val %.9g\n", (double)Q_Points.params.max_val);
      return 0;
    } else {
      fprintf(stderr, "ref_driver: qsearch needs -data_type float\n");
      return 2;
    }
  }

  if (cmd == "full") {
    // frontier and visited lists of each query, as (id u32, dist f32) pairs
    // layout: per query  u32 nf, u32 nv, u32 cmps, then nf pairs, then nv pairs
    int self = atoi(arg(argc, argv, "-self", "0"));  // 1: queries are base rows (build-mode same_as)
    std::vector<std::vector<uint>> rows(nq);
    parlay::parallel_for(0, nq, [&](size_t i) {
      parlay::sequence<uint> starts = {start};
      auto r = self ? beam_search(Points[i], G, Points, starts, QP)
                    : beam_search(Q[i], G, Points, starts, QP);
      auto& fr = r.first.first; auto& vis = r.first.second;
      auto& o = rows[i];
      o.push_back(fr.size()); o.push_back(vis.size()); o.push_back((uint)r.second);
      auto put = [&](std::pair<uint, float> e) { uint b; memcpy(&b, &e.second, 4); o.push_back(e.first); o.push_back(b); };
      for (auto e : fr) put(e);
      for (auto e : vis) put(e);
    });
    std::vector<uint> flat;
    for (auto& r : rows) flat.insert(flat.end(), r.begin(), r.end());
    dump(out + ".full", flat.data(), flat.size() * 4);
    return 0;
  }

  if (cmd == "recall") {
    // -settings "k:beam:cut:limit:degree_limit;..." runs several QueryParams in one process (limit/degree_limit < 0 = n / maxDeg)
    groundTruth<uint> GT((char*)need(argc, argv, "-gt"));
    int reps = atoi(arg(argc, argv, "-reps", "1"));
    std::vector<QueryParams> qps;
    if (const char* st = arg(argc, argv, "-settings")) {
      std::string all = st;
      size_t pos = 0;
      while (pos < all.size()) {
        size_t e = all.find(';', pos);
        if (e == std::string::npos) e = all.size();
        long a, b, l, d; double c;
        if (sscanf(all.substr(pos, e - pos).c_str(), "%ld:%ld:%lf:%ld:%ld", &a, &b, &c, &l, &d) != 5) { fprintf(stderr, "bad -settings\n"); return 2; }
        qps.push_back(QueryParams(a, b, c, l < 0 ? (long)G.size() : l, d < 0 ? (long)G.max_degree() : d));
        pos = e + 1;
      }
    } else {
      qps.push_back(QP);
    }
    for (auto& qp : qps)
      for (int r = 0; r < reps; r++) {
        nn_result N = checkRecall<PR, PR, PR, uint>(G, Points, Q, Points, Q, Points, Q, GT, false, start, qp.k, qp, false);
        printf("REF_RECALL %.9f QPS %.3f avg_cmps %u tail_cmps %u avg_visited %u tail_visited %u k %ld beam %ld limit %ld dl %ld\n",
               N.recall, (double)N.QPS, N.avg_cmps, N.tail_cmps, N.avg_visited, N.tail_visited, qp.k, qp.beamSize, qp.limit, qp.degree_limit);
        fflush(stdout);
      }
    return 0;
  }
  fprintf(stderr, "ref_driver: unknown command %s\n", cmd.c_str());
  return 2;
}

int main(int argc, char** argv) {
  if (argc < 2) { fprintf(stderr, "usage: ref_driver <perm|build|search|full|recall|gt> ...\n"); return 2; }
  std::string cmd = argv[1];
  if (cmd == "selftest") { printf("REF_DRIVER_OK threads=%zu\n", parlay::num_workers()); return 0; }
  if (cmd == "perm") {
    int n = atoi(need(argc, argv, "-n"));
    auto p = parlay::random_permutation<int>(n);
    dump(arg(argc, argv, "-out", "/tmp/ref_perm.bin"), p.begin(), 4 * (size_t)n);
    return 0;
  }
  std::string dt = need(argc, argv, "-data_type"), df = arg(argc, argv, "-dist_func", "Euclidian");
  bool l2 = (df == "Euclidian");
  if (!l2 && df != "mips") { fprintf(stderr, "bad -dist_func\n"); return 2; }
  if (dt == "uint8") return l2 ? run<PointRange<Euclidian_Point<uint8_t>>>(cmd, argc, argv) : run<PointRange<Mips_Point<uint8_t>>>(cmd, argc, argv);
  if (dt == "int8")  return l2 ? run<PointRange<Euclidian_Point<int8_t>>>(cmd, argc, argv)  : run<PointRange<Mips_Point<int8_t>>>(cmd, argc, argv);
  if (dt == "float") return l2 ? run<PointRange<Euclidian_Point<float>>>(cmd, argc, argv)   : run<PointRange<Mips_Point<float>>>(cmd, argc, argv);
  fprintf(stderr, "bad -data_type\n");
  return 2;
}
