// oracle/ref_hnsw.cpp — TEST INFRASTRUCTURE, not product code.
//
// A thin command-line driver around the reference's HNSW (algorithms/HNSW/HNSW.hpp, included from a build-time copy
// under oracle/_ref/patched/ — see oracle/Makefile: the copy differs from /root/reference in ONE line, the memory-safety
// fix SURVEY.md §8c documents: beamSearch.h:74 sizes unvisited_frontier max(beamSize, frontier.size()) instead of
// beamSize, which the ef = 1 descent overflows as soon as there are two entrances).  No reference source is committed.
//
//   build  -> ANN::HNSW<Desc_HNSW<T,Point>>(begin, end, dim, m_l, m, efc, alpha) + save     (HNSW.hpp:746-824, :1615-1686)
//   search -> HNSW(model, getter) + search(q, k, ef, ctrl) per query                         (HNSW.hpp:631-744, :1570-1613)
//
// With PARLAY_NUM_THREADS=1 the build is deterministic: levels come from mt19937{worker_id = 0} in node order.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "utils/euclidian_point.h"
#include "utils/mips_point.h"
#include "utils/point_range.h"
#include "utils/types.h"
#include "HNSW/HNSW.hpp"

using namespace parlayANN;

static double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
static const char* arg(int argc, char** argv, const char* name, const char* dflt = nullptr) {
  for (int i = 2; i + 1 < argc; i++) if (!strcmp(argv[i], name)) return argv[i + 1];
  return dflt;
}
static const char* need(int argc, char** argv, const char* name) {
  const char* v = arg(argc, argv, name);
  if (!v) { fprintf(stderr, "ref_hnsw: missing %s\n", name); exit(2); }
  return v;
}
static void dump(const std::string& path, const void* p, size_t bytes) {
  FILE* f = fopen(path.c_str(), "wb");
  if (!f) { perror(path.c_str()); exit(2); }
  if (bytes && fwrite(p, 1, bytes, f) != bytes) { perror("fwrite"); exit(2); }
  fclose(f);
}

template <typename T, typename Point>
static int run(const std::string& cmd, int argc, char** argv) {
  using desc = Desc_HNSW<T, Point>;
  PointRange<Point> Points((char*)need(argc, argv, "-base"));
  if (cmd == "build") {
    uint32_t m = atoi(need(argc, argv, "-m")), efc = atoi(need(argc, argv, "-efc"));
    float m_l = atof(arg(argc, argv, "-m_l", "0.36")), alpha = atof(arg(argc, argv, "-alpha", "1.0"));
    auto ps = parlay::delayed_seq<Point>(Points.size(), [&](size_t i) { return Points[i]; });   // python/builder.cpp:225-233
    double t0 = now_s();
    auto G = ANN::HNSW<desc>(ps.begin(), ps.end(), Points.dimension(), m_l, m, efc, alpha);
    double t1 = now_s();
    printf("REF_HNSW_BUILD_SECONDS %.6f\n", t1 - t0);
    G.save(need(argc, argv, "-out"));
    return 0;
  }
  if (cmd == "search") {
    PointRange<Point> Q((char*)need(argc, argv, "-query"));
    uint32_t k = atoi(arg(argc, argv, "-k", "10")), ef = atoi(arg(argc, argv, "-ef", "50"));
    long limit = atol(arg(argc, argv, "-limit", "-1"));
    int reps = atoi(arg(argc, argv, "-reps", "1"));
    ANN::HNSW<desc> G(need(argc, argv, "-model"), [&](unsigned int i) { return Points[i]; });   // python/graph_index.cpp:103-109
    size_t nq = Q.size();
    std::vector<std::vector<std::pair<uint32_t, float>>> res(nq);
    std::vector<uint32_t> cmps(nq);
    double best = 1e30;
    for (int r = 0; r < reps; r++) {
      double t0 = now_s();
      parlay::parallel_for(0, nq, [&](size_t i) {
        uint32_t c = 0;
        search_control ctrl{};
        if (limit > 0) ctrl.limit_eval = (uint32_t)limit;
        ctrl.count_cmps = &c;
        auto R = G.search(Q[i], k, ef, ctrl);
        res[i].assign(R.begin(), R.end());
        cmps[i] = c;
      });
      best = std::min(best, now_s() - t0);
    }
    printf("REF_HNSW_SEARCH_SECONDS %.6f\nREF_HNSW_SEARCH_QPS %.3f\n", best, nq / best);
    // layout: per query u32 count, then count (id u32, dist f32 bits) pairs
    std::vector<uint32_t> flat;
    for (size_t i = 0; i < nq; i++) {
      flat.push_back((uint32_t)res[i].size());
      for (auto& e : res[i]) { uint32_t b; memcpy(&b, &e.second, 4); flat.push_back(e.first); flat.push_back(b); }
    }
    std::string out = need(argc, argv, "-out");
    dump(out + ".res", flat.data(), flat.size() * 4);
    dump(out + ".cmps", cmps.data(), cmps.size() * 4);
    return 0;
  }
  fprintf(stderr, "ref_hnsw: unknown command %s\n", cmd.c_str());
  return 2;
}

int main(int argc, char** argv) {
  if (argc < 2) { fprintf(stderr, "usage: ref_hnsw <build|search|selftest> ...\n"); return 2; }
  std::string cmd = argv[1];
  if (cmd == "selftest") { printf("REF_HNSW_OK threads=%zu\n", parlay::num_workers()); return 0; }
  std::string dt = need(argc, argv, "-data_type"), df = arg(argc, argv, "-dist_func", "Euclidian");
  bool l2 = (df == "Euclidian");
  if (dt == "uint8") return l2 ? run<uint8_t, Euclidian_Point<uint8_t>>(cmd, argc, argv) : run<uint8_t, Mips_Point<uint8_t>>(cmd, argc, argv);
  if (dt == "int8") return l2 ? run<int8_t, Euclidian_Point<int8_t>>(cmd, argc, argv) : run<int8_t, Mips_Point<int8_t>>(cmd, argc, argv);
  if (dt == "float") return l2 ? run<float, Euclidian_Point<float>>(cmd, argc, argv) : run<float, Mips_Point<float>>(cmd, argc, argv);
  fprintf(stderr, "bad -data_type\n");
  return 2;
}
