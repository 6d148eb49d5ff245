#!/usr/bin/env python
"""bench.py — QPS at recall@10 >= 0.95 of the batched beam search (BASELINE.json metric), one JSON line.

Workload (config.workload): the configuration the metric is quoted on — Vamana search over a synthetic BIGANN-shaped
10M x 128 uint8 set (BASELINE.json configs[3]; 3.9 GB of points + graph, fits one GPU), graph R=64 L=128 alpha=1.2 built
ON THE GPU by pann_vamana_build (the build time is reported too, with its own roofline and CPU baseline).  A step is one
pass of the search over a batch of 1 000 000 queries per GPU (the 10 000-query set the reference ships is
two waves of warps on one B200; throughput and scaling are quoted on the 1M batch, recall and parity on the 10 000-query
batch).  QueryParams are chosen by the reference's own sweep (check_nn_recall.h:214-257) on the 10 000-query recall batch
against exact ground truth from K4: the cheapest setting whose tie-aware recall@10 is >= 0.95.
`--n 1000000 --nq 10000` is configs[1].

  value     whole-job QPS with queries resident in HBM (device time from CUDA events on the library's stream)
  e2e       the same through pann_search_batch with pinned HOST buffers: H2D queries + search + D2H ids/dists/counters
  roofline  algorithmic bytes of the batch (counted on the device) / K1s' device time, vs measured HBM
  parity    the reference itself (oracle/_ref/ref_driver -> qsearchAll) on the 10 000-query batch at the chosen QueryParams:
            ids, visited and distance-comparison counters compared PER QUERY with the GPU's
  timed_batch_recall  tie-aware recall@10 of a 10 000-row slice of the TIMED 1M batch (exact ground truth from K4)
  recall_batch   the same search over the 10 000-query batch alone, device-resident and end to end
  strong    a FIXED batch (1M queries, and the 10 000-query batch) cut over the N ranks through search_sharded(gather=True):
            H2D + search + NCCL all-gather of the answer + D2H on every rank, max over ranks
  build     pann_vamana_build seconds, phase split, roofline (B_ins counted on the device), and the reference's CPU build
            timed next to the GPU build at the largest n that fits ~1 CPU-minute
  groundtruth   K4 over the recall batch (tensor roofline, CPU baseline on a 1 000-query sample)
  groundtruth_c3  BASELINE configs[2]: 10 000 x 10M x 96 f32, k=100, base rows sharded over the N ranks, NCCL merge
  hnsw_c5   BASELINE configs[4]: HNSW build + search, 1M x 200 f32 inner product, with the reference's HNSW on the host cores
  cpu_baseline  the unmodified reference's qsearchAll on this box's cores, same graph / queries / QueryParams

`--impl reference` times only the reference's CPU search (10 000-query sample per step).  Up to 2M points it also builds
the graph and the ground truth with the reference's own code; at 10M those would take many minutes of CPU, so the inputs
are prepared on the GPU (untimed) when one is present.  `--workload groundtruth` prints the C3 line alone.
Multi-GPU (torchrun): one rank per GPU, replicated index, every rank searches its own 1M-query batch
(weak scaling, no data-path collective); time = max over ranks.
"""
import argparse
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

K = 10
CUT_F = float(np.float32(1.35))   # the sweep passes the cut through a float (check_nn_recall.h:231)
BEAMS = [10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 22, 24, 26, 28, 30, 32, 34, 36, 38, 40, 45, 50, 55, 60, 65, 70, 80, 90, 100,
         120, 140, 160, 180, 200, 225, 250, 275, 300, 375, 500, 750, 1000]
LIMITS = [10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21, 22, 23, 24, 25, 26, 28, 30, 35]


def sweep_settings(maxdeg):
    """(kind, k, beam, cut, limit, degree_limit) in the order of search_and_parse (check_nn_recall.h:214-254)."""
    out = [("beam", K, b, CUT_F, -1, -1) for b in BEAMS if b >= K]
    out += [("limit", K, max(l, K), 1.35, l, min(maxdeg, 5 * l)) for l in LIMITS]
    return out


def recall_with_ties(ids, gt_ids, gt_d, k):
    """check_nn_recall.h:83-109, vectorised."""
    kk = gt_ids.shape[1]
    col = np.arange(kk)[None, :]
    truth = (col < k) | (gt_d == gt_d[:, k - 1:k])
    hit = ((ids[:, :k, None] == gt_ids[:, None, :]) & truth[:, None, :]).any(axis=2)
    return float(np.float32(hit.sum()) / np.float32(k * ids.shape[0]))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.gpu, self.lines, self.p = gpu, [], None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.gpu)],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=lambda: [self.lines.append(l) for l in self.p.stdout], daemon=True).start()
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        sm, mx, reasons = [], [], set()
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------------------------------------------
# the reference's CPU path (oracle/_ref/ref_driver = the unmodified reference headers); used for cpu_baseline, parity and
# --impl reference.  The only place bench.py touches oracle/.
def ref_tools():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    drv = O.ref_driver_path()
    if drv is None:
        raise RuntimeError("oracle/_ref/ref_driver missing or not runnable on this CPU (build it where /root/reference exists)")
    return O, drv


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def ref_recall_runs(O, files, settings, reps, threads=None):
    """runs checkRecall (timed qsearchAll + tie-aware recall) for each setting; returns list of dict per (setting, rep)"""
    st = ";".join(f"{k}:{b}:{c!r}:{l}:{d}" for (_, k, b, c, l, d) in settings)
    out = O.ref_run(["recall", "-data_type", "uint8", "-dist_func", "Euclidian", "-base", files["base"], "-graph", files["graph"],
                     "-query", files["query"], "-gt", files["gt"], "-settings", st, "-reps", reps], threads=threads)
    rows = []
    for m in re.finditer(r"REF_RECALL (\S+) QPS (\S+) avg_cmps (\d+) tail_cmps (\d+) avg_visited (\d+) tail_visited (\d+) k (\d+) beam (\d+) limit (\d+) dl (\d+)", out):
        rows.append(dict(recall=float(m.group(1)), qps=float(m.group(2)), avg_cmps=int(m.group(3)), avg_visited=int(m.group(5)),
                         beam=int(m.group(8)), limit=int(m.group(9)), dl=int(m.group(10))))
    assert len(rows) == len(settings) * reps, out[-2000:]
    return rows


def ref_search_dump(O, files, s, tmp, nq):
    """the reference's qsearchAll at one setting with its raw per-query outputs (ids, stats.visited, stats.distances)"""
    kind, k, beam, cut, limit, dl = s
    out = os.path.join(tmp, "refsearch")
    O.ref_run(["search", "-data_type", "uint8", "-dist_func", "Euclidian", "-base", files["base"], "-graph", files["graph"],
               "-query", files["query"], "-k", k, "-beam", beam, "-cut", repr(cut), "-limit", limit, "-degree_limit", dl, "-out", out])
    return (np.fromfile(out + ".ids", dtype=np.uint32).reshape(nq, k), np.fromfile(out + ".visited", dtype=np.uint32),
            np.fromfile(out + ".distances", dtype=np.uint32))


def choose_setting(settings, evaluate):
    """first beam setting and first limit setting with recall >= 0.95 (both lists ascend in cost); the cheaper one
    by average distance comparisons wins.  evaluate(setting) -> (recall, avg_cmps)."""
    best = None
    for kind in ("beam", "limit"):
        for s in [x for x in settings if x[0] == kind]:
            r, c = evaluate(s)
            if r >= 0.95:
                if best is None or c < best[2]:
                    best = (s, r, c)
                break
    if best is None:
        raise RuntimeError("no sweep setting reaches recall 0.95")
    return best


def device_sweep(idx, settings, n, maxdeg):
    """every setting of the sweep through pann_search_sweep (resident queries + resident ground truth) -> ({setting: (recall, avg_cmps)}, rows)"""
    from parlayann_b200 import harness
    rows = harness.search_sweep(idx, [dict(k=k, beam=b, cut=c, limit=n if l < 0 else l, degree_limit=maxdeg if d < 0 else d) for (_, k, b, c, l, d) in settings])
    bad = [r for r in rows if r["status"] != 0]
    if bad:
        raise RuntimeError(f"sweep setting failed: {bad[0]}")
    return {s: (r["recall"], r["avg_cmps"]) for s, r in zip(settings, rows)}, rows


def gpu_prepared_inputs(args, tmp, local_rank=0):
    """points / graph / recall queries / ground truth for the reference arm at sizes where the reference's own CPU build
    and ground truth would take many minutes (10M points): made on the GPU, untimed, written in the reference's file
    formats.  The graph is the same bytes the reference builds (tests/test_build_gpu.py, tests/test_fullsize_gpu.py)."""
    import torch
    import parlayann_b200 as P
    from parlayann_b200 import fileio
    torch.cuda.set_device(local_rank)
    idx = P.GraphIndex(data_type="uint8", metric="Euclidian", max_degree=64, device=local_rank,
                       synthetic=dict(n=args.n, d=128, kind="uint8", seed=1234, stream=0, r=24, nc=256))
    idx.build(R=64, L=128, alpha=1.2, num_passes=1, sort_neighbors=True)
    idx.synthetic_queries(args.nq_recall, kind="uint8", seed=1234, stream=1)
    qs = idx.download_queries()
    idx.groundtruth_resident(k=100)
    gt_ids, gt_d = idx.fetch_groundtruth()
    cache, _ = device_sweep(idx, sweep_settings(64), args.n, 64)
    s, rec, cm = choose_setting(sweep_settings(64), lambda x: cache[x])
    files = {k: os.path.join(tmp, k) for k in ("base", "query", "graph", "gt")}
    fileio.write_bin(files["base"], idx.points()); fileio.write_bin(files["query"], qs)
    fileio.write_graph(files["graph"], idx.graph()); fileio.write_ibin(files["gt"], gt_ids, gt_d)
    idx.close()
    return files, s


def run_reference(args, rank, world):
    """--impl reference: the reference's own qsearchAll on the host cores, timed by its own checkRecall.  Each step is a
    bounded sample of the workload: the 10 000-query recall batch (same generator, same QueryParams)."""
    if rank != 0:
        return
    O, drv = ref_tools()
    from parlayann_b200 import fileio
    tmp = tempfile.mkdtemp(prefix="pann_ref_")
    try:
        t0 = time.time()
        build_s = gt_s = None
        gpu_setup = False
        if args.n > 2000000:
            try:
                import torch
                gpu_setup = torch.cuda.is_available()
            except Exception:
                gpu_setup = False
        if gpu_setup:
            files, s = gpu_prepared_inputs(args, tmp)
            inputs = "points, graph, ground truth and the QueryParams choice prepared on the GPU (untimed); the timed search is the reference's"
        else:
            base = O.gen_points("uint8", args.n, 128)
            qs = O.gen_points("uint8", args.nq_recall, 128, stream=1)
            files = {k: os.path.join(tmp, k) for k in ("base", "query", "graph", "gt")}
            fileio.write_bin(files["base"], base); fileio.write_bin(files["query"], qs)
            common = ["-data_type", "uint8", "-dist_func", "Euclidian", "-base", files["base"]]
            out = O.ref_run(["build"] + common + ["-R", 64, "-L", 128, "-alpha", 1.2, "-out", os.path.join(tmp, "b")])
            build_s = float(re.search(r"REF_BUILD_SECONDS (\S+)", out).group(1))
            os.replace(os.path.join(tmp, "b.graph"), files["graph"])
            out = O.ref_run(["gt"] + common + ["-query", files["query"], "-k", 100, "-out", os.path.join(tmp, "g")])
            gt_s = float(re.search(r"REF_GT_SECONDS (\S+)", out).group(1))
            os.replace(os.path.join(tmp, "g.ibin"), files["gt"])
            cache = {}

            def ev(s):
                if s not in cache:
                    r = ref_recall_runs(O, files, [s], 1)[0]
                    cache[s] = (r["recall"], r["avg_cmps"])
                return cache[s]
            s = choose_setting(sweep_settings(64), ev)[0]
            inputs = "everything (points file, graph, ground truth, sweep) made by the reference's own code on the CPU"
        rows = ref_recall_runs(O, files, [s], args.warmup + args.steps)[args.warmup:]
        rec, cm = rows[0]["recall"], rows[0]["avg_cmps"]
        ms = [1e3 * args.nq_recall / r["qps"] for r in rows]
        ms_per_step = float(np.mean(ms))
        qps = args.nq_recall / (ms_per_step * 1e-3)
        cores = os.cpu_count()
        line = {"impl": "reference", "metric": "QPS@recall10>=0.95", "value": qps, "unit": "queries/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                "config": workload_config(args, s, rec, cm),
                "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": cores, "kind": "reference", "cpu": cpu_model(), "binary": O.ref_driver_info(),
                                 "sample": f"{args.nq_recall} queries of the batch per step, {args.steps} steps, qsearchAll via checkRecall"},
                "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0, "inputs": inputs,
                "reference_build_seconds": build_s, "reference_groundtruth_seconds": gt_s, "setup_seconds": time.time() - t0}
        print(json.dumps(line), flush=True)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def workload_config(args, s, rec, cm):
    """identical in both arms: the recall is printed to 6 decimals (the arms compute it in different float widths)"""
    kind, k, beam, cut, limit, dl = s
    return {"workload": f"vamana_search_{args.n}x128_u8_{args.nq}q_per_gpu", "n": args.n, "dims": 128, "queries_per_gpu_per_step": args.nq,
            "recall_batch": args.nq_recall, "graph": "Vamana R=64 L=128 alpha=1.2 1 pass", "k": k, "beam": beam, "cut": round(cut, 6),
            "visit_limit": limit if limit >= 0 else args.n, "degree_limit": dl if dl >= 0 else 64, "recall10": round(float(rec), 6), "avg_cmps": int(cm),
            "parallelism": f"replicated index, queries sharded x{args.gpus}",
            "l2": f"index {(args.n * (128 + 260)) >> 20} MB and each step's gathers >> 126 MB L2; 256 MiB flush between steps as well"}


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


# ----------------------------------------------------------------------------------------------------------------
class Dist:
    """torch.distributed plumbing: barrier + max over ranks"""

    def __init__(self, local_rank, world):
        import torch
        self.torch, self.world, self.d = torch, world, None
        torch.cuda.set_device(local_rank)
        if world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            self.d = dist

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.d is not None:
            self.d.barrier()
        self.torch.cuda.synchronize()

    def allmax(self, x):
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        if self.d is not None:
            self.d.all_reduce(t, op=self.d.ReduceOp.MAX)
        return float(t.item())

    def close(self):
        if self.d is not None:
            self.d.barrier()
            self.d.destroy_process_group()


def groundtruth_c3(D, rank, local_rank, world, n=10_000_000, d=96, nq=10_000, k=100, reps=4):
    """BASELINE configs[2]: brute-force ground truth k=100 over synthetic 10M x 96 float (Deep-shaped), base rows sharded over
    the ranks, NCCL all-gather of the local top-k + K5 merge on every rank (compute_groundtruth.cpp:22-100)."""
    import parlayann_b200 as P
    from parlayann_b200.sharded import shard_bounds, groundtruth_sharded
    lo, hi = shard_bounds(n, rank, world)
    idx = P.GraphIndex(data_type="float", metric="Euclidian", max_degree=1, device=local_rank,
                       synthetic=dict(n=hi - lo, d=d, kind="float_norm", seed=1234, stream=0, row0=lo))
    idx.synthetic_queries(nq, kind="float_norm", seed=1234, stream=1)
    times, local_ms = [], []
    gi = gd = None
    for _ in range(reps):
        D.barrier()
        t0 = time.perf_counter()
        gi, gd = groundtruth_sharded(idx, None, k=k, base_offset=lo)
        D.torch.cuda.synchronize()
        times.append(D.allmax(time.perf_counter() - t0))
        local_ms.append(D.allmax(sum(getattr(idx, '_gt_leg_ms', None) or [idx.last_timing()[0]])))
    idx.close()
    best = min(times[1:])
    peaks = load_peaks()
    bf16 = float(peaks.get("bf16_tflops", 1600.0))
    tf32_peak = bf16 / 2.0     # kind::tf32 runs at half the bf16 rate (no measured tf32 number in MEASURED_PEAKS.json)
    flops = 2.0 * nq * n * d
    return {"what": f"ground truth {nq} x {n} x {d} f32, k={k}, base rows sharded x{world} (NCCL all-gather of the local top-k + K5 merge)",
            "n_gpus": world, "seconds": best, "local_device_ms": min(local_ms[1:]), "all_seconds": times,
            "roofline": {"bound": "tensor", "achieved": flops / best / 1e12 / world, "peak": tf32_peak, "unit": "TFLOP/s", "per": "GPU",
                         "frac": flops / best / 1e12 / world / tf32_peak, "peak_source": "MEASURED_PEAKS.json bf16_tflops / 2 (tf32 pipe)",
                         "algorithmic_flop": flops, "kernel": "pann_gt_mma_kernel<false> (tcgen05 kind::tf32) + pann_gt_shrink_kernel"},
            "checksum": int(gi.astype(np.int64).sum() % 1000003), "dist_checksum": float(gd.astype(np.float64).sum())}


def hnsw_c5(args, local_rank, with_cpu=True, n=1_000_000, d=200, nq=10_000, m=32, efc=128, m_l=0.36, alpha=1.2):
    """BASELINE configs[4]: HNSW build + search, synthetic 1M x 200 float inner product (Text2Image-shaped: latent r = 32, no
    clusters).  Build on the GPU (pann_hnsw_build, HNSW.hpp:746-1014), exact ground truth from K4, the smallest ef of a fixed
    list whose recall@10 is >= 0.95, search timed device-side and end to end; next to it the reference's own HNSW on the host
    cores: its search over the SAME model file (ids compared per query) and its build on a bounded prefix of the points."""
    import parlayann_b200 as P
    from parlayann_b200 import fileio
    from parlayann_b200.hnsw import HNSW
    peaks = load_peaks()
    peak = float(peaks.get("hbm_gbs", 6650.0))
    gen = dict(d=d, kind="float", seed=1234, stream=0, r=32, nc=1)
    h = HNSW(None, "float", "mips", m=m, efc=efc, m_l=m_l, alpha=alpha, device=local_rank, synthetic=dict(n=n, **gen))
    h.build()                                  # warm-up (allocations, module load)
    t0 = time.time()
    sec = h.build()
    build_s = time.time() - t0
    M = h.model()
    h.index.synthetic_queries(nq, kind="float", seed=1234, stream=1, r=32, nc=1)
    qs = h.index.download_queries()
    h.index.groundtruth_resident(k=100)
    gi, gd = h.index.fetch_groundtruth()
    chosen = None
    for ef in (10, 12, 14, 16, 20, 24, 28, 32, 40, 50, 64, 80, 100, 128, 160, 200):
        ids, dists, cnt, cmps = h.search(qs, k=10, ef=ef)
        rec = recall_with_ties(ids, gi.astype(np.uint32), gd, K)
        if rec >= 0.95:
            chosen = (ef, rec)
            break
    if chosen is None:
        chosen = (ef, rec)
    ef, rec = chosen
    dev, e2e = [], []
    for i in range(3 + 10):
        t0 = time.perf_counter()
        ids, dists, cnt, cmps = h.search(qs, k=10, ef=ef)
        if i >= 3:
            e2e.append((time.perf_counter() - t0) * 1e3); dev.append(h.index.last_timing()[0])
    dev_ms, e2e_ms = float(np.mean(dev)), float(np.mean(e2e))
    abytes = float(cmps.astype(np.int64).sum()) * d * 4 + nq * (d * 4 + 4 * K)
    out = {"what": f"HNSW {n} x {d} f32 inner product, m={m} efc={efc} m_l={m_l} alpha={alpha}; {nq} queries, k=10",
           "build": {"seconds": build_s, "phase_seconds": {"layer_searches": sec[1], "prune_heuristic+forward": sec[2], "reverse_edges": sec[3]},
                     "levels": np.bincount(M.level).tolist(), "entrances": int(len(M.entrance)), "edges": int(len(M.edges))},
           "search": {"ef": int(ef), "recall10": round(rec, 6), "value": nq / (dev_ms * 1e-3), "unit": "queries/s", "device_ms": dev_ms, "avg_cmps": float(cmps.mean()),
                      "e2e": {"value": nq / (e2e_ms * 1e-3), "unit": "queries/s", "ms": e2e_ms, "h2d_bytes_per_step": int(qs.nbytes),
                              "d2h_bytes_per_step": int(2 * nq * ef * 4 + 8 * nq)},
                      "roofline": {"bound": "hbm", "achieved": abytes / (dev_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "frac": abytes / (dev_ms * 1e-3) / 1e9 / peak,
                                   "what": "distance comparisons x row bytes (all layers) + query + result per query / device time of the 5 layer launches"}}}
    if with_cpu:
        try:
            O, _ = ref_tools()
            if O.ref_hnsw_path() is None:
                raise RuntimeError("oracle/_ref/ref_hnsw missing")
            tmp = tempfile.mkdtemp(prefix="pann_hnsw_")
            try:
                bp, qp, mp = (os.path.join(tmp, x) for x in ("base.fbin", "q.fbin", "model.hnsw"))
                fileio.write_bin(bp, h.index.points()); fileio.write_bin(qp, qs); h.save(mp)
                common = ["-data_type", "float", "-dist_func", "mips", "-base", bp]
                o = O.ref_hnsw_run(["search"] + common + ["-model", mp, "-query", qp, "-k", 10, "-ef", ef, "-reps", 3, "-out", os.path.join(tmp, "s")])
                ref_qps = float(re.search(r"REF_HNSW_SEARCH_QPS (\S+)", o).group(1))
                rows = O.parse_hnsw_results(os.path.join(tmp, "s.res"), nq)
                rc = np.fromfile(os.path.join(tmp, "s.cmps"), dtype=np.uint32)
                same = sum(int(len(a) >= 10 and np.array_equal(a[:10], ids[i])) for i, (a, b) in enumerate(rows))
                out["search"]["cpu_baseline"] = {"value": ref_qps, "unit": "queries/s", "cores": os.cpu_count(), "kind": "reference",
                                                 "sample": f"HNSW::search of the unmodified reference (one-line unvisited_frontier fix) over the GPU-built model file, {nq} queries, best of 3"}
                out["search"]["parity"] = {"queries": nq, "queries_with_equal_ids": int(same), "cmps_equal": bool(np.array_equal(rc, cmps))}
                nb = min(n, args.n_hnsw_cpu)
                pb = os.path.join(tmp, "prefix.fbin")
                fileio.write_bin(pb, h.index.points(0, nb))
                o = O.ref_hnsw_run(["build", "-data_type", "float", "-dist_func", "mips", "-base", pb, "-m", m, "-efc", efc, "-m_l", m_l, "-alpha", alpha, "-out", os.path.join(tmp, "pm")])
                ref_b = float(re.search(r"REF_HNSW_BUILD_SECONDS (\S+)", o).group(1))
                hp = HNSW(None, "float", "mips", m=m, efc=efc, m_l=m_l, alpha=alpha, device=local_rank, synthetic=dict(n=nb, **gen))
                hp.build()
                t0 = time.time(); hp.build(); gpu_b = time.time() - t0
                hp.close()
                out["build"]["cpu_baseline"] = {"value": ref_b, "unit": "s", "cores": os.cpu_count(), "kind": "reference", "n": int(nb), "gpu_seconds_same_n": gpu_b,
                                                "sample": f"the reference's HNSW constructor over the first {nb} points, all host cores"}
            finally:
                shutil.rmtree(tmp, ignore_errors=True)
        except Exception as e:
            out["cpu_baseline_failed"] = str(e)[:300]
    h.close()
    return out


def run_hnsw_workload(args, rank, local_rank, world):
    D = Dist(local_rank, world)
    if rank == 0:
        r = hnsw_c5(args, local_rank, with_cpu=not args.no_cpu_baseline)
        line = {"metric": "HNSW QPS@recall10>=0.95", "value": r["search"]["value"], "unit": "queries/s", "n_gpus": 1, "steps": 10, "warmup": 3,
                "ms_per_step": r["search"]["device_ms"], "higher_is_better": True, "scaling": "replicas only", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "config": {"workload": "hnsw_1000000x200_f32_ip_10000q"}, "e2e": r["search"]["e2e"], "roofline": r["search"]["roofline"],
                "cpu_baseline": r["search"].get("cpu_baseline"), "detail": r}
        print(json.dumps(line), flush=True)
    D.close()


def run_groundtruth_workload(args, rank, local_rank, world):
    D = Dist(local_rank, world)
    r = groundtruth_c3(D, rank, local_rank, world, reps=max(2, args.steps) + 1)
    if rank == 0:
        line = {"metric": "groundtruth_seconds", "value": r["seconds"], "unit": "s", "n_gpus": world, "steps": args.steps, "warmup": 1,
                "ms_per_step": r["seconds"] * 1e3, "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "tf32 filter + f32 exact",
                "data": "synthetic", "config": {"workload": "groundtruth_10000000x96_f32_k100_10000q", "parallelism": f"base rows sharded x{world}"},
                "roofline": r["roofline"], "detail": r}
        print(json.dumps(line), flush=True)
    D.close()


def run_ours(args, rank, local_rank, world):
    import torch
    import parlayann_b200 as P
    from parlayann_b200 import fileio
    from parlayann_b200.sharded import search_sharded
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (parlayann_b200 has no CPU fallback)")
    D = Dist(local_rank, world)
    barrier, allmax = D.barrier, D.allmax
    peaks = load_peaks()
    peak = float(peaks.get("hbm_gbs", 6650.0))

    n, nq, nqr, R, L = args.n, args.nq, args.nq_recall, 64, 128
    # ---- setup (untimed): replicated index, points synthesised on the device, graph built on the device
    t_setup = time.time()
    idx = P.GraphIndex(data_type="uint8", metric="Euclidian", max_degree=R, device=local_rank,
                       synthetic=dict(n=n, d=128, kind="uint8", seed=1234, stream=0, r=24, nc=256))
    t0 = time.time()
    idx.build(R=R, L=L, alpha=1.2, num_passes=1, sort_neighbors=True)
    build_s = time.time() - t0
    build_launches = idx.last_timing()[1]
    phases = idx.build_phase_seconds
    bstats = idx.build_stats()
    # the recall batch (the reference's standard 10 000 queries, generator stream 1): exact ground truth from K4, then the
    # reference's sweep picks the QueryParams; every rank does the same, so all ranks time the same setting
    idx.synthetic_queries(nqr, kind="uint8", seed=1234, stream=1)
    rqueries = idx.download_queries()
    idx.groundtruth_resident(k=100)          # warm-up (allocations)
    t0 = time.time()
    idx.groundtruth_resident(k=100)
    gt_ms = idx.last_timing()[0]
    gt_ids, gt_d = idx.fetch_groundtruth()
    gt_s = time.time() - t0

    settings = sweep_settings(R)

    def qp_of(s):
        kind, k, beam, cut, limit, dl = s
        return dict(k=k, beam=beam, cut=cut, limit=None if limit < 0 else limit, degree_limit=None if dl < 0 else dl)

    # the reference's sweep (check_nn_recall.h:214-254: 43 beams + 20 visit limits) ON THE DEVICE: search, tie-aware recall and
    # counter statistics of every setting stay there (pann_search_sweep), one copy brings the 63 rows back
    t0 = time.time()
    cache, sweep_rows = device_sweep(idx, settings, n, R)
    sweep_s = time.time() - t0
    s, rec, cm = choose_setting(settings, lambda x: cache[x])
    qp = qp_of(s)
    idx.search_resident(**qp)
    r_ids, _, r_vis, r_cmps = idx.fetch_results()     # the recall batch's answer at the chosen setting (parity below)

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def l2_flush():
        flush.zero_()
        torch.cuda.synchronize()

    def pinned_out(m):
        return (torch.empty((m, K), dtype=torch.int32).pin_memory().numpy().view(np.uint32), torch.empty((m, K), dtype=torch.float32).pin_memory().numpy(),
                torch.empty(m, dtype=torch.int32).pin_memory().numpy().view(np.uint32), torch.empty(m, dtype=torch.int32).pin_memory().numpy().view(np.uint32))

    # the same setting over the 10 000-query recall batch alone: about two waves of warps on one B200, so this is the
    # latency-bound end of the curve (SURVEY.md §8d asks for both batch sizes); device-resident, then end to end
    sm, sb = 0.0, 0
    for i in range(3 + 10):
        l2_flush(); idx.search_resident(**qp)
        if i >= 3:
            sm += idx.last_timing()[0]; sb += idx.last_bytes()
    small_ms = allmax(sm / 10)
    rq_pin = torch.from_numpy(rqueries).pin_memory().numpy()
    rout = pinned_out(nqr)
    se = 0.0
    for i in range(3 + 10):
        l2_flush()
        t0 = time.perf_counter()
        idx.search(rq_pin, out=rout, **qp)
        if i >= 3:
            se += (time.perf_counter() - t0) * 1e3
    small_e2e_ms = allmax(se / 10)
    assert np.array_equal(rout[0], r_ids)
    setup_s = time.time() - t_setup

    # ---- the step batch: nq queries per GPU from the same generator (stream 1000 + rank)
    idx.synthetic_queries(nq, kind="uint8", seed=1234, stream=1000 + rank)
    hq = torch.from_numpy(idx.download_queries()).pin_memory()
    out = pinned_out(nq)
    hq_np = hq.numpy()

    # ---- device-resident arm: W warm-up steps, then exactly K timed steps
    for _ in range(args.warmup):
        l2_flush(); idx.search_resident(**qp)
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    dev_ms, launches, abytes = 0.0, 0, 0
    t_wall0 = time.perf_counter()
    if args.profile:
        torch.cuda.profiler.start()        # ncu --profile-from-start off: only the timed steps are captured
    for _ in range(args.steps):
        l2_flush()
        idx.search_resident(**qp)          # synchronous; device time from CUDA events on the library's stream
        ms, nl = idx.last_timing()
        dev_ms += ms; launches += nl; abytes += idx.last_bytes()
    if args.profile:
        torch.cuda.profiler.stop()
    barrier()
    wall_ms = (time.perf_counter() - t_wall0) * 1e3
    # ---- end-to-end arm: host queries in, host results out, through the public batch call
    for _ in range(args.warmup):
        l2_flush(); idx.search(hq_np, out=out, **qp)
    barrier()
    e2e_ms = 0.0
    for _ in range(args.steps):
        l2_flush()
        t0 = time.perf_counter()
        idx.search(hq_np, out=out, **qp)   # H2D + K1 + D2H, returns when the results are in host memory
        e2e_ms += (time.perf_counter() - t0) * 1e3
    barrier()
    clocks = sampler.stop()
    ms_per_step = allmax(dev_ms / args.steps)
    e2e_ms_per_step = allmax(e2e_ms / args.steps)
    launches_all = int(allmax(launches))  # per rank (identical on every rank)

    # ---- the timed batch itself is checked: exact ground truth (K4) for its first rows, tie-aware recall of the timed answer
    nsl = min(nqr, nq)
    idx.upload_queries(hq_np[:nsl])
    idx.groundtruth_resident(k=100)
    tg_ids, tg_d = idx.fetch_groundtruth()
    timed_recall = recall_with_ties(out[0][:nsl], tg_ids, tg_d, K)

    # ---- strong scaling: ONE fixed batch for the whole job (the same rows on every rank: generator stream 999) cut over the
    # ranks; every rank ends with the whole answer (NCCL all-gather from the device buffers).  Timed end to end, max over ranks.
    strong = {}
    for name, m, reps in (("batch_1M", nq, max(3, args.steps // 4)), ("batch_10K", nqr, 10)):
        if name == "batch_1M":
            idx.synthetic_queries(m, kind="uint8", seed=1234, stream=999)
            sq = torch.from_numpy(idx.download_queries()).pin_memory().numpy()
        else:
            sq = rq_pin
        for _ in range(2):
            search_sharded(idx, sq, gather=True, **qp)
        barrier()
        tt, each = 0.0, []
        for _ in range(reps):
            l2_flush()
            barrier()
            t0 = time.perf_counter()
            res = search_sharded(idx, sq, gather=True, **qp)
            each.append(allmax((time.perf_counter() - t0) * 1e3))
            tt += each[-1]
        strong[name] = {"queries_total": int(m), "ms": tt / reps, "ms_each": [round(x, 3) for x in each], "value": m / (tt / reps * 1e-3), "unit": "queries/s",
                        "what": "fixed batch cut over the ranks, H2D + search + all-gather + D2H on every rank (search_sharded gather=True)"}
        if name == "batch_10K":
            strong[name]["ids_equal_single_gpu"] = bool(np.array_equal(res[0], r_ids))
        if world > 1:   # the same with the answer assembled on rank 0 only (a single driving process; NCCL gather instead of all-gather)
            for _ in range(2):
                search_sharded(idx, sq, gather=True, root=0, **qp)
            barrier()
            tt = 0.0
            for _ in range(reps):
                l2_flush()
                barrier()
                t0 = time.perf_counter()
                res0 = search_sharded(idx, sq, gather=True, root=0, **qp)
                tt += allmax((time.perf_counter() - t0) * 1e3)
            strong[name + "_root"] = {"queries_total": int(m), "ms": tt / reps, "value": m / (tt / reps * 1e-3), "unit": "queries/s",
                                      "what": "as above, answer gathered to rank 0 only (search_sharded root=0)"}
            if rank == 0:
                strong[name + "_root"]["ids_equal_all_gather"] = bool(np.array_equal(res0[0], res[0]))

    # ---- ground truth at BASELINE configs[2] (10M x 96 f32, k=100), base rows sharded over the ranks
    c3 = None
    if not args.no_c3:
        try:
            c3 = groundtruth_c3(D, rank, local_rank, world)
        except Exception as e:   # reported, never fatal for the headline
            c3 = {"failed": str(e)[:300]}

    # ---- HNSW at BASELINE configs[4] (1M x 200 f32 inner product): one GPU (the build does not shard), rank 0
    c5 = None
    if rank == 0 and not args.no_c5:
        try:
            c5 = hnsw_c5(args, local_rank, with_cpu=(world == 1 and not args.no_cpu_baseline))
        except Exception as e:
            c5 = {"failed": str(e)[:300]}
    barrier()

    if rank == 0:
        achieved = (abytes / args.steps) / (dev_ms / args.steps * 1e-3) / 1e9
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json"))).get(f"k1s_dram_bytes_per_launch_{n}x{nq}")
        except Exception:
            pass
        bf16 = float(peaks.get("bf16_tflops", 1600.0))
        int8_peak = 2.0 * bf16   # kind::i8 runs at twice the bf16 rate; MEASURED_PEAKS.json holds no int8 number
        gt_ops = 2.0 * nqr * n * 128
        line = {"metric": "QPS@recall10>=0.95", "value": world * nq / (ms_per_step * 1e-3), "unit": "queries/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": workload_config(args, s, rec, cm),
                "e2e": {"value": world * nq / (e2e_ms_per_step * 1e-3), "unit": "queries/s", "ms_per_step": e2e_ms_per_step,
                        "h2d_bytes_per_step": int(hq_np.nbytes), "d2h_bytes_per_step": int(sum(o.nbytes for o in out))},
                "gpu_launches": launches_all,
                "clocks": clocks,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": traffic, "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback",
                             "algorithmic_bytes_per_launch": abytes / args.steps, "kernel": "pann_beam_search_small_kernel<u8,L2> (beam <= 32) / pann_beam_search_kernel<u8,L2>"},
                "wall_ms_per_step_incl_l2_flush": wall_ms / args.steps,
                "timed_batch_recall": {"queries": int(nsl), "recall10": round(timed_recall, 6),
                                       "what": "first rows of the TIMED batch against exact ground truth from K4, tie-aware recall (check_nn_recall.h:83-109)"},
                "recall_batch": {"queries_per_gpu": nqr, "value": world * nqr / (small_ms * 1e-3), "unit": "queries/s", "ms_per_step": small_ms,
                                 "roofline_frac": (sb / 10) / (sm / 10 * 1e-3) / 1e9 / peak,
                                 "e2e": {"value": world * nqr / (small_e2e_ms * 1e-3), "unit": "queries/s", "ms_per_step": small_e2e_ms,
                                         "h2d_bytes_per_step": int(rq_pin.nbytes), "d2h_bytes_per_step": int(sum(o.nbytes for o in rout))},
                                 "what": "the same search over the 10 000-query recall batch alone (about two waves of warps: latency-bound)"},
                "strong": strong,
                "build": {"seconds": build_s, "phase_seconds": {"beam_search+prune": phases[0], "bidirect": phases[1], "reverse_prune": phases[2]},
                          "launches": build_launches, "what": f"pann_vamana_build {n}x128 u8 R=64 L=128 alpha=1.2, final sort incl.",
                          "roofline": {"bound": "hbm", "achieved": bstats["algorithmic_bytes"] / build_s / 1e9, "peak": peak, "unit": "GB/s",
                                       "frac": bstats["algorithmic_bytes"] / build_s / 1e9 / peak, "algorithmic_bytes": bstats["algorithmic_bytes"],
                                       "search_bytes": bstats["search_bytes"], "prune_bytes": bstats["prune_bytes"], "prune_calls": bstats["prune_calls"],
                                       "pair_distances": bstats["pair_distances"],
                                       "what": "SURVEY.md §8d B_ins summed on the device over all insertions / whole build wall time (final sort incl.)"}},
                "groundtruth": {"seconds": gt_s, "device_ms": gt_ms, "what": f"{nqr} x {n} x 128 u8, k=100",
                                "roofline": {"bound": "tensor", "achieved": gt_ops / (gt_ms * 1e-3) / 1e12, "peak": int8_peak, "unit": "TOP/s",
                                             "frac": gt_ops / (gt_ms * 1e-3) / 1e12 / int8_peak, "peak_source": "2 x MEASURED_PEAKS.json bf16_tflops (int8 pipe)",
                                             "algorithmic_op": gt_ops, "kernel": "pann_gt_mma_kernel<true> (tcgen05 kind::i8) + pann_gt_shrink_kernel"}},
                "sweep": {"settings": len(sweep_rows), "seconds": sweep_s, "device_ms_total": float(sum(r["device_ms"] for r in sweep_rows)),
                          "what": "the reference's 63-setting search sweep (check_nn_recall.h:214-254) with recall and counter statistics computed on the device (pann_search_sweep)",
                          "first_beam_rows": [{k_: r[k_] for k_ in ("beam", "recall", "avg_cmps", "avg_visited", "QPS")} for r in sweep_rows[:21:4]]},
                "groundtruth_c3": c3,
                "hnsw_c5": c5,
                "setup_seconds": setup_s}
        if world == 1 and not args.no_cpu_baseline:
            try:
                cb, parity, gtb = cpu_baseline(args, idx, rqueries, gt_ids, gt_d, s, (r_ids, r_vis, r_cmps))
                line["cpu_baseline"] = cb
                line["parity"] = parity
                line["groundtruth"]["cpu_baseline"] = gtb
            except Exception as e:  # the reference binary may be missing on a box that never saw /root/reference
                line["cpu_baseline"] = {"value": None, "unit": "queries/s", "cores": os.cpu_count(), "kind": "reference", "sample": f"failed: {e}"}
            try:
                line["build"]["cpu_baseline"] = build_cpu_baseline(args, local_rank)
            except Exception as e:
                line["build"]["cpu_baseline"] = {"value": None, "sample": f"failed: {e}"}
        print(json.dumps(line), flush=True)
    D.close()


def cpu_baseline(args, idx, queries, gt_ids, gt_d, s, gpu_answer):
    """the unmodified reference's qsearchAll on the host cores, same points / graph / queries / QueryParams: timed (checkRecall),
    and its raw per-query output compared with the GPU's (parity).  Also the reference's compute_groundtruth on a query sample."""
    O, drv = ref_tools()
    from parlayann_b200 import fileio
    tmp = tempfile.mkdtemp(prefix="pann_cpu_")
    try:
        files = {k: os.path.join(tmp, k) for k in ("base", "query", "graph", "gt")}
        fileio.write_bin(files["base"], idx.points()); fileio.write_bin(files["query"], queries)
        fileio.write_graph(files["graph"], idx.graph()); fileio.write_ibin(files["gt"], gt_ids, gt_d)
        rows = ref_recall_runs(O, files, [s], 4)[1:]
        qps = float(np.mean([r["qps"] for r in rows]))
        cb = {"value": qps, "unit": "queries/s", "cores": os.cpu_count(), "kind": "reference", "cpu": cpu_model(), "binary": O.ref_driver_info(),
              "recall10": rows[0]["recall"], "avg_cmps": rows[0]["avg_cmps"],
              "sample": f"the {args.nq_recall}-query recall batch x 3 timed runs (1 warm-up) of the reference's checkRecall/qsearchAll on the GPU-built graph"}
        # one thread (PARLAY_NUM_THREADS=1) on the first 1 000 queries: the per-core rate
        n1 = min(1000, args.nq_recall)
        f1 = dict(files, query=os.path.join(tmp, "query1"), gt=os.path.join(tmp, "gt1"))
        fileio.write_bin(f1["query"], queries[:n1]); fileio.write_ibin(f1["gt"], gt_ids[:n1], gt_d[:n1])
        r1 = ref_recall_runs(O, f1, [s], 3, threads=1)[1:]
        cb["one_thread"] = {"value": float(np.mean([r["qps"] for r in r1])), "unit": "queries/s", "queries": int(n1)}
        ref_ids, ref_vis, ref_cmps = ref_search_dump(O, files, s, tmp, args.nq_recall)
        g_ids, g_vis, g_cmps = gpu_answer
        bad = (ref_ids != g_ids).any(axis=1)
        parity = {"queries": int(args.nq_recall), "ids_equal": bool(not bad.any()), "visited_equal": bool(np.array_equal(ref_vis, g_vis)),
                  "cmps_equal": bool(np.array_equal(ref_cmps, g_cmps)), "queries_with_different_ids": int(bad.sum()),
                  "what": "ref_driver search (the reference's qsearchAll + stats) vs pann_search_resident, per query, same graph / points / QueryParams"}
        # ground truth: the reference's compute_groundtruth on the first rows of the query file, against K4's rows
        nsamp = min(1000, args.nq_recall) if args.n > 2000000 else args.nq_recall
        qsub = os.path.join(tmp, "qsub")
        fileio.write_bin(qsub, queries[:nsamp])
        out = O.ref_run(["gt", "-data_type", "uint8", "-dist_func", "Euclidian", "-base", files["base"], "-query", qsub, "-k", 100, "-out", os.path.join(tmp, "g")])
        ref_gt_s = float(re.search(r"REF_GT_SECONDS (\S+)", out).group(1))
        ri, rd = fileio.read_ibin(os.path.join(tmp, "g.ibin"))
        gtb = {"value": ref_gt_s, "unit": "s", "cores": os.cpu_count(), "kind": "reference", "queries": int(nsamp),
               "distance_evals_per_s": nsamp * args.n / ref_gt_s, "dists_equal_gpu": bool(np.array_equal(rd, gt_d[:nsamp])),
               "sample": f"compute_groundtruth over {nsamp} of the {args.nq_recall} queries x {args.n} base rows, k=100 (scale linearly in queries)"}
        return cb, parity, gtb
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def build_cpu_baseline(args, local_rank):
    """the reference's knn_index::build_index (timed as vamana/neighbors.h:59-61 times it) at the largest n that fits about a
    CPU-minute, next to pann_vamana_build at the SAME n — and the two graph files compared byte for byte"""
    O, drv = ref_tools()
    import parlayann_b200 as P
    from parlayann_b200 import fileio
    nb = min(args.n, args.n_build_cpu)
    tmp = tempfile.mkdtemp(prefix="pann_bld_")
    try:
        idx = P.GraphIndex(data_type="uint8", metric="Euclidian", max_degree=64, device=local_rank,
                           synthetic=dict(n=nb, d=128, kind="uint8", seed=1234, stream=0, r=24, nc=256))
        idx.build(R=64, L=128, alpha=1.2, num_passes=1, sort_neighbors=True)   # warm-up
        t0 = time.time()
        idx.build(R=64, L=128, alpha=1.2, num_passes=1, sort_neighbors=True)
        gpu_s = time.time() - t0
        bs = idx.build_stats()
        base = os.path.join(tmp, "base")
        fileio.write_bin(base, idx.points())
        g_gpu = idx.graph()
        idx.close()
        out = O.ref_run(["build", "-data_type", "uint8", "-dist_func", "Euclidian", "-base", base, "-R", 64, "-L", 128, "-alpha", 1.2, "-out", os.path.join(tmp, "b")])
        ref_s = float(re.search(r"REF_BUILD_SECONDS (\S+)", out).group(1))
        g_ref = fileio.read_graph(os.path.join(tmp, "b.graph"))
        m = np.arange(64)[None, :] < g_ref[:, 0][:, None]
        same = bool(np.array_equal(g_ref[:, 0], g_gpu[:, 0]) and np.array_equal(g_ref[:, 1:][m], g_gpu[:, 1:][m]))
        return {"value": ref_s, "unit": "s", "cores": os.cpu_count(), "kind": "reference", "cpu": cpu_model(), "binary": O.ref_driver_info(), "n": int(nb),
                "gpu_seconds_same_n": gpu_s, "gpu_algorithmic_bytes_same_n": bs["algorithmic_bytes"], "graph_identical_to_reference": same,
                "sample": f"knn_index::build_index over the first {nb} points of the same set (R=64 L=128 alpha=1.2, 1 pass, final sort), 'Graph built in' time"}
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="search", choices=["search", "groundtruth", "hnsw"])
    ap.add_argument("--n", type=int, default=10000000, help="base points (BASELINE.json metric: 10M x 128 uint8; 1000000 = configs[1])")
    ap.add_argument("--nq", type=int, default=1000000, help="queries per GPU per step")
    ap.add_argument("--nq-recall", dest="nq_recall", type=int, default=10000, help="recall / sweep batch (the reference's standard query set size)")
    ap.add_argument("--n-build-cpu", dest="n_build_cpu", type=int, default=1000000, help="points of the reference's CPU build baseline")
    ap.add_argument("--profile", action="store_true", help="cudaProfilerStart/Stop around the timed device-resident steps")
    ap.add_argument("--no-cpu-baseline", dest="no_cpu_baseline", action="store_true")
    ap.add_argument("--no-c3", dest="no_c3", action="store_true", help="skip the configs[2] ground-truth measurement")
    ap.add_argument("--no-c5", dest="no_c5", action="store_true", help="skip the configs[4] HNSW measurement")
    ap.add_argument("--n-hnsw-cpu", dest="n_hnsw_cpu", type=int, default=100000, help="points of the reference's CPU HNSW build baseline")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    elif args.workload == "groundtruth":
        run_groundtruth_workload(args, rank, local_rank, world)
    elif args.workload == "hnsw":
        run_hnsw_workload(args, rank, local_rank, world)
    else:
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
