"""C5's data shape (1M x 200 f32, inner product) through the Vamana path the Python surface offers for it
(python/builder.cpp + graph_index.cpp): rows normalised, graph built over the Quantized_Mips_Point<8,true> copy (int8 MIPS
kernels), search on the copy + exact float re-rank (K7), against exact ground truth (K4) on the normalised rows.
(The HNSW half of C5 is not built.)  Host-side normalisation / quantisation times are printed too: they are numpy."""
import os, sys, time, json, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import parlayann_b200 as P
from parlayann_b200.index import normalize_rows, quant_params_mips8, quantize_mips8, normalize_rows_device, quant_params_mips8_device, quantize_mips8_device
n = int(os.environ.get("N", "1000000")); d = int(os.environ.get("D", "200")); nq = 10000
gen = P.GraphIndex(data_type="float", metric="mips", max_degree=64, device=0, synthetic=dict(n=n, d=d, kind="float", seed=1234, stream=0, r=24, nc=256))
gen.synthetic_queries(nq, kind="float", seed=1234, stream=1)
base, qs = gen.points(), gen.download_queries(); gen.close()
t0 = time.time(); nb, nqs = normalize_rows(base), normalize_rows(qs); t_norm = time.time() - t0
t0 = time.time(); mv = quant_params_mips8(nb); qb, qq = quantize_mips8(nb, mv), quantize_mips8(nqs, mv); t_quant = time.time() - t0
normalize_rows_device(qs)   # warm-up (context, kernels)
t0 = time.time(); nb_d, nqs_d = normalize_rows_device(base), normalize_rows_device(qs); t_norm_d = time.time() - t0
t0 = time.time(); mv_d = quant_params_mips8_device(nb_d); qb_d, qq_d = quantize_mips8_device(nb_d, mv_d), quantize_mips8_device(nqs_d, mv_d); t_quant_d = time.time() - t0
assert np.array_equal(nb_d, nb) and mv_d == mv and np.array_equal(qb_d, qb) and np.array_equal(qq_d, qq)
del base
qidx = P.GraphIndex(qb, None, "int8", "mips", max_degree=64)
alpha = float(np.float32(1.0 if d < 100 else .98))
qidx.build(R=64, L=128, alpha=alpha)
t0 = time.time(); qidx.build(R=64, L=128, alpha=alpha); t_build = time.time() - t0
exact = P.GraphIndex(nb, qidx.graph(), "float", "mips")
gt_ids, gt_d = exact.groundtruth(nqs, k=100)
rec_of = lambda ids: float((((ids[:, :, None] == gt_ids[:, None, :]) & ((np.arange(100)[None, :] < 10) | (gt_d == gt_d[:, 9:10]))[:, None, :]).any(axis=2)).mean())
out = {"config": f"{n} x {d} f32 MIPS, normalised, Quantized_Mips_Point<8,true> copy, Vamana R=64 L=128 alpha={alpha:.2f}", "max_val": float(mv),
       "host_normalise_s": round(t_norm, 2), "host_quantise_s": round(t_quant, 2), "device_normalise_s": round(t_norm_d, 3), "device_quantise_s": round(t_quant_d, 3), "device_equals_host": True, "build_s": round(t_build, 3), "build_phases": [round(x, 3) for x in qidx.build_phase_seconds], "search": []}
for beam in (10, 16, 24, 32, 64):
    exact.search_rerank(qidx, nqs, qq, k=10, beam=beam)
    ts = []
    for _ in range(3):
        t0 = time.time(); ids, dists, vis, cmps = exact.search_rerank(qidx, nqs, qq, k=10, beam=beam); ts.append(time.time() - t0)
    t0 = time.time(); ids_f, _, _, cmps_f = exact.search(nqs, k=10, beam=beam); t_f = time.time() - t0
    out["search"].append({"beam": beam, "recall10_quant_rerank": round(rec_of(ids), 4), "qps_quant_rerank_host_buffers": round(nq / min(ts)),
                          "recall10_float": round(rec_of(ids_f), 4), "qps_float_host_buffers": round(nq / t_f), "avg_cmps": float(cmps.mean())})
print(json.dumps(out))
os.makedirs("gpurun_out", exist_ok=True); json.dump(out, open("gpurun_out/c5_mips_quant.json", "w"), indent=1)
