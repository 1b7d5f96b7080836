"""Wall time of run_fhe_gpu_task for the small BASELINE graphs on a warm handle (development aid): cfg1 (1 BFV mult_relin),
cfg3 (degree-7 polynomial, 39 nodes), cfg4 (logistic regression, 15 nodes) at n=16384 and n=65536, the toy bootstrap graph."""
import os, sys, time, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import graph_eval as GE
cases = sys.argv[1:] or ["cfg1_bfv_mult", "cfg3_bfv_poly7", "cfg4_ckks_logreg", "cfg4_ckks_logreg_n65536", "ckks16k_euclidean_distance"]
for case in cases:
    path = os.path.join(GE.GRAPHS, case)
    if not os.path.exists(os.path.join(path, "mega_ag.json")):
        path = GE.unpack_case(case, tempfile.mkdtemp())
    gc = GE.GraphCase(path)
    h = GE.harness(); task = h.th_open(path.encode(), 1)
    for key_cache in (0, 1):
        GE.runner_lib().lsb_gpu_task_set_key_cache(task, key_cache)
        for _ in range(3): gc.run(task=task, reuse_args=True)
        ts = []
        for _ in range(20):
            t0 = time.perf_counter(); gc.run(task=task, reuse_args=True); ts.append(time.perf_counter() - t0)
        ts.sort()
        print("%-28s key cache %d: median %.3f ms, best %.3f ms  (stats %s, h2d %.1f MiB)" % (case, key_cache, 1e3 * ts[len(ts) // 2], 1e3 * ts[0], gc.stats, gc.transfer["h2d"] / 2**20), flush=True)
    h.th_close(task)
