"""Runs the 38 400-node convolution graph through run_fhe_gpu_task twice on one task handle and prints wall times,
runner statistics and the output digest (development aid; the parity test is in tests/test_runner_gpu.py)."""
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import graph_eval as GE  # noqa: E402

if __name__ == "__main__":
    with tempfile.TemporaryDirectory() as tmp:
        t = time.time()
        gc = GE.GraphCase(GE.unpack_case("conv_4x4_5x5_stress", tmp), seed=7, pt_pool=16)
        print("inputs %.1f s" % (time.time() - t), flush=True)
        h = GE.harness()
        t = time.time()
        task = h.th_open(gc.path.encode(), 1)
        print("task load %.2f s" % (time.time() - t), flush=True)
        for rep in range(2):
            t = time.time()
            out = gc.run(task=task)
            dt = time.time() - t
            print("run %d: %.3f s, %.0f rotations/s, stats %s, digest %s" % (rep, dt, 12799 / dt, gc.stats, GE.digest(out)),
                  flush=True)
        h.th_close(task)
