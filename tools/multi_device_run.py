"""cfg3 x 64 (BFV degree-7 polynomial, 64 independent evaluations in one mega_ag graph) through ONE run_fhe_gpu_task
call on 1, 2, 4, 8 devices (lsb_gpu_task_set_devices): wall time of the call on a warm task handle, evaluations/s, and
equality of the outputs across device counts.  Development aid; parity is in tests/test_runner_gpu.py."""
import os
import sys
import tempfile
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import graph_eval as GE  # noqa: E402

if __name__ == "__main__":
    ndev = torch.cuda.device_count()
    with tempfile.TemporaryDirectory() as tmp:
        gc = GE.GraphCase(GE.unpack_case("cfg3_bfv_poly7_x64", tmp), seed=3)
        ref = None
        for n in (1, 2, 4, 8):
            if n > ndev:
                break
            h = GE.harness()
            task = h.th_open(gc.path.encode(), 1)
            best = 1e9
            for rep in range(4):
                if rep == 3 and os.environ.get("MD_TRACE"):
                    os.environ["LSB_TRACE"] = "1"
                t = time.time()
                out = gc.run(task=task, devices=list(range(n)))
                best = min(best, time.time() - t)
                os.environ.pop("LSB_TRACE", None)
            d = GE.digest(out)
            ref = ref or d
            print("devices %d: %.1f ms per call, %.0f polynomial evaluations/s, stats %s, outputs %s" %
                  (n, best * 1e3, 64 / best, gc.stats, "equal" if d == ref else "DIFFER"), flush=True)
            h.th_close(task)
