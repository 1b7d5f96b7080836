"""Evaluates the 38 400-node convolution stress graph (tests/golden/graphs/conv_4x4_5x5_stress, emitted by the
reference frontend via make_graphs.py conv) with the CPU oracle on seeded inputs and records the SHA-256 of the
output limbs in tests/golden/conv_stress_digest.json.  One-off (about a quarter of an hour on one core); the GPU test
replays the same seeded inputs through run_fhe_gpu_task and compares digests.

    python tests/golden/make_conv_digest.py
"""
import json
import os
import sys
import tempfile
import time

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import graph_eval as GE  # noqa: E402

SEED, POOL = 7, 16

if __name__ == "__main__":
    with tempfile.TemporaryDirectory() as tmp:
        gc = GE.GraphCase(GE.unpack_case("conv_4x4_5x5_stress", tmp), seed=SEED, pt_pool=POOL)
        t = time.time()
        out = gc.expected(free=True)
        rec = {"case": "conv_4x4_5x5_stress", "seed": SEED, "pt_pool": POOL, "nodes": len(gc.compute),
               "sha256": GE.digest(out), "oracle_seconds": round(time.time() - t, 1)}
    with open(os.path.join(HERE, "conv_stress_digest.json"), "w") as f:
        json.dump(rec, f, indent=1)
    print(rec)
