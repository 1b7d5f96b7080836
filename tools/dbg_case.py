"""Run one golden graph several times through the runner and list which outputs differ from the oracle."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import graph_eval as GE
case = sys.argv[1]
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
gc = GE.GraphCase(os.path.join(GE.GRAPHS, case))
want = gc.expected()
for rep in range(reps):
    got = gc.run()
    bad = []
    for i, (a, b) in enumerate(zip(got, want)):
        if not np.array_equal(a, b):
            d = a != b
            bad.append((i, [int(d[p].sum()) for p in range(d.shape[0])], [int(x) for x in d.reshape(d.shape[0]*d.shape[1], -1).sum(axis=1)]))
    print("rep", rep, "stats", gc.stats, "bad", bad)
