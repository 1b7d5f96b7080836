"""Records, from the REFERENCE frontend (frontend/custom_task.py + bootstrap_params.py, imported from
/root/reference), the rotation structure of the default bootstrapping parameters: per-matrix diagonal index sets
of CoeffsToSlots / SlotsToCoeffs and the Galois-key rotation list a bootstrap node is given.

    PYTHONPATH=/root/reference python tests/golden/make_btp_fixture.py
"""
import json
import os
import sys

sys.path.insert(0, os.environ.get("LATTISENSE_REFERENCE", "/root/reference"))
from frontend.custom_task import CkksBtpParam  # noqa: E402

if __name__ == "__main__":
    p = CkksBtpParam.create_default_param()
    rots = sorted(p.rotations_for_bootstrapping())
    rec = {"n": p.n, "slots": p.slots, "rotations": rots}
    for name, ep in (("cts", p.cts_params), ("stc", p.stc_params)):
        m = ep.compute_bootstrapping_dft_index_map()
        rec[name] = {"depth": ep.depth(actual=False), "level_start": ep.level_start, "bsgs_ratio": ep.bsgs_ratio,
                     "index": [sorted(m[i].keys()) for i in sorted(m)], "rotations": sorted(set(ep.rotations()))}
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "btp_rotations.json")
    with open(out, "w") as f:
        json.dump(rec, f)
    print({k: (v if not isinstance(v, (dict, list)) else "...") for k, v in rec.items()}, len(rots))
