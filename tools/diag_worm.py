"""Diagnostic (GPU): worm scene, fused steps vs oracle under different switches."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import common
P, O = common.product_lib(), common.oracle_lib()
cfg = common.config_worm(P)
ref = common.create(O, cfg)
states = []
for s in range(2):
    ref.set_muscles(common.muscle_signal(s)); ref.step()
    states.append({b: common.canon(b, ref.read(b)) for b in common.ALL_STATE})
for env in ({}, {"SPH_PDL": "0"}, {"SPH_PDL": "1", "SPH_GRAPH": "0"}, {"SPH_PDL": "0", "SPH_GRAPH": "0"}, {"SPH_GRAPH": "0", "SPH_PDL": "1", "SPH_LANES": "4"},
            {"SPH_PDL": "1", "SPH_GRAPH": "1", "SPH_NEIGHBOR_PATH": "generic"}):
    for k in ("SPH_PDL", "SPH_GRAPH", "SPH_LANES", "SPH_NEIGHBOR_PATH"):
        os.environ.pop(k, None)
    os.environ.update(env)
    a = common.create(P, cfg)
    out = []
    for s in range(2):
        a.set_muscles(common.muscle_signal(s)); a.step()
        for b in common.ALL_STATE:
            mm = common.first_mismatch(common.canon(b, a.read(b)), states[s][b])
            if mm is not None:
                out.append((s, b, mm["rows"], mm["first_row"]))
    a.close()
    print(env, "OK" if not out else out[:8], flush=True)
