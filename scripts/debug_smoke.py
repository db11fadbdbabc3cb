import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from mri_nanoparticle_simulation_b200 import api
import oracle_lib as ol
p = api.low_permeability(api.ParameterStruct(num_water=512, num_cells=40, bound=10, water_start_bound=6, hashDim=8, border=3, avoid_intracellular_throw=0))
box = api.BacteriaBox(p, 3).populateSimulation()
waters = box.getWaters().copy(); steps = 2 * p.pfreq
ctx = api.WalkerContext(p, 0); ctx.set_scene(box); ctx.set_waters(waters, 0xC0FFEE)
sig, tr = ctx.trace_sprint(steps); got, grng = ctx.get_waters(), ctx.get_rng()
orc = ol.oracle_for_box(box, p); want, wrng = waters.copy(), orc.seed(len(waters), 0xC0FFEE)
osig, otr = orc.run_sprint(want, wrng, 0, steps, trace=True)
for k in ("x","y","z","in_cell","leaf"):
    bad = np.argwhere(tr[k] != otr[k]); print("trace", k, len(bad), bad[:3].tolist())
    if len(bad):
        s,w = bad[0]; print("   gpu", tr[k][s,w], "cpu", otr[k][s,w], "prev gpu", tr[k][s-1,w], "prev cpu", otr[k][s-1,w])
        for kk in ("x","y","z","in_cell"): print("     ", kk, tr[kk][s,w], otr[kk][s,w])
for k in ("x","y","z","phase","in_cell","nearest"):
    bad = np.nonzero(got[k] != want[k])[0]; print("state", k, len(bad), bad[:5], got[k][bad[:3]], want[k][bad[:3]])
print("bytes equal", got.tobytes()==want.tobytes(), "rng", grng.tobytes()==wrng.tobytes())
gb = np.frombuffer(got.tobytes(), np.uint8).reshape(-1,40); wb = np.frombuffer(want.tobytes(), np.uint8).reshape(-1,40)
print("byte columns differing:", np.nonzero((gb!=wb).any(0))[0])
