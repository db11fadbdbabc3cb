"""First contact with the GPU: parity of one short sprint against the oracle + timing."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from mri_nanoparticle_simulation_b200 import api
import oracle_lib as ol

def check(name, params, seed, steps, trace=True, field_mode=0):
    box = api.BacteriaBox(params, seed).populateSimulation()
    waters = box.getWaters().copy()
    ctx = api.WalkerContext(params)
    ctx.set_field_mode(field_mode)
    ctx.set_scene(box)
    ctx.set_waters(waters, 0xC0FFEE)
    t = time.time()
    if trace:
        sig, tr = ctx.trace_sprint(steps)
    else:
        sig = ctx.run_sprint(steps)
    dt = time.time() - t
    gw = ctx.get_waters(); grng = ctx.get_rng()
    orc = ol.oracle_for_box(box, params)
    ow = waters.copy(); orng = orc.seed(len(ow), 0xC0FFEE)
    t = time.time()
    osig, otr = orc.run_sprint(ow, orng, 0, steps, trace=trace)
    dto = time.time() - t
    ok = True
    for f in ("x", "y", "z", "phase", "in_cell", "nearest"):
        same = np.array_equal(gw[f].view(np.uint8 if gw[f].dtype.itemsize == 1 else ('u8' if gw[f].dtype.itemsize == 8 else 'u4')),
                              ow[f].view(np.uint8 if ow[f].dtype.itemsize == 1 else ('u8' if ow[f].dtype.itemsize == 8 else 'u4')))
        if not same:
            ok = False
            bad = np.nonzero(gw[f] != ow[f])[0]
            print(f"  {f}: {len(bad)} walkers differ, first {bad[:5]}, gpu {gw[f][bad[:3]]} cpu {ow[f][bad[:3]]}")
    for f in ("s0", "s1", "spare"):
        if not np.array_equal(grng[f].view('u8'), orng[f].view('u8')):
            ok = False; print("  rng", f, "differs")
    if trace:
        for f in ("x", "y", "z", "in_cell", "leaf"):
            if not np.array_equal(tr[f], otr[f]):
                ok = False
                bad = np.argwhere(tr[f] != otr[f])
                print(f"  trace {f}: {len(bad)} entries differ, first at (step,walker) {bad[:3].tolist()}")
    err = np.max(np.abs(sig - osig)) / len(ow)
    print(f"{name}: n={len(ow)} steps={steps} bit-exact={ok} max|dS|/N={err:.3e} gpu {dt*1e3:.1f} ms (kernel {ctx.last_sprint_kernel_ms():.2f} ms) oracle {dto:.1f} s"
          f" -> gpu {len(ow)*steps/ max(ctx.last_sprint_kernel_ms(),1e-9)*1e3:.3e} steps/s, cpu {len(ow)*steps/dto:.3e} steps/s")
    print("  signal gpu", sig[:3], "cpu", osig[:3])
    return ok and err < 1e-6

if __name__ == "__main__":
    import torch
    print(torch.cuda.get_device_name(0))
    print("fp64 peak TFLOP/s:", api.measure_fp64_tflops(0))
    P = api.ParameterStruct
    r = []
    r.append(check("tiny", P(num_water=256, num_cells=40, bound=10, water_start_bound=6, hashDim=8, border=3), 3, 2000))
    r.append(check("default-short", P(), 12345, 2000))
    r.append(check("noavoid-lowperm", api.low_permeability(P(avoid_intracellular_throw=0)), 99, 3000))
    r.append(check("default-sprint", P(), 12345, 20000, trace=False))
    r.append(check("default-sprint-fast", P(), 12345, 20000, trace=False, field_mode=1))
    print("ALL OK" if all(r) else "FAILURES", r)
