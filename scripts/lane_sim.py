"""Lane-efficiency simulation of the fused kernel's resident loop on an oracle trace (CPU only).

The walkers of a workload are traced with the oracle for two launches of `steps` steps; at the start of
the second they are ordered as the kernel orders them (stable sort by leaf rank: leaves ranked by
resident count, heaviest first, then by leaf index).  Per warp and step the script counts how many
resident evaluations the WARP executes under several loop structures and reports walker-level work /
warp-level work (x32 = active lanes as ncu reports them).

  python scripts/lane_sim.py [workload] [walkers] [steps]
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
from mri_nanoparticle_simulation_b200 import api  # noqa: E402
import oracle_lib as ol  # noqa: E402
import ctypes as C  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else bench.DEFAULT_WORKLOAD
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 20
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 125
    p, seed, _ = bench.make_params(api, name)
    path = f"/tmp/lane_sim_{name}.scn"
    t = time.time()
    if os.path.exists(path):
        box = api.BacteriaBox.load(path, p)
    else:
        q = p.copy(); q.num_water = 64
        box = api.BacteriaBox(q, seed).populateSimulation()
        box.save(path)
    box.rethrow_waters(n, 7000, True, p)
    waters = box.getWaters().copy()
    print(f"scene + {n} waters in {time.time() - t:.1f} s", flush=True)

    p = p.copy(); p.pfreq = steps            # the oracle samples every pfreq steps; the walk does not depend on it
    orc = ol.oracle_for_box(box, p)
    rng = orc.seed(n, bench.WALKER_SEED)
    nodes = box.getOctree().nodes
    is_leaf = (nodes["mc"] >> np.uint64(63)).astype(bool)
    cnt_of_node = nodes["n_res"].astype(np.int32)
    # leaf index (kind_index) per flat node and the kernel's leaf rank
    leaf_index = np.cumsum(is_leaf) - 1
    leaf_cnt = cnt_of_node[is_leaf]
    order = np.argsort(-leaf_cnt.astype(np.int64), kind="stable")
    leaf_rank = np.empty_like(order); leaf_rank[order] = np.arange(len(order))

    def trace(time0):
        leaf = np.zeros((steps, n), np.uint32)
        inc = np.zeros((steps, n), np.uint8)
        tr = ol.OracleTrace(None, None, None, inc.ctypes.data, leaf.ctypes.data, None)
        sig = np.zeros(max(1, steps // p.pfreq))
        rc = ol.lib().oracle_run_sprint(C.byref(p), C.byref(orc.sc), waters.ctypes.data, rng.ctypes.data, n, time0, steps,
                                        sig.ctypes.data, C.byref(tr), os.cpu_count())
        assert rc == 0
        return leaf, inc

    t = time.time()
    leaf0, _ = trace(0)
    print(f"launch 1 traced in {time.time() - t:.1f} s", flush=True)
    key = leaf_rank[leaf_index[leaf0[-1]]]
    perm = np.argsort(key, kind="stable")
    leaf1, inc1 = trace(steps)
    cnt = cnt_of_node[leaf1][:, perm].astype(np.int32)         # [step][thread]
    nw = n // 32
    cw = cnt[:, :nw * 32].reshape(steps, nw, 32)
    total = cw.sum()
    mx = cw.max(axis=2)
    print(f"walker-level residents per step: {total / (steps * nw * 32):.3f}; warp-level ideal (max count): "
          f"{mx.mean():.3f} -> lanes {32 * total / (32 * mx.sum()):.2f}")

    # shipped structure: loop over groups of 4 while j + 4 <= cnt (lanes leave at different trips), then the remainder:
    # a pair if rem >= 2, a single if rem is odd.  warp cost = (max full groups)*4 + 2*any(rem>=2) + 1*any(rem odd)
    groups = (cw // 4).max(axis=2) * 4
    rem = cw % 4
    cost_shipped = groups + 2 * (rem >= 2).any(axis=2) + 1 * ((rem & 1) == 1).any(axis=2)
    print(f"shipped structure (4-unrolled + remainder): {cost_shipped.mean():.3f} per warp-step -> lanes "
          f"{32 * total / (32 * cost_shipped.sum()):.2f}")
    cost_pair = ((mx + 1) // 2) * 2
    print(f"uniform loop over pairs (second predicated): {cost_pair.mean():.3f} -> lanes {32 * total / (32 * cost_pair.sum()):.2f}")
    # decay over the launch
    for s in (0, steps // 4, steps // 2, steps - 1):
        print(f"  step {s:4d}: ideal lanes {32 * cw[s].sum() / (32 * mx[s].sum()):.2f}, shipped {32 * cw[s].sum() / (32 * cost_shipped[s].sum()):.2f}")
    # CTA-level distribution: how uneven are CTAs (128 threads)
    ncta = nw // 4
    cta_cost = cost_shipped[:, :ncta * 4].reshape(steps, ncta, 4).sum(axis=(0, 2))
    print(f"CTA cost (warp-residents per launch): first {cta_cost[:8].mean():.0f}, median {np.median(cta_cost):.0f}, "
          f"last {cta_cost[-8:].mean():.0f}, max {cta_cost.max()}")
    # membrane test occupancy: lanes whose leaf is near a cell are not in the trace; in_cell share instead
    print(f"in_cell share {inc1.mean():.3f}")
    np.save("/tmp/lane_sim_cnt.npy", cnt)

    # risk-aware order: key = the largest resident count a walker STARTING in this leaf reached within a launch
    # (estimated from launch 1), ties by own count, then leaf index
    li0 = leaf_index[leaf0]                                   # [step][walker] leaf index, launch 1
    c0 = leaf_cnt[li0]
    reach = c0.max(axis=0)                                     # per walker: max count over launch 1
    risk = np.zeros(len(leaf_cnt), np.int64)
    np.maximum.at(risk, li0[0], reach)
    risk = np.maximum(risk, leaf_cnt)
    start_leaf = leaf_index[leaf0[-1]]
    for name, key in (("risk, own count, leaf", (risk[start_leaf] * 256 + leaf_cnt[start_leaf]) * (1 << 20) + (len(leaf_cnt) - 1 - start_leaf)),
                      ("risk, leaf", risk[start_leaf] * (1 << 20) + (len(leaf_cnt) - 1 - start_leaf))):
        perm2 = np.argsort(-key, kind="stable")
        c2 = cnt_of_node[leaf1][:, perm2].astype(np.int32)[:, :nw * 32].reshape(steps, nw, 32)
        g2 = (c2 // 4).max(axis=2) * 4; r2 = c2 % 4
        cost2 = g2 + 2 * (r2 >= 2).any(axis=2) + 1 * ((r2 & 1) == 1).any(axis=2)
        print(f"order by ({name}): shipped structure {cost2.mean():.3f} per warp-step -> lanes {32 * total / (32 * cost2.sum()):.2f}")
        wc = cost2.sum(axis=0)                                 # per warp cost over the launch
        for lo in (0, nw // 2, nw * 3 // 4, nw * 7 // 8, nw - 1024):
            seg = wc[lo:lo + 1024]
            print(f"   warps {lo:6d}+1024: cost mean {seg.mean():7.0f} p50 {np.median(seg):7.0f} p99 {np.percentile(seg, 99):7.0f} max {seg.max():7d}")
    wc = cost_shipped.sum(axis=0)
    print("shipped order:")
    for lo in (0, nw // 2, nw * 3 // 4, nw * 7 // 8, nw - 1024):
        seg = wc[lo:lo + 1024]
        print(f"   warps {lo:6d}+1024: cost mean {seg.mean():7.0f} p50 {np.median(seg):7.0f} p99 {np.percentile(seg, 99):7.0f} max {seg.max():7d}")


if __name__ == "__main__":
    main()
