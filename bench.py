#!/usr/bin/env python
"""bench.py -- walker-steps/s of the sprint hot path on N B200s (one process per GPU).

    python bench.py --gpus 1 --steps K --warmup W
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...        (CPU arm: the oracle port on all host cores)

A "step" is ONE SPRINT: `sprintSteps` (20 000) timesteps of every walker on this rank --
Gaussian step, membrane test, octree field lookup, phase kick, 20 signal samples -- plus,
for N > 1, the one all-reduce of the 20 signal partials (NCCL).  Workload (config.workload):
BASELINE.json configs[3], the configuration the 1/2/4/8-GPU metric is quoted on: 1 Mi waters
per GPU in the dense 2000-cell box (SURVEY.md 8d "4a": hashDim 40, maxNeighbors 16, 89 448
octree nodes).  Walkers are independent, so ranks own disjoint walker ranges of one global
index space (weak scaling: 1 Mi walkers per GPU), the scene is replicated.

Timing: W untimed warm-up sprints, then K sprints, each bracketed by CUDA events on the
stream the kernels are launched on (torch's current stream, handed to the C ABI), with an
L2 flush (256 MiB write) between sprints outside the event pairs; barrier +
torch.cuda.synchronize() on both sides of the timed region; max over ranks.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (knob overrides, scene seed, walkers per GPU)
    "cfg4a_dense2000cells_1Mi_waters_per_gpu": (dict(num_cells=2000, hashDim=40, maxNeighbors=16), 11, 1 << 20),
    "cfg1_default_4032_waters": (dict(), 12345, 4032),
    "cfg2_8000_waters_intracellular_kick": (dict(num_water=8000, avoid_intracellular_throw=0), 21, 8000),
    "cfg3_lowperm_100k_waters": (dict(), 12345, 100000),
    "default_scene_1Mi_waters_per_gpu": (dict(), 12345, 1 << 20),
    # SURVEY.md 8d "4b": 8000 cells (21 385 dipoles, 455 008 octree nodes, up to 104 residents per leaf)
    "cfg4b_dense8000cells_1Mi_waters_per_gpu": (dict(num_cells=8000, hashDim=40, maxNeighbors=32), 11, 1 << 20),
    # BASELINE configs[4]: 16 Mi waters IN TOTAL, split over the ranks (strong scaling), shipped scene
    "cfg5_16Mi_waters_total": (dict(), 12345, 16 << 20),
}
STRONG = {"cfg5_16Mi_waters_total"}          # the walker count is the job's, not the rank's
DEFAULT_WORKLOAD = "cfg4a_dense2000cells_1Mi_waters_per_gpu"
SMALL_CONFIGS = ("cfg1_default_4032_waters", "cfg2_8000_waters_intracellular_kick")     # the reference's shipped sizes
WALKER_SEED = 0xC0FFEE


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner to
# fd 1 from C whatever NCCL_DEBUG_FILE says on some boxes), so fd 1 is pointed at stderr for the whole
# run and the result line goes to a private duplicate of the original stdout.
_RESULT_OUT = None


def claim_stdout():
    global _RESULT_OUT
    if _RESULT_OUT is None:
        sys.stdout.flush()
        _RESULT_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _RESULT_OUT if _RESULT_OUT is not None else sys.stdout
    print(json.dumps(line), file=out, flush=True)


# ----------------------------------------------------------------------------------------
def make_params(api, name):
    over, seed, n = WORKLOADS[name]
    p = api.ParameterStruct(**over)
    if name.startswith("cfg3"):
        api.low_permeability(p)
    return p, seed, n


def get_scene(api, name, rank, world, barrier, device=None):
    """The workload's scene.  device=None (the CPU reference arm): rank 0 builds it with the host
    threads once (~3 s for 2000 cells on 16 cores) and saves it, the other ranks load the file.
    device=k (our arm): every rank builds it on its own GPU (csrc/octree_build.cu, byte-identical
    to the host build, ~50 ms) -- no file, no barrier."""
    p, seed, n = make_params(api, name)
    if device is not None:
        t = time.time()
        q = p.copy(); q.num_water = 64
        box = api.BacteriaBox(q, seed).populateSimulation(device=device)
        if rank == 0:
            log(f"[bench] scene {name} built on the GPU in {time.time() - t:.2f} s")
        return box, p, n
    path = f"/tmp/mrisim_bench_{name}_seed{seed}.scn"
    if rank == 0 and not os.path.exists(path):
        t = time.time()
        q = p.copy(); q.num_water = 64
        box = api.BacteriaBox(q, seed).populateSimulation()
        box.save(path + ".tmp"); os.replace(path + ".tmp", path)
        log(f"[bench] scene {name} built in {time.time() - t:.1f} s -> {path}")
    barrier()
    box = api.BacteriaBox.load(path, p)
    return box, p, n


def throw_waters(box, p, n, rank, device=None):
    """This rank's walkers: BacteriaBox::init_waters + sortWaters from XORShift(7000 + rank);
    on the GPU for our arm (csrc/scene_gpu.cu, same bytes)."""
    box.rethrow_waters(n, 7000 + rank, True, p, device=device)
    return box.getWaters().copy()


def workload_stats(box, p, waters, sample=4000):
    """C (lookup candidates), L (levels below min_depth), R (leaf residents) at the walkers'
    current positions -- the variables of the algorithmic op count F = 73 + 9C + 9L + 24R
    (SURVEY.md 8d).  Pure numpy/python on the product scene arrays."""
    oct_ = box.getOctree()
    nodes, prefix = oct_.nodes, oct_.subtree_prefix
    lut = box.getLookupTable()
    rng = np.random.default_rng(0)
    idx = rng.choice(len(waters), size=min(sample, len(waters)), replace=False)
    Cs, Ls, Rs = [], [], []
    h, cube = p.hashDim, p.bound / p.hashDim
    md = oct_.min_depth
    for i in idx:
        x, y, z = float(waters["x"][i]), float(waters["y"][i]), float(waters["z"][i])
        row = (int(z / cube) * h + int(y / cube)) * h + int(x / cube)
        Cs.append(int((lut[row] != -1).sum()))
        m = [int(np.floor(v / p.bound * (1 << 21))) for v in (x, y, z)]
        code = 0
        for b in range(md):
            sh = 20 - b
            code = (code << 3) | (((m[0] >> sh) & 1) << 2) | (((m[1] >> sh) & 1) << 1) | ((m[2] >> sh) & 1)
        n, d = int(prefix[code]), md
        while not (int(nodes["mc"][n]) >> 63):
            sh = 20 - d
            c = (((m[0] >> sh) & 1) << 2) | (((m[1] >> sh) & 1) << 1) | ((m[2] >> sh) & 1)
            n += int(nodes["child"][n][c]); d += 1
        Ls.append(d - md); Rs.append(int(nodes["n_res"][n]))
    return float(np.mean(Cs)), float(np.mean(Ls)), float(np.mean(Rs))


#: FP64 operations of the contract's Gaussian transform per walker-step (half a Box-Muller
#: pair: ln 27, sqrt 1, sincos 33, glue 5 -> 66 per pair) + 7 uniform conversions (1 sub each
#: incl. the 1-u1) -- counted from include/mrisim_math.h, one op per + - * / sqrt fma.
RNG_FP64_OPS_PER_STEP = 66 / 2 + 6


def algorithmic_ops(Cbar, Lbar, Rbar):
    return 73 + 9 * Cbar + 9 * Lbar + 24 * Rbar + RNG_FP64_OPS_PER_STEP


class ClockSampler:
    """nvidia-smi clocks/throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------------------
def cpu_arm(make_oracle, p, waters, budget_s, n_threads, steps):
    """Time the oracle port (CPU restatement of the reference's device functions; the
    reference has no CPU walk of its own) on a bounded walker sample."""
    orc = make_oracle()
    calib_n = min(len(waters), 64 * n_threads)
    w = waters[:calib_n].copy(); rng = orc.seed(calib_n, WALKER_SEED)
    t = time.perf_counter(); orc.run_sprint(w, rng, 0, p.pfreq, n_threads=n_threads); dt = time.perf_counter() - t
    rate = calib_n * p.pfreq / max(dt, 1e-9)
    n = int(min(len(waters), max(n_threads, rate * budget_s / steps)))
    n -= n % n_threads or 0
    n = max(n, n_threads)
    return orc, n, rate


def shipped_params(cabi, **over):
    """The shipped src/parameters.h as an mrisim_params, filled in Python: the reference arm must not load
    the product library.  tests/test_bench_contract.py checks it against mrisim_params_default byte for byte."""
    import math
    p = cabi.Params()
    vals = dict(g=42.5781e6, scale=2.0, border=6.0, num_threads=16, sprintSteps=20000, num_water=4032, num_cells=270,
                cell_r=.55, mmoment=3.5e-17, kick_mode=cabi.KICK_RANDOM, avoid_intracellular_throw=1,
                phase_stdev=2 * 3.14 * 1.5 * 1e-3 * 42 * 12 * 10e-3, phase_k=5.0, bound=30.0, water_start_bound=10.0,
                hashDim=20, maxNeighbors=13, D_cell=.5547, D_extra=1.6642, P_expr=0.2, tau=1e-6, totaltime=40, taucp=5.5,
                reflectIO=0.0, reflectOI=0.0)
    vals.update(over)
    for k, v in vals.items():
        setattr(p, k, v)
    p.t = int(p.totaltime / p.tau); p.tcp = int(p.taucp / p.tau); p.pfreq = int(1e-3 / p.tau)       # parameters.h:93,95; cuda_sim.cu:44-45
    p.in_stdev = math.sqrt(math.pi * p.D_cell * p.tau); p.out_stdev = math.sqrt(math.pi * p.D_extra * p.tau)   # :120-121
    return p


#: workload -> (oracle/_ref binary = the reference's own host scene code compiled in place, its seed)
REF_SCENES = {"cfg4a_dense2000cells_1Mi_waters_per_gpu": ("ref_scene_dense2000_32k", 11)}


def reference_scene(workload):
    """Scene + waters for the CPU arm from the REFERENCE'S OWN host code (oracle/_ref/ref_scene_*: BacteriaBox,
    SimulationBox, octree .cpp compiled where they lie), read with the pure-Python MRISCN01 reader.  Returns
    None when the binary is not there (then the arm falls back to the product's byte-identical builder)."""
    name = REF_SCENES.get(workload)
    exe = name and os.path.join(ROOT, "oracle", "_ref", name[0])
    if not exe or not os.path.exists(exe):
        return None
    from mri_nanoparticle_simulation_b200 import scene_io, _cabi as cabi
    path = f"/tmp/mrisim_refarm_{name[0]}_seed{name[1]}.scn"
    if not os.path.exists(path):
        t = time.time()
        subprocess.run([exe, str(name[1]), path + ".tmp"], check=True, stdout=subprocess.DEVNULL)
        os.replace(path + ".tmp", path)
        log(f"[bench] reference scene by {name[0]} (the reference's own host code) in {time.time() - t:.1f} s")
    sc = scene_io.load(path)
    over, _, _ = WORKLOADS[workload]
    p = shipped_params(cabi, **over)
    assert (sc.num_cells, sc.hash_dim, sc.max_neighbors) == (p.num_cells, p.hashDim, p.maxNeighbors)
    waters = np.zeros(sc.num_water, dtype=cabi.WATER_DTYPE)
    for i, k in enumerate(("x", "y", "z", "phase")):
        waters[k] = sc.waters[:, i]
    waters["nearest"] = -1
    return sc, p, waters


def run_reference_arm(args, api_loader, ol):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    n_threads = os.cpu_count() or 1
    ref = reference_scene(args.workload)
    if ref is not None:
        sc, p, waters = ref
        scene_by = "oracle/_ref (the reference's own BacteriaBox/SimulationBox/octree .cpp compiled in place)"
        make_oracle = lambda: ol.Oracle(sc, p)
    else:
        api = api_loader()
        box, p, n_per_gpu = get_scene(api, args.workload, 0, 1, lambda: None)
        waters = throw_waters(box, p, min(n_per_gpu, 1 << 18), 0)
        scene_by = "product host builder (byte-identical to the reference's, tests/test_scene_parity.py); oracle/_ref binary absent"
        make_oracle = lambda: ol.oracle_for_box(box, p)
    steps = p.sprintSteps
    per_step_budget = max(2.0, min(20.0, 150.0 / max(args.steps + args.warmup, 1)))
    if os.environ.get("MRISIM_REFARM_BUDGET_S"):                 # tests: a short sample
        per_step_budget = float(os.environ["MRISIM_REFARM_BUDGET_S"])
    orc, n, _ = cpu_arm(make_oracle, p, waters, per_step_budget, n_threads, steps)
    w = waters[:n].copy(); rng = orc.seed(n, WALKER_SEED)
    time0 = 0
    for _ in range(args.warmup):
        orc.run_sprint(w, rng, time0, steps, n_threads=n_threads); time0 += steps
    t = time.perf_counter()
    for _ in range(args.steps):
        orc.run_sprint(w, rng, time0, steps, n_threads=n_threads); time0 += steps
    dt = time.perf_counter() - t
    value = n * steps * args.steps / dt
    sample = f"{n} of the workload's walkers x {steps} timesteps per step (one full sprint), {n_threads} pthreads over walkers"
    line = {
        "impl": "reference", "metric": "walker-steps/sec", "value": value, "unit": "walker-steps/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, "sprintSteps": steps, "walkers_in_sample": n, "host_cores": n_threads,
                   "scene_by": scene_by,
                   "note": "reference has no CPU walk (SURVEY.md D1): CPU restatement of its device functions (oracle port, "
                           "bit-equal to the reference's kernels: tests/test_ref_gpu_bitpin.py)"},
        "cpu_baseline": {"value": value, "unit": "walker-steps/s", "cores": n_threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "walker-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ----------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--walkers-per-gpu", type=int, default=0, help="override the workload's walker count")
    ap.add_argument("--field-mode", default="exact", choices=["exact", "fast"])
    ap.add_argument("--path-mode", default="auto", choices=["auto", "fused", "pipeline"],
                    help="fused: one thread per walker for the whole sprint; pipeline: the few-walker path; auto: by walker count")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-small-configs", action="store_true", help="skip the few-walker block (cfg1/cfg2) of the default line")
    ap.add_argument("--cpu-budget-s", type=float, default=12.0)
    args = ap.parse_args()
    claim_stdout()
    if args.warmup < 3 and args.impl == "ours":
        log("[bench] note: W < 3 warm-up steps does not meet the timing rules")

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    if args.impl == "reference":
        import oracle_lib as ol          # the one other place bench.py may execute oracle/

        def api_loader():                # only if oracle/_ref is absent: the product's host scene builder
            from mri_nanoparticle_simulation_b200 import api
            return api
        return run_reference_arm(args, api_loader, ol)
    from mri_nanoparticle_simulation_b200 import api

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference for the CPU arm)")
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")     # NCCL's version banner must not precede the JSON line on stdout
        dist.init_process_group("nccl", device_id=dev)
    barrier = (lambda: dist.barrier()) if world > 1 else (lambda: None)

    box, p, n_per_gpu = get_scene(api, args.workload, rank, world, barrier, device=local)
    strong = args.workload in STRONG
    first_id = rank * n_per_gpu
    if strong:                                   # the job's walkers, cut into contiguous rank ranges
        from mri_nanoparticle_simulation_b200 import dist as shard
        lo, hi = shard.shard_range(n_per_gpu, rank, world)
        first_id, n_per_gpu = lo, hi - lo
    if args.walkers_per_gpu:
        n_per_gpu = args.walkers_per_gpu; first_id = rank * n_per_gpu
    waters = throw_waters(box, p, n_per_gpu, rank, device=local)
    steps = p.sprintSteps
    n_int = steps // p.pfreq
    ctx = api.WalkerContext(p, local)
    ctx.set_field_mode(api.FIELD_FAST if args.field_mode == "fast" else api.FIELD_EXACT)
    ctx.set_path_mode({"auto": api.PATH_AUTO, "fused": api.PATH_FUSED, "pipeline": api.PATH_PIPELINE}[args.path_mode])
    ctx.set_scene(box)
    ctx.set_waters(waters, WALKER_SEED, first_id)
    if world > 1:
        # the path's one exchange runs inside libmrisim.so: the context owns an NCCL communicator (rank 0's
        # unique id travels through torch.distributed, which stays the plumbing: barriers, max over ranks)
        ids = [api.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0, device=dev)
        ctx.comm_init_rank(ids[0], world, rank)
        if rank == 0:
            log(f"[bench] NCCL bound by libmrisim.so: {api.nccl_origin()}; communicator of {ctx.comm_size} ranks")

    stream = torch.cuda.Stream(device=dev)       # the sprint's kernels run on this stream, the exchange on the context's side stream
    torch.cuda.set_stream(stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def one_step():
        # mrisim_submit_sprint: kernels on `stream`; on the side stream the all-reduce of the 20 samples (N > 1)
        # and their D2H.  The step's event pair closes only after that exchange (stream_wait_sprint), so a
        # step's time includes its own all-reduce (in a production loop it overlaps the next sprint instead).
        ctx.submit_sprint(steps, stream.cuda_stream)
        ctx.stream_wait_sprint(stream.cuda_stream)

    for _ in range(args.warmup):
        one_step()
        ctx.wait_sprint(steps)
    torch.cuda.synchronize(); barrier()

    fp64_peak = api.measure_fp64_tflops(local)
    sampler = ClockSampler(local).start() if rank == 0 else None
    launches0 = ctx.kernel_launches
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    kernel_ms = []
    torch.cuda.synchronize(); barrier()
    wall0 = time.perf_counter()
    for a, b in ev:
        flush.fill_(1)                       # L2 flush between timed iterations (not inside an event pair)
        a.record(stream)
        one_step()
        b.record(stream)
        b.synchronize()
        signal_last = ctx.wait_sprint(steps)          # the all-reduced samples, already on the host
        kernel_ms.append(ctx.last_sprint_kernel_ms())
    torch.cuda.synchronize(); barrier()
    wall = time.perf_counter() - wall0
    clocks = sampler.stop() if sampler else None
    launches = ctx.kernel_launches - launches0
    step_ms = [a.elapsed_time(b) for a, b in ev]
    total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    total_s = float(total_ms.item()) * 1e-3
    n_total = WORKLOADS[args.workload][2] if strong and not args.walkers_per_gpu else n_per_gpu * world
    value = n_total * steps * args.steps / total_s

    # ---- e2e: the same sprint through the host-buffer C-ABI calls a user makes ----------
    host_w = ctx.get_waters()
    h2d = host_w.nbytes; d2h = host_w.nbytes + n_int * 8
    e2e_steps = args.steps
    torch.cuda.synchronize(); barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        ctx.set_waters(host_w, WALKER_SEED, first_id)               # H2D of the walker records (pinned staging)
        ctx.submit_sprint(steps)                                     # kernels + all-reduce (N > 1) + D2H of the 20 samples
        part = ctx.wait_sprint(steps)
        host_w = ctx.get_waters()                                    # D2H of the walker records
    torch.cuda.synchronize()
    e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_value = n_total * steps * e2e_steps / float(e2e_s.item())

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    Cbar, Lbar, Rbar = workload_stats(box, p, host_w)
    F = algorithmic_ops(Cbar, Lbar, Rbar)
    k_ms = float(np.mean(kernel_ms))
    achieved = n_per_gpu * steps * F / (k_ms * 1e-3) / 1e12
    # Instruction mix of the hot kernel from the committed ncu capture -- only if that capture was taken from
    # THIS kernel source (digest of csrc/mrisim.cu + include/mrisim_math.h, scripts/make_instr_mix.py).
    traffic, issue = None, None
    mix = load_instr_mix()
    fused = not ctx.pipeline_sprints
    if mix is not None and mix.get("stale"):
        issue = {"stale": True, "note": "profiles/instr_mix_r02.json was captured from another version of the kernel source; "
                                        "no issue-slot view for this line (re-run scripts/make_instr_mix.py on an ncu capture of HEAD)"}
    elif mix is not None and args.field_mode == "exact" and args.workload == DEFAULT_WORKLOAD and not args.walkers_per_gpu and fused:
        traffic = mix.get("dram_bytes_per_launch")
        if clocks and clocks.get("sm_mhz"):
            # Issue-slot view: an FP64 warp instruction holds the SMSP issue port for two cycles on
            # B200, so the binding resource is issue slots = 2 x FP64 + other instructions; the rate and the clock are live.
            n_sm = torch.cuda.get_device_properties(local).multi_processor_count
            warp_steps_per_s = n_per_gpu / 32.0 * steps / (k_ms * 1e-3)
            cyc_per_s = n_sm * 4 * clocks["sm_mhz"] * 1e6
            issue = {"fp64_inst_per_warp_step": mix["fp64_inst_per_warp_step"], "other_inst_per_warp_step": mix["other_inst_per_warp_step"],
                     "issue_slots_per_warp_step": mix["issue_slots_per_warp_step"],
                     "issue_slot_frac": mix["issue_slots_per_warp_step"] * warp_steps_per_s / cyc_per_s,
                     "fp64_pipe_frac": 2 * mix["fp64_inst_per_warp_step"] * warp_steps_per_s / cyc_per_s,
                     "instr_mix_from": f"profiles/instr_mix_r02.json (ncu capture {mix.get('tag')}, kernel source digest {mix.get('source_digest', '')[:12]} == this tree's)",
                     "note": "issue slots = 2*FP64 + other warp instructions (an FP64 instruction holds the SMSP issue port for 2 cycles, "
                             "profiles/fp64_issue_probe_r01.txt); frac = slots issued per SMSP cycle at the sampled SM clock. "
                             "Explains roofline.frac; does not replace it."}
    F_survey = F - RNG_FP64_OPS_PER_STEP          # SURVEY.md 8(d)'s count: the contract's Gaussian transform excluded
    props = torch.cuda.get_device_properties(local)
    sm_max = (clocks or {}).get("sm_max_mhz") or 1965.0
    peak_theoretical = props.multi_processor_count * 64 * 2 * sm_max * 1e6 / 1e12      # 64 FP64 FMA lanes per SM
    line = {
        "metric": "walker-steps/sec", "value": value, "unit": "walker-steps/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_s / args.steps * 1e3,
        "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, "walkers_per_gpu": n_per_gpu, "walkers_total": n_total,
                   "sprintSteps": steps, "signal_samples_per_step": n_int, "num_cells": p.num_cells,
                   "octree_nodes": int(len(box.getOctree().nodes)), "field_mode": args.field_mode,
                   "path_mode": args.path_mode, "pipeline_sprints": ctx.pipeline_sprints,
                   "l2": "flushed between timed steps (256 MiB write outside the event pairs)",
                   "parallelism": f"walker-sharded x{world}, scene replicated, 1 ncclAllReduce of {n_int} f64 per step issued by libmrisim.so on a side stream"},
        "e2e": {"value": e2e_value, "unit": "walker-steps/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "steps": e2e_steps, "call": "mrisim_set_waters + mrisim_submit_sprint/mrisim_wait_sprint + mrisim_get_waters (host buffers)"},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "fp64", "achieved": achieved, "peak": fp64_peak, "unit": "TFLOP/s", "frac": achieved / fp64_peak,
                     "frac_survey_F": achieved * F_survey / F / fp64_peak, "peak_theoretical": peak_theoretical,
                     "frac_of_theoretical": achieved / peak_theoretical,
                     "traffic": traffic, "kernel_ms": k_ms,
                     "kernel": "walker_sprint_kernel" if not ctx.pipeline_sprints else
                               "pipelined path (walker_draws + walker_path + step_field + walker_phase kernels; kernel_ms is their sum)",
                     "algorithmic_fp64_ops_per_walker_step": F, "algorithmic_fp64_ops_survey_8d": F_survey, "C_L_R": [Cbar, Lbar, Rbar], "issue": issue,
                     "peak_source": "measured live: mrisim_measure_fp64_tflops (8 independent DFMA chains/thread, all SMs); "
                                    "MEASURED_PEAKS.json has no FP64 entry",
                     "note": "path is FP64-pipe bound, not HBM/tensor (SURVEY.md 8d): per-step HBM traffic is ~0"},
        "signal_last_step": [float(v) for v in signal_last[:3]],
        "wall_s_timed_region": wall,
    }
    if not args.no_cpu_baseline and world == 1:
        import oracle_lib as ol          # cpu_baseline leg: the oracle as the thing timed on the host cores
        n_threads = os.cpu_count() or 1
        orc, n, _ = cpu_arm(lambda: ol.oracle_for_box(box, p), p, waters, args.cpu_budget_s, n_threads, steps)
        w = waters[:n].copy(); rng = orc.seed(n, WALKER_SEED)
        t = time.perf_counter(); orc.run_sprint(w, rng, 0, steps, n_threads=n_threads); dt = time.perf_counter() - t
        line["cpu_baseline"] = {"value": n * steps / dt, "unit": "walker-steps/s", "cores": n_threads, "kind": "port",
                                "sample": f"first {n} walkers of rank 0 x {steps} timesteps (one sprint), {dt:.1f} s, "
                                          f"{n_threads} pthreads over walkers"}
    if world == 1 and args.workload == DEFAULT_WORKLOAD and not args.no_small_configs:
        ctx.close()
        line["small_configs"] = small_configs(api, torch, local, args)
        if not args.no_cpu_baseline:
            try:                                                   # informational: must never cost the contract line
                line["reference_gpu"] = reference_gpu_block(api, torch, local)
            except Exception as e:
                line["reference_gpu"] = {"unavailable": f"{type(e).__name__}: {e}"}
    emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


def reference_gpu_block(api, torch, local):
    """Informational (part of the reported-baseline leg, like cpu_baseline): the reference's OWN GPU code -- cuda_sim.cu +
    gpu_random.cu (cuRAND) + cuBLAS dgemv + thrust::reduce, compiled for sm_100 where it lies by `make -C oracle refgpu`
    (oracle/_ref/ref_gpu_short: the SHIPPED parameters.h -- 4032 waters, sprints of 20 000 steps -- cut to the first 2 of its
    40 ms) -- run on this GPU, next to this library on the same configuration.  Its walker randomness is cuRAND seeded from the clock, so the two runs agree
    statistically only (tests/test_ref_gpu_crosscheck.py); the bit-level pin is tests/test_ref_gpu_bitpin.py."""
    import re
    import subprocess
    import tempfile
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_gpu_short")
    if not os.path.exists(exe):
        return {"unavailable": "oracle/_ref/ref_gpu_short not built (make -C oracle refgpu needs the reference tree)"}
    n, sprint, total_ms = 4032, 20000, 2
    steps_total = total_ms * 1_000_000                        # tau = 1e-6 ms per step (parameters.h)
    try:
        with tempfile.TemporaryDirectory() as tmp:
            out = subprocess.run([exe, os.path.join(tmp, "ref.csv")], capture_output=True, text=True, timeout=300, cwd=tmp)
        m = re.search(r"Elapsed time: ([0-9.eE+]+) ms", out.stdout)
        if out.returncode != 0 or not m:
            return {"unavailable": "reference GPU binary failed: " + (out.stderr or out.stdout)[-200:]}
        ref_ms = float(m.group(1))
    except Exception as e:                                     # informational: never fail the bench line
        return {"unavailable": f"reference GPU binary: {e}"}
    p = api.ParameterStruct(num_water=n, sprintSteps=sprint, totaltime=total_ms)
    assert p.t == steps_total
    box = api.BacteriaBox(p, 12345).populateSimulation(device=local)
    ctx = api.WalkerContext(p, local)
    ctx.set_scene(box)
    ctx.set_waters(box.getWaters().copy(), WALKER_SEED)
    ctx.run_sprint(sprint)                                     # warm-up (buffers, jump tables)
    ctx.set_waters(box.getWaters().copy(), WALKER_SEED)
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(steps_total // sprint):
        ctx.run_sprint(sprint)
    torch.cuda.synchronize()
    our_ms = (time.perf_counter() - t) * 1e3
    ctx.close()
    return {"config": f"shipped parameters.h (BASELINE configs[0]) cut to {total_ms} of its 40 ms: {n} waters, {steps_total} steps in sprints of {sprint} (oracle/Makefile GCFG_short)",
            "reference_gpu_code": {"value": n * steps_total / (ref_ms * 1e-3), "unit": "walker-steps/s", "ms": ref_ms,
                                   "what": "the reference's cuda_sim.cu + cuRAND + cuBLAS + thrust recompiled for sm_100, its own 'Elapsed time' (sprint loop)"},
            "this_library": {"value": n * steps_total / (our_ms * 1e-3), "unit": "walker-steps/s", "ms": our_ms,
                             "what": f"mrisim_run_sprint x {steps_total // sprint}, wall clock incl. the signal samples read back after every sprint"},
            "ratio": ref_ms / our_ms}


def kernel_source_digest():
    import hashlib
    h = hashlib.sha256()
    for rel in ("mri_nanoparticle_simulation_b200/csrc/mrisim.cu", "include/mrisim_math.h"):
        with open(os.path.join(ROOT, rel), "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def load_instr_mix():
    path = os.path.join(ROOT, "profiles", "instr_mix_r02.json")
    if not os.path.exists(path):
        return None
    try:
        mix = json.load(open(path))
    except Exception:
        return None
    if mix.get("source_digest") != kernel_source_digest():
        return {"stale": True}
    return mix


def small_configs(api, torch, local, args, sprints=5, warm=2):
    """The reference's own shipped sizes (BASELINE configs[0-1]: 4032 / 8000 waters, the few-walker path),
    device-timed with CUDA events on the launch stream, L2 flushed between sprints: one entry per config."""
    out = {}
    stream = torch.cuda.current_stream()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=torch.device("cuda", local))
    for name in SMALL_CONFIGS:
        box, p, n = get_scene(api, name, 0, 1, lambda: None, device=local)
        waters = throw_waters(box, p, n, 0, device=local)
        ctx = api.WalkerContext(p, local)
        ctx.set_scene(box)
        ctx.set_waters(waters, WALKER_SEED, 0)
        steps = p.sprintSteps
        for _ in range(warm):
            ctx.submit_sprint(steps, stream.cuda_stream); ctx.wait_sprint(steps)
        ms = []
        for _ in range(sprints):
            flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            ctx.submit_sprint(steps, stream.cuda_stream); ctx.stream_wait_sprint(stream.cuda_stream)
            b.record(stream); b.synchronize()
            ctx.wait_sprint(steps)
            ms.append(a.elapsed_time(b))
        out[name] = {"value": n * steps * sprints / (sum(ms) * 1e-3), "unit": "walker-steps/s", "walkers": n, "sprints": sprints,
                     "ms_per_sprint": sum(ms) / sprints, "path": "pipeline" if ctx.pipeline_sprints else "fused"}
        ctx.close()
    return out


if __name__ == "__main__":
    sys.exit(main())
