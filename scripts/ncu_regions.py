"""Per-region issue-slot accounting of an ncu report of walker_sprint_kernel.
   python scripts/ncu_regions.py <report.ncu-rep> <warp-steps of the profiled launch>
Regions are found by source line ranges (needs -lineinfo): each SASS instruction is charged to the
function its CUDA source line belongs to.  Issue slots = 2 x FP64 instructions + 1 x the others
(an FP64 warp instruction occupies the SMSP issue port for two cycles on B200, profiles/fp64_issue_probe_r01.txt)."""
import collections, csv, io, re, subprocess, sys
rep, N = sys.argv[1], float(sys.argv[2])
src_path = __file__.rsplit("/scripts/", 1)[0] + "/mri_nanoparticle_simulation_b200/csrc/mrisim.cu"
src = open(src_path).read().split("\n")
# map each line of mrisim.cu to the enclosing top-level function name
fn_of, cur = {}, "?"
for i, l in enumerate(src, 1):
    m = re.match(r"^(?:template.*>\s*)?(?:__device__|__global__|static|inline).*?\b(\w+)\s*\(", l) or re.match(r"^(\w+)\(const __grid_constant__", l)
    if m: cur = m.group(1)
    fn_of[i] = cur
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
agg = collections.defaultdict(lambda: [0, 0, 0]); fname = None; hdr = None; curfn = None; seen = set()
FP64 = ("DFMA", "DMUL", "DADD", "DSETP")
for row in csv.reader(io.StringIO(txt)):
    if not row: continue
    if row[0] == "File Path": fname = row[1].split("/")[-1]; hdr = None; continue
    if row[0] == "Line No": hdr = row; ie = hdr.index("Instructions Executed"); ism = hdr.index("# Samples"); continue
    if hdr is None or len(row) <= ism: continue
    if row[0] != "":
        try: ln = int(row[0])
        except ValueError: continue
        curfn = fn_of.get(ln, "?") if fname == "mrisim.cu" else fname
        continue
    try: e = int(row[ie])
    except ValueError: continue
    if row[2] in seen: continue          # an inlined instruction is listed under every source line of its inline stack
    seen.add(row[2])
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_]+)", row[3]); op = m.group(2) if m else "?"
    a = agg[curfn]; a[0] += e; a[2] += int(row[ism])
    if op in FP64: a[1] += e
T = sum(a[0] for a in agg.values()); F = sum(a[1] for a in agg.values()); S = sum(a[2] for a in agg.values())
print(f"per warp-step: {T/N:.1f} instructions, {F/N:.1f} FP64 -> {(T+F)/N:.1f} issue slots")
print(f"{'function':34s} {'inst':>7s} {'fp64':>7s} {'slots':>7s} {'share':>6s} {'samples':>7s}")
for k, a in sorted(agg.items(), key=lambda kv: -(kv[1][0] + kv[1][1])):
    if a[0] / N < 0.05: continue
    print(f"{k:34s} {a[0]/N:7.1f} {a[1]/N:7.1f} {(a[0]+a[1])/N:7.1f} {100*(a[0]+a[1])/(T+F):5.1f}% {100*a[2]/max(S,1):6.1f}%")
