"""Aggregate an ncu report by CUDA source line and by SASS opcode.
   python scripts/ncu_lines.py report.ncu-rep <warps*steps of the profiled launch> [rows]      (SORT=samples: order by stall samples)"""
import collections, csv, io, os, re, subprocess, sys
rep, N = sys.argv[1], float(sys.argv[2])
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
lines = collections.defaultdict(lambda: [0, 0, 0, ""]); ops = collections.Counter(); opthr = collections.Counter()
fname = None; hdr = None; seen = set(); cur = None
for row in csv.reader(io.StringIO(txt)):
    if not row: continue
    if row[0] == "File Path": fname = row[1].split("/")[-1]; hdr = None; continue
    if row[0] == "Function Name": continue
    if row[0] == "Line No": hdr = row; ie = hdr.index("Instructions Executed"); it = hdr.index("Thread Instructions Executed"); ism = hdr.index("# Samples"); continue
    if hdr is None or len(row) <= ism: continue
    try: e = int(row[ie]); t = int(row[it]); s = int(row[ism])
    except ValueError: continue
    if row[0] != "":                       # aggregate row of one source line (its totals double-count inlined code)
        cur = lines[(fname, int(row[0]))]; cur[3] = row[1].strip()[:90]
    else:                                  # one SASS instruction attributed to that line; charged once
        if row[2] in seen: continue
        seen.add(row[2])
        cur[0] += e; cur[1] += t; cur[2] += s
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_]+)", row[3]); op = m.group(2) if m else "?"
        ops[op] += e; opthr[op] += t
T = sum(v[0] for v in lines.values()); S = sum(v[2] for v in lines.values())
print(f"total warp-instr per warp-step: {T/N:.1f}   FP64-pipe (DFMA+DMUL+DADD+DSETP): {sum(ops[o] for o in ('DFMA','DMUL','DADD','DSETP'))/N:.1f}")
print("top opcodes:", ", ".join(f"{o} {c/N:.0f} ({opthr[o]/max(c,1):.0f}ln)" for o, c in ops.most_common(16)))
print(f"{'file:line':28s} {'inst/step':>9s} {'share':>6s} {'lanes':>5s} {'stall%':>6s}  source")
for (f, ln), (e, t, s, src) in sorted(lines.items(), key=lambda kv: -kv[1][2 if os.environ.get("SORT") == "samples" else 0])[:int(sys.argv[3]) if len(sys.argv) > 3 else 45]:
    print(f"{f+':'+str(ln):28s} {e/N:9.1f} {100*e/T:5.1f}% {t/max(e,1):5.1f} {100*s/max(S,1):5.1f}%  {src}")
