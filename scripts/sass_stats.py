"""Static SASS statistics of one kernel of a built library/object:  python scripts/sass_stats.py <file> [name substring]"""
import collections, re, subprocess, sys
path = sys.argv[1]; want = sys.argv[2] if len(sys.argv) > 2 else "walker_sprint_kernelILb0ELi0ELi0E"
out = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
cur = None; ops = collections.Counter()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m: cur = m.group(1); continue
    if cur and want in cur:
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m: ops[m.group(1).split(".")[0]] += 1
tot = sum(ops.values())
fp64 = sum(ops[k] for k in ("DFMA", "DMUL", "DADD", "DSETP"))
print(f"{want}: {tot} instructions, FP64 {fp64}, LDL {ops['LDL']}, STL {ops['STL']}")
print("  " + "  ".join(f"{k} {v}" for k, v in ops.most_common(24)))
