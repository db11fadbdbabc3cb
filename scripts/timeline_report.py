"""Text report of the CTA timelines behind the 'overlapping walker groups' change (profiles/timeline_r02_groups_and_sort.txt).
   python scripts/timeline_report.py <cub_1group.bin> <cub_2groups.bin> <own_sort_Ngroups.bin> <N>
Each .bin is a MRISIM_TIMELINE dump: the last 8 launches, per CTA {SM, start ns, end ns}."""
import sys
import numpy as np

a, b, c, gc = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4])
print("""CTA timelines of walker_sprint_kernel (MRISIM_TIMELINE: %smid + %globaltimer at CTA start, atomicMax of the warps' ends),
B200, cfg4a dense 2000-cell scene, 1 Mi walkers, 125-step launches; times in us from the first CTA of the first recorded launch.

The question: two walker groups with separate (sort -> kernel) chains on separate streams were meant to fill each other's
tails, and measured NO gain in round 1 and again early in round 2.  Why?
""")


def show(path, g, title):
    t = np.fromfile(path, dtype=np.uint64).reshape(8, -1, 3)
    t0 = t[:, :, 1][t[:, :, 1] > 0].min()
    per = -(-(-(-(t.shape[1] * 128) // g)) // 128)          # group_ranges(): ceil(ceil(n / g) / 128) CTAs per group, the last takes the rest
    sizes = [per] * g
    print(title)
    for f in range(8):
        lo = 0
        for k in range(g):
            s = t[f, lo:lo + sizes[k]] if k < g - 1 else t[f, lo:]
            lo += sizes[k]
            s = s[s[:, 1] > 0]
            st = (s[:, 1].astype(np.int64) - t0) / 1e3
            en = (s[:, 2].astype(np.int64) - t0) / 1e3
            print(f"  launch {f} group {k}: first CTA starts {st.min():8.0f}   last CTA starts {st.max():8.0f}   kernel ends {en.max():8.0f}")
    span = (t[:, :, 2].max() - t0) / 1e3
    print(f"  -> {span / 8:.0f} us per 125 steps of all walkers\n")


show(a, 1, "A. one group, CUB radix sort between launches (the shipped path of round 1): every launch ends in a ~650 us tail")
show(b, 2, "B. two groups, CUB radix sort: group 1 fills group 0's tail INSIDE a launch pair, but group 0's next kernel starts only when\n"
           "   group 1's kernel has ENDED (5093 after 5020, 10166 after 10095, ...) although its own kernel ended a millisecond earlier")
show(c, gc, f"C. {gc} groups, the library's own small-shared-memory sort (shipped): a group's next kernel starts while the others' are still\n"
            "   running -- the chains overlap, no SM waits for a tail")
print("""Reading: an SM runs CTAs of kernels with different L1/shared-memory splits one split at a time.  CUB's sort kernels want tens of
KB of shared memory per CTA, the walker kernel wants all of L1; so the sort of group 0 cannot start on any SM that still holds a
walker CTA of group 1, i.e. not before group 1's kernel has drained.  Cross-checks (same session): with the sort taken out (stale
order) two groups gained 8 % over one; with the walker kernel forced to a 64 / 132 / 196 KB shared-memory split two groups gained
11 % over one (at a 17 % loss from the smaller L1).  The fix is a sort that fits the walker kernel's split: csrc/mrisim.cu
sort_*_kernel (1-6 KB per CTA), every kernel between two walker launches pinned to the same split (kCarveoutPercent).""")
