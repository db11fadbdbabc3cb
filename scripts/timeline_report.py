"""Text report of the CTA timelines behind the 'overlapping walker groups' change (profiles/timeline_r02_groups_and_sort.txt).
   python scripts/timeline_report.py <cub_1group.bin> <cub_2groups.bin> <own_sort_Ngroups.bin> <N>
Each .bin is a MRISIM_TIMELINE dump: the last 8 launches, per CTA {SM, start ns, end ns}."""
import sys
import numpy as np

a, b, c, gc = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4])
print("""CTA timelines of walker_sprint_kernel (MRISIM_TIMELINE: %smid + %globaltimer at CTA start, atomicMax of the warps' ends),
B200, cfg4a dense 2000-cell scene, 1 Mi walkers, 125-step launches; times in us from the first CTA of the first recorded launch.

The question: walker groups with separate (sort -> kernel) chains on separate streams fill each other's tails -- how well,
and what does the sort between two launches of a group cost them?
(Round 1 and the first sessions of round 2 recorded "two groups: no gain".  That measurement was void: the scripts passed the
group count as `GROUPS=2 python ...`, and bash keeps GROUPS to itself -- the assignment never reached the process, every run
used one group.  The scripts now say WGROUPS.)
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
show(b, 2, "B. two groups, CUB radix sort (commit 232218b): the chains overlap, but every kernel starts ~280-300 us after the other group's\n"
           "   last CTA was dispatched (2959 after 2682, 4678 after 4387, ...): the sort before it had to wait for SMs to drain")
show(c, gc, f"C. {gc} groups, the library's own small-shared-memory sort (shipped): a group's next kernel starts ~70-170 us after the\n"
            "   previous group's last CTA was dispatched -- sort and tail both hide behind the other groups' CTAs")
print("""Reading: an SM runs CTAs of kernels with different L1/shared-memory splits one split at a time.  CUB's sort kernels want tens of
KB of shared memory per CTA, the walker kernel wants all of L1; so a group's sort cannot start on an SM that still holds a walker
CTA of another group and waits until some SMs have drained completely.  Whole sprints, ms (20 000 steps, same session):
  CUB sort:            1 group 802.6   2 groups 768.7 (+4.4 %)   3 groups 764.1 (+5.0 %)
  own sort (shipped):  1 group 807.7   2 groups 723.0 (+11.7 %)  3 groups 722.8          4 groups 723.2
Cross-checks: with the sort taken out (stale order) two groups gained 8 % over one; with the walker kernel forced to a 64 / 132 /
196 KB shared-memory split and CUB's sort, two groups gained 11-12 % over one (at a 17 % loss from the smaller L1).  Hence a sort
that fits the walker kernel's split: csrc/mrisim.cu sort_*_kernel (1-6 KB per CTA), every kernel between two walker launches
pinned to the same split (kCarveoutPercent).""")
