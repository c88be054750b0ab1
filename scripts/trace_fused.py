"""Summarise a DGB_FUSED_TRACE file: start / split-end / first-run-end / end per CTA (ns, relative)."""
import sys, numpy as np
a = np.loadtxt(sys.argv[1], dtype=np.uint64)
t = a[:, 1:5].astype(np.int64); role = (a[:, 5] & 1).astype(int)
base = t[:, 0].min(); t -= base
print("CTAs", len(a), "split-role", role.sum())
print("kernel span us", t[:, 3].max() / 1e3)
print("start spread us", t[:, 0].max() / 1e3)
s = role == 1
print("split-role: split phase end us: min %.1f med %.1f max %.1f" % tuple(np.percentile(t[s, 1], [0, 50, 100]) / 1e3))
print("encode-role: first run end us: min %.1f med %.1f max %.1f" % tuple(np.percentile(t[~s, 2], [0, 50, 100]) / 1e3))
print("all: end us: min %.1f p10 %.1f med %.1f max %.1f" % tuple(np.percentile(t[:, 3], [0, 10, 50, 100]) / 1e3))
if a.shape[1] > 6:
    te = a[:, 6].astype(np.int64) - base
    print("all: encode loop end (warp 0) us: min %.1f p10 %.1f med %.1f p90 %.1f max %.1f" % tuple(np.percentile(te, [0, 10, 50, 90, 100]) / 1e3))
    print("all: drain (compaction) time per CTA us: min %.1f med %.1f max %.1f" % tuple(np.percentile(t[:, 3] - te, [0, 50, 100]) / 1e3))

# per-element timeline of a build with -DDGB_TRACE_ELEMS (make -C dietgpu_fork_b200/csrc trace): <trace>.elems
import os
if os.path.exists(sys.argv[1] + ".elems"):
    e = np.loadtxt(sys.argv[1] + ".elems", dtype=np.uint64).reshape(-1, 7)
    v = e[:, 1:6].astype(np.int64) - base
    wait = e[:, 6].astype(np.int64)
    print("element: ready | first run start | last block end | first compaction start | last compaction end | summed wait of its first-run warps (us)")
    step = max(1, len(e) // 32)
    for i in list(range(0, len(e), step)) + [len(e) - 1]:
        print("  %4d  %7.1f %7.1f %7.1f %7.1f %7.1f   wait %8.1f" % ((int(e[i, 0]),) + tuple(v[i] / 1e3) + (wait[i] / 1e3,)))
    d = np.diff(v[:, 0]) / 1e3
    print("ready-to-ready interval us: med %.2f p90 %.2f max %.2f" % tuple(np.percentile(d, [50, 90, 100])))
    print("ready -> first run start us: med %.1f max %.1f" % tuple(np.percentile((v[:, 1] - v[:, 0]) / 1e3, [50, 100])))
    print("first run start -> last block end us: med %.1f min %.1f max %.1f" % tuple(np.percentile((v[:, 2] - v[:, 1]) / 1e3, [50, 0, 100])))
    print("last block end -> first compaction start us: med %.1f max %.1f" % tuple(np.percentile((v[:, 3] - v[:, 2]) / 1e3, [50, 100])))
    print("compaction duration us: med %.1f max %.1f" % tuple(np.percentile((v[:, 4] - v[:, 3]) / 1e3, [50, 100])))
    print("total wait (warp-us): %.0f" % (wait.sum() / 1e3))
