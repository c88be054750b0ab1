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
