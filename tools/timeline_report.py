#!/usr/bin/env python
"""Summary of a per-CTA timeline dump of the container scan (tools/k1_experiment.py, -DSB_K1_TIMELINE build)."""
import sys
import numpy as np
b = np.load(sys.argv[1]).astype(np.int64)
sm, t1, t2, t3, t4, t5 = b[:, 0], b[:, 1], b[:, 2], b[:, 3], b[:, 4], b[:, 5]
nl, nu = b[:, 6] >> 32, b[:, 6] & 0xffffffff
T0 = t1.min()
us = lambda x: x / 1e3
dur = us(t5 - t1)
print("CTAs", len(b), "kernel span us %.1f" % us(t5.max() - T0))
print("CTA life us: mean %.2f median %.2f p90 %.2f max %.2f" % (dur.mean(), np.median(dur), np.percentile(dur, 90), dur.max()))
for name, x in (("startup", t2 - t1), ("stream", t3 - t2), ("barrier", t4 - t3), ("flush", t5 - t4)):
    print("  %-8s mean %.2f median %.2f p90 %.2f  (share of CTA time %.0f%%)" % (name, us(x).mean(), np.median(us(x)), np.percentile(us(x), 90), 100 * x.sum() / (t5 - t1).sum()))
npost = nu - nl
X = np.stack([np.ones(len(b)), npost, nl], 1)
coef = np.linalg.lstsq(X, us(t4 - t2), rcond=None)[0]
print("stream+barrier ~ %.2f + %.3f * posting chunks + %.3f * large units   (chunks %d, large units %d)" % (*coef, npost.sum(), nl.sum()))
ends = np.array([us(t5[sm == s].max() - T0) for s in np.unique(sm)])
print("SM end times: min %.1f mean %.1f max %.1f" % (ends.min(), ends.mean(), ends.max()))
ev = np.concatenate([np.stack([t1, np.ones_like(t1)], 1), np.stack([t5, -np.ones_like(t5)], 1)])
ev = ev[np.argsort(ev[:, 0])]
c, tt = np.cumsum(ev[:, 1]), us(ev[:, 0] - T0)
print("resident CTAs by 4 us window:", [int(c[(tt >= q) & (tt < q + 4)].mean()) for q in range(0, int(tt.max()) + 1, 4) if ((tt >= q) & (tt < q + 4)).any()])
