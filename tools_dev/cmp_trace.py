#!/usr/bin/env python3
"""dev helper: compare two traces (reference vs engine), report the first mismatching step/vehicle."""
import sys
import numpy as np
sys.path.insert(0, __file__.rsplit('/', 2)[0])
from sumo_b200.flatfile import load_trace, load_flat, names

def main():
    a = load_trace(sys.argv[1]); b = load_trace(sys.argv[2])
    flat = load_flat(sys.argv[3]) if len(sys.argv) > 3 else None
    lanes = names(flat, "lane") if flat else None
    vnames = None
    if flat:
        vn = names(flat, "veh"); vnames = {int(n): s for n, s in zip(flat["veh_numid"], vn)}
    n = min(len(a), len(b))
    for i in range(n):
        ta, ida, la, pa, sa = a[i]; tb, idb, lb, pb, sb = b[i]
        if ta != tb:
            print("time mismatch at step", i, ta, tb); return 1
        if len(ida) != len(idb) or (ida != idb).any():
            sa_, sb_ = set(ida.tolist()), set(idb.tolist())
            print("t=%d: vehicle sets differ: only ref %s only engine %s" % (ta, [vnames[x] if vnames else x for x in sorted(sa_ - sb_)][:10], [vnames[x] if vnames else x for x in sorted(sb_ - sa_)][:10])); return 1
        bad = (la != lb) | (pa.view(np.int64) != pb.view(np.int64)) | (sa.view(np.int64) != sb.view(np.int64))
        if bad.any():
            k = np.nonzero(bad)[0]
            print("t=%d: %d of %d vehicles differ" % (ta, len(k), len(ida)))
            for j in k[:8]:
                print("  veh %s ref lane=%s pos=%.17g speed=%.17g | engine lane=%s pos=%.17g speed=%.17g" % (
                    vnames[int(ida[j])] if vnames else ida[j], lanes[la[j]] if lanes else la[j], pa[j], sa[j], lanes[lb[j]] if lanes else lb[j], pb[j], sb[j]))
            return 1
    print("identical for %d steps (ref %d, engine %d); vehicle-steps compared: %d" % (n, len(a), len(b), sum(len(x[1]) for x in a[:n])))
    return 0 if len(a) == len(b) else 2

sys.exit(main())
