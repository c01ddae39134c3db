import sys; sys.path.insert(0,'.')
import numpy as np
from tests import helpers
from oracle.binding import Oracle
O, corp, ali = helpers.trained_oracle(dim=25, n_utts=24, splits=2)
ctx = helpers.gpu_context(O, corp, 25)
ctx.set_alignment(ali)
cnt, sx, sxx, ssum = ctx.accumulate(False)
dens,_ = ctx.last_frame_results()
P = Oracle(25, O.n_mix, "none", True, O.tdp, O.silence); P.load_tables(O.tables())
wd = P.accumulate(ali, corp["x"]); t = P.tables()
wc, wsx = helpers.packed_stats(t, 25)
print(cnt.shape, wc.shape, (cnt!=wc).sum(), np.nonzero(cnt!=wc)[0][:10], cnt[cnt!=wc][:10], wc[cnt!=wc][:10])
off = t["mix_offsets"]; g = off[ali] + dens
bc = np.bincount(g, minlength=len(wc)); print((bc!=wc).sum(), (bc!=cnt).sum(), cnt.sum(), wc.sum(), len(ali))
