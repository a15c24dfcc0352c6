import os, sys, time
import numpy as np, torch
sys.path.insert(0, "/root/repo")
import __graft_entry__ as graft, bench
pkg = graft.load_package(); ctx = pkg.Context(0)
os.environ["PM_KD_ONLY"] = "1"
cloud = bench.make_cloud(10_000_000)
d_tree = torch.from_numpy(cloud.view(np.uint8)).cuda()
pmap = pkg.PhotonMap.build(ctx, d_tree)
q = bench.make_queries(200_000, 0)
# far queries: outside the cloud
qf = (q + np.float32(60.0)).astype(np.float32)
for name, qq in (("inside", q), ("far", qf)):
    d_q = torch.from_numpy(qq).cuda()
    for it in range(3):
        ids, d2, r2 = pmap.knn(d_q, 64, 100.0)
    print(name, "kernels ms", ctx.last_kernel_ms, "fallbacks", ctx.last_fallback_count, flush=True)
