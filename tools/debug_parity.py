import sys, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sparknet_b200 import zoo
from sparknet_b200.library import load_library
from sparknet_b200.net import CaffeNet, MinibatchSource
from tests.harness import load_oracle, synthetic_batches, rel_err
oracle, product = load_oracle(), load_library()
name, batch = sys.argv[1], int(sys.argv[2])
sol = zoo.solver_with_net(name, train_bs=batch, test_bs=batch, random_seed=1701, display=0)
o, p = CaffeNet.from_solver(sol, oracle), CaffeNet.from_solver(sol, product)
product.sparknet_set_precision(p.state, int(os.environ.get("PREC", "0")))
imgs, labs = synthetic_batches(name, 1, batch)
o.setTrainData(MinibatchSource(imgs, labs)); p.setTrainData(MinibatchSource(imgs, labs))
p.setWeights(o.getWeights())
o.forward(); p.forward()
do, dp = o.getData(), p.getData()
for k in do: print("fwd", k, do[k].shape, "%.2e" % rel_err(dp[k], do[k]))
o.backward(); p.backward()
go, gp = o.getData(diff=True), p.getData(diff=True)
for k in go:
    a, b = go[k].astype(np.float64), gp[k].astype(np.float64)
    sc = max(np.abs(a).max(), 1e-30)
    bad = np.abs(a-b) > 1e-5*sc
    print("bwd", k, "rel %.2e" % rel_err(gp[k], go[k]), "bad frac %.2e" % bad.mean(), "nnz o/p", (a!=0).sum(), (b!=0).sum())
    if bad.any() and k in ("conv1", "pool1"):
        idx = np.argwhere(bad)[:8]
        for i in idx: print("   ", tuple(i), a[tuple(i)], b[tuple(i)])
wo, wp = o.getWeightDiffs(), p.getWeightDiffs()
for l in wo:
    for j,(a,b) in enumerate(zip(wo[l], wp[l])): print("wdiff", l, j, "%.2e" % rel_err(b, a))
