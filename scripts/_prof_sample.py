import cProfile, pstats, sys, numpy as np, torch
sys.path.insert(0, ".")
import tianshou_fork_b200 as ts
N, m = 131072, 64
rng = np.random.default_rng(5)
f = lambda a: torch.from_numpy(a.astype(np.float32)).cuda()
adv, ret, vs = f(rng.standard_normal(N)), f(rng.standard_normal(N)), f(rng.standard_normal(N))
lo = f(rng.normal(-8, 0.3, N))
mb = torch.from_numpy(rng.permutation(N)[:m]).cuda()
lp = (lo[mb] + 0.1 * torch.randn(m, device="cuda")).requires_grad_(True)
val = (vs[mb] + 0.3 * torch.randn(m, device="cuda")).requires_grad_(True)
def fused():
    cl, vl, _ = ts.ppo_loss(lp, val, lo, adv, ret, vs, mb, 0.2, None, True, True)
    (cl + 0.5 * vl).backward()
for _ in range(200): fused()
torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable()
for _ in range(2000): fused()
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
