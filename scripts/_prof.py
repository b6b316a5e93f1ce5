import cProfile, pstats, sys, numpy as np, torch
sys.path.insert(0, "."); sys.path.insert(0, "scripts")
import bench_rows as br
import tianshou_fork_b200 as ts
E, T = 64, 2048
rng = np.random.default_rng(9)
buf = ts.VectorReplayBuffer(E * T, E, rew_dtype=torch.float32)
br.fill_vec(buf, E, T, rng, 1 / 200, 1 / 1000, rew3=False)
batch, indices = buf.sample(0)
v_s = torch.randn(E * T, device="cuda"); v_n = torch.randn(E * T, device="cuda")
f = lambda: ts.BasePolicy.compute_episodic_return(batch, buf, indices, v_n, v_s, 0.99, 0.95)
for _ in range(200): f()
torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable()
for _ in range(2000): f()
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(18)
idx2 = indices.clone(); idx2[0], idx2[1] = indices[1], indices[0]  # not canonical
b2 = buf[idx2]
g = lambda: ts.BasePolicy.compute_episodic_return(b2, buf, idx2, v_n, v_s, 0.99, 0.95)
for _ in range(200): g()
torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable()
for _ in range(2000): g()
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(18)
