import cProfile, pstats, sys, numpy as np, torch
sys.path.insert(0, ".")
import tianshou_fork_b200 as ts
E, T = 16, 1 << 12
buf = ts.PrioritizedVectorReplayBuffer(E * T, E, alpha=0.6, beta=0.4, stack_num=4, ignore_obs_next=True, save_only_last_obs=True)
rng = np.random.default_rng(0)
for t in range(64):
    buf.add(ts.Batch(obs=np.zeros((E, 4, 84, 84), np.uint8), act=np.zeros(E, np.int64), rew=np.zeros(E),
                     terminated=rng.random(E) < 0.05, truncated=np.zeros(E, bool), info={}))
for _ in range(200): buf.sample(32)
torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable()
for _ in range(2000): buf.sample(32)
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(35)
