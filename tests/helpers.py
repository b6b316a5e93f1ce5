"""Shared test helpers: replay of the add-streams that tests/golden/make_golden.py used."""
import numpy as np


def golden_stream(seed, steps, n_env, p_term, p_trunc, ragged=False):
    """Yields (ids, batch_dict) in exactly the RNG order of make_golden.fill_vecbuf."""
    rng = np.random.default_rng(seed)
    return rng, _stream(rng, steps, n_env, p_term, p_trunc, ragged)


def _stream(rng, steps, n_env, p_term, p_trunc, ragged):
    for _ in range(steps):
        ids = np.arange(n_env)
        if ragged:
            ids = ids[rng.random(n_env) < 0.7]
            if ids.size == 0:
                continue
        m = ids.size
        term = rng.random(m) < p_term
        trunc = (rng.random(m) < p_trunc) & ~term
        yield ids, dict(
            obs=rng.standard_normal((m, 3)).astype(np.float32),
            act=rng.integers(0, 4, m),
            rew=rng.standard_normal(m),
            terminated=term,
            truncated=trunc,
            obs_next=rng.standard_normal((m, 3)).astype(np.float32),
            info={},
        )


def fill(buf, Batch, seed, steps, n_env, p_term, p_trunc, ragged=False):
    rng, it = golden_stream(seed, steps, n_env, p_term, p_trunc, ragged)
    for ids, d in it:
        buf.add(Batch(**d), buffer_ids=ids)
    return rng


GAE_SPECS = [
    # name, n_env, steps, cap, gamma, lam, p_term, p_trunc, ragged  (make_golden.gen_gae)
    ("small", 4, 37, 64, 0.99, 0.95, 0.1, 0.05, False),
    ("wrap", 3, 150, 100, 0.99, 0.95, 0.05, 0.02, False),
    ("ragged", 5, 60, 80, 0.9, 0.8, 0.1, 0.05, True),
    ("long4096", 2, 4096, 4096, 0.99, 0.95, 1 / 500, 1 / 2000, False),
    ("mc4096", 2, 4096, 4096, 1.0, 1.0, 1 / 1000, 0.0, False),
    ("noend", 2, 512, 512, 0.99, 0.95, 0.0, 0.0, False),
]

INDEX_SPECS = [("v_partial", 4, 20, 11, True), ("v_wrap", 4, 20, 57, False),
               ("v_big", 16, 1024, 700, True), ("v_one", 1, 16, 40, False)]

NSTEP_SPECS = [("n1", 4, 64, 40, 1, 1), ("n3", 4, 64, 100, 3, 1), ("n3_d51", 4, 64, 100, 3, 51),
               ("n10", 8, 256, 300, 10, 1), ("n5_partial", 6, 600, 37, 5, 3)]


def c2_rollout(E, T, seed=0, p_term=1 / 200, p_trunc=1 / 1000):
    """Synthetic rollout of SURVEY.md 8d / BASELINE.json configs[1] (f32 values, bool flags)."""
    rng = np.random.default_rng(seed)
    rew = rng.random((E, T), dtype=np.float32)
    v_s = rng.standard_normal((E, T), dtype=np.float32)
    v_next = np.empty_like(v_s)
    v_next[:, :-1] = v_s[:, 1:]
    v_next[:, -1] = rng.standard_normal(E, dtype=np.float32)
    term = rng.random((E, T)) < p_term
    trunc = (rng.random((E, T)) < p_trunc) & ~term
    return rew, v_s, v_next, term, trunc


STACK_SPECS = [
    # name, n_env, total, steps, stack_num, ragged, frame shape (make_golden.gen_stack)
    ("atari_like", 4, 80, 13, 4, False, (6,)),
    ("wrapped", 3, 30, 47, 4, True, (2, 3)),
    ("deep", 2, 64, 25, 7, False, (4,)),
]


def stack_stream(rng, steps, n_env, ragged, frame_shape, hist):
    """The add-stream of make_golden.stack_stream, in the same RNG order."""
    for _ in range(steps):
        ids = np.arange(n_env)
        if ragged:
            ids = ids[rng.random(n_env) < 0.7]
            if ids.size == 0:
                continue
        m = ids.size
        term = rng.random(m) < 0.2
        trunc = (rng.random(m) < 0.1) & ~term
        yield ids, dict(
            obs=rng.integers(0, 256, (m, hist) + frame_shape).astype(np.uint8),
            act=rng.integers(0, 4, m), rew=rng.standard_normal(m), terminated=term,
            truncated=trunc, info={})


def random_rings(rng, B, cap, p_done=0.02):
    """A plausible manager state: (offset[B+1], done[B*cap], last_index[B], lengths[B])."""
    offset = np.arange(B + 1, dtype=np.int64) * cap
    lengths = rng.integers(cap // 2, cap + 1, B).astype(np.int64)
    lengths[0] = cap
    last = (offset[:-1] + rng.integers(0, lengths)).astype(np.int64)
    done = rng.random(B * cap) < p_done
    return offset, done, last, lengths


# ---- PPO update goldens (make_golden.gen_ppo_update) -------------------------------------------
PPO_UPDATE_SPECS = [
    # name, kwargs of PPOPolicy, batch_size, repeat, number of update() calls
    ("clip_norm", dict(eps_clip=0.2, value_clip=True, advantage_normalization=True, vf_coef=0.5,
                       ent_coef=0.01, max_grad_norm=0.5, gae_lambda=0.95, discount_factor=0.99),
     64, 2, 1),
    ("rewnorm_recompute_dual", dict(eps_clip=0.1, dual_clip=3.0, value_clip=False,
                                    advantage_normalization=False, recompute_advantage=True,
                                    reward_normalization=True, vf_coef=0.25, ent_coef=0.0,
                                    max_grad_norm=None, gae_lambda=0.9, discount_factor=0.97,
                                    max_batchsize=300),
     128, 3, 2),
]


def ppo_update_nets(torch):
    """Actor / critic of the PPO-update goldens (shared with tests/test_gpu_ppo.py)."""

    class Actor(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.net = torch.nn.Sequential(torch.nn.Linear(3, 64), torch.nn.Tanh(),
                                           torch.nn.Linear(64, 4))

        def forward(self, obs, state=None, info=None):
            obs = torch.as_tensor(obs, dtype=torch.float32,
                                  device=self.net[0].weight.device)
            return self.net(obs), state

    class Critic(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.net = torch.nn.Sequential(torch.nn.Linear(3, 64), torch.nn.Tanh(),
                                           torch.nn.Linear(64, 1))

        def forward(self, obs, **kw):
            obs = torch.as_tensor(obs, dtype=torch.float32,
                                  device=self.net[0].weight.device)
            return self.net(obs)

    return Actor, Critic


DQN_UPDATE_SPECS = [
    # name, DQNPolicy kwargs, PER kwargs, sample size, updates
    ("double_n3", dict(discount_factor=0.97, estimation_step=3, target_update_freq=2,
                       is_double=True), dict(alpha=0.6, beta=0.4), 64, 6),
    ("nature_n1_huber", dict(discount_factor=0.9, estimation_step=1, target_update_freq=0,
                             is_double=False, clip_loss_grad=True),
     dict(alpha=0.8, beta=1.0), 48, 4),
]




# ---- BASELINE.json configs[3] at full size (make_golden.gen_c4) ----------------------------------
C4_E, C4_SUB, C4_EXTRA = 16, 65536, 1234       # 16 sub-buffers x 65 536 = 2^20 slots, rings wrapped
C4_BATCHES = (32, 256, 4096)
C4_DIMS = (1, 51)


def c4_stream(seed=4000):
    """SURVEY.md 8d C4 rollout: rew in {-1, 0, 1} with P = (.05, .9, .05), terminated ~ B(1/1000),
    truncated ~ B(1/10000) & ~terminated; ``[steps, E]`` arrays, steps = 65 536 + 1234 so every
    ring has wrapped."""
    rng = np.random.default_rng(seed)
    steps = C4_SUB + C4_EXTRA
    rew = rng.choice(np.array([-1.0, 0.0, 1.0]), size=(steps, C4_E), p=[0.05, 0.9, 0.05])
    term = rng.random((steps, C4_E)) < 1 / 1000
    trunc = (rng.random((steps, C4_E)) < 1 / 10000) & ~term
    return rew, term, trunc


def c4_leaves(seed=4001):
    rng = np.random.default_rng(seed)
    return (np.abs(rng.standard_normal(C4_E * C4_SUB)) + np.finfo(np.float32).eps) ** 0.6


def c4_target_q(idx, D):
    """Deterministic ``target_q_fn`` table: exact in f32 on any library (integer hash / 2^16)."""
    idx = np.asarray(idx, dtype=np.int64)
    h = (idx[:, None] * 2654435761 + np.arange(D, dtype=np.int64)[None] * 40503) % 65536
    return (h / 65536.0 - 0.5).astype(np.float32)


def c4_final_storage(rew, term, trunc):
    """What the rings hold after the stream: slot j of a sub-buffer keeps the LAST step t with
    t % sub == j.  Returns ``[E * sub]`` arrays in storage order plus (ring index, last_index)."""
    steps = rew.shape[0]
    t_of_slot = np.arange(C4_SUB)
    t_of_slot = np.where(t_of_slot + C4_SUB < steps, t_of_slot + C4_SUB, t_of_slot)
    pick = lambda a: np.ascontiguousarray(a[t_of_slot].T).reshape(-1)  # noqa: E731
    ring_index = steps % C4_SUB
    last = np.arange(C4_E, dtype=np.int64) * C4_SUB + (ring_index - 1) % C4_SUB
    return pick(rew), pick(term), pick(trunc), ring_index, last


# ---- BASELINE.json configs[2]: Ant-shaped PPO update (make_golden.gen_ppo_ant) -------------------
ANT = dict(E=64, T=2048, obs=27, act=8, batch_size=1024, repeat=1)
ANT_KW = dict(eps_clip=0.2, value_clip=True, advantage_normalization=True, vf_coef=0.25,
              ent_coef=0.0, max_grad_norm=0.5, gae_lambda=0.95, discount_factor=0.99,
              max_batchsize=131072)


def ant_rollout(seed=4100):
    """``[T, E, ...]`` arrays of an Ant-shaped rollout (obs 27, act 8, 64 envs x 2048 steps)."""
    rng = np.random.default_rng(seed)
    T, E = ANT["T"], ANT["E"]
    obs = rng.standard_normal((T + 1, E, ANT["obs"]), dtype=np.float32)
    act = rng.standard_normal((T, E, ANT["act"]), dtype=np.float32)
    rew = rng.standard_normal((T, E))
    term = rng.random((T, E)) < 1 / 500
    trunc = (rng.random((T, E)) < 1 / 1000) & ~term
    return obs, act, rew, term, trunc


def ant_nets(torch):
    """Gaussian actor with a state-independent log-std (examples/mujoco ActorProb shape) and a
    critic, two 64 x 64 tanh MLPs."""

    class Actor(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.mu = torch.nn.Sequential(
                torch.nn.Linear(ANT["obs"], 64), torch.nn.Tanh(), torch.nn.Linear(64, 64),
                torch.nn.Tanh(), torch.nn.Linear(64, ANT["act"]))
            self.sigma = torch.nn.Parameter(torch.full((ANT["act"],), -0.5))

        def forward(self, obs, state=None, info=None):
            obs = torch.as_tensor(obs, dtype=torch.float32, device=self.sigma.device)
            mu = self.mu(obs)
            return (mu, self.sigma.exp().expand_as(mu)), state

    class Critic(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.net = torch.nn.Sequential(
                torch.nn.Linear(ANT["obs"], 64), torch.nn.Tanh(), torch.nn.Linear(64, 64),
                torch.nn.Tanh(), torch.nn.Linear(64, 1))

        def forward(self, obs, **kw):
            obs = torch.as_tensor(obs, dtype=torch.float32, device=self.net[0].weight.device)
            return self.net(obs)

    return Actor, Critic


def ant_dist(torch):
    return lambda mu, sigma: torch.distributions.Independent(
        torch.distributions.Normal(mu, sigma), 1)
