#!/usr/bin/env python
"""Times the REFERENCE ITSELF (imported from /root/reference under the four attribute-only stub
modules of tests/golden/make_golden.py) on this container's host cores: the CPU denominator that
BASELINE.md section 3 and SURVEY.md 8(d) name.  Runs in the build container only -- the GPU box
has no /root/reference.

    python scripts/time_reference.py            # writes profiles/r02_reference_cpu.json

Timed, numba warmed (``BasePolicy._compile`` / ``SegmentTree._compile``), median of the passes:
  C2  ``BasePolicy.compute_episodic_return`` over 4096 streams x 2048 steps (policy/base.py:337-384)
  C3  the same over 64 x 2048 = 131 072 transitions
  C4  ``compute_nstep_return`` n = 3, batch 4096, D = 1 / 51 on a 2^20-slot
      ``PrioritizedVectorReplayBuffer`` (policy/base.py:386-440), and the PER step
      ``sample_indices`` + ``get_weight`` / max-normalisation + ``update_weight`` (prio.py:59-102)
each on ONE core (the reference as shipped: serial njit loops, single-threaded Python) and as
``os.cpu_count()`` processes that each own 1/cores of the env streams (GAE) or an independent
tree (PER) -- the reference's best case.  Beside every reference figure the same workload through
oracle/oracle_nb.py (the numba restatement that travels to the GPU box as bench.py's
``cpu_baseline``) and oracle/ts_oracle.c (the C parity oracle), so the stand-ins are calibrated
against the real thing on the same cores.
"""
from __future__ import annotations

import json
import multiprocessing as mp
import os
import platform
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

GAMMA, LAMBDA = 0.99, 0.95


def cpu_model() -> str:
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return platform.processor() or "unknown"


def median_time(fn, min_passes=5, seconds=3.0):
    fn()  # warm
    times, t0 = [], time.perf_counter()
    while len(times) < min_passes or time.perf_counter() - t0 < seconds:
        a = time.perf_counter()
        fn()
        times.append(time.perf_counter() - a)
    return float(np.median(times)), len(times)


def rollout(E, T, seed):
    import helpers
    return helpers.c2_rollout(E, T, seed=seed)


def full_vector_buffer(ts, cls, E, T, rew, term, trunc, **kw):
    """A reference (Prioritized)VectorReplayBuffer whose E rings each hold T = capacity steps,
    built by ``set_batch`` + the bookkeeping ``add`` would have left (manager.py:136-146,
    base.py:195-214) -- 8.4 M ``add`` calls through the Python loop would take minutes."""
    from tianshou.data import Batch
    buf = cls(E * T, E, **kw)
    n = E * T
    z = np.zeros(n, np.uint8)
    buf.set_batch(Batch(obs=z, act=z, rew=rew.reshape(-1).astype(np.float64),
                        terminated=term.reshape(-1), truncated=trunc.reshape(-1),
                        done=(term | trunc).reshape(-1), obs_next=z, info=Batch(),
                        policy=Batch()))
    for k, b in enumerate(buf.buffers):
        b._size, b._index = T, 0
        b.last_index = np.array([T - 1])
    buf.last_index = buf._offset + T - 1
    buf._lengths = np.full(E, T)
    return buf


# ---- GAE ----------------------------------------------------------------------------------------
def gae_case(E, T, seed=0):
    ts = import_reference()
    from tianshou.data import Batch, VectorReplayBuffer
    from tianshou.policy import BasePolicy
    BasePolicy._compile()
    rew, v_s, v_n, term, trunc = rollout(E, T, seed)
    buf = full_vector_buffer(ts, VectorReplayBuffer, E, T, rew, term, trunc)
    indices = buf.sample_indices(0)
    batch = Batch(rew=buf.rew[indices], terminated=buf.terminated[indices],
                  truncated=buf.truncated[indices])
    vs, vn = v_s.reshape(-1), v_n.reshape(-1)
    return BasePolicy, batch, buf, indices, vs, vn, (rew, v_s, v_n, term, trunc)


def time_gae_reference(E, T, seconds=3.0):
    BasePolicy, batch, buf, indices, vs, vn, _ = gae_case(E, T)
    f = lambda: BasePolicy.compute_episodic_return(batch, buf, indices, vn, vs, GAMMA, LAMBDA)  # noqa: E731
    return median_time(f, 3, seconds)


def _gae_worker(args):
    E, T, seconds = args
    med, n = time_gae_reference(E, T, seconds)
    return E * T / med


def time_gae_stand_ins(E, T):
    from oracle import oracle as orc
    from oracle import oracle_nb as nb
    nb.warm()
    _, batch, buf, indices, vs, vn, (rew, v_s, v_n, term, trunc) = gae_case(E, T)
    unf = buf.unfinished_index()
    bt = np.asarray(buf.terminated)
    r64 = np.asarray(batch.rew)
    f_nb = lambda: nb.compute_episodic_return(r64, batch.terminated, batch.truncated, bt,  # noqa: E731
                                              indices, unf, vn, vs, GAMMA, LAMBDA)
    f_c = lambda: orc.episodic_return_rows(rew, v_s, v_n, term, trunc, GAMMA, LAMBDA, 1)  # noqa: E731
    return median_time(f_nb, 3, 2.0)[0], median_time(f_c, 3, 2.0)[0]


# ---- C4: n-step + PER ------------------------------------------------------------------------------
def c4_buffer():
    import helpers as H
    ts = import_reference()
    from tianshou.data import PrioritizedVectorReplayBuffer
    from tianshou.policy import BasePolicy
    BasePolicy._compile()
    rew, term, trunc = H.c4_stream()
    s_rew, s_term, s_trunc, _, _ = H.c4_final_storage(rew, term, trunc)
    E, T = H.C4_E, H.C4_SUB
    buf = full_vector_buffer(ts, PrioritizedVectorReplayBuffer, E, T, s_rew.reshape(E, T),
                             s_term.reshape(E, T), s_trunc.reshape(E, T), alpha=0.6, beta=0.4)
    buf.weight[np.arange(E * T)] = H.c4_leaves()
    return buf, BasePolicy, H


def time_c4_reference(seconds=3.0, bsz=4096):
    import torch
    buf, BasePolicy, H = c4_buffer()  # installs the stub modules before tianshou is imported
    from tianshou.data import Batch
    rng = np.random.default_rng(3)
    out = {}
    idx = buf.sample_indices(bsz)
    for D in (1, 51):
        table = {}

        def target_q_fn(buffer, terminal, D=D):
            t = torch.from_numpy(H.c4_target_q(terminal, D))
            return t if D > 1 else t.flatten()

        f = lambda: BasePolicy.compute_nstep_return(Batch(), buf, idx, target_q_fn, 0.99, 3)  # noqa: E731
        out[f"nstep_d{D}_s"] = median_time(f, 5, seconds)[0]
        del table
    tds = [rng.standard_normal(bsz).astype(np.float32) for _ in range(8)]
    state = {"i": 0}

    def per_step():
        i = buf.sample_indices(bsz)                      # prio.py:59-64
        w = buf.get_weight(i)                            # :66-76
        w = w / np.max(w)                                # :101
        buf.update_weight(i, tds[state["i"] % 8])        # :78-89
        state["i"] += 1
        return w

    out["per_step_s"] = median_time(per_step, 20, seconds)[0]
    return out


def _per_worker(seconds):
    return 4096 / time_c4_reference(seconds)["per_step_s"]


def time_c4_stand_ins(bsz=4096):
    import helpers as H
    from oracle import oracle as orc
    from oracle import oracle_nb as nb
    nb.warm()
    size = H.C4_E * H.C4_SUB
    rng = np.random.default_rng(3)
    tds = [rng.standard_normal(bsz).astype(np.float32) for _ in range(8)]
    per = nb.PER(size, 0.6, 0.4)
    per.set_leaves(np.arange(size), H.c4_leaves())
    state = {"i": 0}

    def nb_step():
        i = per.sample_indices(bsz)
        w = per.batch_weight(i)
        per.update_weight(i, tds[state["i"] % 8])
        state["i"] += 1
        return w

    tree = orc.SegTree(size)
    tree[np.arange(size)] = H.c4_leaves()
    mm = [1.0, 1.0]

    def c_step():
        u = rng.random(bsz)
        _, _, mm[0], mm[1] = orc.per_step_c(tree, u, tds[state["i"] % 8], 0.6, 0.4, True, *mm)
        state["i"] += 1

    return median_time(nb_step, 20, 2.0)[0], median_time(c_step, 20, 2.0)[0]


def import_reference():
    import make_golden
    return make_golden.import_reference()


def main():
    import numba
    ts = import_reference()
    cores = os.cpu_count()
    out = {"cpu_model": cpu_model(), "cores": cores, "python": platform.python_version(),
           "numpy": np.__version__, "numba": numba.__version__, "tianshou": ts.__version__,
           "where": "build container (the GPU box has no /root/reference)",
           "how": "time.perf_counter, numba warmed by _compile(), median of >= 3 passes"}
    ctx = mp.get_context("spawn")
    # ---- C2 / C3 GAE -------------------------------------------------------------------------------
    for tag, E, T in (("c2", 4096, 2048), ("c3", 64, 2048)):
        med, n = time_gae_reference(E, T)
        nb_s, c_s = time_gae_stand_ins(E, T)
        rec = {"transitions": E * T, "reference_1core_s": med, "passes": n,
               "reference_1core_transitions_per_s": E * T / med,
               "oracle_nb_1core_s": nb_s, "oracle_c_1core_s": c_s,
               "oracle_nb_over_reference": nb_s / med, "oracle_c_over_reference": c_s / med}
        if tag == "c2":
            with ctx.Pool(cores) as pool:  # one reference process per core, E / cores streams each
                rates = pool.map(_gae_worker, [(E // cores, T, 3.0)] * cores)
            rec["reference_allcore_transitions_per_s"] = float(sum(rates))
            rec["allcore_processes"] = cores
        out[tag + "_gae"] = rec
        print(tag, json.dumps(rec), flush=True)
    # ---- C4 n-step + PER --------------------------------------------------------------------------
    c4 = time_c4_reference()
    nb_s, c_s = time_c4_stand_ins()
    with ctx.Pool(cores) as pool:  # independent trees, one per core
        rates = pool.map(_per_worker, [3.0] * cores)
    out["c4"] = {"batch": 4096, "leaves": 1 << 20, **c4,
                 "per_samples_per_s_1core": 4096 / c4["per_step_s"],
                 "per_samples_per_s_allcore": float(sum(rates)), "allcore_processes": cores,
                 "oracle_nb_per_step_s": nb_s, "oracle_c_per_step_s": c_s,
                 "oracle_nb_over_reference": nb_s / c4["per_step_s"],
                 "oracle_c_over_reference": c_s / c4["per_step_s"]}
    print("c4", json.dumps(out["c4"]), flush=True)
    path = os.path.join(ROOT, "profiles", "r02_reference_cpu.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", path)


if __name__ == "__main__":
    main()
