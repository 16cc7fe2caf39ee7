"""GPU parity of additional_params['sort_vehicles'] (sort_pre / sort_obs kernels around the
step kernel) and env_params.evaluate, against the reference-generated golden
tests/golden/ring_sorted.npz and the oracle."""
import os

import numpy as np
import pytest

from tests import scenarios as S
from tests.test_oracle_golden import SORT_GROUPS
from oracle import sim as OS

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), 'golden')


def close(a, b, floor=1.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float((np.abs(a - b) / np.maximum(np.abs(b), floor)).max())


def test_sort_vehicles_accel_env_golden():
    """Observation rows in absolute-position order and RL action slots handed out in that
    order, fused launches and single steps, vs the reference's own sorted_ids machinery."""
    import torch
    z = np.load(os.path.join(G, 'ring_sorted.npz'))
    E, warm = 6, 30
    eng = S.make_engine([S.vt(c, n) for c, n in SORT_GROUPS], E, 'f64', length=120,
                        env_kind=OS.ENV_ACCEL, act_lo=-3.0, act_hi=3.0, sort_vehicles=1,
                        warmup_steps=warm)
    acts = z['actions_0'][warm:]
    ref_obs, ref_rew = z['obs_0'], z['reward_0']
    obs = eng.reset()
    assert close(obs.cpu().numpy()[2], ref_obs[warm]) <= 1e-9
    n, s = len(acts), 0
    for K in (1, 3, 50, 200, n):
        K = min(K, n - s)
        act = torch.tensor(np.repeat(acts[s:s + K, :, None], E, 2), device='cuda')
        obs_seq = torch.empty(K, E, 24, device='cuda', dtype=torch.float64)
        rew_seq = torch.empty(K, E, device='cuda', dtype=torch.float64)
        eng.step(act, k=K, obs_seq=obs_seq, reward_seq=rew_seq)
        o, r = obs_seq.cpu().numpy(), rew_seq.cpu().numpy()
        for e in (0, 5):
            assert close(o[:, e], ref_obs[warm + 1 + s:warm + 1 + s + K]) <= 1e-9, (s, K)
            assert close(r[:, e], ref_rew[warm + s:warm + s + K]) <= 1e-9, (s, K)
        assert close(eng.obs.cpu().numpy()[3], ref_obs[warm + s + K]) <= 1e-9
        assert close(eng.pos.cpu().numpy()[:, 1], z['pos_0'][warm + s + K]) <= 1e-9
        s += K
    assert s == n


def test_sort_vehicles_second_episode_uses_stale_keys_like_the_oracle():
    import torch
    E = 3
    g = [S.vt(c, n) for c, n in SORT_GROUPS]
    kw = dict(env_kind=OS.ENV_ACCEL, act_lo=-3.0, act_hi=3.0, sort_vehicles=1, warmup_steps=0)
    eng = S.make_engine(g, E, 'f64', length=120, **kw)
    o = S.make_oracle(g, 120, **dict(kw, sort_vehicles=True))
    o.abs_position = o.prev_pos = None
    assert close(eng.reset().cpu().numpy()[0], o.reset(120)) <= 1e-12       # id order
    rng = np.random.default_rng(0)
    for ep in range(2):
        for s in range(260):
            a = rng.uniform(-0.3, 0.3, 2)
            # keep the RL cars from ramming: brake when close
            for q, i in enumerate([j for j in o.sorted_idx() if j in o.rl_idx]):
                if o.headway[i] < 6 + 1.5 * o.vel[i]:
                    a[q] = -3.0
            obs, rew, done = eng.step(torch.tensor(np.repeat(a[:, None], E, 1), device='cuda'))
            ob, r, d = o.step(list(a))
            assert close(obs.cpu().numpy()[1], ob) <= 1e-9, (ep, s)
            assert close(rew.cpu().numpy()[1], r) <= 1e-9 and bool(done[1]) == d
            assert not d
        assert o.sorted_idx() != list(range(12))
        # reset: observation sorted with the keys the episode left behind
        assert close(eng.reset().cpu().numpy()[2], o.reset(120)) <= 1e-9


def test_sort_vehicles_lane_change_env():
    import torch
    z = np.load(os.path.join(G, 'ring_sorted.npz'))
    E = 4
    eng = S.make_engine([S.vt('IDM', 9), S.vt('RL', 3)], E, 'f64', length=230,
                        env_kind=OS.ENV_LANE_CHANGE_ACCEL, n_lanes=3, act_lo=-3.0, act_hi=3.0,
                        x0=200.0, sort_vehicles=1, warmup_steps=20)
    obs = eng.reset()
    assert close(obs.cpu().numpy()[0], z['lc_obs'][19]) <= 1e-9
    acts = z['lc_actions'][20:]
    K = len(acts)
    obs_seq = torch.empty(K, E, 36, device='cuda', dtype=torch.float64)
    rew_seq = torch.empty(K, E, device='cuda', dtype=torch.float64)
    eng.step(torch.tensor(np.repeat(acts[:, :, None], E, 2), device='cuda'), k=K,
             obs_seq=obs_seq, reward_seq=rew_seq)
    assert close(obs_seq.cpu().numpy()[:, 3], z['lc_obs'][20:]) <= 1e-9
    assert close(rew_seq.cpu().numpy()[:, 3], z['lc_reward'][20:]) <= 1e-9
    assert (eng.lane.cpu().numpy()[:, 0] == z['lc_lane'][-1]).all()


@pytest.mark.parametrize('dtype,tol', [('f64', 1e-9), ('f32', 2e-4)])
def test_evaluate_reward(dtype, tol):
    """env_params.evaluate: AccelEnv reward = mean speed (accel.py:111-112); generic
    kernel (the IDM fast path is bypassed) and the multi-lane kernel."""
    import torch
    z = np.load(os.path.join(G, 'ring_sorted.npz'))
    td = torch.float64 if dtype == 'f64' else torch.float32
    eng = S.make_engine(S.c1_groups(), 2, dtype, length=230, bunching=20.0, evaluate=1)
    eng.reset()
    rew_seq = torch.empty(120, 2, device='cuda', dtype=td)
    eng.step(None, k=120, reward_seq=rew_seq)
    assert close(rew_seq.cpu().numpy()[:, 1], z['eval_reward'], floor=0.1) <= tol
    ml = S.make_engine([S.vt('IDM', 12)], 2, dtype, length=230, n_lanes=3, evaluate=1)
    ml.reset()
    o = S.make_oracle([S.vt('IDM', 12)], 230, n_lanes=3, evaluate=True)
    r2 = torch.empty(60, 2, device='cuda', dtype=td)
    ml.step(None, k=60, reward_seq=r2)
    assert close(r2.cpu().numpy()[:, 0], [o.step(None)[1] for _ in range(60)], floor=0.1) <= tol


def test_sort_vehicles_rejected_where_the_reference_has_no_such_option():
    from flow_b200.engine import FlowStepError
    with pytest.raises(FlowStepError):
        S.make_engine(S.c2_groups(), 2, 'f32', env_kind=OS.ENV_WAVE_ATTENUATION_PO,
                      sort_vehicles=1, **S.C2)


def test_sims_per_step_counters_and_trajectory():
    """EnvParams.sims_per_step (flow/envs/base.py:325-383): the reference's own checks
    (tests/fast_tests/test_environment_base_class.py:256-303: reset advances
    warmup_steps x sims_per_step, a step advances sims_per_step) and the trajectory /
    observation / reward against the oracle, fused and stepwise."""
    import torch
    S5 = 5
    kw = dict(env_kind=OS.ENV_WAVE_ATTENUATION_PO, length_lo=230., length_hi=230.,
              sims_per_step=S5, **dict(S.C2, warmup_steps=7, horizon=40))
    E = 4
    eng = S.make_engine(S.c2_groups(), E, 'f64', **kw)
    okw = dict(kw); okw.pop('length_lo'); okw.pop('length_hi')
    o = S.make_oracle(S.c2_groups(), 230, **okw)
    obs = eng.reset()
    assert int(eng.time_counter[0]) == 7 * S5 == o.time_counter
    assert close(obs.cpu().numpy()[0], o.get_state()) <= 1e-9
    rng = np.random.default_rng(2)
    acts = rng.uniform(-1.2, 0.6, (40, 1))
    obs, rew, done = eng.step(torch.tensor(np.repeat(acts[0][:, None], E, 1), device='cuda'))
    ob, r, d = o.step(list(acts[0]))
    assert int(eng.time_counter[1]) == 8 * S5 == o.time_counter
    assert close(obs.cpu().numpy()[1], ob) <= 1e-9 and close(rew.cpu().numpy()[1], r) <= 1e-9
    K = 39
    obs_seq = torch.empty(K, E, 3, device='cuda', dtype=torch.float64)
    rew_seq = torch.empty(K, E, device='cuda', dtype=torch.float64)
    done_seq = torch.empty(K, E, device='cuda', dtype=torch.uint8)
    eng.step(torch.tensor(np.repeat(acts[1:, :, None], E, 2), device='cuda'), k=K,
             obs_seq=obs_seq, reward_seq=rew_seq, done_seq=done_seq)
    for k in range(K):
        ob, r, d = o.step(list(acts[1 + k]))
        assert close(obs_seq.cpu().numpy()[k, 2], ob) <= 1e-9, k
        assert close(rew_seq.cpu().numpy()[k, 2], r) <= 1e-9, k
        assert bool(done_seq[k, 2]) == d, k
    assert d and int(eng.time_counter[3]) == o.time_counter      # stopped where the oracle stopped
    assert o.time_counter % S5 == 0 or o.crash
    assert close(eng.pos.cpu().numpy()[:, 0], o.pos) <= 1e-9


@pytest.mark.parametrize('lanes', [1, 3])
def test_ballistic_update_vs_oracle(lanes):
    """SumoParams(use_ballistic=True): x += (v_old + v_new)/2 dt in the generic ring kernel
    (the IDM fast paths are bypassed) and the multi-lane kernel; differs from the Euler run."""
    import torch
    g = [S.vt('IDM', 14), S.vt('OVM', 4)] if lanes == 1 else [S.vt('IDM', 18)]
    kw = dict(env_kind=OS.ENV_ACCEL, n_lanes=lanes, use_ballistic=1)
    eng = S.make_engine(g, 2, 'f64', length=200, **kw)
    eul = S.make_engine(g, 2, 'f64', length=200, **dict(kw, use_ballistic=0))
    o = S.make_oracle(g, 200, **dict(kw, use_ballistic=True))
    eng.reset(); eul.reset()
    K = 150
    obs_seq = torch.empty(K, 2, 36, device='cuda', dtype=torch.float64)
    rew_seq = torch.empty(K, 2, device='cuda', dtype=torch.float64)
    eng.step(None, k=K, obs_seq=obs_seq, reward_seq=rew_seq)
    eul.step(None, k=K)
    for k in range(K):
        ob, r, d = o.step(None)
        assert close(obs_seq.cpu().numpy()[k, 1], ob) <= 1e-9, k
        assert close(rew_seq.cpu().numpy()[k, 1], r) <= 1e-9, k
    assert close(eng.pos.cpu().numpy()[:, 0], o.pos) <= 1e-9
    assert np.abs(eng.pos.cpu().numpy() - eul.pos.cpu().numpy()).max() > 1e-3
