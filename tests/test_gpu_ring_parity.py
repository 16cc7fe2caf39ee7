"""GPU parity: CUDA single-lane ring path (through the C-ABI) vs the CPU oracle.

Tolerances (BASELINE.json north_star): fp64 1e-9 relative per step, fp32 1e-4
relative per step (re-synchronised from the oracle state each step), 3000-step
aggregates within 1 %; leader / follower indices bit-exact."""
import numpy as np
import pytest

from tests import scenarios as S
from oracle import sim as OS

pytestmark = pytest.mark.gpu

TOL64, TOL32 = 1e-9, 1e-4


def close(a, b, tol, floor=1.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    err = np.abs(a - b) / np.maximum(np.abs(b), floor)
    return float(err.max()) if err.size else 0.0


def state_of(eng):
    eng.sync()
    return eng.pos.cpu().numpy().astype(np.float64), eng.vel.cpu().numpy().astype(np.float64)


def test_c1_fp64_stepwise():
    g = S.c1_groups()
    eng = S.make_engine(g, 3, 'f64', length=230, bunching=20.0, record_views=1)
    ora = S.make_oracle(g, 230, bunching=20.0)
    obs = eng.reset().cpu().numpy()
    assert close(obs[0], ora.get_state(), TOL64) <= TOL64
    for step in range(400):
        o_obs, o_rew, o_done = ora.step(None)
        obs, rew, done = eng.step()
        pos, vel = state_of(eng)
        for e in range(3):
            assert close(pos[:, e], ora.pos, TOL64) <= TOL64, step
            assert close(vel[:, e], ora.vel, TOL64) <= TOL64, step
        assert close(eng.accel.cpu().numpy()[:, 1], ora.accel, TOL64) <= TOL64
        assert close(eng.headway.cpu().numpy()[:, 2], ora.headway, TOL64) <= TOL64
        assert (eng.leader.cpu().numpy()[:, 0] == np.array(ora.leader)).all()
        assert (eng.follower.cpu().numpy()[:, 0] == np.array(ora.follower)).all()
        assert close(obs.cpu().numpy()[1], o_obs, TOL64) <= TOL64
        assert close(rew.cpu().numpy()[1], o_rew, TOL64) <= TOL64
        assert bool(done.cpu().numpy()[0]) == o_done


def test_c2_fp64_warmup_and_actions():
    g = S.c2_groups()
    lengths = [220, 233, 250, 260, 270]
    E = len(lengths)
    eng = S.make_engine(g, E, 'f64', env_kind=OS.ENV_WAVE_ATTENUATION_PO, length_lo=220.,
                        length_hi=270., **S.C2)
    obs = eng.reset(lengths=lengths).cpu().numpy()
    oras = []
    for e, L in enumerate(lengths):
        o = S.make_oracle(g, L, env_kind=OS.ENV_WAVE_ATTENUATION_PO, bunching=50.0,
                          warmup_steps=750, horizon=3000, act_lo=-1, act_hi=1,
                          obs_max_length=270.0)
        oras.append(o)
        assert close(obs[e], o.get_state(), TOL64) <= TOL64
    pos, vel = state_of(eng)
    for e, o in enumerate(oras):
        assert close(pos[:, e], o.pos, TOL64) <= TOL64
        assert close(vel[:, e], o.vel, TOL64) <= TOL64
    assert (eng.time_counter.cpu().numpy() == 750).all()
    import torch
    rng = np.random.default_rng(7)
    for step in range(150):
        act = rng.uniform(-1.5, 1.5, size=(1, E))
        obs, rew, done = eng.step(torch.tensor(act, device='cuda'))
        obs, rew = obs.cpu().numpy(), rew.cpu().numpy()
        for e, o in enumerate(oras):
            o_obs, o_rew, o_done = o.step([act[0, e]])
            assert close(obs[e], o_obs, TOL64) <= TOL64, (step, e)
            assert close(rew[e], o_rew, TOL64) <= TOL64, (step, e)


def test_fused_equals_single_step_bitwise():
    import torch
    g = S.c2_groups()
    kw = dict(env_kind=OS.ENV_WAVE_ATTENUATION_PO, length_lo=220., length_hi=270., seed=3, **S.C2)
    K, E = 64, 256
    for dt in ('f32', 'f64'):
        a = S.make_engine(g, E, dt, **kw)
        b = S.make_engine(g, E, dt, **kw)
        a.reset(); b.reset()
        tdt = torch.float32 if dt == 'f32' else torch.float64
        act = (torch.rand(K, 1, E, device='cuda', dtype=tdt) * 2 - 1)
        obs_seq = torch.empty(K, E, 3, device='cuda', dtype=tdt)
        rew_seq = torch.empty(K, E, device='cuda', dtype=tdt)
        done_seq = torch.empty(K, E, device='cuda', dtype=torch.uint8)
        a.step(act, k=K, obs_seq=obs_seq, reward_seq=rew_seq, done_seq=done_seq)
        for k in range(K):
            obs, rew, done = b.step(act[k])
            assert torch.equal(obs, obs_seq[k]) and torch.equal(rew, rew_seq[k])
            assert torch.equal(done, done_seq[k])
        assert torch.equal(a.pos, b.pos) and torch.equal(a.vel, b.vel)
        # host-buffer entry point gives the same numbers
        c = S.make_engine(g, E, dt, **kw)
        c.reset()
        h_obs, h_rew, h_done = c.step_host(act.cpu().numpy(), k=K)
        assert np.array_equal(h_obs, obs_seq.cpu().numpy())
        assert np.array_equal(h_rew, rew_seq.cpu().numpy())
        assert np.array_equal(h_done, done_seq.cpu().numpy())


def test_reset_lengths_in_range_and_invariant_to_sharding():
    g = S.c2_groups()
    kw = dict(env_kind=OS.ENV_WAVE_ATTENUATION_PO, length_lo=220., length_hi=270., seed=11, **S.C2)
    whole = S.make_engine(g, 512, 'f32', **kw)
    whole.reset()
    L = whole.length.cpu().numpy() - 0.4
    assert L.min() >= 220 - 1e-3 and L.max() <= 270 + 1e-3
    assert np.allclose(L, np.round(L), atol=1e-3) and len(np.unique(np.round(L))) > 30
    half = S.make_engine(g, 256, 'f32', env_id_base=256, **kw)
    half.reset()
    assert np.array_equal(half.length.cpu().numpy(), whole.length.cpu().numpy()[256:])
    assert np.array_equal(half.pos.cpu().numpy(), whole.pos.cpu().numpy()[:, 256:])


def test_fp32_resync_per_step_and_aggregates():
    g = S.c1_groups()
    eng = S.make_engine(g, 2, 'f32', length=230, bunching=20.0, record_views=1)
    ora = S.make_oracle(g, 230, bunching=20.0)
    eng.reset()
    for step in range(300):
        ora.step(None)
        # re-synchronise: engine takes the oracle state rounded to fp32, the
        # oracle is evaluated on that same rounded state
        p32 = np.asarray(ora.pos, np.float32); v32 = np.asarray(ora.vel, np.float32)
        eng.set_state(np.stack([p32, p32], 1), np.stack([v32, v32], 1))
        chk = S.make_oracle(g, 230, bunching=20.0)
        chk.circ = float(np.float32(ora.circ))
        chk.set_state(p32.astype(np.float64), v32.astype(np.float64))
        chk.step(None)
        eng.step()
        eng.sync()
        assert close(eng.accel.cpu().numpy()[:, 0], chk.accel, TOL32, floor=0.1) <= TOL32, step
        assert close(eng.vel.cpu().numpy()[:, 0], chk.vel, TOL32, floor=1e-2) <= TOL32, step
        assert close(eng.pos.cpu().numpy()[:, 0], chk.pos, TOL32) <= TOL32, step
    # free-running 3000-step aggregates within 1 %
    eng = S.make_engine(g, 2, 'f32', length=230, bunching=20.0)
    ora = S.make_oracle(g, 230, bunching=20.0)
    eng.reset()
    import torch
    K = 3000
    obs_seq = torch.empty(K, 2, 44, device='cuda')
    rew_seq = torch.empty(K, 2, device='cuda')
    eng.step(None, k=K, obs_seq=obs_seq, reward_seq=rew_seq)
    v_eng = obs_seq[:, 0, :22].cpu().numpy() * 30.0
    o_v, o_r = [], []
    for _ in range(K):
        _, r, _ = ora.step(None)
        o_v.append(list(ora.vel)); o_r.append(r)
    o_v = np.array(o_v)
    m_e, m_o = v_eng[-1000:].mean(), o_v[-1000:].mean()
    amp_e = (v_eng[-1000:].max(1) - v_eng[-1000:].min(1)).mean()
    amp_o = (o_v[-1000:].max(1) - o_v[-1000:].min(1)).mean()
    r_e, r_o = float(rew_seq[:, 0].sum()), float(np.sum(o_r))
    assert abs(m_e - m_o) <= 0.01 * abs(m_o)
    assert abs(amp_e - amp_o) <= 0.01 * max(abs(amp_o), 0.05)
    assert abs(r_e - r_o) <= 0.01 * abs(r_o)


def test_sims_per_step_statistics_count_env_steps_not_sub_steps():
    """with sims_per_step = 3 the rollout statistics follow Experiment.run's
    bookkeeping - one reward and one env-step per Env.step (sub-steps still count their
    vehicle-steps), also for the Env.step that ends in a collision on an early sub-step."""
    import torch
    g = [S.vt('IDM', 21, sim_min_gap=0.0), S.vt('RL', 1)]       # mode 0: the RL car can ram
    E, K, SPS = 64, 40, 3
    eng = S.make_engine(g, E, 'f64', env_kind=OS.ENV_WAVE_ATTENUATION_PO, length_lo=220.,
                        length_hi=270., seed=3, sims_per_step=SPS, **dict(S.C2, warmup_steps=10))
    eng.reset()
    eng.stats(reset=True)
    act = torch.full((K, 1, E), 1.0, device='cuda', dtype=torch.float64)    # full throttle
    rew = torch.empty(K, E, device='cuda', dtype=torch.float64)
    done = torch.empty(K, E, device='cuda', dtype=torch.uint8)
    eng.step(act, k=K, reward_seq=rew, done_seq=done)
    st = eng.stats()
    d = done.cpu().numpy().astype(bool)
    steps_alive = (~d).sum() + d.any(0).sum()       # Env.steps taken, the terminal one included
    assert d.any(), 'the scenario is meant to collide'
    assert st['env_steps'] == steps_alive
    assert st['crashes'] == eng.crash.sum().item() > 0
    assert abs(st['sum_reward'] - rew.sum().item()) <= 1e-9 * max(1.0, abs(rew.sum().item()))
    tc = eng.time_counter.cpu().numpy() - SPS * 10          # simulation sub-steps since the reset
    assert st['vehicle_steps'] == 22.0 * tc.sum()


def test_engine_follows_the_callers_stream_and_rejects_host_tensors():
    """Engine.step runs on torch's CURRENT stream (re-read per call) and refuses
    tensors that do not live on its GPU or have the wrong size / dtype."""
    import torch
    g = S.c2_groups()
    kw = dict(env_kind=OS.ENV_WAVE_ATTENUATION_PO, length_lo=220., length_hi=270., seed=3, **S.C2)
    a, b = S.make_engine(g, 128, 'f32', **kw), S.make_engine(g, 128, 'f32', **kw)
    a.reset(); b.reset()
    act = torch.rand(30, 1, 128, device='cuda') * 2 - 1
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        scaled = act * 1.0                      # produced on the side stream ...
        a.step(scaled, k=30)                    # ... consumed there too, no race
        pa = a.pos.clone()
    side.synchronize()
    b.step(act, k=30)
    assert torch.equal(pa, b.pos)
    with pytest.raises(ValueError):
        a.step(act.cpu(), k=30)
    with pytest.raises(ValueError):
        a.step(act, k=30, obs_seq=torch.empty(30, 128, 2, device='cuda'))
    with pytest.raises(ValueError):
        a.step(act, k=30, reward_seq=torch.empty(30, 128, device='cuda', dtype=torch.float64))
    a.close()
    assert a.pos is None and a.obs is None
