"""Seeded differential fuzz: random ring configurations (vehicle counts around the
thread-mapping boundaries, controller mixes, fail-safes, RL vehicles anywhere, env fronts,
lanes, placement parameters, time steps, option flags) stepped by the CUDA engine in fp64
and by the oracle; trajectories, observations, rewards and done flags must agree to 1e-9
until the oracle's episode ends. Every case is a few dozen steps of a handful of vehicles."""
import numpy as np
import pytest

from tests import scenarios as S
from oracle import sim as OS

pytestmark = pytest.mark.gpu

HUMANS = ['IDM', 'OVM', 'LINEAR_OVM', 'BCM', 'CFM', 'LAC', 'GIPPS', 'FOLLOWER_STOPPER',
          'NONLOCAL_FOLLOWER_STOPPER', 'PI_SATURATION', 'SIM']
V_CHOICES = [1, 2, 3, 7, 8, 9, 16, 21, 22, 23, 31, 32, 33, 40, 64, 65, 95, 96, 97, 130]


def close(a, b, floor=1.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float((np.abs(a - b) / np.maximum(np.abs(b), floor)).max()) if a.size else 0.0


def draw_case(seed):
    rng = np.random.default_rng(1000 + seed)
    lanes = int(rng.choice([1, 1, 1, 2, 3, 4]))
    V = int(rng.choice(V_CHOICES))
    if lanes > 1:
        V = min(V, 97)
    env = rng.choice(['accel', 'wave', 'wave_po', 'ma_wave', 'ma_accel'] if lanes == 1
                     else ['accel', 'lc', 'lc_po'])
    n_rl = 0 if env == 'accel' and rng.random() < 0.5 else int(rng.integers(1, min(V, 4) + 1))
    if env == 'wave_po':
        n_rl = 1
    only_idm = rng.random() < 0.4
    kinds = ['IDM'] if only_idm else [str(k) for k in rng.choice(HUMANS, size=int(rng.integers(1, 4)), replace=False)]
    if lanes > 1:
        kinds = [k for k in kinds if k != 'PI_SATURATION'] or ['IDM']
    # a random order of type groups: humans and RL vehicles interleaved
    n_h = V - n_rl
    groups = []
    if n_h > 0:
        cuts = sorted(rng.integers(0, n_h + 1, size=len(kinds) - 1)) if len(kinds) > 1 else []
        sizes = np.diff([0] + list(cuts) + [n_h])
        for k, n in zip(kinds, sizes):
            if n > 0:
                fs = None
                if k in ('IDM', 'OVM', 'LINEAR_OVM', 'CFM', 'GIPPS') and rng.random() < 0.4:
                    fs = str(rng.choice(['instantaneous', 'safe_velocity']))
                groups.append(S.vt(k, int(n), fail_safe=fs))
    for _ in range(n_rl):
        groups.insert(int(rng.integers(0, len(groups) + 1)), S.vt('RL', 1))
    # SUMO speed modes (restated clamp on commanded speeds): drawn from a generator of their own
    # so that the cases keep the configurations they had before the modes existed
    rng_sm = np.random.default_rng(50_000 + seed)
    if rng_sm.random() < 0.55:
        uniform_mode = int(rng_sm.choice([1, 7, 25, 31])) if rng_sm.random() < 0.5 else None
        for g in groups:
            g['speed_mode'] = uniform_mode if uniform_mode is not None else \
                int(rng_sm.choice([0, 1, 2, 4, 7, 25, 31]))
    per_lane = -(-V // lanes)
    L = float(int(per_lane * rng.uniform(9.0, 16.0)) + 8)
    dt = float(rng.choice([0.1, 0.1, 0.2, 0.5]))
    if any(g['controller'] == 'LAC' for g in groups):
        dt = 0.1   # LAC integrates da/dt = (u - a)/tau with tau = 0.1 explicitly: dt > 0.2 explodes
                   # (speeds of 1e6 m/s; the engine's single-subtraction ring wrap assumes a
                   # vehicle moves less than one lap per step)
    common = dict(
        n_lanes=lanes, x0=float(rng.choice([0.0, 0.0, 3.5, L / 4 + 0.05])),
        bunching=float(rng.choice([0.0, 0.0, 10.0])), warmup_steps=int(rng.choice([0, 0, 4])),
        horizon=int(rng.choice([25, 1000])), target_velocity=float(rng.choice([4.0, 10.0, 20.0])),
        act_lo=-float(rng.choice([1.0, 3.0])), act_hi=float(rng.choice([1.0, 3.0])),
        clip_actions=bool(rng.random() < 0.5), sims_per_step=int(rng.choice([1, 1, 1, 3])),
        use_ballistic=bool(rng.random() < 0.15))
    if lanes > 1:
        common['lanes_distribution'] = float(rng.choice([lanes, float('inf'), max(1, lanes - 1)]))
    flags = {}
    if env == 'accel':
        flags = dict(evaluate=bool(rng.random() < 0.3), sort_vehicles=bool(rng.random() < 0.3))
    if env == 'lc':
        flags = dict(sort_vehicles=bool(rng.random() < 0.3))
    kind = dict(accel=OS.ENV_ACCEL, wave=OS.ENV_WAVE_ATTENUATION, wave_po=OS.ENV_WAVE_ATTENUATION_PO,
                ma_wave=OS.ENV_MA_WAVE_ATTENUATION_PO, ma_accel=OS.ENV_MA_ACCEL_PO,
                lc=OS.ENV_LANE_CHANGE_ACCEL, lc_po=OS.ENV_LANE_CHANGE_ACCEL_PO)[env]
    if env in ('wave_po', 'ma_wave') and rng.random() < 0.5:
        common['obs_max_length'] = L + 30.0
    return dict(groups=groups, L=L, dt=dt, kind=kind, env=env, common=common, flags=flags,
                n_rl=n_rl, V=V, rng=rng)


@pytest.mark.parametrize('seed', range(120))
def test_fuzz_engine_vs_oracle_fp64(seed):
    import torch
    c = draw_case(seed)
    groups, L, common, flags = c['groups'], c['L'], c['common'], c['flags']
    okw = dict(common, dt=c['dt'], env_kind=c['kind'], **flags)
    try:
        o = S.make_oracle(groups, L, **okw)
    except ValueError:
        pytest.skip('vehicles do not fit this ring')
    ekw = dict(common, sim_step=c['dt'], env_kind=c['kind'])
    ekw['clip_actions'] = int(ekw['clip_actions'])
    ekw['use_ballistic'] = int(ekw['use_ballistic'])
    ekw.update({k: int(v) for k, v in flags.items()})
    E = 3
    eng = S.make_engine(groups, E, 'f64', length=L, **ekw)
    obs = eng.reset()
    assert close(eng.pos.cpu().numpy()[:, 1], o.pos) <= 1e-9
    if not o.crash:     # (a collision during the warm-up: see `reordered` below)
        assert close(obs.cpu().numpy()[1], o.get_state()) <= 1e-9, 'reset observation'
    lc = c['env'] in ('lc', 'lc_po')
    A = c['n_rl'] * (2 if lc else 1)
    K = 36
    rng = c['rng']
    acts = rng.uniform(-1.5, 1.0, (K, max(A, 1)))
    if lc:
        acts[:, 1::2] = rng.integers(-1, 2, (K, c['n_rl']))
    ref = []
    for k in range(K):
        try:
            ob, r, d = o.step(list(acts[k][:A]) if A else None)
        except ValueError:      # the reference raises here too (Gipps: sqrt of a negative number)
            break
        ref.append((ob, r, d, list(o.pos), list(o.vel)))
        if d:
            break
    n = len(ref)
    if n == 0:
        pytest.skip('the reference raises on the first step')
    # a collision in which a vehicle went through its leader leaves the ring out of index
    # order; what is observed afterwards is undefined (SUMO would teleport), so the
    # observation of that terminal step is not compared
    V = c['V']
    reordered = o.crash and common['n_lanes'] == 1 and V > 1 and \
        list(o.leader) != [(i + 1) % V for i in range(V)]
    td = torch.float64
    obs_seq = torch.empty(n, E, eng.obs_dim, device='cuda', dtype=td)
    rew_seq = torch.empty(n, E, device='cuda', dtype=td)
    done_seq = torch.empty(n, E, device='cuda', dtype=torch.uint8)
    a_dev = torch.tensor(np.repeat(acts[:n, :A, None], E, 2), device='cuda') if A else None
    half = n // 2
    if half:      # two launches: a fused one and the rest
        eng.step(a_dev[:half] if A else None, k=half, obs_seq=obs_seq[:half],
                 reward_seq=rew_seq[:half], done_seq=done_seq[:half])
    eng.step(a_dev[half:] if A else None, k=n - half, obs_seq=obs_seq[half:],
             reward_seq=rew_seq[half:], done_seq=done_seq[half:])
    og, rg, dg = obs_seq.cpu().numpy(), rew_seq.cpu().numpy(), done_seq.cpu().numpy()
    for k in range(n):
        ob, r, d, _, _ = ref[k]
        assert bool(dg[k, 2]) == d, (k, 'done')
        if not np.isnan(r):
            assert close(rg[k, 2], r) <= 1e-9, (k, 'reward', rg[k, 2], r)
        if not (reordered and k == n - 1):
            assert close(og[k, 2], ob) <= 1e-9, (k, 'obs')
    assert close(eng.pos.cpu().numpy()[:, 0], ref[-1][3]) <= 1e-9
    assert close(eng.vel.cpu().numpy()[:, 0], ref[-1][4]) <= 1e-9


# ---------------------------------------------------------------- open highway
HW_HUMANS = ['IDM', 'OVM', 'LINEAR_OVM', 'BCM', 'CFM', 'GIPPS', 'FOLLOWER_STOPPER',
             'NONLOCAL_FOLLOWER_STOPPER', 'SIM']


def draw_highway(seed):
    rng = np.random.default_rng(5000 + seed)
    kinds = [str(k) for k in rng.choice(HW_HUMANS, size=int(rng.integers(1, 4)), replace=False)]
    groups = []
    for k in kinds:
        fs = None
        if k in ('IDM', 'OVM', 'LINEAR_OVM', 'CFM') and rng.random() < 0.4:
            fs = str(rng.choice(['instantaneous', 'safe_velocity']))
        groups.append(S.vt(k, 0, fail_safe=fs))
    rng_sm = np.random.default_rng(55_000 + seed)     # SUMO speed modes, own generator (see draw_case)
    if rng_sm.random() < 0.55:
        for g in groups:
            g['speed_mode'] = int(rng_sm.choice([0, 1, 7, 25, 25, 31]))
    cap = int(rng.choice([32, 64, 96, 160]))
    length = float(rng.choice([300.0, 600.0, 1000.0]))
    n_init = int(rng.integers(0, min(cap, int(length / 12)) // 2 + 1))
    init_types = [int(t) for t in rng.integers(0, len(groups), n_init)]
    kw = dict(length=length, dt=float(rng.choice([0.1, 0.1, 0.2, 0.5])),
              inflow_vehs_per_hour=float(rng.choice([0.0, 600.0, 2000.0, 3500.0])),
              inflow_type=int(rng.integers(0, len(groups))),
              inflow_depart_speed=float(rng.choice([0.0, 10.0, 20.0])),
              inflow_begin=float(rng.choice([0.0, 1.0, 2.5])),
              x0=float(rng.choice([0.0, 7.5])), bunching=float(rng.choice([0.0, 20.0])),
              horizon=int(rng.choice([60, 3000])), warmup_steps=int(rng.choice([0, 5])),
              target_velocity=float(rng.choice([10.0, 25.0])),
              evaluate=bool(rng.random() < 0.25), use_ballistic=bool(rng.random() < 0.25),
              sims_per_step=int(rng.choice([1, 1, 2])))
    return groups, init_types, cap, kw


@pytest.mark.parametrize('seed', range(40))
def test_fuzz_highway_vs_oracle_fp64(seed):
    import torch
    from oracle import highway as OH
    from flow_b200 import engine as EN
    groups, init_types, cap, kw = draw_highway(seed)
    try:
        h = OH.HighwaySim(S.oracle_types(groups), init_types, capacity=cap, **kw)
    except ValueError:
        pytest.skip('initial vehicles do not fit')
    ekw = dict(kw)
    L = ekw.pop('length')
    ekw['sim_step'] = ekw.pop('dt')
    ekw['evaluate'], ekw['use_ballistic'] = int(ekw['evaluate']), int(ekw['use_ballistic'])
    E = 3
    eng = EN.Engine(S.engine_types(groups), init_types, E, dtype='f64', capacity=cap,
                    topology=EN.TOPO_HIGHWAY, env_kind=EN.ENV_HIGHWAY, length_lo=L, length_hi=L,
                    junction_len=0.0, record_views=1, **ekw)
    obs = eng.reset()
    n = len(h.pos)
    assert close(eng.pos.cpu().numpy()[:n, 1], h.pos) <= 1e-9
    assert close(obs.cpu().numpy()[1], h.get_state()) <= 1e-9
    K = 150
    ref = []
    for k in range(K):
        try:
            ob, r, d = h.step()
        except ValueError:
            break
        ref.append((ob, r, d, len(h.pos), h.departed, h.arrived))
        if d:
            break
    if not ref:
        pytest.skip('the reference raises on the first step')
    m = len(ref)
    obs_seq = torch.empty(m, E, eng.obs_dim, device='cuda', dtype=torch.float64)
    rew_seq = torch.empty(m, E, device='cuda', dtype=torch.float64)
    done_seq = torch.empty(m, E, device='cuda', dtype=torch.uint8)
    half = m // 2
    if half:
        eng.step(None, k=half, obs_seq=obs_seq[:half], reward_seq=rew_seq[:half], done_seq=done_seq[:half])
    eng.step(None, k=m - half, obs_seq=obs_seq[half:], reward_seq=rew_seq[half:], done_seq=done_seq[half:])
    og, rg, dg = obs_seq.cpu().numpy(), rew_seq.cpu().numpy(), done_seq.cpu().numpy()
    # NonLocalFollowerStopper feeds np.mean of all speeds (pairwise summation; the warp sums in
    # another order) into (v_cmd - v)/dt and the 2h/dt of its safe-velocity fail-safe: last-bit
    # differences grow by up to 2/dt per step (seed 7: 2e-9 at step 46, 1e-7 at step 50), so
    # those cases are compared over their first 30 steps
    nonlocal_mean = any(g['controller'] == 'NONLOCAL_FOLLOWER_STOPPER' for g in groups)
    tol = 1e-7 if nonlocal_mean else 1e-9
    for k in range(min(m, 30) if nonlocal_mean else m):
        ob, r, d = ref[k][:3]
        assert bool(dg[k, 2]) == d, (k, 'done')
        assert close(og[k, 0], ob) <= tol, (k, 'obs')
        if np.isnan(r):
            assert np.isnan(rg[k, 0]), (k, 'reward nan')
        else:
            assert close(rg[k, 0], r) <= tol, (k, 'reward', rg[k, 0], r)
    c = eng.counters.cpu().numpy()
    if not nonlocal_mean:
        assert (c[:3, 1] == list(ref[-1][3:])).all(), (c[:, 1], ref[-1][3:])


# ------------------------------------------------- multi-agent open highway
def draw_highway_ma(seed):
    rng = np.random.default_rng(9000 + seed)
    kinds = [str(k) for k in rng.choice(['IDM', 'OVM', 'LINEAR_OVM', 'BCM', 'CFM', 'GIPPS', 'SIM'],
                                        size=int(rng.integers(1, 3)), replace=False)]
    groups = []
    for k in kinds:
        fs = None
        if k in ('IDM', 'OVM', 'LINEAR_OVM', 'CFM') and rng.random() < 0.4:
            fs = str(rng.choice(['instantaneous', 'safe_velocity']))
        groups.append(S.vt(k, 0, fail_safe=fs))
    groups.append(S.vt('RL', 0, max_accel=float(rng.choice([1.0, 2.6])),
                       sim_min_gap=float(rng.choice([0.0, 2.5]))))
    rl_t = len(groups) - 1
    rng_sm = np.random.default_rng(59_000 + seed)     # SUMO speed modes, own generator (see draw_case)
    if rng_sm.random() < 0.55:
        for g in groups:
            g['speed_mode'] = int(rng_sm.choice([0, 1, 7, 25, 25, 31]))
    cap = int(rng.choice([32, 64, 96, 160]))
    length = float(rng.choice([300.0, 600.0, 1000.0]))
    n_init = int(rng.integers(0, min(cap, int(length / 12)) // 2 + 1))
    init_types = [int(t) for t in rng.integers(0, len(groups), n_init)]
    flows = []
    for _ in range(int(rng.integers(1, 5))):
        flows.append((int(rl_t if rng.random() < 0.4 else rng.integers(0, len(groups))),
                      float(rng.choice([300.0, 900.0, 2000.0])),
                      float(rng.choice([0.0, 8.0, 15.0])), float(rng.choice([0.0, 1.0, 2.5]))))
    kw = dict(length=length, dt=float(rng.choice([0.1, 0.2, 0.5])),
              x0=float(rng.choice([0.0, 7.5, 50.0])), horizon=int(rng.choice([80, 3000])),
              warmup_steps=int(rng.choice([0, 5])), target_velocity=float(rng.choice([10.0, 25.0])),
              evaluate=bool(rng.random() < 0.25), use_ballistic=bool(rng.random() < 0.25))
    A = int(rng.choice([1, 4, 33])) if cap > 32 else int(rng.choice([1, 4, 32]))
    return groups, init_types, cap, flows, min(A, cap), kw, rng


@pytest.mark.parametrize('seed', range(30))
def test_fuzz_highway_multiagent_vs_oracle_fp64(seed):
    """Random open highways with up to four inflows (RL ones included), random agent commands
    (some missing = NaN), more agents than action rows in some cases, collisions allowed: agent
    ids, observations, per-agent rewards, slot occupancy and counters vs the oracle each step."""
    import torch
    from oracle import highway as OH
    from flow_b200 import engine as EN
    groups, init_types, cap, flows, A, kw, rng = draw_highway_ma(seed)
    try:
        h = OH.HighwaySim(S.oracle_types(groups), init_types, capacity=cap,
                          inflows=[dict(type=t, vehs_per_hour=q, depart_speed=v, begin=b)
                                   for t, q, v, b in flows], **kw)
    except ValueError:
        pytest.skip('initial vehicles do not fit')
    E = 2
    eng = EN.Engine(S.engine_types(groups), init_types, E, dtype='f64', capacity=cap,
                    topology=EN.TOPO_HIGHWAY, env_kind=EN.ENV_MA_HIGHWAY_PO, length_lo=kw['length'],
                    length_hi=kw['length'], junction_len=0.0, record_views=1, sim_step=kw['dt'],
                    x0=kw['x0'], horizon=kw['horizon'], warmup_steps=kw['warmup_steps'],
                    target_velocity=kw['target_velocity'], evaluate=int(kw['evaluate']),
                    use_ballistic=int(kw['use_ballistic']), max_agents=A, inflows=flows)
    eng.reset()
    n_agents_seen = 0
    for k in range(160):
        slots = h.rl_slots()
        ids = [h.ids[s] for s in slots]
        got = eng.agent_id.cpu().numpy()
        want = (ids + [-1] * A)[:A]
        assert (got == want).all(), (k, got, want)
        st = h.ma_state()
        ob = eng.agent_obs.cpu().numpy()
        for a, vid in enumerate(ids[:A]):
            assert close(ob[1, a], st[vid], floor=1e-3) <= 1e-9, (k, a)
        assert (ob[:, len(ids[:A]):] == 0).all()
        n_agents_seen = max(n_agents_seen, len(ids))
        if h.done:
            assert eng.done.cpu().numpy().all(), k
            break
        # commands for the agents that have an action row; ~20 % missing
        acts, row = {}, np.full((A, E), np.nan)
        for a, (vid, s) in enumerate(zip(ids[:A], slots[:A])):
            if rng.random() < 0.8:
                acts[vid] = float(np.clip(0.2 * (h.headway[s] - 4 - 1.2 * h.vel[s]), -4, 1.5)
                                  + rng.normal(0, 0.4)) if s > 0 else float(rng.normal(0.3, 0.5))
                row[a] = acts[vid]
        use = bool(acts) or rng.random() < 0.5
        eng.step(torch.as_tensor(row, dtype=torch.float64, device='cuda') if use else None)
        h.step(acts if use else None)
        n = len(h.pos)
        c = eng.counters.cpu().numpy()
        assert (c[:, 0] == [n, h.departed, h.arrived]).all(), (k, c[:, 0], n, h.departed, h.arrived)
        assert (eng.veh_id.cpu().numpy()[0, :n] == h.ids).all(), k
        assert close(eng.pos.cpu().numpy()[:n, 1], h.pos) <= 1e-9, k
        assert close(eng.vel.cpu().numpy()[:n, 1], h.vel) <= 1e-9, k
        assert bool(eng.crash.cpu().numpy()[0]) == h.crash, k
        if not h.crash:
            rw = h.ma_reward(False)
            er = eng.agent_reward.cpu().numpy()
            now = [h.ids[s] for s in h.rl_slots()][:A]
            for a, vid in enumerate(now):
                assert close(er[0, a], rw[vid]) <= 1e-9, (k, a, er[0, a], rw[vid])


# ------------------------------------------------- inflow kinds (periodic / end / probabilistic)
@pytest.mark.parametrize('seed', range(16))
def test_fuzz_highway_inflow_kinds_vs_oracle_fp64(seed):
    """One to four flows drawn from {periodic, periodic with an end, probabilistic, probabilistic
    with an end}, random seeds and env_id_base: per env, arrivals / occupancy / trajectories equal
    the oracle's over two episodes (the Philox stream is keyed by seed, GLOBAL env id, episode,
    flow and step)."""
    from oracle import highway as OH
    from flow_b200 import engine as EN
    rng = np.random.default_rng(12000 + seed)
    groups = [S.vt('IDM', 0), S.vt(str(rng.choice(['OVM', 'SIM', 'IDM'])), 0)]
    flows = []
    for _ in range(int(rng.integers(1, 5))):
        kind = int(rng.integers(0, 4))
        end = float(rng.choice([8.0, 15.0])) if kind & 1 else 0.0
        if kind & 2:
            flows.append((int(rng.integers(0, 2)), 0.0, float(rng.choice([5.0, 12.0])),
                          float(rng.choice([0.0, 1.0])), end, float(rng.choice([0.1, 0.3, 1.0]))))
        else:
            flows.append((int(rng.integers(0, 2)), float(rng.choice([400.0, 1200.0])),
                          float(rng.choice([5.0, 12.0])), float(rng.choice([0.0, 1.0])), end, 0.0))
    dt = float(rng.choice([0.1, 0.2]))
    sd, base = int(rng.integers(0, 2 ** 40)), int(rng.choice([0, 7, 2 ** 33]))
    E = 2
    eng = EN.Engine(S.engine_types(groups), [], E, dtype='f64', capacity=64, topology=EN.TOPO_HIGHWAY,
                    env_kind=EN.ENV_HIGHWAY, length_lo=400.0, length_hi=400.0, sim_step=dt, seed=sd,
                    env_id_base=base, horizon=3000, inflows=flows, record_views=1)
    sims = [OH.HighwaySim(S.oracle_types(groups), [], capacity=64, length=400.0, dt=dt,
                          inflows=[dict(type=t, vehs_per_hour=q, depart_speed=v, begin=b, end=en,
                                        probability=p) for t, q, v, b, en, p in flows],
                          seed=sd, env_gid=base + g) for g in range(E)]
    for episode in range(2):
        eng.reset()
        if episode:
            for h in sims:
                h.reset()
        for k in range(150):
            eng.step()
            for h in sims:
                h.step()
            c = eng.counters.cpu().numpy()
            for g, h in enumerate(sims):
                assert (c[:, g] == [len(h.pos), h.departed, h.arrived]).all(), (episode, k, g)
            if k % 10 == 9:
                for g, h in enumerate(sims):
                    n = len(h.pos)
                    assert close(eng.pos.cpu().numpy()[:n, g], h.pos) <= 1e-9, (episode, k, g)
                    assert bool(eng.crash.cpu().numpy()[g]) == h.crash
            if any(h.done for h in sims):
                break
