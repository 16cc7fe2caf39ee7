"""Open highway with inflow (c4). CPU: invariants of the oracle restatement
(no reference fixture exists for SUMO's insertion/arrival: parity unpinned).
GPU: CUDA path vs the oracle — fp64 1e-9 per step, slot occupancy, departure
and arrival counters bit-exact; fp32 aggregates within 1 %."""
import numpy as np
import pytest

from oracle import controllers as OC
from oracle import highway as OH
from tests import scenarios as S


def make_oracle(n_init=10, cap=160, **kw):
    return OH.HighwaySim([OC.VType(OC.IDM, OC.idm_params())], [0] * n_init, capacity=cap, **kw)


def test_oracle_invariants_and_inflow_rate():
    h = make_oracle(0)
    for s in range(3000):
        _, r, d = h.step()
        assert len(h.pos) == h.departed - h.arrived
        assert all(h.pos[i] > h.pos[i + 1] for i in range(len(h.pos) - 1))     # head first
        assert all(0 <= x < 1000 for x in h.pos)
    # 2000 veh/h from t = 1 s, one per 1.8 s: floor((300 - 1) / 1.8) + 1
    assert h.departed == int((300 - 1) / 1.8) + 1
    assert h.arrived > 100 and not h.crash


def test_oracle_blocked_entry_waits():
    h = make_oracle(10)            # stationary vehicle at x = 0 blocks the entry
    for _ in range(30):
        h.step()
    assert h.spawned == 0
    for _ in range(300):
        h.step()
    assert h.spawned > 0


def engine(E, dtype, n_init=10, cap=160, **kw):
    from flow_b200 import engine as EN
    types = [EN.make_vtype(EN.CTRL_IDM, [30, 1, 1, 1.5, 4, 2])]
    return EN.Engine(types, [0] * n_init, E, dtype=dtype, capacity=cap, topology=EN.TOPO_HIGHWAY,
                     env_kind=EN.ENV_HIGHWAY, length_lo=1000.0, length_hi=1000.0,
                     inflow_vehs_per_hour=2000.0, inflow_depart_speed=20.0, inflow_type=0,
                     horizon=3000, **kw)


def close(a, b, floor=1.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float((np.abs(a - b) / np.maximum(np.abs(b), floor)).max()) if a.size else 0.0


@pytest.mark.gpu
@pytest.mark.parametrize('n_init', [0, 10])
def test_gpu_highway_fp64_vs_oracle(n_init):
    eng = engine(3, 'f64', n_init, record_views=1)
    eng.reset()
    h = make_oracle(n_init)
    assert close(eng.pos.cpu().numpy()[:n_init, 1], h.pos) <= 1e-12
    for s in range(1200):
        o_obs, o_rew, o_done = h.step()
        obs, rew, done = eng.step()
        n = len(h.pos)
        c = eng.counters.cpu().numpy()
        assert (c[:, 2] == [n, h.departed, h.arrived]).all(), s
        assert (eng.active.cpu().numpy()[:, 0] == np.array([1] * n + [0] * (160 - n))).all(), s
        assert close(eng.pos.cpu().numpy()[:n, 0], h.pos) <= 1e-9, s
        assert close(eng.vel.cpu().numpy()[:n, 1], h.vel) <= 1e-9, s
        assert close(eng.headway.cpu().numpy()[:n, 2], h.headway) <= 1e-9, s
        assert close(obs.cpu().numpy()[1], o_obs) <= 1e-9, s
        assert close(rew.cpu().numpy()[2], o_rew) <= 1e-9, s
    assert h.arrived > 20


@pytest.mark.gpu
def test_gpu_highway_fused_fp32_aggregates():
    import torch
    K = 3000
    eng = engine(40, 'f32', 10)
    eng.reset()
    rew_seq = torch.empty(K, 40, device='cuda')
    eng.step(None, k=K, reward_seq=rew_seq)
    h = make_oracle(10)
    r = 0.0
    for _ in range(K):
        _, rr, _ = h.step()
        r += rr
    c = eng.counters.cpu().numpy()
    assert (c[1] == h.departed).all() and (c[2] == h.arrived).all() and (c[0] == len(h.pos)).all()
    n = len(h.pos)
    assert abs(eng.vel.cpu().numpy()[:n, 39].mean() - np.mean(h.vel)) <= 0.01 * np.mean(h.vel)
    assert abs(float(rew_seq[:, 7].sum()) - r) <= 0.01 * max(abs(r), 1.0)
    st = eng.stats()
    assert st['env_steps'] == K * 40 and st['crashes'] == 0


@pytest.mark.gpu
def test_gpu_highway_start_positions_that_wrap_past_the_end():
    """InitialConfig(x0 > 0): gen_even_start_pos wraps the last vehicles to the start of the edge
    (`% length`, network/base.py:380); the slots must still be head first."""
    eng = engine(2, 'f64', 30, x0=400.0, record_views=1)
    eng.reset()
    h = make_oracle(30, x0=400.0)
    assert max(h.pos) == h.pos[0] and min(h.pos) < 400.0 < max(h.pos)      # it did wrap
    assert close(eng.pos.cpu().numpy()[:30, 1], h.pos) <= 1e-12
    for s in range(300):
        o_obs, o_rew, o_done = h.step()
        obs, rew, done = eng.step()
        n = len(h.pos)
        assert close(eng.pos.cpu().numpy()[:n, 0], h.pos) <= 1e-9, s
        assert close(rew.cpu().numpy()[1], o_rew) <= 1e-9, s
    assert not h.crash and h.arrived > 5


# ---------------------------------------------------------------- probabilistic inflows
def prob_oracle(gid, seed=77, p=0.2, **kw):
    return OH.HighwaySim([OC.VType(OC.IDM, OC.idm_params())], [], capacity=96, length=600.0,
                         inflows=[dict(type=0, probability=p, depart_speed=15.0, begin=1.0, end=120.0),
                                  dict(type=0, vehs_per_hour=300.0, depart_speed=15.0, begin=2.0,
                                       end=30.0)],
                         seed=seed, env_gid=gid, **kw)


def test_oracle_philox_known_answer():
    """Random123's published known-answer vectors for Philox4x32-10."""
    assert OH.philox4x32_10([0, 0, 0, 0], [0, 0]) == \
        [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert OH.philox4x32_10([0xffffffff] * 4, [0xffffffff] * 2) == \
        [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert OH.philox4x32_10([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344],
                            [0xa4093822, 0x299f31d0]) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_oracle_probabilistic_inflow_rate_and_end():
    """p = 0.2 / s over [1 s, 120 s): about 24 vehicles (binomial, sd 4.8), none generated after
    `end`; the periodic flow (one per 12 s from 2 s) stops before its end = 30 s: 2, 14, 26."""
    tot = []
    for gid in range(8):
        h = prob_oracle(gid)
        for _ in range(1500):
            h.step()
        assert h.spawned_by_flow[1] == 3
        assert h.pending[0] == 0 and not h.crash
        tot.append(h.spawned_by_flow[0])
    assert len(set(tot)) > 1                      # keyed by the env id
    assert abs(np.mean(tot) - 23.8) < 3 * 4.8 / np.sqrt(8), tot
    a, b = prob_oracle(3), prob_oracle(3)
    for _ in range(300):
        a.step(); b.step()
    assert a.pos == b.pos                          # counter-based: reproducible


@pytest.mark.gpu
def test_gpu_probabilistic_inflow_matches_oracle_bit_for_bit():
    """The engine draws from the same Philox stream (seed, GLOBAL env id, episode, flow, step):
    arrivals, slot occupancy and trajectories equal the oracle's, and a shard that starts at
    env_id_base = 5 reproduces envs 5.. of the unsharded run."""
    from flow_b200 import engine as EN
    types = [EN.make_vtype(EN.CTRL_IDM, [30, 1, 1, 1.5, 4, 2])]
    flows = [(0, 0.0, 15.0, 1.0, 120.0, 0.2), (0, 300.0, 15.0, 2.0, 30.0, 0.0)]

    def mk(E, base):
        return EN.Engine(types, [], E, dtype='f64', capacity=96, topology=EN.TOPO_HIGHWAY,
                         env_kind=EN.ENV_HIGHWAY, length_lo=600.0, length_hi=600.0, seed=77,
                         env_id_base=base, horizon=3000, inflows=flows, record_views=1)
    eng, shard = mk(8, 0), mk(3, 5)
    eng.reset(); shard.reset()
    sims = [prob_oracle(g) for g in range(8)]
    for s in range(1500):
        eng.step(); shard.step()
        for h in sims:
            h.step()
        if s % 50 == 0 or s > 1450:
            c = eng.counters.cpu().numpy()
            for g, h in enumerate(sims):
                assert (c[:, g] == [len(h.pos), h.departed, h.arrived]).all(), (s, g)
                n = len(h.pos)
                assert close(eng.pos.cpu().numpy()[:n, g], h.pos) <= 1e-9, (s, g)
    import torch
    assert torch.equal(eng.pos[:, 5:], shard.pos) and torch.equal(eng.counters[:, 5:], shard.counters)
    # second episode: another stream
    eng.reset()
    for h in sims:
        h.reset()
    for s in range(400):
        eng.step()
        for h in sims:
            h.step()
    c = eng.counters.cpu().numpy()
    assert (c[1] == [h.departed for h in sims]).all()
