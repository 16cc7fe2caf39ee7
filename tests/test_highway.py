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
