"""The merge network under MergePOEnv: the restatement oracle/merge.py against the
golden rollout recorded with the REFERENCE's own MergePOEnv methods (tests/golden/make_golden.py
merge(): _apply_rl_actions, additional_command, get_state, compute_reward of flow/envs/merge.py
run on the restated network), and the geometry against flow/networks/merge.py:162-211."""
import os

import numpy as np
import pytest

from oracle import merge as OM
from tests import scenarios as S

GOLD = os.path.join(os.path.dirname(__file__), 'golden', 'merge.npz')


def test_geometry_matches_specify_edge_starts():
    """MergeNetwork.specify_edge_starts / specify_internal_edge_starts (flow/networks/merge.py:
    196-221) for pre 500, merge 100, post 100: inflow_highway 0, ':left' 100, left 100.1,
    ':center' 600.1, center 622.6, inflow_merge 722.6, ':bottom' 822.6, bottom 822.7. get_x_by_id
    adds the position on a real edge and returns the bare start of an internal one
    (network/traci.py:267-271 with vehicle/traci.py:1001-1007)."""
    g = OM.MergeGeometry(500, 100, 100)
    ax = pytest.approx
    assert g.LH == ax(622.6) and g.LM == ax(222.6) and g.L_total == ax(722.6)
    assert g.net_length == ax(200 + 500 + 100 + 100 + 0.2 + 22.5)
    # highway arm: inflow_highway -> :left -> left -> :center -> center
    assert g.x_abs(0, 50.0) == 50.0
    assert g.x_abs(0, 100.05) == 100.0                    # on ':left'
    assert g.x_abs(0, 110.1) == ax(110.1)                 # 10 m into 'left'
    assert g.x_abs(0, 610.0) == ax(600.1)                 # on ':center'
    assert g.x_abs(0, 632.6) == ax(632.6)                 # 10 m into 'center'
    # ramp arm: inflow_merge -> :bottom -> bottom -> :center -> center
    assert g.x_abs(1, 10.0) == ax(732.6)
    assert g.x_abs(1, 100.05) == ax(822.6)                # on ':bottom'
    assert g.x_abs(1, 110.1) == ax(832.7)                 # 10 m into 'bottom'
    assert g.x_abs(1, 210.0) == ax(600.1)                 # on ':center'
    assert g.x_abs(1, 232.6) == ax(632.6)                 # 10 m into 'center'
    # both arms share the axis on 'center': same u -> same x
    u = g.LH + 30.0
    assert g.x_abs(0, u - g.off[0]) == ax(g.x_abs(1, u - g.off[1]))


def test_oracle_reproduces_reference_env_methods():
    z = np.load(GOLD)
    sim = S.merge_sim()
    n_gone = 0
    for k in range(len(z['actions'])):
        obs, rew, done = sim.step(z['actions'][k])
        np.testing.assert_array_equal(obs, z['obs'][k], err_msg='step %d' % k)
        assert rew == z['reward'][k], k
        assert done == bool(z['done'][k]), k
        ctrl = list(sim.rl_veh) + [-1] * (sim.num_rl - len(sim.rl_veh))
        assert ctrl == list(z['ctrl'][k]), k
        assert [sim.num_vehicles, sim.departed, sim.arrived] == list(z['counts'][k]), k
        n_gone += int((obs[0::5] < -30).sum())
    assert n_gone >= 1                       # a vehicle that left was observed (the -1001 rows)
    assert z['counts'][-1][2] > 50           # the network really flows
    np.testing.assert_array_equal(np.array(sim.pos), z['pos_final'])
    np.testing.assert_array_equal(np.array(sim.ids), z['ids_final'])


def test_leader_rules_on_the_merged_axis():
    """Before the zone a ramp vehicle and a highway vehicle do not see each other; inside it the
    one ahead on the axis leads; past the junction everything is one road."""
    g = OM.MergeGeometry(500, 100, 100)
    types = S.oracle_types(S.merge_groups())

    def sim_with(initial):
        return OM.MergeSim(types, g, initial=initial, inflows=[], dt=0.2, num_rl=1)
    far = g.LH - 300.0
    s = sim_with([(0, 0, far - g.off[0]), (0, 1, far + 20.0 - g.off[1])])
    assert s.leader == [-1, -1]
    near = g.LH - 10.0
    s = sim_with([(0, 0, near - g.off[0]), (0, 1, near + 7.0 - g.off[1])])
    assert s.leader == [1, -1] and s.headway[0] == pytest.approx(2.0)
    assert not s.collided()                  # side by side on different roads is not a crash
    s = sim_with([(0, 0, near - g.off[0]), (0, 1, g.LH + 5.0 - g.off[1])])
    assert s.leader == [1, -1]               # a vehicle on the merged road leads both arms
    s = sim_with([(0, 0, g.LH + 1.0 - g.off[0]), (0, 1, g.LH + 4.0 - g.off[1])])
    assert s.collided()                      # overlapping on the shared road


def test_rl_rows_close_up_and_fill_in_arrival_order():
    sim = S.merge_sim(num_rl=2, inflows=[(1, 0, 3600.0, 10.0, 1.0)], initial=[(0, 0, 400.0)])
    seen = []
    for k in range(120):
        sim.step(np.zeros(2))
        if sim.rl_veh not in seen[-1:]:
            seen.append(list(sim.rl_veh))
    assert seen[0] == [] or len(seen[0]) <= 2
    flat = [v for row in seen for v in row]
    assert all(len(r) <= 2 for r in seen)
    firsts = []
    for v in flat:
        if v not in firsts:
            firsts.append(v)
    assert firsts == sorted(firsts)          # ids of one flow grow with departure order
    assert len(firsts) >= 3                  # rows were handed on to later vehicles
