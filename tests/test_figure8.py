"""The figure-eight network under AccelEnv: the restatement oracle/figure8.py against
the golden rollout recorded with the REFERENCE's own IDMController.get_action (junction guard),
AccelEnv._apply_rl_actions / get_state / compute_reward and TraCIKernelNetwork.get_x on the
network's own edge-start tables (tests/golden/make_golden.py figure8()), the loop geometry against
flow/networks/figure_eight.py:217-263, and the host placement against the reference's
gen_even_start_pos on that network (tests/golden/start_pos_figure8.json)."""
import json
import math
import os

import numpy as np
import pytest

from flow_b200 import placement as PL
from oracle import figure8 as F8
from tests import scenarios as S

G = os.path.join(os.path.dirname(__file__), 'golden')
CASES = json.load(open(os.path.join(G, 'start_pos_figure8.json')))


def test_geometry_matches_specify_edge_starts():
    """radius 30: ring_edgelen 15 pi, intersection 60, junction 6.2, inner 0.28
    (figure_eight.py:66-73); edge starts :217-238, internal starts :240-263."""
    g = F8.Figure8Geometry(30)
    rel, ax = 30 * math.pi / 2, pytest.approx
    starts = {n: s for n, s, _, _ in g.segments}
    assert starts['bottom'] == ax(0.28) and starts['top'] == ax(30 + 6.2 + 0.28)
    assert starts['upper_ring'] == ax(60 + 6.2 + 0.56)
    assert starts['right'] == ax(60 + 3 * rel + 6.2 + 0.84)
    assert starts['left'] == ax(90 + 3 * rel + 12.4 + 0.84)
    assert starts['lower_ring'] == ax(120 + 3 * rel + 12.4 + 1.12)
    assert starts[':center_1'] == ax(30.28) and starts[':center_0'] == ax(90 + 3 * rel + 6.2 + 0.84)
    assert g.L == ax(120 + 6 * rel + 12.4 + 1.12)
    assert g.cB - g.cA == ax(g.L / 2)                       # the loop is symmetric
    # get_x: node slots give their start, junction passes and edges start + position
    assert g.x_abs(0.1) == 0.0 and g.x_abs(10.28) == ax(10.28) and g.x_abs(33.0) == ax(33.0)
    assert g.x_abs(starts[':top'] + 0.2) == starts[':top']
    assert g.on_internal(31.0) and not g.on_internal(40.0) and g.on_internal(starts[':left'] + 0.1)
    assert g.delta(20.28) == (0, ax(10.0)) and g.delta(g.cB - 5.0) == (1, ax(5.0))


@pytest.mark.parametrize('c', CASES, ids=['r%d-n%d' % (c['radius'], c['n']) for c in CASES])
def test_even_placement_on_the_figure_eight(c):
    real, internal = PL.figure8_edges(c['radius'])
    kw = c['kw']
    pos, lane = PL.even_positions(real, internal, kw.get('x0', 0), kw.get('min_gap', 0),
                                  kw.get('bunching', 0), c['n'])
    assert [[e, p] for e, p in pos] == c['pos']              # bit for bit
    starts = {n: s for n, s, _ in real}
    assert [starts[e] + p for e, p in pos] == c['x']
    for x in c['x']:
        assert PL.figure8_edge_of(x, c['radius'])[0][0] != ':'


def test_oracle_reproduces_reference_controllers_and_env():
    z = np.load(os.path.join(G, 'figure8.npz'))
    sim = S.figure8_sim(positions=CASES[0]['x'])
    np.testing.assert_array_equal(np.array(sim.pos), z['pos'][0])
    crossing = 0
    for k in range(len(z['actions'])):
        obs, rew, done = sim.step(z['actions'][k])
        np.testing.assert_array_equal(obs, z['obs'][k], err_msg='step %d' % k)
        assert rew == z['reward'][k] and done == bool(z['done'][k]), k
        np.testing.assert_array_equal(np.array(sim.pos), z['pos'][k + 1])
        crossing += sum(sim.cross)
    assert crossing > 100                    # vehicles did wait for the other pass
    assert (z['pos'][1:] < z['pos'][:-1]).any()        # and went round the loop


def test_crossing_rules():
    g = F8.Figure8Geometry(30)
    types = S.oracle_types(S.figure8_groups())

    def sim_with(*pos):
        return F8.Figure8Sim(types, [0] * len(pos), 30, list(pos))
    # two vehicles approaching the junction on the two passes: the nearer one leads the other
    s = sim_with(g.cA - 20.0, g.cB - 12.0)
    assert s.leader == [1, 0] and s.cross == [True, False]
    assert s.headway[0] == pytest.approx(8.0 - 5.0)
    # outside the zone (default: the approach edge, 30 m) the other pass is not regarded
    s = sim_with(g.cA - 31.0, g.cB - 12.0)
    assert s.cross == [False, False]
    # a vehicle whose body is still in the box is regarded, one that cleared it is not
    s = sim_with(g.cA - 3.0, g.cB + 6.2 + 4.0)
    assert s.cross[0] and s.headway[0] == pytest.approx(3.0 + 10.2 - 5.0)
    s = sim_with(g.cA - 3.0, g.cB + 6.2 + 5.5)
    assert not s.cross[0]
    # abreast outside the box: no collision, the follower reads a 1 mm gap
    s = sim_with(g.cA - 6.0, g.cB - 4.0)
    assert s.headway[0] < 0 and not s.collided([-1, -1], [0, 0], [False, False])
    assert F8._ControlView(s).headway[0] == F8.MIN_CTRL_GAP
    # both bodies in the box: collision
    s = sim_with(g.cA + 3.0, g.cB + 1.0)
    assert s.collided([-1, -1], [0, 0], [False, False])
