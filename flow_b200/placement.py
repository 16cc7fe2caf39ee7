"""Host-side starting positions that the reference draws from HOST random streams.

The engine's reset kernel generates `gen_even_start_pos` in closed form (csrc/fs_reset.cuh).
The stochastic placements of InitialConfig are drawn by the reference from Python's `random`
and numpy's GLOBAL generator, so they are restated here on the host with the very same calls
in the very same order — a caller that seeds `random` / `np.random` like it did for the
reference gets the reference's placements, number for number (pinned by
tests/golden/start_pos.json) — and handed to the engine between the two halves of its reset
(fs_reset_stage + fs_set_state):

  perturbation > 0     flow/core/kernel/network/base.py:384-389  (np.random.normal per vehicle,
                       clipped to the vehicle's own edge)
  spacing = 'random'   flow/core/kernel/network/base.py:393-487  (random.random per vehicle,
                       sorted, packed edge by edge and lane by lane)
  shuffle = True       flow/envs/base.py:275-277                 (random.shuffle of initial_ids)

Positions are loop / road coordinates x = edge start + position on the edge
(`get_x`, flow/core/kernel/network/traci.py:267-271).
"""
import random

import numpy as np

VEHICLE_LENGTH = 5.0       # flow/core/kernel/network/base.py:10
RING_JUNCTION_LENGTH = 0.1


def ring_edges(length, junction=RING_JUNCTION_LENGTH):
    """[(name, start, edge_length)] of a RingNetwork of `length` (flow/networks/ring.py:194-216):
    four edges of length / 4 with a junction slot after each."""
    q = length / 4.0
    return [('bottom', 0.0, q), ('right', q + junction, q), ('top', 2 * q + 2 * junction, q),
            ('left', 3 * q + 3 * junction, q)]


def highway_edges(length):
    return [('highway_0', 0.0, float(length))]


def edge_of(x, edges):
    """The (non-internal) edge that holds loop coordinate x: the last edge starting at or before x
    (get_edge walks the edge starts backwards, network/traci.py:267-271)."""
    best = edges[0]
    for e in edges:
        if x >= e[1]:
            best = e
    return best


def perturb(xs, edges, sigma, normal=None):
    """network/base.py:384-389: every vehicle, in index order, moves by N(0, sigma) along its
    own edge and is kept on it: pos <- max(0, min(edge_length, pos + perturb))."""
    normal = np.random.normal if normal is None else normal
    out = []
    for x in xs:
        _, start, elen = edge_of(x, edges)
        p = normal(0, sigma)
        out.append(start + max(0, min(elen, (x - start) + p)))
    return out


def random_positions(edges, num_lanes, lanes_distribution, min_gap, bunching, num_vehicles,
                     rand=None):
    """gen_random_start_pos (network/base.py:393-487) for edges_distribution = 'all' on a network
    whose edges all have `num_lanes` lanes. Returns (xs, lanes). Raises ValueError where the
    reference raises FatalFlowError (no room)."""
    rand = random.random if rand is None else rand
    min_gap = max(0, min_gap)
    lanes_distr = min(max(int(min(lanes_distribution, 1 << 30)), 1), num_lanes)
    usable = [e for e in edges if e[2] > min_gap + VEHICLE_LENGTH]
    per_edge = min(num_lanes, lanes_distr)
    available = sum(e[2] * per_edge for e in usable) - lanes_distr * bunching - \
        num_vehicles * (min_gap + VEHICLE_LENGTH)
    if available < 0:
        raise ValueError('There is not enough space to place all vehicles in the network.')
    efs = min_gap + VEHICLE_LENGTH          # extra front space
    for _ in usable:
        available -= efs * per_edge
    pos = [rand() * available for _ in range(num_vehicles)]
    pos.sort()
    for i in range(num_vehicles):
        pos[i] += (VEHICLE_LENGTH + min_gap) * i
    decrement, k = 0, 0
    xs, lanes = [], []
    for i in range(num_vehicles):
        name, start, elen = usable[k]
        p = (pos[i] - decrement) % (elen - efs)
        lane = int(((pos[i] - decrement) - p) / (elen - efs))
        p += efs
        while lane > per_edge - 1:
            decrement += per_edge * (elen - efs)
            k += 1
            name, start, elen = usable[k]
            p = (pos[i] - decrement) % (elen - efs)
            lane = int(((pos[i] - decrement) - p) / (elen - efs))
            p += efs
        xs.append(start + p)
        lanes.append(lane)
    return xs, lanes


def shuffled(ids, shuffle=None):
    """flow/envs/base.py:275-277: random.shuffle of the id list (a copy is returned)."""
    out = list(ids)
    (random.shuffle if shuffle is None else shuffle)(out)
    return out
