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


# ---- networks with several edges of different length (the merge) -----------------------------
MERGE_INFLOW_EDGE_LEN = 100.0     # flow/networks/merge.py:9


def merge_edges(pre, merge, post):
    """([(name, start, length)] real edges, [(name, start)] internal edges) of a MergeNetwork
    (specify_edge_starts / specify_internal_edge_starts, flow/networks/merge.py:196-221), with
    the nominal edge lengths of specify_edges (merge.py:120-160)."""
    I = MERGE_INFLOW_EDGE_LEN
    real = [('inflow_highway', 0.0, I), ('left', I + 0.1, float(pre)),
            ('center', I + pre + 22.6, float(post)), ('inflow_merge', I + pre + post + 22.6, I),
            ('bottom', 2 * I + pre + post + 22.7, float(merge))]
    internal = [(':left', I), (':center', I + pre + 0.1), (':bottom', 2 * I + pre + post + 22.6)]
    return real, internal


def merge_arm_position(edge, p, pre):
    """(arm, metres along the arm) of position p on `edge`: the engine's merge coordinates
    (include/flowstep.h FS_TOPO_MERGE). Arm 0 = inflow_highway -> left -> center,
    arm 1 = inflow_merge -> bottom -> center."""
    I = MERGE_INFLOW_EDGE_LEN
    return {'inflow_highway': (0, p), 'left': (0, I + 0.1 + p), 'center': (0, I + 0.1 + pre + 22.5 + p),
            'inflow_merge': (1, p), 'bottom': (1, I + 0.1 + p)}[edge]


def even_positions(real, internal, x0, min_gap, bunching, num_vehicles, num_lanes=1,
                   lanes_distribution=1 << 30):
    """gen_even_start_pos (network/base.py:256-392) with edges_distribution = 'all' on a network
    whose edges all have `num_lanes` lanes: [(edge, position)] and lanes, in placement order.
    Raises ValueError where the reference raises FatalFlowError."""
    min_gap = max(0, min_gap)
    lanes_distr = min(max(int(min(lanes_distribution, 1 << 30)), 1), num_lanes)
    length = {n: ln for n, _, ln in real}
    usable = [n for n, _, ln in real if ln > min_gap + VEHICLE_LENGTH]
    available = sum(length[n] * min(num_lanes, lanes_distr) for n in usable) - \
        lanes_distr * bunching - num_vehicles * (min_gap + VEHICLE_LENGTH)
    if available < 0:
        raise ValueError('There is not enough space to place all vehicles in the network.')
    if num_vehicles == 0:
        return [], []
    non_internal = sum(length.values())
    total = sorted([(n, s) for n, s, _ in real] + list(internal), key=lambda t: t[1])
    names = [n for n, _ in total]
    internal_names = {n for n, _ in internal}

    def get_edge(x):
        for n, s in reversed(total):
            if x >= s:
                return n, x - s

    increment = available / num_vehicles
    x, count, out, lanes = x0, 0, [], []
    while count < num_vehicles:
        pos = get_edge(x)
        while pos[0] in internal_names:           # never inside a junction: start of the next edge
            i = names.index(pos[0])
            nxt = total[0] if i == len(total) - 1 else total[i + 1]
            x, pos = nxt[1], (nxt[0], 0)
        while pos[0] not in usable:
            x = (x + length[pos[0]]) % non_internal
            pos = get_edge(x)
        for lane in range(min(num_lanes, lanes_distr)):
            count += 1
            out.append(pos)
            lanes.append(lane)
            if count == num_vehicles:
                break
        x = (x + increment + VEHICLE_LENGTH + min_gap) % non_internal
    return out, lanes


def merge_x(arm, p, pre, merge, post):
    """get_x_by_id of a vehicle p metres into its arm: edge start + position on a real edge, the
    bare start of an internal one (network/traci.py:273-289 with the merge tables above)."""
    I, jc = MERGE_INFLOW_EDGE_LEN, MERGE_INFLOW_EDGE_LEN + pre + 0.1
    ramp0 = I + pre + post + 22.6
    LH, LM = I + 0.1 + pre + 22.5, I + 0.1 + merge + 22.5
    if arm == 0:
        if p < I:
            return p
        if p < I + 0.1:
            return I
        if p < I + 0.1 + pre:
            return p
        return jc if p < LH else p
    if p < I:
        return ramp0 + p
    if p < I + 0.1:
        return ramp0 + I
    if p < I + 0.1 + merge:
        return ramp0 + p
    return jc if p < LM else LH + (p - LM)


def merge_edge_of(arm, p, pre, merge, post):
    """(edge name, position on it) of a vehicle p metres into its arm."""
    I = MERGE_INFLOW_EDGE_LEN
    first, second, mid = (('inflow_highway', 'left', pre), ('inflow_merge', 'bottom', merge))[arm][0:3]
    junction = ':left' if arm == 0 else ':bottom'
    if p < I:
        return first, p
    if p < I + 0.1:
        return junction, p - I
    if p < I + 0.1 + mid:
        return second, p - (I + 0.1)
    if p < I + 0.1 + mid + 22.5:
        return ':center', p - (I + 0.1 + mid)
    return 'center', p - (I + 0.1 + mid + 22.5)


# ---- the figure-eight loop -------------------------------------------------------------------
def figure8_edges(radius, placement_table=True):
    """([(name, start, length)] real edges in specify_edges order, [(name, start)] internal edges)
    of a one-lane FigureEightNetwork (flow/networks/figure_eight.py:66-73, 120-176, 217-263).
    placement_table: keep the two entries the network lists "for aimsun" ('bottom_to_top' at the
    start of ':center_1', 'right_to_left' at 6.2 + 3 * 0.28 = 7.04 m). gen_even_start_pos reads
    them through get_edge like any internal edge, so on the reference every x from 7.04 m to the
    end of the first junction pass is moved to the start of 'top'."""
    import math
    r = float(radius)
    rel, il, jl, sp = r * math.pi / 2., 2 * r, 2.9 + 3.3 * 1, 0.28
    real = [('bottom', sp, il / 2), ('top', il / 2 + jl + sp, il / 2),
            ('right', il + 3 * rel + jl + 3 * sp, il / 2),
            ('left', 3 / 2 * il + 3 * rel + 2 * jl + 3 * sp, il / 2),
            ('upper_ring', il + jl + 2 * sp, 3 * r * math.pi / 2.),
            ('lower_ring', 2 * il + 3 * rel + 2 * jl + 4 * sp, 3 * r * math.pi / 2.)]
    internal = [(':bottom', 0), (':center_1', il / 2 + sp), (':top', il + jl + sp),
                (':right', il + 3 * rel + jl + 2 * sp),
                (':center_0', 3 / 2 * il + 3 * rel + jl + 3 * sp),
                (':left', 2 * il + 3 * rel + 2 * jl + 3 * sp)]
    if placement_table:
        internal += [('bottom_to_top', il / 2 + sp), ('right_to_left', + jl + 3 * sp)]
    return real, internal


def figure8_edge_of(s, radius):
    """(edge name as SUMO reports it, position on it) of loop coordinate s."""
    real, internal = figure8_edges(radius, placement_table=False)
    total = sorted([(n, st) for n, st, _ in real] + list(internal), key=lambda t: t[1])
    name, start = total[0]
    for n, st in total:
        if s >= st:
            name, start = n, st
    if name in (':bottom', ':top', ':right', ':left'):
        name += '_0'
    return name, s - start
