"""TEST INFRASTRUCTURE ONLY (see oracle/sim.py): CPU restatement of the figure-eight path
(SURVEY 8f rank 4): FigureEightNetwork (flow/networks/figure_eight.py) stepped under AccelEnv
(flow/envs/ring/accel.py), one lane - the network of
examples/exp_configs/rl/singleagent/singleagent_figure_eight.py.

What is the reference's own arithmetic and is followed line by line:
  * AccelEnv.get_state / compute_reward / _apply_rl_actions (accel.py:100-123) with
    rewards.desired_velocity (flow/core/rewards.py:6-59);
  * get_x_by_id on FigureEightNetwork.specify_edge_starts / specify_internal_edge_starts
    (figure_eight.py:217-263) through TraCIKernelNetwork.get_x (network/traci.py:273-289): the two
    passes through the centre junction are keys of the internal table (':center_1' bottom -> top,
    ':center_0' right -> left: start + position), the four 0.28 m node slots (':bottom', ':top',
    ':right', ':left') are not under SUMO's '<node>_0' names, so their START is returned;
  * BaseController.get_action's guard (base_controller.py:96-99): a vehicle on an internal edge
    (name starts with ':') gets no command - the simulator's model drives it through the junction;
  * controllers, noise, fail-safes (oracle/controllers.py), apply_acceleration.

What SUMO does and is RESTATED ("parity unpinned"). The network is ONE closed
loop of length L = 4 r + 3 pi r + 2 * 6.2 + 4 * 0.28 that passes the centre junction twice:
northbound at c_A = 0.28 + r and westbound at c_B = c_A + L / 2 (loop coordinate s, position 0 =
start of ':bottom'). Along the loop this is a single-lane ring: the leader is the nearest vehicle
ahead (mod L), nobody overtakes, a vehicle that went through its leader inside a step has
collided (oracle/sim.py). The crossing is restated as a merge followed by a diverge - the same
statement as the merge network's (oracle/merge.py): with delta = c - s the distance of a
vehicle's front bumper to the junction box [c, c + 6.2) of its own pass, a vehicle with
-6.2 < delta <= zone (approaching, or its front inside the box) also regards the vehicles of the
OTHER pass whose body has not cleared the box (-(6.2 + length) < delta_j <= zone) and that are
closer to / further through it (delta_j < delta_i; ties: the lower index) as ahead of it, at a
bumper-to-bumper gap (delta_i - delta_j) - length_j. Its leader is the nearer of the loop leader
and that crossing vehicle. Collision <=> negative headway towards the loop leader, or towards a
crossing leader while both bodies are inside the box; a negative gap towards a crossing leader
outside the box reads as MIN_CTRL_GAP for the controllers (bumper to bumper, oracle/merge.py).
"""
import math

import numpy as np

from . import controllers as C

JUNCTION_LEN = 2.9 + 3.3 * 1       # figure_eight.py:72, one lane
INNER_LEN = 0.28                   # figure_eight.py:73
F32_EPS = float(np.finfo(np.float32).eps)
MIN_CTRL_GAP = 1e-3


class Figure8Geometry:
    """Loop layout and the x coordinate of FigureEightNetwork (figure_eight.py:217-263)."""

    def __init__(self, radius):
        r = self.r = float(radius)
        self.ring_edgelen = r * math.pi / 2.0          # figure_eight.py:70
        il, jl, sp, re3 = 2 * r, JUNCTION_LEN, INNER_LEN, 3 * self.ring_edgelen
        rl = 3 * r * math.pi / 2.      # the edge LENGTH of specify_edges (figure_eight.py:124)
        # (name, start on the loop, length, internal?) in driving order
        self.segments = [
            (':bottom', 0.0, sp, True), ('bottom', sp, r, False),
            (':center_1', il / 2 + sp, jl, True), ('top', il / 2 + jl + sp, r, False),
            (':top', il + jl + sp, sp, True), ('upper_ring', il + jl + 2 * sp, rl, False),
            (':right', il + re3 + jl + 2 * sp, sp, True),
            ('right', il + re3 + jl + 3 * sp, r, False),
            (':center_0', 3 / 2 * il + re3 + jl + 3 * sp, jl, True),
            ('left', 3 / 2 * il + re3 + 2 * jl + 3 * sp, r, False),
            (':left', 2 * il + re3 + 2 * jl + 3 * sp, sp, True),
            ('lower_ring', 2 * il + re3 + 2 * jl + 4 * sp, rl, False)]
        self.L = 2 * il + 2 * re3 + 2 * jl + 4 * sp
        self.cA = il / 2 + sp                            # box of the northbound pass
        self.cB = 3 / 2 * il + re3 + jl + 3 * sp         # box of the westbound pass
        self.J = jl
        # junction slots whose SUMO edge id is not a key of the internal table: x = their start
        self.start_only = (':bottom', ':top', ':right', ':left')

    def segment(self, s):
        """(name, start, internal?) of the edge that holds loop coordinate s."""
        best = self.segments[0]
        for seg in self.segments:
            if s >= seg[1]:
                best = seg
        return best[0], best[1], best[3]

    def on_internal(self, s):
        return self.segment(s)[2]

    def x_abs(self, s):
        name, start, _ = self.segment(s)
        return start if name in self.start_only else start + (s - start)

    def delta(self, s):
        """(pass, distance of the front bumper to the box of that pass): the northbound pass for
        the quarter loop either side of c_A, else the westbound one (c_B = c_A + L / 2). Vehicles
        in the last quarter before s = L read a delta below -(L / 4), i.e. no role at all."""
        dA = self.cA - s
        if dA > -0.25 * self.L:
            return 0, dA
        return 1, self.cB - s


class _ControlView:
    """The simulator as the controllers see it (headways floored for crossing leaders that
    overlap outside the box)."""

    def __init__(self, sim):
        self._sim = sim
        self.headway = [MIN_CTRL_GAP if (h < 0 and sim.cross[i] and not sim.both_in_box(i, j)) else h
                        for i, (h, j) in enumerate(zip(sim.headway, sim.leader))]

    def __getattr__(self, name):
        return getattr(self._sim, name)


class Figure8Sim:
    def __init__(self, types, type_of, radius, positions, dt=0.1, speed_limit=30.0,
                 target_velocity=20.0, horizon=1500, warmup_steps=0, sims_per_step=1, zone=None,
                 act_lo=-3.0, act_hi=3.0, clip_actions=True, evaluate=False, noise_fn=None):
        self.types, self.type_of = types, list(type_of)
        self.g = Figure8Geometry(radius)
        self.initial = [float(p) for p in positions]
        self.dt, self.speed_limit, self.target_velocity = dt, speed_limit, target_velocity
        self.horizon, self.warmup_steps, self.sims_per_step = horizon, warmup_steps, int(sims_per_step)
        self.zone = float(radius if zone is None else zone)     # default: the approach edge
        self.act_lo, self.act_hi, self.clip_actions = act_lo, act_hi, clip_actions
        self.evaluate, self.noise_fn = evaluate, noise_fn
        self.rl_idx = [i for i, t in enumerate(self.type_of) if types[t].controller == C.RL]
        self.reset()

    # ------------------------------------------------------------------ state
    @property
    def num_vehicles(self):
        return len(self.pos)

    def mean_speed(self):
        return float(np.mean(self.vel))

    def reset(self):
        V = len(self.initial)
        self.pos, self.vel, self.accel = list(self.initial), [0.0] * V, [0.0] * V
        self.time_counter = 0
        self.crash = self.done = False
        self.update()
        obs = self.get_state()
        for _ in range(self.warmup_steps):
            obs, _, _ = self.step(None)
        return obs

    def set_state(self, pos=None, vel=None):
        if pos is not None:
            self.pos = [float(x) for x in pos]
        if vel is not None:
            self.vel = [float(x) for x in vel]
        self.update()

    def length_of(self, i):
        return self.types[self.type_of[i]].length

    def roles(self, i):
        """(pass, delta, regards the other pass?, is regarded by the other pass?)"""
        p, d = self.g.delta(self.pos[i])
        return p, d, -self.g.J < d <= self.zone, -(self.g.J + self.length_of(i)) < d <= self.zone

    def both_in_box(self, i, j):
        if j < 0:
            return False
        _, di, _, _ = self.roles(i)
        _, dj, _, _ = self.roles(j)
        return -(self.g.J + self.length_of(i)) < di < 0 and -(self.g.J + self.length_of(j)) < dj < 0

    def update(self):
        V, L = self.num_vehicles, self.g.L
        self.leader, self.headway, self.follower = [-1] * V, [1e3] * V, [-1] * V
        self.cross = [False] * V          # leader is a vehicle of the other pass
        role = [self.roles(i) for i in range(V)]
        for i in range(V):
            best, bd = -1, None
            for j in range(V):            # the loop leader: nearest ahead, ties to the lower index
                if j == i:
                    continue
                d = (self.pos[j] - self.pos[i]) % L
                if bd is None or d < bd:
                    best, bd = j, d
            cross = False
            pi_, di, regards, _ = role[i]
            if regards:
                for j in range(V):
                    pj, dj, _, regarded = role[j]
                    if j == i or pj == pi_ or not regarded:
                        continue
                    if not (dj < di or (dj == di and j < i)):
                        continue
                    d = di - dj
                    if bd is None or d < bd:
                        best, bd, cross = j, d, True
            if best >= 0:
                self.leader[i], self.cross[i] = best, cross
                self.headway[i] = bd - self.length_of(best)
        fh = [None] * V
        for i in range(V):
            j = self.leader[i]
            if j >= 0 and (fh[j] is None or self.headway[i] < fh[j]):
                self.follower[j], fh[j] = i, self.headway[i]

    # ------------------------------------------------------------------ Env.step
    def step(self, rl_actions=None):
        if self.done:
            return self.get_state(), 0.0, True
        for _ in range(self.sims_per_step):
            self._sim_step(rl_actions)
            if self.crash:
                break
        self.done = self.crash or self.time_counter >= self.sims_per_step * (
            self.warmup_steps + self.horizon)
        return self.get_state(), self.compute_reward(), self.done

    def _sim_step(self, rl_actions, commanded=None):
        """commanded: hook of tests/golden/ref_harness.RefFigure8 ({index: acceleration} from the
        REFERENCE's controllers and _apply_rl_actions; indices missing = no command)."""
        dt, V, L = self.dt, self.num_vehicles, self.g.L
        self.time_counter += 1
        view = _ControlView(self)
        nv = [0.0] * V
        for i in range(V):
            vt = self.types[self.type_of[i]]
            if commanded is not None:
                cmd = i in commanded
                a = commanded[i] if cmd else C.sim_model_accel(view, i, vt, self.speed_limit)
            else:
                internal = self.g.on_internal(self.pos[i])
                if vt.controller == C.RL and rl_actions is not None:
                    # _apply_rl_actions: straight to apply_acceleration, whatever the edge
                    a = float(rl_actions[self.rl_idx.index(i)])
                    if self.clip_actions:
                        a = min(max(a, self.act_lo), self.act_hi)
                    cmd = True
                elif vt.controller in (C.SIM, C.RL) or internal:
                    a, cmd = C.sim_model_accel(view, i, vt, self.speed_limit), False
                else:
                    z = self.noise_fn(self.time_counter - 1, i) if (self.noise_fn and vt.noise > 0) else 0.0
                    a, cmd = C.get_action(view, i, vt, {}, dt, self.speed_limit, z), True
            self.accel[i] = a
            if cmd and vt.speed_mode:
                a = C.speed_mode_accel(view, i, vt, a, self.speed_limit)
            nv[i] = max(self.vel[i] + a * dt, 0.0)
        old_leader, old_headway, old_cross = list(self.leader), list(self.headway), list(self.cross)
        for i in range(V):
            self.vel[i] = nv[i]
            self.pos[i] = (self.pos[i] + nv[i] * dt) % L
        self.update()
        self.crash = self.collided(old_leader, old_headway, old_cross)

    def collided(self, old_leader, old_headway, old_cross):
        L = self.g.L
        for i, h in enumerate(self.headway):
            j = self.leader[i]
            if j >= 0 and h < 0 and (not self.cross[i] or self.both_in_box(i, j)):
                return True
        for i, j in enumerate(old_leader):     # went through its loop leader inside the step
            if j >= 0 and not old_cross[i]:
                d_new = (self.pos[j] - self.pos[i]) % L - self.length_of(j)
                if d_new - old_headway[i] > 0.5 * L:
                    return True
        return False

    # ------------------------------------------------------------------ AccelEnv
    def get_state(self):
        """accel.py:116-123 (sort_vehicles False: get_ids order)."""
        return np.array([v / self.speed_limit for v in self.vel] +
                        [self.g.x_abs(s) / self.g.L for s in self.pos])

    def compute_reward(self):
        """accel.py:108-114 with rewards.desired_velocity (rewards.py:6-59)."""
        vel = np.array(self.vel)
        if self.evaluate:
            return float(np.mean(vel))
        if self.crash or len(vel) == 0:
            return 0.0
        max_cost = np.linalg.norm(np.array([self.target_velocity] * len(vel)))
        cost = np.linalg.norm(vel - self.target_velocity)
        return max(max_cost - cost, 0) / (max_cost + F32_EPS)
