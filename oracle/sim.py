"""CPU oracle — TEST INFRASTRUCTURE ONLY (never imported by flow_b200/).

One-environment fp64 restatement of the reference's `Env.reset/step` path on a
ring road: initial placement, k.update (leader / headway / follower),
controllers + fail-safes (oracle/controllers.py), apply_acceleration, the
SUMO Euler move, collision flag, observation and reward.

What is restated from the reference (pinned, see tests/golden/):
  gen_even_start_pos      flow/core/kernel/network/base.py:263-391, :515-608
  ring edge-start table   flow/networks/ring.py:194-216
  apply_acceleration      flow/core/kernel/vehicle/traci.py:952-963
  headway/leader/follower flow/core/kernel/vehicle/traci.py:218-250
  Env.step order + done   flow/envs/base.py:294-412
  desired_velocity        flow/core/rewards.py:6-59
  WaveAttenuation reward  flow/envs/ring/wave_attenuation.py:113-139
  WaveAttenuationPO obs   flow/envs/ring/wave_attenuation.py:248-269
  Accel/WaveAtt full obs  flow/envs/ring/accel.py:116-123
  sorted_ids / abs. pos.  flow/envs/ring/accel.py:100-183
  LaneChangeAccel(PO)Env  flow/envs/ring/lane_change_accel.py:100-263
  per-lane neighbours     flow/core/kernel/vehicle/traci.py:699-937
  multi-agent fronts      flow/envs/multiagent/ring/wave_attenuation.py:195-247,
                          flow/envs/multiagent/ring/accel.py:154-206
  sims_per_step, evaluate flow/envs/base.py:325-383, flow/envs/ring/accel.py:109-114

What stands in for SUMO (third-party C++, not in /root/reference; pinned only
by tests/fast_tests/test_files/ring_230_emission.csv for the first steps,
otherwise "parity unpinned"):
  * the ring is one continuous loop of circumference C = length + 4*junction
    (junction = 0.1 m, flow/networks/ring.py:197); controllers never see a
    ':' edge;
  * Euler move x' = (x + v'*dt) mod C with v' = the commanded speed;
  * leader = next vehicle ahead in the same lane, headway = front-to-front
    distance mod C minus the leader's length; alone in lane -> none / 1000;
  * collision <=> some headway < 0 after the move, or (single lane) a vehicle went through
    its leader inside the step;
  * SumoParams(use_ballistic=True): x' = x + (v + v')/2 * dt;
  * vehicles without a command (RL vehicle when rl_actions is None,
    SimCarFollowingController) follow IDM with their SumoCarFollowingParams.
"""
import math

import numpy as np

from . import controllers as C

VEHICLE_LENGTH = 5.0  # flow/core/kernel/network/base.py:10

ENV_ACCEL, ENV_WAVE_ATTENUATION, ENV_WAVE_ATTENUATION_PO, \
    ENV_LANE_CHANGE_ACCEL, ENV_LANE_CHANGE_ACCEL_PO, ENV_HIGHWAY, \
    ENV_MA_WAVE_ATTENUATION_PO, ENV_MA_ACCEL_PO = range(8)

F32_EPS = float(np.finfo(np.float32).eps)


def ring_edge_tables(length, junction=0.1):
    """(edgestarts, internal_edgestarts) of RingNetwork, ring.py:194-216."""
    edgestarts = [("bottom", 0.0),
                  ("right", 0.25 * length + junction),
                  ("top", 0.5 * length + 2 * junction),
                  ("left", 0.75 * length + 3 * junction)]
    internal = [(":right_0", 0.25 * length),
                (":top_0", 0.5 * length + junction),
                (":left_0", 0.75 * length + 2 * junction),
                (":bottom_0", length + 3 * junction)]
    return edgestarts, internal


def gen_even_start_pos(length, num_vehicles, n_lanes=1, x0=0.0, bunching=0.0,
                       min_gap=0.0, lanes_distribution=float('inf'),
                       junction=0.1):
    """Even placement on a ring; returns (x[V] in loop coordinates, lane[V]).

    Follows gen_even_start_pos / _get_start_pos_util
    (flow/core/kernel/network/base.py:263-391, :515-608) specialised to the
    four equal ring edges; a position that falls into a junction slot is
    moved to the start of the next edge (:339-354)."""
    min_gap = max(0.0, min_gap)
    bunching = max(0.0, bunching)
    lanes_distr = int(min(max(lanes_distribution, 1), n_lanes))
    edge_len = length / 4.0
    if edge_len > min_gap + VEHICLE_LENGTH:
        distribution_length = 4 * (edge_len * min(n_lanes, lanes_distr))
    else:
        distribution_length = 0.0
    available = distribution_length - lanes_distr * bunching - \
        num_vehicles * (min_gap + VEHICLE_LENGTH)
    if available < 0:
        raise ValueError('There is not enough space to place all vehicles '
                         'in the network.')
    if num_vehicles == 0:
        return [], []
    increment = available / num_vehicles
    edgestarts, internal = ring_edge_tables(length, junction)
    total = sorted(edgestarts + internal, key=lambda t: t[1])
    xs, lanes = [], []
    x = x0
    count = 0
    while count < num_vehicles:
        # get_edge(x): flow/core/kernel/network/traci.py:267-271
        idx = max(i for i, (_, s) in enumerate(total) if x >= s)
        if total[idx][0].startswith(':'):
            idx = 0 if idx == len(total) - 1 else idx + 1
            x = total[idx][1]
        for lane in range(min(n_lanes, lanes_distr)):
            count += 1
            xs.append(x)
            lanes.append(lane)
            if count == num_vehicles:
                break
        x = (x + increment + VEHICLE_LENGTH + min_gap) % length
    return xs, lanes


class RingSim:
    """One ring environment. Vehicle index == position in `initial_ids`."""

    def __init__(self, types, type_of_vehicle, length=230.0, n_lanes=1,
                 dt=0.1, junction=0.1, speed_limit=30.0, env_kind=ENV_ACCEL,
                 x0=0.0, bunching=0.0, init_min_gap=0.0,
                 lanes_distribution=float('inf'), warmup_steps=0,
                 horizon=3000, target_velocity=10.0, act_lo=-4.5, act_hi=2.6,
                 clip_actions=False, obs_max_length=None,
                 lane_change_duration=5, noise_fn=None, evaluate=False,
                 sort_vehicles=False, sims_per_step=1, use_ballistic=False):
        self.types = types
        self.type_of = list(type_of_vehicle)
        self.num_vehicles = len(self.type_of)
        self.n_lanes = n_lanes
        self.dt = dt
        self.junction = junction
        self.speed_limit = speed_limit
        self.env_kind = env_kind
        self.init = dict(x0=x0, bunching=bunching, min_gap=init_min_gap,
                         lanes_distribution=lanes_distribution)
        self.warmup_steps = warmup_steps
        self.horizon = horizon
        self.target_velocity = target_velocity
        self.act_lo, self.act_hi = act_lo, act_hi
        self.clip_actions = clip_actions
        self.obs_max_length = obs_max_length
        self.lane_change_duration = lane_change_duration
        self.noise_fn = noise_fn  # (step, vehicle) -> N(0,1) sample
        self.lc_bug_compat = True  # get_last_lc returns the headway (vehicle/traci.py:614)
        self.sims_per_step = int(sims_per_step)   # EnvParams.sims_per_step (base.py:325)
        self.use_ballistic = use_ballistic        # SumoParams.use_ballistic (simulation/traci.py:95-96)
        self.evaluate = evaluate            # EnvParams.evaluate (accel.py:111-112)
        self.sort_vehicles = sort_vehicles  # additional_params['sort_vehicles'] (accel.py:150-169)
        # AccelEnv.absolute_position / prev_pos: dicts made once in __init__ (accel.py:77-78),
        # so they persist across resets; None = no entry yet (key -1001)
        self.abs_position = None
        self.prev_pos = None
        self.rl_idx = [i for i, t in enumerate(self.type_of)
                       if types[t].controller == C.RL]
        self.reset(length)

    # -- reset ------------------------------------------------------------
    def reset(self, length, run_warmup=True):
        """Env.reset (flow/envs/base.py:414-560): place, update, warm up."""
        V = self.num_vehicles
        self.net_length = float(length)
        self.circ = float(length) + 4 * self.junction
        xs, lanes = gen_even_start_pos(
            length, V, self.n_lanes, junction=self.junction, **self.init)
        self.pos = [float(x) for x in xs]
        self.vel = [0.0] * V
        self.lane = list(lanes)
        self.accel = [0.0] * V
        self.ctrl_state = [dict() for _ in range(V)]
        self.last_lc = [-float('inf')] * V   # vehicle/traci.py:178-180
        self.time_counter = 0
        self.step_index = 0
        self.crash = False
        self.done = False
        self.update()
        obs = self.get_state()
        if run_warmup:
            for _ in range(self.warmup_steps):
                obs, _, _ = self.step(None)
        if self.env_kind in (ENV_ACCEL, ENV_LANE_CHANGE_ACCEL, ENV_LANE_CHANGE_ACCEL_PO):
            # AccelEnv.reset tail (accel.py:178-181), after the observation was taken
            self.abs_position = list(self.pos)
            self.prev_pos = list(self.pos)
        return obs

    def sorted_idx(self):
        """AccelEnv.sorted_ids (accel.py:150-169): stable sort by absolute position."""
        V = self.num_vehicles
        if not self.sort_vehicles or self.abs_position is None:
            return list(range(V))
        return sorted(range(V), key=lambda i: self.abs_position[i])

    def additional_command(self):
        """AccelEnv.additional_command (accel.py:125-148), called before the simulator
        moves (base.py:365). WaveAttenuationEnv is not an AccelEnv and LaneChangeAccelEnv
        overrides it without the bookkeeping (lane_change_accel.py:156-161)."""
        if self.env_kind != ENV_ACCEL:
            return
        V = self.num_vehicles
        first = self.abs_position is None
        if first:
            self.abs_position, self.prev_pos = [0.0] * V, [0.0] * V
        for i in range(V):
            this = self.pos[i]
            change = 0.0 if first else this - self.prev_pos[i]
            self.abs_position[i] = ((this if first else self.abs_position[i]) + change) % self.circ
            self.prev_pos[i] = this

    def set_state(self, pos=None, vel=None, lane=None):
        if pos is not None:
            self.pos = [float(x) for x in pos]
        if vel is not None:
            self.vel = [float(x) for x in vel]
        if lane is not None:
            self.lane = [int(x) for x in lane]
        self.update()

    # -- k.update ---------------------------------------------------------
    def update(self):
        """Leader / headway / follower of every vehicle
        (flow/core/kernel/vehicle/traci.py:218-250 around SUMO's getLeader)."""
        V = self.num_vehicles
        if self.n_lanes == 1 and V > 1:
            # single lane, no overtaking: if the vehicles are still in index
            # order round the loop (forward distances sum to one lap) the
            # leader of i is i+1 and the follower i-1 — the same answer the
            # general search below gives, without the O(V^2) scan.
            C_, pos, types, type_of = self.circ, self.pos, self.types, self.type_of
            d = [(pos[i + 1 - V] - pos[i]) % C_ for i in range(V)]
            if abs(sum(d) - C_) < 1e-6 * C_ and min(d) > 0:
                self.leader = [(i + 1) % V for i in range(V)]
                self.follower = [(i - 1) % V for i in range(V)]
                self.headway = [d[i] - types[type_of[i + 1 - V]].length
                                for i in range(V)]
                return
        self.leader = [-1] * V
        self.follower = [-1] * V
        self.headway = [1e3] * V
        fol_h = [None] * V
        for i in range(V):
            best, best_d = -1, None
            for j in range(V):
                if j == i or self.lane[j] != self.lane[i]:
                    continue
                d = (self.pos[j] - self.pos[i]) % self.circ
                if best_d is None or d < best_d:
                    best, best_d = j, d
            if best >= 0:
                h = best_d - self.types[self.type_of[best]].length
                self.leader[i] = best
                self.headway[i] = h
        for i in range(V):
            j = self.leader[i]
            if j >= 0 and (fol_h[j] is None or self.headway[i] < fol_h[j]):
                self.follower[j] = i
                fol_h[j] = self.headway[i]

    def lane_neighbours(self, i):
        """Per lane: (headway, tailway, leader, follower) of vehicle i, the result
        of TraCIVehicle._multi_lane_headways on a closed loop
        (flow/core/kernel/vehicle/traci.py:699-937). In each lane the leader is
        the vehicle with the smallest forward distance d = x_j - x_i (+ circ when
        negative; a vehicle of ANOTHER lane at the same x counts as ahead, d = 0),
        the follower the one with the smallest backward distance b = x_i - x_j
        for x_j < x_i, else x_i - x_j + circ. A vehicle alone in its own lane is
        its own leader and follower one lap away (the edge walk comes back to the
        vehicle's own edge); an empty lane gives 1000 / -1. Ties: leader = lowest
        index, follower = highest index (stable sort by position + bisect_left)."""
        p, C_ = self.pos[i], self.circ
        own_len = self.types[self.type_of[i]].length
        out = []
        for lane in range(self.n_lanes):
            members = [j for j in range(self.num_vehicles) if self.lane[j] == lane]
            others = [j for j in members if j != i]
            if not members:
                out.append((1000, 1000, -1, -1))
                continue
            cand = others if others else [i]
            lead, best = -1, None
            for j in cand:
                d = self.pos[j] - p
                if d < 0 or j == i:
                    d = d + C_
                if best is None or d < best:
                    lead, best = j, d
            head = best - self.types[self.type_of[lead]].length
            foll, bestb = -1, None
            for j in cand:
                b = p - self.pos[j]
                if not self.pos[j] < p:
                    b = b + C_
                if bestb is None or b <= bestb:
                    foll, bestb = j, b
            out.append((head, bestb - own_len, lead, foll))
        return out

    def mean_speed(self):
        return float(np.mean(self.vel))

    # -- Env.step ---------------------------------------------------------
    def step(self, rl_actions):
        """One Env.step (flow/envs/base.py:294-412): sims_per_step simulation steps with
        the same actions (stopping at a collision, :378-383), then observation, done and
        reward once. Returns (obs, reward, done). A finished env is frozen."""
        if self.done:
            return self.get_state(), 0.0, True
        clipped = None
        for _ in range(self.sims_per_step):
            clipped = self._sim_step(rl_actions)
            if self.crash:
                break
        obs = self.get_state()
        self.done = (self.time_counter >= self.sims_per_step *
                     (self.warmup_steps + self.horizon)) or self.crash
        acts = rl_actions
        if rl_actions is not None and self.clip_actions:
            acts = clipped
        reward = self.compute_reward(acts, self.crash)
        return obs, reward, self.done

    def _sim_step(self, rl_actions):
        """Body of the `for _ in range(sims_per_step)` loop; returns the clipped actions."""
        V, dt = self.num_vehicles, self.dt
        self.time_counter += 1
        next_vel = [0.0] * V
        clipped = None
        lc_env = self.env_kind in (ENV_LANE_CHANGE_ACCEL, ENV_LANE_CHANGE_ACCEL_PO)
        directions = None
        # _apply_rl_actions hands action r to the r-th RL vehicle of sorted_ids
        # (accel.py:100-106, lane_change_accel.py:138-154)
        rl_order = [i for i in self.sorted_idx() if i in self.rl_idx]
        if rl_actions is not None:
            if lc_env:
                # LaneChangeAccelEnv._apply_rl_actions (lane_change_accel.py:132-154):
                # interleaved [acc, dir] per RL vehicle; Box clip first (base.py:614)
                acc_part = [min(max(float(a), self.act_lo), self.act_hi)
                            for a in rl_actions[0::2]]
                directions = [min(max(float(d), -1.0), 1.0) for d in rl_actions[1::2]]
                clipped = acc_part
                for q, i in enumerate(rl_order):
                    last = self.headway[i] if self.lc_bug_compat else self.last_lc[i]
                    if self.time_counter <= self.lane_change_duration + last:
                        directions[q] = 0.0
                if any(d not in (-1.0, 0.0, 1.0) for d in directions):
                    # vehicle/traci.py:973-975
                    raise ValueError("Direction values for lane changes may only "
                                     "be: -1, 0, or 1.")
            else:
                clipped = [min(max(float(a), self.act_lo), self.act_hi)
                           for a in rl_actions]
        for i in range(V):
            vt = self.types[self.type_of[i]]
            if vt.controller == C.RL and clipped is not None:
                a = clipped[rl_order.index(i)]
            else:
                z = self.noise_fn(self.step_index, i) \
                    if (self.noise_fn and vt.noise > 0) else 0.0
                a = C.get_action(self, i, vt, self.ctrl_state[i], dt,
                                 self.speed_limit, z)
            self.accel[i] = a
            # a speed commanded through slowDown passes SUMO's speed mode (restated); vehicles
            # the simulator drives itself (SimCarFollowingController, un-commanded RL) do not
            commanded = vt.controller != C.SIM and not (vt.controller == C.RL and clipped is None)
            if commanded and vt.speed_mode:
                a = C.speed_mode_accel(self, i, vt, a, self.speed_limit)
            # apply_acceleration, vehicle/traci.py:961-962
            next_vel[i] = max(self.vel[i] + a * dt, 0.0)
        # apply_lane_change (vehicle/traci.py:965-997) + restated execution: the
        # change happens in this simulation step unless the vehicle would overlap
        # its new leader or new follower in the target lane (pre-move positions)
        if directions is not None:
            new_lane = list(self.lane)
            for q, i in enumerate(rl_order):
                d = int(directions[q])
                if d == 0:
                    continue
                target = min(max(self.lane[i] + d, 0), self.n_lanes - 1)
                if target == self.lane[i]:
                    continue
                ok = True
                for j in range(V):
                    if j == i or self.lane[j] != target:
                        continue
                    fwd = (self.pos[j] - self.pos[i]) % self.circ
                    back = (self.pos[i] - self.pos[j]) % self.circ
                    if fwd - self.types[self.type_of[j]].length < 0 or \
                            back - self.types[self.type_of[i]].length < 0:
                        ok = False
                if ok:
                    new_lane[i] = target
                    self.last_lc[i] = self.time_counter
            self.lane = new_lane
        self.additional_command()      # base.py:365, before simulation_step
        # simulator move (restated SUMO Euler step)
        for i in range(V):
            if self.use_ballistic:      # restated SUMO ballistic update
                x = self.pos[i] + 0.5 * (self.vel[i] + next_vel[i]) * dt
            else:
                x = self.pos[i] + next_vel[i] * dt
            self.vel[i] = next_vel[i]
            self.pos[i] = x % self.circ
        self.step_index += 1
        old_leader, old_headway = list(self.leader), list(self.headway)
        self.update()
        self.crash = any(h < 0 for h in self.headway)
        if self.n_lanes == 1:
            # restated SUMO: on a single lane a vehicle cannot get past its leader; one that
            # went THROUGH it inside the step (the forward distance to that leader wraps by a
            # lap) has collided even if the two no longer overlap afterwards
            for i, j in enumerate(old_leader):
                if j >= 0:
                    d_new = (self.pos[j] - self.pos[i]) % self.circ - self.types[self.type_of[j]].length
                    if d_new - old_headway[i] > 0.5 * self.circ:
                        self.crash = True
        return clipped

    # -- observations / rewards -------------------------------------------
    def get_state(self):
        k = self.env_kind
        if k == ENV_ACCEL:
            # accel.py:116-123, rows in sorted_ids order
            order = self.sorted_idx()
            return np.array([self.vel[i] / self.speed_limit for i in order] +
                            [self.pos[i] / self.circ for i in order])
        if k == ENV_WAVE_ATTENUATION:
            # wave_attenuation.py:141-148 (get_ids order)
            return np.array([v / self.speed_limit for v in self.vel] +
                            [x / self.circ for x in self.pos])
        if k == ENV_WAVE_ATTENUATION_PO:
            # wave_attenuation.py:248-269
            rl = self.rl_idx[0]
            lead = self.leader[rl] if self.leader[rl] >= 0 else rl
            max_length = self.obs_max_length or self.circ
            return np.array([
                self.vel[rl] / 15.,
                (self.vel[lead] - self.vel[rl]) / 15.,
                (self.pos[lead] - self.pos[rl]) % self.circ / max_length])
        if k == ENV_MA_WAVE_ATTENUATION_PO:
            # multiagent/ring/wave_attenuation.py:195-213, one row per RL vehicle
            max_length = self.obs_max_length or self.circ
            out = []
            for rl in self.rl_idx:
                lead = self.leader[rl] if self.leader[rl] >= 0 else rl
                out += [self.vel[rl] / 15., (self.vel[lead] - self.vel[rl]) / 15.,
                        self.headway[rl] / max_length]
            return np.array(out)
        if k == ENV_MA_ACCEL_PO:
            # MultiAgentAccelPOEnv.get_state, multiagent/ring/accel.py:163-206: rows in
            # get_rl_ids order; lead_head is x_lead - x_rl - length(rl), NOT wrapped
            out = []
            for rl in self.rl_idx:
                lead, fol = self.leader[rl], self.follower[rl]
                if lead < 0:
                    lead_speed, lead_head = self.speed_limit, self.circ
                else:
                    lead_speed = self.vel[lead]
                    lead_head = self.pos[lead] - self.pos[rl] - self.types[self.type_of[rl]].length
                if fol < 0:
                    follow_speed, follow_head = 0, self.circ
                else:
                    follow_speed, follow_head = self.vel[fol], self.headway[fol]
                out += [self.pos[rl] / self.circ, self.vel[rl] / self.speed_limit,
                        (lead_speed - self.vel[rl]) / self.speed_limit, lead_head / self.circ,
                        (self.vel[rl] - follow_speed) / self.speed_limit, follow_head / self.circ]
            return np.array(out)
        if k == ENV_LANE_CHANGE_ACCEL:
            # lane_change_accel.py:114-130
            order = self.sorted_idx()
            return np.array([self.vel[i] / self.speed_limit for i in order] +
                            [self.pos[i] / self.circ for i in order] +
                            [self.lane[i] / self.n_lanes for i in order])
        if k == ENV_LANE_CHANGE_ACCEL_PO:
            # LaneChangeAccelPOEnv.get_state (lane_change_accel.py:215-263) as it
            # behaves: the `return` sits inside the loop over RL vehicles, so only
            # the FIRST RL vehicle's block is filled and one ego speed appended
            # (length 4*n_rl*lanes + 1); headways / tailways are copied before the
            # in-place normalisation, i.e. they stay in metres (1000 = empty lane);
            # neighbour speeds / max_speed; ego speed un-normalised.
            L = self.n_lanes
            obs = [0.0] * (4 * len(self.rl_idx) * L)
            if not self.lc_bug_compat:
                # corrected form (the declared observation_space, 4*n_rl*L + n_rl): every
                # RL vehicle's block, distances / network length (1 for an empty lane),
                # ego speeds / max_speed
                for q, rl in enumerate(self.rl_idx):
                    nb = self.lane_neighbours(rl)
                    obs[4 * L * q:4 * L * (q + 1)] = \
                        [t[0] / self.circ if t[2] >= 0 else 1.0 for t in nb] + \
                        [t[1] / self.circ if t[3] >= 0 else 1.0 for t in nb] + \
                        [self.vel[t[2]] / self.speed_limit if t[2] >= 0 else 0.0 for t in nb] + \
                        [self.vel[t[3]] / self.speed_limit if t[3] >= 0 else 0.0 for t in nb]
                return np.array(obs + [self.vel[rl] / self.speed_limit for rl in self.rl_idx])
            rl = self.rl_idx[0]
            nb = self.lane_neighbours(rl)
            obs[0:4 * L] = [float(t[0]) for t in nb] + [float(t[1]) for t in nb] + \
                [self.vel[t[2]] / self.speed_limit if t[2] >= 0 else 0.0 for t in nb] + \
                [self.vel[t[3]] / self.speed_limit if t[3] >= 0 else 0.0 for t in nb]
            obs.append(self.vel[rl])
            return np.array(obs)
        raise NotImplementedError(k)

    def desired_velocity(self, fail=False):
        """flow/core/rewards.py:6-59."""
        vel = np.array(self.vel)
        n = len(vel)
        if any(vel < -100) or fail or n == 0:
            return 0.
        target = self.target_velocity
        max_cost = np.linalg.norm(np.array([target] * n))
        cost = np.linalg.norm(vel - target)
        return max(max_cost - cost, 0) / (max_cost + F32_EPS)

    def compute_reward(self, rl_actions, fail):
        k = self.env_kind
        if k == ENV_MA_ACCEL_PO:    # multiagent/ring/accel.py:154-161, shared by the agents
            return self.desired_velocity(fail)
        if k == ENV_ACCEL:
            if self.evaluate:     # accel.py:111-112
                return float(np.mean(self.vel))
            return self.desired_velocity(fail)
        if k in (ENV_LANE_CHANGE_ACCEL, ENV_LANE_CHANGE_ACCEL_PO):
            # lane_change_accel.py:100-112 (penalty compares get_last_lc, i.e.
            # the headway under the reference's bug, with the time counter)
            reward = self.desired_velocity(fail)
            for i in self.rl_idx:
                last = self.headway[i] if self.lc_bug_compat else self.last_lc[i]
                if last == self.time_counter:
                    reward -= 0.1
            return reward
        if k in (ENV_WAVE_ATTENUATION, ENV_WAVE_ATTENUATION_PO, ENV_MA_WAVE_ATTENUATION_PO):
            # wave_attenuation.py:113-139; multiagent/ring/wave_attenuation.py:221-247 (shared)
            if rl_actions is None:
                return 0
            vel = np.array(self.vel)
            if any(vel < -100) or fail:
                return 0.
            reward = 4. * np.mean(vel) / 20
            mean_actions = np.mean(np.abs(np.array(rl_actions)))
            if mean_actions > 0:
                reward += 4 * (0 - mean_actions)
            return float(reward)
        raise NotImplementedError(k)
