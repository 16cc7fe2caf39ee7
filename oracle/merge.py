"""TEST INFRASTRUCTURE ONLY (see oracle/sim.py): CPU restatement of the merge network path
(SURVEY 8f rank 4): MergeNetwork (flow/networks/merge.py) stepped under MergePOEnv
(flow/envs/merge.py:28-231), single-lane highway + single-lane on-ramp.

What is the reference's own arithmetic and is followed line by line:
  * MergePOEnv.get_state / compute_reward / _apply_rl_actions / additional_command
    (flow/envs/merge.py:112-231): the `rl_veh` list of controlled vehicles (FIFO from `rl_queue`,
    at most num_rl, a slot is freed when its vehicle left), 5 numbers per controlled vehicle,
    desired_velocity + time-headway penalty; including what it does for a controlled vehicle
    that left the network in the step just taken (get_speed -> -1001, no leader / follower);
  * get_x_by_id on MergeNetwork.specify_edge_starts / specify_internal_edge_starts
    (flow/networks/merge.py:196-221) with TraCIKernelNetwork.get_x's treatment of junction lanes
    (network/traci.py:273-289: ':center_0' is not a key, so the junction's START is returned);
  * controllers, noise, fail-safes (oracle/controllers.py), apply_acceleration.

What SUMO does and is RESTATED ("parity unpinned"): both arms are laid on
one axis u = distance travelled on the arm + (LH - L_arm), so that the merge point (start of
edge 'center') is u = LH for both. A vehicle's leader is the nearest vehicle ahead on that axis
among: vehicles of its own arm, vehicles past the merge point, and - once both are within
`zone` metres of the merge point - vehicles of the other arm (a zipper by position stands in
for SUMO's junction right-of-way). Ties in u: the lower slot is ahead. Collision <=> headway < 0
towards a leader that shares the road: same arm, or both inside the junction / on 'center'
(u >= LH - 22.5); two vehicles side by side on the two approach roads are no collision: the
one behind on the axis brakes for the other as if bumper to bumper - its controller and the
simulator model read a gap of MIN_CTRL_GAP instead of the negative projected one (with a negative
gap an IDM accelerates again once |gap| > s*, and the pair would drive into the junction abreast).
Observations and rewards keep the geometric headway.
Vehicles leave at the end of 'center'; a <flow> emits at the start of its arm, at most one
vehicle per flow and step, the flow with the earliest scheduled departure first, once the gap to the arm's last vehicle reaches
the simulator model's desired gap s* for the depart speed (SUMO delays an insertion it deems
unsafe for the given departSpeed).
"""
import math

import numpy as np

from . import controllers as C

INFLOW_EDGE_LEN = 100.0        # flow/networks/merge.py:8
J_SMALL, J_CENTER = 0.1, 22.5  # junction lengths implied by specify_edge_starts (0.1 / 22.6 - 0.1)
F32_EPS = float(np.finfo(np.float32).eps)
ARM_HIGHWAY, ARM_RAMP = 0, 1
MIN_CTRL_GAP = 1e-3            # what a controller reads for an overlapping cross-arm leader


class _ControlView:
    """The simulator as the controllers see it: everything of MergeSim, headways floored for
    overlapping leaders on the other approach road (module docstring)."""

    def __init__(self, sim):
        self._sim = sim
        zc = sim.g.LH - J_CENTER
        self.headway = [
            MIN_CTRL_GAP if (h < 0 and j >= 0 and sim.arm[i] != sim.arm[j] and
                             not (sim.u(i) >= zc and sim.u(j) >= zc)) else h
            for i, (h, j) in enumerate(zip(sim.headway, sim.leader))]

    def __getattr__(self, name):
        return getattr(self._sim, name)


class MergeGeometry:
    """Lengths and the x coordinate of MergeNetwork (edge starts of flow/networks/merge.py:196-221)."""

    def __init__(self, pre, merge, post):
        self.pre, self.merge, self.post = float(pre), float(merge), float(post)
        self.LH = INFLOW_EDGE_LEN + J_SMALL + self.pre + J_CENTER     # highway arm, merge point
        self.LM = INFLOW_EDGE_LEN + J_SMALL + self.merge + J_CENTER   # ramp arm, merge point
        self.off = (0.0, self.LH - self.LM)                           # arm coordinate -> axis u
        self.L_total = self.LH + self.post                            # u of the exit
        # k.network.length(): every edge and junction (network/traci.py:210-212)
        self.net_length = 2 * INFLOW_EDGE_LEN + self.pre + self.merge + self.post + \
            2 * J_SMALL + J_CENTER
        self.x_ramp0 = INFLOW_EDGE_LEN + self.pre + self.post + J_SMALL + J_CENTER   # 'inflow_merge'

    def x_abs(self, arm, p):
        """get_x_by_id of a vehicle p metres into its arm."""
        j_center = INFLOW_EDGE_LEN + self.pre + J_SMALL          # start of ':center'
        if arm == ARM_HIGHWAY:
            if p < INFLOW_EDGE_LEN:
                return p
            if p < INFLOW_EDGE_LEN + J_SMALL:
                return INFLOW_EDGE_LEN                            # ':left': start, no position added
            if p < INFLOW_EDGE_LEN + J_SMALL + self.pre:
                return p
            if p < self.LH:
                return j_center
            return p
        if p < INFLOW_EDGE_LEN:
            return self.x_ramp0 + p
        if p < INFLOW_EDGE_LEN + J_SMALL:
            return self.x_ramp0 + INFLOW_EDGE_LEN                 # ':bottom'
        if p < INFLOW_EDGE_LEN + J_SMALL + self.merge:
            return self.x_ramp0 + p
        if p < self.LM:
            return j_center
        return self.LH + (p - self.LM)                            # on 'center'


class MergeSim:
    def __init__(self, types, geometry, initial=(), inflows=(), dt=0.2, speed_limit=30.0,
                 capacity=128, num_rl=5, target_velocity=25.0, horizon=600, warmup_steps=0,
                 sims_per_step=1, zone=None, evaluate=False, noise_fn=None, seed=0, env_gid=0):
        """initial: [(type index, arm, p)]; inflows: [dict(type, arm, vehs_per_hour, depart_speed,
        begin)] served in definition order."""
        self.types, self.g = types, geometry
        self.initial, self.inflows = list(initial), [dict(f) for f in inflows]
        self.dt, self.speed_limit, self.capacity = dt, speed_limit, capacity
        self.num_rl, self.target_velocity = int(num_rl), target_velocity
        self.horizon, self.warmup_steps, self.sims_per_step = horizon, warmup_steps, int(sims_per_step)
        # look-ahead zone before the merge point: by default the ramp edge 'bottom' + the junction
        self.zone = float(geometry.merge + J_CENTER if zone is None else zone)
        self.evaluate, self.noise_fn = evaluate, noise_fn
        self.seed, self.env_gid = seed, env_gid
        self.episode = 0
        self.reset()

    # ------------------------------------------------------------------ state
    def reset(self):
        self.episode += 1
        self.pos = [float(p) for _, _, p in self.initial]
        self.arm = [int(a) for _, a, _ in self.initial]
        self.type_of = [int(t) for t, _, _ in self.initial]
        self.vel = [0.0] * len(self.pos)
        self.accel = [0.0] * len(self.pos)
        self.ids = list(range(len(self.pos)))                 # initial vehicle i -> id i
        self.seq = list(range(len(self.pos)))                 # departure order
        self.next_seq = len(self.pos)
        self.spawned_by_flow = [0] * len(self.inflows)
        self.time_counter = 0
        self.departed, self.arrived = len(self.pos), 0     # initial vehicles count as departed
        self.crash = self.done = False
        self.rl_veh = []                                      # ids of the controlled vehicles
        self.hi_seq = -1                                      # highest seq ever moved into rl_veh
        self.update()
        obs = self.get_state()
        for _ in range(self.warmup_steps):
            obs, _, _ = self.step(None)
        return obs

    @property
    def num_vehicles(self):
        return len(self.pos)

    def mean_speed(self):
        return float(np.mean(self.vel)) if self.vel else 0.0

    def u(self, i):
        return self.pos[i] + self.g.off[self.arm[i]]

    def is_rl(self, i):
        return self.types[self.type_of[i]].controller == C.RL

    def update(self):
        """Leader / headway / follower on the merged axis (see the module docstring)."""
        n, LH, z0 = len(self.pos), self.g.LH, self.g.LH - self.zone
        us = [self.u(i) for i in range(n)]
        self.leader, self.headway, self.follower = [-1] * n, [1e3] * n, [-1] * n
        for i in range(n):
            best, bd = -1, None
            for j in range(n):
                if j == i:
                    continue
                if not (us[j] > us[i] or (us[j] == us[i] and j < i)):
                    continue
                ok = self.arm[j] == self.arm[i] or us[j] >= LH or (us[i] >= z0 and us[j] >= z0)
                if not ok:
                    continue
                d = us[j] - us[i]
                if bd is None or d < bd:
                    best, bd = j, d
            if best >= 0:
                self.leader[i] = best
                self.headway[i] = bd - self.types[self.type_of[best]].length
        fh = [None] * n
        for i in range(n):
            j = self.leader[i]
            if j >= 0 and (fh[j] is None or self.headway[i] < fh[j]):
                self.follower[j], fh[j] = i, self.headway[i]

    # ------------------------------------------------------------------ Env.step
    def step(self, actions=None):
        """One Env.step: sims_per_step simulation steps on the same action vector
        (flow/envs/base.py:325-383), then observation, done, reward."""
        if self.done:
            return self.get_state(), 0.0, True
        for _ in range(self.sims_per_step):
            self._sim_step(actions)
            if self.crash:
                break
        self.done = self.crash or self.time_counter >= self.sims_per_step * (
            self.warmup_steps + self.horizon)
        return self.get_state(), self.compute_reward(), self.done

    def additional_command(self):
        """MergePOEnv.additional_command (merge.py:191-220): vehicles that left free their slot
        (the list closes up), waiting RL vehicles fill it in order of arrival. (The reference
        removes from rl_veh while iterating over it, which skips the element after a removed
        one for a step; not reproduced.)"""
        present = {self.ids[i]: i for i in range(len(self.ids))}
        self.rl_veh = [v for v in self.rl_veh if v in present]
        waiting = sorted((self.seq[i], self.ids[i]) for i in range(len(self.ids))
                         if self.is_rl(i) and self.seq[i] > self.hi_seq)
        for sq, vid in waiting:
            if len(self.rl_veh) >= self.num_rl:
                break
            self.rl_veh.append(vid)
            self.hi_seq = sq

    def _sim_step(self, actions, commanded_ids=None, additional_command=None):
        """commanded_ids / additional_command: hooks of tests/golden/ref_harness.RefMerge, which
        runs the REFERENCE's _apply_rl_actions and additional_command in their place."""
        dt = self.dt
        self.time_counter += 1
        n = len(self.pos)
        slot_of = {self.ids[i]: i for i in range(n)}
        commanded = {}
        if commanded_ids is not None:
            commanded = {slot_of[v]: float(a) for v, a in commanded_ids.items() if v in slot_of}
        elif actions is not None:                     # _apply_rl_actions (merge.py:106-112)
            for k, vid in enumerate(self.rl_veh):
                if vid in slot_of:
                    commanded[slot_of[vid]] = float(actions[k])
        nv = [0.0] * n
        view = _ControlView(self)
        for s in range(n):
            vt = self.types[self.type_of[s]]
            z = self.noise_fn(self.time_counter - 1, s) if (self.noise_fn and vt.noise > 0) else 0.0
            a = C.get_action(view, s, vt, {}, dt, self.speed_limit, z)
            cmd = vt.controller not in (C.SIM, C.RL)
            if s in commanded:
                a, cmd = commanded[s], True           # clipped by the caller's Box (base.py:614)
            self.accel[s] = a
            if cmd and vt.speed_mode:
                a = C.speed_mode_accel(view, s, vt, a, self.speed_limit)
            nv[s] = max(self.vel[s] + a * dt, 0.0)
        (additional_command or self.additional_command)()   # base.py:365, before simulation_step
        for s in range(n):
            self.pos[s] = self.pos[s] + nv[s] * dt
            self.vel[s] = nv[s]
        keep = [s for s in range(n) if self.u(s) < self.g.L_total]
        self.arrived += n - len(keep)
        for name in ('pos', 'vel', 'arm', 'type_of', 'accel', 'ids', 'seq'):
            setattr(self, name, [getattr(self, name)[s] for s in keep])
        t = self.time_counter * dt
        dues = [(fl.get('begin', 1.0) + self.spawned_by_flow[f] * (3600.0 / fl['vehs_per_hour']), f)
                for f, fl in enumerate(self.inflows) if fl.get('vehs_per_hour', 0) > 0]
        for due, f in sorted(dues):            # earliest scheduled departure first (ties: flow order)
            fl = self.inflows[f]
            if not (t >= due and len(self.pos) < self.capacity):
                continue
            nt = self.types[fl['type']]
            same = [s for s in range(len(self.pos)) if self.arm[s] == fl['arm']]
            free = True
            if same:
                # restated SUMO insertion with a given departSpeed: delayed until the gap to the
                # last vehicle of the arm is the simulator model's desired gap s* at that speed
                last = min(same, key=lambda s: self.pos[s])
                vd = float(fl['depart_speed'])
                s_star = nt.sim_min_gap + max(0.0, vd * nt.sim_tau + vd * (vd - self.vel[last]) /
                                              (2 * math.sqrt(nt.max_accel * nt.max_decel)))
                free = self.pos[last] - self.types[self.type_of[last]].length >= s_star
            if free:
                self.pos.append(0.0); self.vel.append(float(fl['depart_speed']))
                self.arm.append(int(fl['arm'])); self.type_of.append(int(fl['type']))
                self.accel.append(0.0)
                self.ids.append(((f + 1) << 24) | self.spawned_by_flow[f])
                self.seq.append(self.next_seq)
                self.next_seq += 1
                self.spawned_by_flow[f] += 1
                self.departed += 1
        self.update()
        self.crash = self.collided()

    def collided(self):
        zc = self.g.LH - J_CENTER
        for i, h in enumerate(self.headway):
            j = self.leader[i]
            if j >= 0 and h < 0 and (self.arm[i] == self.arm[j] or
                                     (self.u(i) >= zc and self.u(j) >= zc)):
                return True
        return False

    # ------------------------------------------------------------------ MergePOEnv
    def get_state(self):
        """merge.py:114-159; a controlled vehicle that is gone reads as the kernel's error values."""
        ms, ml = self.speed_limit, self.g.net_length
        slot_of = {self.ids[i]: i for i in range(len(self.ids))}
        obs = [0.0] * (5 * self.num_rl)
        for k, vid in enumerate(self.rl_veh):
            s = slot_of.get(vid, -1)
            this_speed = self.vel[s] if s >= 0 else -1001
            lead = self.leader[s] if s >= 0 else -1
            fol = self.follower[s] if s >= 0 else -1
            if lead < 0:
                lead_speed, lead_head = ms, ml
            else:
                lead_speed = self.vel[lead]
                lead_head = self.g.x_abs(self.arm[lead], self.pos[lead]) - \
                    self.g.x_abs(self.arm[s], self.pos[s]) - self.types[self.type_of[s]].length
            if fol < 0:
                follow_speed, follow_head = 0, ml
            else:
                follow_speed, follow_head = self.vel[fol], self.headway[fol]
            obs[5 * k:5 * k + 5] = [this_speed / ms, (lead_speed - this_speed) / ms, lead_head / ml,
                                    (this_speed - follow_speed) / ms, follow_head / ml]
        return np.array(obs)

    def compute_reward(self):
        """merge.py:161-189 with rewards.desired_velocity (flow/core/rewards.py:6-59)."""
        if self.evaluate:
            return float(np.mean(self.vel)) if self.vel else float('nan')
        if self.crash:
            return 0
        vel = np.array(self.vel)
        n = len(vel)
        if n == 0:
            cost1 = 0
        else:
            max_cost = np.linalg.norm(np.array([self.target_velocity] * n))
            cost = np.linalg.norm(vel - self.target_velocity)
            cost1 = max(max_cost - cost, 0) / (max_cost + F32_EPS)
        slot_of = {self.ids[i]: i for i in range(len(self.ids))}
        cost2 = 0
        for vid in self.rl_veh:
            s = slot_of.get(vid, -1)
            if s >= 0 and self.leader[s] >= 0 and self.vel[s] > 0:
                t_headway = max(self.headway[s] / self.vel[s], 0)
                cost2 += min((t_headway - 1) / 1, 0)
        return max(1.00 * cost1 + 0.10 * cost2, 0)
