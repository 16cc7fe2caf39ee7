"""CPU oracle — TEST INFRASTRUCTURE ONLY (never imported by flow_b200/).

Single-lane open highway with an inflow (config c4): one-environment fp64
restatement. Controllers, fail-safes, apply_acceleration and the
desired_velocity reward are the reference's (oracle/controllers.py,
flow/core/rewards.py:6-59). Everything SUMO does on this network is restated
and is "parity unpinned" (SURVEY.md §8c items 5, 6): no reference test or
fixture holds numbers for inflow insertion or arrival.

  * network: HighwayNetwork(length, lanes=1, num_edges=1) — one edge
    'highway_0' of `length` metres, no junctions (flow/networks/highway.py:88-104);
  * vehicles are kept HEAD FIRST: slot 0 is the vehicle furthest downstream;
    the leader of slot s is slot s-1, the head has no leader (headway 1000,
    flow/core/kernel/vehicle/traci.py:234-238); follower of s is s+1;
  * arrival: a vehicle whose position after the move is >= length leaves the
    network in that step (SUMO VAR_ARRIVED -> TraCIVehicle.remove,
    vehicle/traci.py:145-160) and the slots behind it close up;
  * inflow `<flow vehsPerHour=q begin=b departSpeed=v0 departPos=0>`
    (flow/core/kernel/network/traci.py:801-833): vehicle k is due at time
    b + k*3600/q; at the end of a step at most ONE due vehicle is inserted at
    x = 0 with speed v0 if a slot is free and the bumper gap to the last
    vehicle (x_last - length_last) is at least the new type's minGap;
    otherwise it waits (SUMO's insertion backlog);
  * initial vehicles: gen_even_start_pos on the single edge
    (flow/core/kernel/network/base.py:263-391), speed 0;
  * collision <=> some headway < 0 after the move; observation = speeds /
    speed_limit and positions / length per slot (0 for empty slots);
  * several inflows (`inflows=[dict(type, vehs_per_hour, depart_speed, begin)]`)
    are served in definition order, each needing the entry gap on its own; a flow
    with `end` emits only vehicles due before it; a flow with `probability` p
    (per second) instead makes one more vehicle due with probability p * dt in
    every step inside [begin, end) - SUMO scales the per-second probability by
    the step length - drawn from Philox4x32-10 keyed by (seed, global env id,
    episode, flow, step) exactly as the engine does, so the arrival pattern is
    reproducible bit for bit (SUMO's own generator is not: parity unpinned);
    vehicle k of inflow f carries the id ((f + 1) << 24) | k ('flow_<f>.<k>' in
    the reference, flow/core/params.py:1163-1213), initial vehicle i the id i.

Multi-agent front (`ma_state` / `ma_reward`): MultiAgentHighwayPOEnv.get_state
and compute_reward (flow/envs/multiagent/highway.py:97-172) over the RL vehicles
present, head first; pinned on the reference's own methods by
tests/golden/make_golden.py::highway_ma (tests/golden/highway_ma.npz). RL
vehicles take `actions[id]` UNCLIPPED (MultiEnv.clip_actions returns its input,
multiagent/base.py:372-395) and fall back to the restated simulator model when
no action is given for them.
"""
import numpy as np

from . import controllers as C


def philox4x32_10(counter, key):
    """Philox4x32-10 (Salmon et al., SC'11), the counter-based generator the engine uses;
    counter: 4 x u32, key: 2 x u32 -> 4 x u32."""
    c = [int(x) & 0xFFFFFFFF for x in counter]
    k = [int(x) & 0xFFFFFFFF for x in key]
    for _ in range(10):
        p0, p1 = 0xD2511F53 * c[0], 0xCD9E8D57 * c[2]
        c = [((p1 >> 32) ^ c[1] ^ k[0]) & 0xFFFFFFFF, p1 & 0xFFFFFFFF,
             ((p0 >> 32) ^ c[3] ^ k[1]) & 0xFFFFFFFF, p0 & 0xFFFFFFFF]
        k = [(k[0] + 0x9E3779B9) & 0xFFFFFFFF, (k[1] + 0xBB67AE85) & 0xFFFFFFFF]
    return c

F32_EPS = float(np.finfo(np.float32).eps)


class HighwaySim:
    def __init__(self, types, initial_types, capacity=160, length=1000.0, dt=0.1,
                 speed_limit=30.0, inflow_vehs_per_hour=2000.0, inflow_type=0,
                 inflow_depart_speed=20.0, inflow_begin=1.0, x0=0.0, bunching=0.0,
                 init_min_gap=0.0, horizon=3000, warmup_steps=0, target_velocity=10.0,
                 noise_fn=None, evaluate=False, use_ballistic=False, sims_per_step=1,
                 inflows=None, seed=0, env_gid=0):
        self.types = types
        self.initial_types = list(initial_types)
        self.capacity = capacity
        self.length = float(length)
        self.dt = dt
        self.speed_limit = speed_limit
        self.q = inflow_vehs_per_hour
        self.inflow_type = inflow_type
        self.depart_speed = inflow_depart_speed
        self.begin = inflow_begin
        if inflows is None:
            inflows = [dict(type=inflow_type, vehs_per_hour=inflow_vehs_per_hour,
                            depart_speed=inflow_depart_speed, begin=inflow_begin)] \
                if inflow_vehs_per_hour > 0 else []
        self.inflows = [dict(f) for f in inflows]
        self.seed, self.env_gid, self.episode = int(seed), int(env_gid), 0
        self.init = dict(x0=x0, bunching=bunching, min_gap=init_min_gap)
        self.horizon, self.warmup_steps = horizon, warmup_steps
        self.target_velocity = target_velocity
        self.noise_fn = noise_fn
        self.evaluate = evaluate              # AccelEnv.compute_reward, accel.py:111-112
        self.use_ballistic = use_ballistic    # SumoParams.use_ballistic
        self.sims_per_step = int(sims_per_step)   # EnvParams.sims_per_step, base.py:325
        self.reset()

    # getters the controller functions use ---------------------------------
    @property
    def num_vehicles(self):
        return len(self.pos)

    def mean_speed(self):
        return float(np.mean(self.vel))

    def reset(self):
        n = len(self.initial_types)
        min_gap = max(0.0, self.init['min_gap'])
        available = self.length - max(0.0, self.init['bunching']) - n * (min_gap + 5.0)
        if available < 0:
            raise ValueError('There is not enough space to place all vehicles in the network.')
        xs = []
        if n:
            increment = available / n
            x = self.init['x0']
            for _ in range(n):
                xs.append(x)
                x = (x + increment + 5.0 + min_gap) % self.length
        order = sorted(range(n), key=lambda i: -xs[i])        # head first
        self.pos = [float(xs[i]) for i in order]
        self.vel = [0.0] * n
        self.type_of = [self.initial_types[i] for i in order]
        self.ids = list(order)
        self.accel = [0.0] * n
        self.time_counter = 0
        self.spawned_by_flow = [0] * len(self.inflows)
        self.pending = [0] * len(self.inflows)
        self.episode += 1
        self.departed = n
        self.arrived = 0
        self.crash = False
        self.done = False
        self.update()
        for _ in range(self.warmup_steps):
            self.step()
        return self.get_state()

    def update(self):
        n = len(self.pos)
        self.leader = [s - 1 for s in range(n)]
        self.follower = [s + 1 if s + 1 < n else -1 for s in range(n)]
        self.headway = [1e3 if s == 0 else
                        self.pos[s - 1] - self.pos[s] - self.types[self.type_of[s - 1]].length
                        for s in range(n)]

    @property
    def spawned(self):
        return sum(self.spawned_by_flow)

    def step(self, actions=None):
        """One Env.step: sims_per_step simulation steps (base.py:325-383), then obs / reward.
        actions: {vehicle id: acceleration} for RL vehicles (multi-agent front) or None."""
        if self.done:
            return self.get_state(), 0.0, True
        for _ in range(self.sims_per_step):
            self._sim_step(actions)
            if self.crash:
                break
        self.done = self.crash or self.time_counter >= self.sims_per_step * (
            self.warmup_steps + self.horizon)
        if self.evaluate:
            rew = float(np.mean(self.vel)) if self.vel else float('nan')
        else:
            rew = self.desired_velocity(self.crash)
        return self.get_state(), rew, self.done

    def _sim_step(self, actions=None):
        dt = self.dt
        self.time_counter += 1
        n = len(self.pos)
        nv = [0.0] * n
        for s in range(n):
            vt = self.types[self.type_of[s]]
            z = self.noise_fn(self.time_counter - 1, s) if (self.noise_fn and vt.noise > 0) else 0.0
            a = C.get_action(self, s, vt, {}, dt, self.speed_limit, z)
            commanded = vt.controller not in (C.SIM, C.RL)
            if vt.controller == C.RL and actions and self.ids[s] in actions:
                raw = float(actions[self.ids[s]])
                if raw == raw:          # NaN = no command for this agent
                    a = raw             # multiagent/highway.py:86-95, not clipped
                    commanded = True
            self.accel[s] = a
            if commanded and vt.speed_mode:     # SUMO's speed mode on a commanded speed (restated)
                a = C.speed_mode_accel(self, s, vt, a, self.speed_limit)
            nv[s] = max(self.vel[s] + a * dt, 0.0)
        for s in range(n):
            if self.use_ballistic:
                self.pos[s] = self.pos[s] + 0.5 * (self.vel[s] + nv[s]) * dt
            else:
                self.pos[s] = self.pos[s] + nv[s] * dt
            self.vel[s] = nv[s]
        # arrivals
        keep = [s for s in range(n) if self.pos[s] < self.length]
        self.arrived += n - len(keep)
        self.pos = [self.pos[s] for s in keep]
        self.vel = [self.vel[s] for s in keep]
        self.type_of = [self.type_of[s] for s in keep]
        self.accel = [self.accel[s] for s in keep]
        self.arrived_ids = [self.ids[s] for s in range(n) if s not in set(keep)]
        self.ids = [self.ids[s] for s in keep]
        # inflows, in definition order
        for f, fl in enumerate(self.inflows):
            t = self.time_counter * dt
            end = fl.get('end', 0.0)
            prob = 0.0 if fl.get('vehs_per_hour', 0) > 0 else fl.get('probability', 0.0)
            if prob > 0:
                if t >= fl['begin'] and (end <= 0 or t < end):
                    r = philox4x32_10(
                        [self.time_counter, 0x40000000 | f, self.env_gid, self.env_gid >> 32],
                        [(self.seed & 0xFFFFFFFF) ^ ((self.episode * 0x9E3779B9) & 0xFFFFFFFF),
                         self.seed >> 32])
                    if r[0] * (1.0 / 4294967296.0) < prob * dt:
                        self.pending[f] += 1
                want = self.pending[f] > 0
            elif fl.get('vehs_per_hour', 0) > 0:
                due = fl['begin'] + self.spawned_by_flow[f] * (3600.0 / fl['vehs_per_hour'])
                want = t >= due and (end <= 0 or due < end)
            else:
                continue
            if want and len(self.pos) < self.capacity:
                nt = self.types[fl['type']]
                free = (not self.pos) or \
                    (self.pos[-1] - self.types[self.type_of[-1]].length >= nt.sim_min_gap)
                if free:
                    self.pos.append(0.0)
                    self.vel.append(float(fl['depart_speed']))
                    self.type_of.append(fl['type'])
                    self.ids.append(((f + 1) << 24) | self.spawned_by_flow[f])
                    self.accel.append(0.0)
                    self.spawned_by_flow[f] += 1
                    self.departed += 1
                    if prob > 0:
                        self.pending[f] -= 1
        self.update()
        self.crash = any(h < 0 for h in self.headway)

    def get_state(self):
        cap = self.capacity
        v = [x / self.speed_limit for x in self.vel] + [0.0] * (cap - len(self.vel))
        p = [x / self.length for x in self.pos] + [0.0] * (cap - len(self.pos))
        return np.array(v + p)

    def desired_velocity(self, fail=False):
        """flow/core/rewards.py:6-59 over the vehicles currently in the network."""
        vel = np.array(self.vel)
        n = len(vel)
        if n == 0 or fail or any(vel < -100):
            return 0.
        target = self.target_velocity
        max_cost = np.linalg.norm(np.array([target] * n))
        cost = np.linalg.norm(vel - target)
        return max(max_cost - cost, 0) / (max_cost + F32_EPS)

    # ---- multi-agent front: flow/envs/multiagent/highway.py --------------------
    def rl_slots(self):
        """RL vehicles present, head first (= the engine's agent rows)."""
        return [s for s in range(len(self.pos)) if self.types[self.type_of[s]].controller == C.RL]

    def ma_state(self):
        """get_state, highway.py:97-136: {id: 5 numbers}; note `lead_head` is the LEADER's
        headway (highway.py:118) and the follower's headway is the gap behind the agent."""
        max_speed, max_length = self.speed_limit, self.length
        obs = {}
        n = len(self.pos)
        for s in self.rl_slots():
            v = self.vel[s]
            if s == 0:
                lead_speed, lead_head = max_speed, max_length
            else:
                lead_speed, lead_head = self.vel[s - 1], self.headway[s - 1]
            if s + 1 >= n:
                follow_speed, follow_head = 0, max_length
            else:
                follow_speed, follow_head = self.vel[s + 1], self.headway[s + 1]
            obs[self.ids[s]] = np.array([v / max_speed, (lead_speed - v) / max_speed,
                                         lead_head / max_length, (v - follow_speed) / max_speed,
                                         follow_head / max_length])
        return obs

    def ma_reward(self, fail=False):
        """compute_reward, highway.py:138-172 (rl_actions not None)."""
        rewards = {}
        for s in self.rl_slots():
            v = self.vel[s]
            if self.evaluate:
                r = v
            elif fail:
                r = 0
            else:
                cost1 = self.desired_velocity(fail)
                cost2 = 0
                if s > 0 and v > 0:
                    t_headway = max(self.headway[s] / v, 0)
                    cost2 += min((t_headway - 1) / 1, 0)
                r = max(1.00 * cost1 + 0.10 * cost2, 0)
            rewards[self.ids[s]] = r
        return rewards
