"""CPU oracle — TEST INFRASTRUCTURE ONLY (never imported by flow_b200/).

Scalar fp64 restatement of the reference's acceleration controllers and
fail-safes. Each function cites the reference lines it follows
(paths relative to the parthjaggi/flow checkout). `k` is the oracle's
vehicle-kernel view (oracle/sim.py: RingSim / HighwaySim) which offers the
same getters as flow/core/kernel/vehicle/traci.py:541-602 but keyed by integer
vehicle index; "no leader" is index -1 (reference: None / '').

Parity status: pinned — tests/test_oracle_vs_reference.py drives the
reference's own controller classes (imported from /root/reference when it is
present) on identical states, and tests/golden/*.json holds the vectors of
tests/fast_tests/test_controllers.py plus reference-generated trajectories.
"""
import math

SIM, IDM, OVM, LINEAR_OVM, BCM, CFM, LAC, GIPPS, FOLLOWER_STOPPER, \
    NONLOCAL_FOLLOWER_STOPPER, PI_SATURATION, RL = range(12)
FS_NONE, FS_INSTANTANEOUS, FS_SAFE_VELOCITY = 0, 1, 2
NO_LEADER_SPEED = -1001.0   # get_speed(None): vehicle/traci.py:541-545


class VType:
    """One row of the per-type table (mirrors fs_vtype in include/flowstep.h)."""

    def __init__(self, controller=IDM, p=(), fail_safe=FS_NONE, noise=0.0,
                 delay=0.0, max_accel=2.6, max_decel=4.5, length=5.0,
                 sim_tau=1.0, sim_min_gap=2.5, sim_max_speed=30.0,
                 lc_controller=0, speed_mode=0):
        self.controller = controller
        self.p = list(p) + [0.0] * (8 - len(p))
        self.fail_safe = fail_safe
        self.noise = noise
        self.delay = delay
        self.max_accel = max_accel
        self.max_decel = abs(max_decel)
        self.length = length
        self.sim_tau = sim_tau
        self.sim_min_gap = sim_min_gap
        self.sim_max_speed = sim_max_speed
        self.lc_controller = lc_controller
        self.speed_mode = int(speed_mode)   # SUMO speed-mode bits (flow/core/params.py:12-18)


def idm_params(v0=30, T=1, a=1, b=1.5, delta=4, s0=2):
    """Defaults: car_following_models.py:437-447."""
    return [v0, T, a, b, delta, s0]


def _speed(k, j):
    return NO_LEADER_SPEED if j < 0 else k.vel[j]


def idm_accel(v, v_lead, h, has_leader, v0, T, a, b, delta, s0):
    """car_following_models.py:464-482."""
    if abs(h) < 1e-3:
        h = 1e-3
    if not has_leader:
        s_star = 0.0
    else:
        s_star = s0 + max(0.0, v * T + v * (v - v_lead) / (2 * math.sqrt(a * b)))
    return a * (1 - (v / v0) ** delta - (s_star / h) ** 2)


def get_accel(k, i, vt, st, dt, speed_limit):
    """Dispatch to the controller law of vehicle i. `st` is the per-vehicle
    controller state dict (LAC `a`, PISaturation history / v_cmd)."""
    c = vt.controller
    p = vt.p
    v = k.vel[i]
    lead = k.leader[i]
    h = k.headway[i]
    if c == IDM:
        return idm_accel(v, _speed(k, lead), h, lead >= 0, *p[:6])
    if c == SIM or c == RL:
        # Restated simulator default (parity unpinned): IDM
        # with the type's SumoCarFollowingParams (flow/core/params.py:842-846).
        v0 = min(vt.sim_max_speed, speed_limit)
        return idm_accel(v, _speed(k, lead), h, lead >= 0, v0, vt.sim_tau,
                         vt.max_accel, vt.max_decel, 4, vt.sim_min_gap)
    if c == OVM:
        # car_following_models.py:308-328
        alpha, beta, h_st, h_go, v_max = p[:5]
        if lead < 0:
            return vt.max_accel
        h_dot = k.vel[lead] - v
        if h <= h_st:
            v_h = 0
        elif h_st < h < h_go:
            v_h = v_max / 2 * (1 - math.cos(math.pi * (h - h_st) / (h_go - h_st)))
        else:
            v_h = v_max
        return alpha * (v_h - v) + beta * h_dot
    if c == LINEAR_OVM:
        # car_following_models.py:383-397
        v_max, adaptation, h_st = p[:3]
        alpha = 1.689
        if h < h_st:
            v_h = 0
        elif h_st <= h <= h_st + v_max / alpha:
            v_h = alpha * (h - h_st)
        else:
            v_h = v_max
        return (v_h - v) / adaptation
    if c == CFM:
        # car_following_models.py:76-88
        k_d, k_v, k_c, d_des, v_des = p[:5]
        if lead < 0:
            return vt.max_accel
        return k_d * (h - d_des) + k_v * (k.vel[lead] - v) + k_c * (v_des - v)
    if c == BCM:
        # car_following_models.py:152-176
        k_d, k_v, k_c, d_des, v_des = p[:5]
        if lead < 0:
            return vt.max_accel
        trail = k.follower[i]
        trail_vel = _speed(k, trail)
        footway = k.headway[trail] if trail >= 0 else -1001.0
        return k_d * (h - footway) + \
            k_v * ((k.vel[lead] - v) - (v - trail_vel)) + k_c * (v_des - v)
    if c == LAC:
        # car_following_models.py:232-245
        k_1, k_2, hh, tau = p[:4]
        a_prev = st.setdefault('a', p[4])
        ex = h - vt.length - hh * v
        ev = _speed(k, lead) - v
        u = k_1 * ex + k_2 * ev
        a_dot = -(a_prev / tau) + (u / tau)
        st['a'] = a_dot * dt + a_prev
        return st['a']
    if c == GIPPS:
        # car_following_models.py:567-582
        v0, acc, b, b_l, s0, tau = p[:6]
        v_l = _speed(k, lead)
        v_acc = v + (2.5 * acc * tau * (1 - (v / v0)) * math.sqrt(0.025 + (v / v0)))
        inner = ((tau ** 2) * (b ** 2)) - (b * ((2 * (h - s0)) - (tau * v) - ((v_l ** 2) / b_l)))
        # the reference takes np.sqrt: a negative argument gives NaN (a RuntimeWarning, no
        # exception), and Python's min() then never selects it because every comparison with
        # NaN is false and v_acc comes first
        v_safe = (tau * b) + (math.sqrt(inner) if inner >= 0 else float('nan'))
        v_next = min(v_acc, v_safe, v0)
        return (v_next - v) / dt
    if c in (FOLLOWER_STOPPER, NONLOCAL_FOLLOWER_STOPPER):
        # velocity_controllers.py:75-116 / :122-164
        if c == NONLOCAL_FOLLOWER_STOPPER:
            v_des = k.mean_speed()
        else:
            v_des = p[0]
        lead_vel = _speed(k, lead)
        if lead < 0:
            v_cmd = v_des
        else:
            dx = h
            dv_minus = min(lead_vel - v, 0)
            dx_1 = 4.5 + 1 / (2 * 1.5) * dv_minus ** 2
            dx_2 = 5.25 + 1 / (2 * 1.0) * dv_minus ** 2
            dx_3 = 6.0 + 1 / (2 * 0.5) * dv_minus ** 2
            vv = min(max(lead_vel, 0), v_des)
            if dx <= dx_1:
                v_cmd = 0
            elif dx <= dx_2:
                v_cmd = vv * (dx - dx_1) / (dx_2 - dx_1)
            elif dx <= dx_3:
                v_cmd = vv + (v_des - v) * (dx - dx_2) / (dx_3 - dx_2)
            else:
                v_cmd = v_des
        return (v_cmd - v) / dt
    if c == PI_SATURATION:
        # velocity_controllers.py:208-240
        lead_vel = _speed(k, lead)
        dx = h
        dv = lead_vel - v
        dx_s = max(2 * dv, 4)
        hist = st.setdefault('v_history', [])
        hist.append(v)
        if len(hist) == int(38 / dt):
            del hist[0]
        v_des = sum(hist) / len(hist)
        v_target = v_des + 1 * min(max((dx - 7) / (30 - 7), 0), 1)
        alpha = min(max((dx - dx_s) / 2, 0), 1)
        beta = 1 - 0.5 * alpha
        v_cmd = beta * (alpha * v_target + (1 - alpha) * lead_vel) \
            + (1 - beta) * st.get('v_cmd', 0.0)
        st['v_cmd'] = v_cmd
        return min((v_cmd - v) / dt, vt.max_accel)
    raise ValueError('unknown controller %r' % c)


def safe_action_instantaneous(k, i, action, dt):
    """base_controller.py:120-169."""
    if k.num_vehicles == 1:
        return action
    if k.leader[i] < 0:
        return action
    this_vel = k.vel[i]
    next_vel = this_vel + action * dt
    h = k.headway[i]
    if next_vel > 0:
        if h < dt * next_vel + this_vel * 1e-3 + 0.5 * this_vel * dt:
            return -this_vel / dt
        return action
    return action


def safe_velocity_action(k, i, action, dt, delay):
    """base_controller.py:171-236."""
    if k.num_vehicles == 1:
        return action
    lead_vel = _speed(k, k.leader[i])
    this_vel = k.vel[i]
    h = k.headway[i]
    dv = lead_vel - this_vel
    v_safe = 2 * h / dt + dv - this_vel * (2 * delay)
    if this_vel + action * dt > v_safe:
        if v_safe > 0:
            return (v_safe - this_vel) / dt
        return -this_vel / dt
    return action


def sim_model_accel(k, i, vt, speed_limit):
    """Acceleration of the restated simulator model for vehicle i: IDM fed with the type's
    SumoCarFollowingParams (carFollowModel "IDM" is the default of
    flow/core/params.py:833)."""
    lead = k.leader[i]
    v0 = min(vt.sim_max_speed, speed_limit)
    return idm_accel(k.vel[i], _speed(k, lead), k.headway[i], lead >= 0, v0, vt.sim_tau,
                     vt.max_accel, vt.max_decel, 4, vt.sim_min_gap)


def speed_mode_accel(k, i, vt, accel, speed_limit):
    """What SUMO makes of a speed commanded through slowDown (apply_acceleration,
    flow/core/kernel/vehicle/traci.py:952-963) under the vehicle's speed mode
    (setSpeedMode at departure, :343-347; modes flow/core/params.py:12-18). RESTATED SUMO,
    "parity unpinned": bit 0 (regard safe speed) caps the commanded speed by the car-following
    model's own next speed - here the restated simulator model; bit 1 caps the acceleration at
    the type's `accel`, bit 2 floors it at -`decel`. Expressed on accelerations:
    v' = max(v + a dt, 0) with a = this function's result equals min / max of the speeds."""
    m = vt.speed_mode
    if m & 1:
        accel = min(accel, sim_model_accel(k, i, vt, speed_limit))
    if m & 2:
        accel = min(accel, vt.max_accel)
    if m & 4:
        accel = max(accel, -vt.max_decel)
    return accel


def get_action(k, i, vt, st, dt, speed_limit, noise_sample=0.0):
    """base_controller.py:70-118 (guards for '' / ':' edges never fire on the
    restated continuous ring/highway). Order: accel -> noise -> fail-safe."""
    accel = get_accel(k, i, vt, st, dt, speed_limit)
    if vt.controller in (SIM, RL):
        # get_accel -> None in the reference (car_following_models.py:485-497):
        # the simulator's model drives; no Flow noise / fail-safe is applied.
        return accel
    if vt.noise > 0:
        accel += vt.noise * noise_sample
    if vt.fail_safe == FS_INSTANTANEOUS:
        accel = safe_action_instantaneous(k, i, accel, dt)
    elif vt.fail_safe == FS_SAFE_VELOCITY:
        accel = safe_velocity_action(k, i, accel, dt, vt.delay)
    return accel
