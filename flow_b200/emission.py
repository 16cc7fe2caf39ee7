"""Emission-trace writer: the on-disk format the
reference's visualisers and `Experiment.run(convert_to_csv=True)` consume.

In the reference SUMO writes `<name>-emission.xml` and `emission_to_csv`
(flow/core/util.py:36-99) turns it into a CSV with the columns below, sorted by
vehicle id; `flow/visualize/time_space_diagram.py:39-71` reads `time, id,
edge_id, relative_position, speed, lane_number` from it. Here the rows come from
the engine's state after every step. The pollutant columns (CO, CO2, NOx, fuel,
HC, PMx, electricity, noise) are SUMO's HBEFA emission model — third-party, out
of scope — and are written as 0 so that the header keeps the reference's shape.
"""
import csv
import math
import os

import numpy as np

# header of the reference's own fixture tests/fast_tests/test_files/ring_230_emission.csv
COLUMNS = ['speed', 'CO', 'electricity', 'x', 'NOx', 'id', 'fuel', 'angle', 'time', 'edge_id',
           'eclass', 'route', 'waiting', 'CO2', 'lane_number', 'PMx', 'type', 'noise',
           'relative_position', 'HC', 'y']
_ZERO = ('CO', 'electricity', 'NOx', 'fuel', 'CO2', 'PMx', 'noise', 'HC')


class EmissionRecorder:
    """Collects one row per vehicle per step from a `flow_api` env (env index
    `env.view_env` of a batch) and writes the CSV."""

    def __init__(self, env, decimals=2):
        self.env, self.decimals, self.rows = env, decimals, []
        self._waiting = {}

    def _xy(self, x_loop):
        """Cartesian position on the reference's ring geometry (flow/networks/ring.py:70-94:
        nodes at radius L / 2 pi, `bottom` starts at (0, -r)); highway: along the x axis."""
        env = self.env
        if env.network.TOPOLOGY != 0:
            return x_loop, 0.0, 90.0
        C = env.k.network.length()
        r = env.k.network.non_internal_length() / (2 * math.pi)
        th = 2 * math.pi * x_loop / C - math.pi / 2
        return r * math.cos(th), r * math.sin(th), (90.0 - math.degrees(th + math.pi / 2)) % 360.0

    def record(self):
        env, kv = self.env, self.env.k.vehicle
        # SUMO's first reported time step (0.1) is the state right after insertion: Env.reset
        # performs one simulation step before anything moves (flow/envs/base.py:497-500)
        t = (env.time_counter + 1) * env.sim_step
        ids = kv.get_ids()
        speeds, xs, lanes = kv.get_speed(ids), kv.get_x_by_id(ids), kv.get_lane(ids)
        for vid, v, x, ln in zip(ids, speeds, xs, lanes):
            edge, start = env._edge_of(x)
            px, py, ang = self._xy(x)
            w = self._waiting.get(vid, 0.0)
            w = w + env.sim_step if v < 0.1 else 0.0      # SUMO: time spent below 0.1 m/s
            self._waiting[vid] = w
            row = dict.fromkeys(_ZERO, 0.0)
            row.update(speed=round(v, self.decimals), x=round(px, self.decimals),
                       y=round(py, self.decimals), id=vid, angle=round(ang, self.decimals),
                       time=round(t, 6), edge_id=edge, eclass='HBEFA3/PC_G_EU4',
                       route='route' + edge.lstrip(':').rsplit('_', 1)[0] if edge.startswith(':')
                       else 'route' + edge, waiting=round(w, self.decimals),
                       lane_number=int(ln if ln is not None and ln >= 0 else 0),
                       type=kv.get_type(vid), relative_position=round(x - start, self.decimals))
            self.rows.append(row)

    def write_csv(self, path):
        d = os.path.dirname(path)
        if d:
            os.makedirs(d, exist_ok=True)
        rows = sorted(self.rows, key=lambda r: r['id'])      # emission_to_csv sorts by id
        with open(path, 'w', newline='') as f:
            w = csv.DictWriter(f, COLUMNS)
            w.writeheader()
            w.writerows(rows)
        return path


def rollout_to_csv(env, n_steps, path, policy=None):
    """Reset `env`, run `n_steps` steps (policy(obs) -> actions, default None =
    no RL command) and write the trace. Returns the list of rewards."""
    rec = EmissionRecorder(env)
    obs = env.reset()
    rec.record()
    rewards = []
    for _ in range(n_steps):
        obs, r, done, _ = env.step(policy(obs) if policy else None)
        rec.record()
        rewards.append(r)
        if np.all(done if not isinstance(done, dict) else done['__all__']):
            break
    rec.write_csv(path)
    return rewards
