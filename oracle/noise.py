"""TEST INFRASTRUCTURE ONLY (see oracle/sim.py): counter-exact restatement of the Gaussian
noise stream of the c2 / c5 kernel (flow_b200/csrc/fs_ring_hot.cuh: PhiloxCK, normal4_ck and the
counter layout of ring1_hot_kernel), so that the TIMED instantiation (idm noise 0.2) can be
stepped side by side with the oracle (RingSim(noise_fn=...)) instead of only statistically.

The reference draws `np.random.normal(0, noise)` from numpy's global MT19937 inside
BaseController.get_action (flow/controllers/base_controller.py:109-110, noise BEFORE the
fail-safe); no counter-based generator can reproduce that stream, so what is pinned here is
the engine's own stream: Philox4x32-10 with the fixed key (0xA4093822, 0x299F31D0), counter
words (c0, c1, c2, c3) = (step block, episode * 256 + vehicle block, global env id low,
global env id high ^ seed low ^ seed high * 0x9E3779B9), Box-Muller on both word pairs.

Which counter a (vehicle, step) pair reads depends on the thread shape the engine picked for
the batch (threads per env G, vehicles per thread VPT = ceil(22 / G)):
  * VPT <= 4  (G = 8, 16, 32): a vehicle consumes the four normals of ONE call over four
    consecutive steps: c0 = tc >> 2, c1 = episode * 256 + vehicle, component tc & 3;
  * VPT > 4   (G = 1, 2, 4): a thread draws for its vehicles four at a time every step:
    c0 = tc, c1 = (episode * 256 + base + 4 * (i >> 2)) | 0x80000000 with base = first vehicle
    of the thread and i = vehicle - base, component i & 3.
tc = the env's time counter BEFORE the step (== RingSim.step_index), episode = number of
resets the env has seen (the engine's counter starts at 0 and fs_reset increments it).
"""
import math

from oracle.highway import philox4x32_10

KEY = (0xA4093822, 0x299F31D0)
V_HOT = 22


def vpt_of_shape(g):
    """vehicles per thread of ring1_hot_kernel's shapes (launch_hot22)."""
    return {1: 22, 2: 11, 4: 6, 8: 3, 16: 2, 32: 1}[g]


def auto_shape(n_envs, f64=False):
    """threads per env the engine picks when FS_RING_SHAPE is unset (flowstep.cu run_steps)."""
    g = 1
    while g < 32 and n_envs * g < 32768:
        g <<= 1
    if f64 and g == 1:
        g = 2
    return g


def normal4_ck(c0, c1, c2, c3):
    """Four standard normals of one call, double-precision Box-Muller as in normal4_ck<double>."""
    r = philox4x32_10([c0, c1, c2, c3], KEY)
    k = 1.0 / 4294967296.0
    ra = math.sqrt(-2.0 * math.log((r[0] + 1.0) * k))
    rb = math.sqrt(-2.0 * math.log((r[2] + 1.0) * k))
    a0, a1 = 6.283185307179586 * (r[1] * k), 6.283185307179586 * (r[3] * k)
    return [ra * math.cos(a0), ra * math.sin(a0), rb * math.cos(a1), rb * math.sin(a1)]


class HotNoise:
    """noise_fn(step, vehicle) of RingSim for one env of the hot kernel."""

    def __init__(self, seed, env_gid, episode, shape):
        self.vpt = vpt_of_shape(shape)
        self.episode = int(episode)
        self.c2 = env_gid & 0xFFFFFFFF
        self.c3 = ((env_gid >> 32) ^ (seed & 0xFFFFFFFF) ^
                   (((seed >> 32) * 0x9E3779B9) & 0xFFFFFFFF)) & 0xFFFFFFFF
        self._memo = {}

    def _call(self, c0, c1):
        key = (c0, c1)
        if key not in self._memo:
            if len(self._memo) > 4096:
                self._memo.clear()
            self._memo[key] = normal4_ck(c0 & 0xFFFFFFFF, c1 & 0xFFFFFFFF, self.c2, self.c3)
        return self._memo[key]

    def __call__(self, step, veh):
        if self.vpt <= 4:
            return self._call(step >> 2, self.episode * 256 + veh)[step & 3]
        base = (veh // self.vpt) * self.vpt
        i = veh - base
        return self._call(step, (self.episode * 256 + base + 4 * (i >> 2)) | 0x80000000)[i & 3]
