"""CPU, build container only: the oracle against the LIVE reference code
(/root/reference, imported through tests/golden/ref_shim). Skipped on machines
without the reference checkout (the GPU box)."""
import numpy as np
import pytest

from tests.golden import ref_shim

pytestmark = pytest.mark.skipif(not ref_shim.available(), reason='no /root/reference here')


def test_c2_random_actions_bit_exact_vs_reference_code():
    from tests.golden.ref_harness import RefRing, ref_ctrl
    from tests import scenarios as S
    from oracle import sim as OS
    ref = RefRing([('human', ref_ctrl.IDMController, {}, {'min_gap': 0}, 21),
                   ('rl', 'rl', {}, {}, 1)], 247, 'wave_po',
                  init_kw=dict(bunching=50, min_gap=0),
                  env_additional={'max_accel': 1, 'max_decel': 1, 'ring_length': [220, 270]})
    o = S.make_oracle(S.c2_groups(), 247, env_kind=OS.ENV_WAVE_ATTENUATION_PO, bunching=50.0,
                      warmup_steps=0, act_lo=-1, act_hi=1, obs_max_length=270.0)
    rng = np.random.default_rng(5)
    for s in range(400):
        a = None if s < 100 else rng.uniform(-1.2, 1.2, 1)
        ob1, r1, crash = ref.step(a)
        ob2, r2, done = o.step(a)
        assert crash == done
        assert np.array_equal(np.array(ref.pos), np.array(o.pos))
        assert np.array_equal(np.array(ref.vel), np.array(o.vel))
        assert np.array_equal(ob1, ob2) and r1 == r2
        if crash:   # the oracle freezes a finished env; the harness does not
            break
