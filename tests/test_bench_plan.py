"""CPU: the planning helpers of bench.py (how the driver's K steps become a >= 100 ms timed
region of whole fused launches, ring buffers larger than L2, episodes with resets between)."""
import bench


def test_reps_fill_the_target_and_make_whole_launches():
    # the driver's call: K = 20 on c2 (0.45 us per step, fuse 500)
    r = bench.round_reps(bench.plan_reps(20, 0.00045, 100.0), 20, 500)
    assert r * 20 * 0.00045 >= 100.0 and (r * 20) % 500 == 0
    assert bench.plan_reps(3000, 0.00045, 100.0) == 75
    assert bench.plan_reps(20, 0.0, 100.0) == 1
    assert bench.plan_reps(20, 1e-9, 100.0, max_total=1000) == 50      # capped
    assert bench.round_reps(7, 3000, 500) == 7 and bench.round_reps(1, 20, 100) == 5


def test_ring_buffers_exceed_l2_and_hold_whole_launches():
    for fuse, act_b, out_b in ((500, 16384, 69632), (100, 262144, 1114112), (100, 0, 1310720),
                               (100, 49152, 1632256)):
        nb = bench.plan_ring(fuse, 20, act_b, out_b)
        assert nb % fuse == 0 and nb >= fuse
        assert nb * (act_b + out_b) >= 2 * bench.L2_BYTES
        if act_b:
            assert nb * act_b > bench.L2_BYTES
    assert bench.plan_ring(100, 20, 1 << 20, 1 << 30, cap_bytes=4e9) == 100   # memory cap wins


def test_episodes_cover_the_total_in_calls_of_at_most_chunk():
    eps = bench.plan_episodes(7000, 3000, 500)
    assert [n for n, _ in eps] == [3000, 3000, 1000]
    assert all(sum(k for _, k in calls) == n and max(k for _, k in calls) <= 500 for n, calls in eps)
    assert bench.plan_episodes(20, 3000, 500) == [(20, [(0, 20)])]


def test_default_arguments_finish_quickly():
    a = bench.parse([])
    assert a.gpus == 1 and a.impl == 'b200' and a.workload == 'c2' and a.target_ms == 100.0
