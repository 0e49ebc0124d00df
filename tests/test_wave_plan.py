"""The host-side planning of a render call through the C ABI hook rt_plan_waves (no GPU): which rays of which passes a call
traces and how they are grouped into waves (cuda_raytracer_b200/csrc/device/wave_plan.cuh, launched by render_impl). The GPU
suite checks the IMAGES of a few such calls; this checks the plan itself over thousands of option sets."""
import random

import pytest

from cuda_raytracer_b200 import core
from cuda_raytracer_b200.multi_gpu import sample_range

PASS = 20  # MAX_RAYS_PER_PIXEL_PER_PASS, raytracing.cu:109


def spp_of(p, spp, pass_spp=PASS):
    return min(spp - p * pass_spp, pass_spp)


def expected_samples(spp, s0, n, pass_spp=PASS):
    """Independent restatement: the set of (pass, sample inside the pass) a sample range covers."""
    return {(s // pass_spp, s % pass_spp) for s in range(max(0, s0), min(spp, s0 + n))}


def covered_samples(waves, pixels, spp, pass_spp=PASS):
    """(pass, sample) pairs of a plan whose segments are whole passes or sample ranges; asserts no pair twice."""
    seen = set()
    for members, _ in waves:
        for p, offset, count, sub_first, sub_count in members:
            assert offset == 0
            k = sub_count if sub_count else spp_of(p, spp, pass_spp)
            assert count == pixels * k
            for s in range(sub_first, sub_first + k):
                assert (p, s) not in seen
                seen.add((p, s))
    return seen


def check_shape(waves, pixels, spp, max_group, pass_spp=PASS):
    for members, partial in waves:
        assert 1 <= len(members) <= max_group
        assert len({spp_of(m[0], spp, pass_spp) for m in members}) == 1          # a wave holds passes of one size
        assert partial in (-1, len(members) - 1)                                # the partial pass is last (or alone)
        for i, (p, offset, count, sub_first, sub_count) in enumerate(members):
            whole = offset == 0 and sub_count == 0 and count == pixels * spp_of(p, spp, pass_spp)
            assert whole == (i != partial)
            if i == partial:
                assert 0 < count < pixels * spp_of(p, spp, pass_spp) or len(members) == 1
                assert count <= pixels * spp_of(p, spp, pass_spp)


def test_known_plans():
    sizes = lambda waves: [(len(m), part) for m, part in waves]
    # teapot as shipped on 8 GPUs (1000 spp = 50 passes, 125 samples per rank), three pipes, waves of at most ceil(7 / 3) passes
    assert sizes(core.plan_waves(1000, 1000, 1000, 3, 3, sample_first=0, sample_count=125)) == [(3, 2), (3, -1), (1, -1)]
    assert sizes(core.plan_waves(1000, 1000, 1000, 3, 3, sample_first=125, sample_count=125)) == [(3, 2), (3, 2), (1, -1)]
    w = core.plan_waves(1000, 1000, 1000, 3, 3, sample_first=0, sample_count=125)
    assert [m[0] for m in w[0][0]] == [0, 1, 6] and w[0][0][2] == (6, 0, 5_000_000, 0, 5)   # samples [120, 125) ride with passes 0, 1
    # the whole frame with room for 16 passes per wave on two pipes: equal waves, a multiple of the pipes
    assert sizes(core.plan_waves(1000, 1000, 1000, 16, 2)) == [(13, -1), (13, -1), (12, -1), (12, -1)]
    # lamp 4K, 640 spp, room for 4: eight waves of four; 160 spp: 4 + 4
    assert sizes(core.plan_waves(3840, 2160, 640, 4, 2)) == [(4, -1)] * 8
    assert sizes(core.plan_waves(3840, 2160, 160, 4, 2)) == [(4, -1)] * 2
    # one pass at a time (the reference's loop): a partial pass cannot ride
    assert sizes(core.plan_waves(64, 48, 70, 1, 1, sample_first=0, sample_count=25)) == [(1, -1), (1, 0)]
    # the last, shorter pass never shares a wave with full ones; a partial piece of it rides nowhere
    w = core.plan_waves(64, 48, 70, 4, 2)
    assert sizes(w) == [(3, -1), (1, -1)] and w[1][0][0][0] == 3
    assert sizes(core.plan_waves(64, 48, 70, 4, 2, sample_first=25, sample_count=40)) == [(2, 1), (1, 0)]
    # nothing to do
    assert core.plan_waves(64, 48, 70, 4, 2, sample_first=70, sample_count=5) == []
    with pytest.raises(core.RtError):
        core.plan_waves(64, 48, 70, 4, 2, sample_first=-1, sample_count=5)
    with pytest.raises(core.RtError):
        core.plan_waves(64, 48, 70, 17, 2)


def test_sample_ranges_cover_exactly_their_samples():
    rng = random.Random(1)
    for _ in range(1500):
        w, h = rng.randint(1, 40), rng.randint(1, 40)
        pass_spp = rng.choice((20, 20, 7, 1))
        spp = rng.randint(1, 400 if pass_spp > 1 else 60)
        s0, n = rng.randint(0, spp), rng.randint(1, spp + 5)
        max_group, pipes = rng.randint(1, 16), rng.randint(1, 4)
        waves = core.plan_waves(w, h, spp, max_group, pipes, pass_spp=pass_spp, sample_first=s0, sample_count=n)
        check_shape(waves, w * h, spp, max_group, pass_spp)
        assert covered_samples(waves, w * h, spp, pass_spp) == expected_samples(spp, s0, n, pass_spp)


def test_ranks_cover_the_frame_once():
    rng = random.Random(2)
    for _ in range(300):
        spp, world = rng.randint(1, 1200), rng.randint(1, 9)
        w, h = rng.randint(1, 30), rng.randint(1, 30)
        union = set()
        for rank in range(world):
            first, count = sample_range(spp, rank, world)
            if count == 0:
                continue
            passes = -(-count // PASS) + 1
            pipes = 3 if passes <= 8 else 2
            max_group = max(1, min(16, -(-passes // pipes)))
            waves = core.plan_waves(w, h, spp, max_group, pipes, sample_first=first, sample_count=count)
            check_shape(waves, w * h, spp, max_group)
            mine = covered_samples(waves, w * h, spp)
            assert not (mine & union)
            union |= mine
        assert union == expected_samples(spp, 0, spp)


def test_ray_ranges_cover_exactly_their_rays():
    rng = random.Random(3)
    for _ in range(1500):
        w, h, spp = rng.randint(1, 30), rng.randint(1, 30), rng.randint(1, 300)
        pixels, total = w * h, w * h * spp
        a = rng.randint(0, total)
        n = rng.randint(1, total + 10)
        max_group, pipes = rng.randint(1, 16), rng.randint(1, 4)
        waves = core.plan_waves(w, h, spp, max_group, pipes, ray_first=a, ray_range=n)
        check_shape(waves, pixels, spp, max_group)
        got = []
        for members, _ in waves:
            for p, offset, count, sub_first, sub_count in members:
                assert sub_count == 0 and sub_first == 0
                start = p * PASS * pixels + offset  # passes before p are full
                got.append((start, start + count))
        got.sort()
        want_lo, want_hi = a, min(total, a + n)
        if want_hi <= want_lo:
            assert got == []
            continue
        assert got[0][0] == want_lo and got[-1][1] == want_hi
        for (l0, h0), (l1, h1) in zip(got, got[1:]):
            assert h0 == l1  # contiguous, no overlap


def test_pass_ranges_and_equal_waves():
    rng = random.Random(4)
    for _ in range(1500):
        w, h, spp = rng.randint(1, 30), rng.randint(1, 30), rng.randint(1, 2000)
        passes = -(-spp // PASS)
        first, count = rng.randint(0, passes), rng.randint(0, passes + 2)
        max_group, pipes = rng.randint(1, 16), rng.randint(1, 4)
        waves = core.plan_waves(w, h, spp, max_group, pipes, pass_first=first, pass_count=count)
        check_shape(waves, w * h, spp, max_group)
        want = list(range(first, min(passes, first + count) if count > 0 else passes))
        assert [m[0] for members, _ in waves for m in members] == want          # whole passes, in order
        assert all(part == -1 for _, part in waves)
        # full-size passes (all but a shorter last one): first fit when one round of waves is enough, equal waves otherwise
        full = [len(m) for m, _ in waves if spp_of(m[0][0], spp) == PASS]
        n_full = sum(full)
        if n_full and -(-n_full // max_group) > pipes:
            assert max(full) - min(full) <= 1
            least = -(-n_full // max_group)
            assert len(full) in (least, -(-least // pipes) * pipes)
        elif n_full:
            assert full == [max_group] * (n_full // max_group) + ([n_full % max_group] if n_full % max_group else [])
