"""Host-side logic of bench.py that the driver's ratios rest on (no GPU): the config object is the same in both arms, the
reference subprocesses get every host core even under torchrun, the reference-GPU rate is a slope whose one-sided outliers
are measured again, and the roofline block only uses an ncu capture of the running build."""
import json
import os
import sys
import types

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def _args(**kw):
    a = types.SimpleNamespace(workload="lamp_4k", sort=0, gpus=1)
    a.__dict__.update(kw)
    return a


def test_config_is_the_same_object_in_both_arms():
    for wl in bench.WORKLOADS:
        for n in (1, 8):
            c1, c2 = bench.config_for(_args(workload=wl, gpus=n)), bench.config_for(_args(workload=wl, gpus=n))
            assert c1 == c2 and json.dumps(c1) == json.dumps(c2)
            assert set(c1) == {"workload", "passes", "pass_spp", "sort", "parallelism", "l2"}
            assert f"x{n}" in c1["parallelism"]
    scene, w, h, spp, bounces = bench.WORKLOADS["lamp_4k"]
    assert (scene, w, h, spp, bounces) == ("lamp", 3840, 2160, 640, 10)  # BASELINE.json configs[4]
    assert bench.config_for(_args())["passes"] == 32


def test_reference_subprocess_env_overrides_torchrun(monkeypatch):
    monkeypatch.setenv("OMP_NUM_THREADS", "1")  # what torch.distributed.run exports to its workers
    env = bench.ref_env()
    assert int(env["OMP_NUM_THREADS"]) == bench.host_threads() >= 1
    assert "RT_LIBRTCORE" not in env or env["RT_LIBRTCORE"] == os.environ.get("RT_LIBRTCORE")


def test_slope_fit_exact_line():
    xs = (20, 40, 60, 80, 100)
    ys = [0.3 + 0.008 * x for x in xs]
    slope, icpt, res, se = bench.slope_fit(xs, ys)
    assert slope == pytest.approx(0.008) and icpt == pytest.approx(0.3)
    assert max(abs(r) for r in res) < 1e-12 and se < 1e-12


def test_one_sided_outlier_is_measured_again():
    xs = (20, 40, 60, 80, 100)
    # the r2f run of the reference GPU build (no_sort): the 60 spp point sat 0.13 s above the line, spread 14.7 %
    ys = [0.475127, 0.672374, 0.994364, 1.00934, 1.1883]
    asked = []

    def again(x):
        asked.append(x)
        return [0.84, 0.99, 0.85]

    slope, icpt, res, se, re_x = bench.fit_with_remeasure(xs, ys, again)
    assert asked == [60] and re_x == [60] and ys[2] == 0.84
    assert se / slope < 0.05
    # the sort-mode series of the next run: least squares alone is pulled up by the bad point (residual 0.114 against a median
    # of 0.043) and would not have flagged it
    ys = [0.503434, 0.649788, 0.780731, 1.1003, 1.08669]
    *_, se2, re_x = bench.fit_with_remeasure(xs, ys, lambda x: [0.94, 0.95])
    assert re_x == [80] and ys[3] == 0.94
    # a clean series asks for nothing
    ys = [0.3 + 0.008 * x + d for x, d in zip(xs, (0.001, -0.002, 0.0015, -0.001, 0.0005))]
    *_, re_x = bench.fit_with_remeasure(xs, ys, lambda x: pytest.fail("should not re-measure"))
    assert re_x == []
    # a point BELOW the line is not an outlier of this timer (it is the better measurement): the points above it are not
    # re-measured unless they stand out by themselves
    ys = [0.3 + 0.008 * x for x in xs]
    ys[2] -= 0.2
    calls = []
    bench.fit_with_remeasure(xs, ys, lambda x: calls.append(x) or [ys[xs.index(x)]])
    assert 60 not in calls


def test_agreeing_runs_keeps_sampling_a_noisy_point():
    it = iter([1.38, 1.21, 1.30, 1.135, 1.14, 9.0])
    runs = bench.agreeing_runs(lambda: next(it))
    assert runs == [1.38, 1.21, 1.30, 1.135, 1.14] and min(runs) == 1.135   # stopped once 1.135 and 1.14 agreed
    it = iter([0.50, 0.505, 0.51, 9.0])
    assert bench.agreeing_runs(lambda: next(it)) == [0.50, 0.505, 0.51]     # quiet point: three runs
    it = iter([None, 0.5, None, 0.7, 0.9, 1.1, 1.3, 1.5, 1.7, 1.9, 2.1])
    assert len(bench.agreeing_runs(lambda: next(it))) <= 5                  # never agrees: bounded


def test_ncu_summary_of_the_running_build_is_preferred(tmp_path, monkeypatch):
    prof = tmp_path / "profiles"
    prof.mkdir()
    for name, build, nodes in (("r2a_f32", "aaaa", "f32"), ("r2b_f32", "bbbb", "f32"), ("r2c_q16", "bbbb", "q16")):
        json.dump({"build_id": build, "workload": f"lamp nodes={nodes}", "trace_rays_kernel": {}}, open(prof / f"{name}_traffic.json", "w"))
    json.dump({"no": "kernels"}, open(prof / "r2z_traffic.json", "w"))
    monkeypatch.setattr(bench, "ROOT", str(tmp_path))
    assert bench.newest_ncu_summary("aaaa")["build_id"] == "aaaa"          # an older capture of THIS build wins
    assert bench.newest_ncu_summary("bbbb")["_file"].endswith("r2b_f32_traffic.json")  # the default node format's capture
    stale = bench.newest_ncu_summary("cccc")
    assert stale["build_id"] == "bbbb" and "f32" in stale["workload"]      # reported as stale by main(): same_build false
    # a host-only change: new source hash, same SASS of the captured kernels -> the capture of those kernels still counts
    json.dump({"build_id": "dddd", "kernel_id": "sass-1", "workload": "lamp nodes=f32", "trace_rays_kernel": {}}, open(prof / "r2a_zz_traffic.json", "w"))
    assert bench.newest_ncu_summary("eeee", "sass-1")["build_id"] == "dddd"
    assert bench.newest_ncu_summary("eeee", "sass-2")["build_id"] == "bbbb"
    assert bench.newest_ncu_summary("eeee", None)["build_id"] == "bbbb"


def test_committed_capture_matches_the_committed_library_when_present():
    """The roofline keys of the driver's bench line need a capture of the shipped build: if librtcore.so has been built here,
    its id must be the one of the newest committed f32 capture (rebuild + re-capture otherwise)."""
    so = os.path.join(ROOT, "cuda_raytracer_b200", "librtcore.so")
    if not os.path.exists(so):
        pytest.skip("librtcore.so not built")
    from cuda_raytracer_b200.core import build_id, kernel_id
    d = bench.newest_ncu_summary(build_id(), kernel_id())
    assert d is not None
    if d["build_id"] != build_id() and (kernel_id() is None or d.get("kernel_id") != kernel_id()):
        pytest.skip(f"capture {d['_file']} is of build {d['build_id']} / kernels {d.get('kernel_id')}, library is {build_id()} / {kernel_id()}: "
                    "bench.py will leave the per-level keys out")
