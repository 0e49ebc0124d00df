#!/usr/bin/env python3
"""bench.py — headline metric of the render path: Mrays/s (all bounces) on lamp.scene, ms/frame.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload lamp_4k|teapot_1k|...] [--impl reference]
  torchrun ... bench.py --gpus N ...        (one rank per GPU, NCCL; launched by the driver for N > 1)

A "step" is one whole frame of the workload: every pass of every bounce (generate -> closest hit -> shade -> compact
[-> reorder]) plus, for N > 1, the NCCL reduce of the float accumulation buffer to rank 0. Every pixel's SAMPLES are
split into equal shares across the ranks (rt_render_opts.sample_first / sample_count: whole passes when the pass count
divides, a few samples of every pixel of one more pass otherwise; strong scaling: the frame is fixed). `value` = live ray-bounces
of the whole frame / device time (CUDA events on the render stream, max over ranks). `e2e` is the same frame through the
reference-facing call with HOST buffers: rt_scene_upload (H2D of the scene) + render + D2H of the sums, every step.

Only the `cpu_baseline` leg and `--impl reference` execute anything under oracle/ (the reference's own CPU path); the
reference arm loads nothing of the product.
"""
import argparse
import glob
import json
import os
import re
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
SCENES = os.path.join(ROOT, "scenes")
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "raytracing_ref")

# name -> (scene, width, height, spp, bounces)   BASELINE.md §3 configs
WORKLOADS = {
    "lamp_4k": ("lamp", 3840, 2160, 640, 10),      # BASELINE.json configs[4]: "lamp.scene at 4K and high rays_per_pixel"
    "lamp_1k": ("lamp", 1000, 1000, 1000, 10),     # lamp.scene as shipped
    "teapot_1k": ("teapot", 1000, 1000, 1000, 10),  # BASELINE.json configs[3]: teapot.scene as shipped, 1/2/4/8 GPUs
    "cornell_1k": ("cornell", 1000, 1000, 1000, 10),
    "cornell_plus_1k": ("cornell_plus", 1000, 1000, 1000, 10),
    "spheres_1k": ("spheres", 1000, 1000, 1000, 10),
    "glass_teapot_1k": ("glass_teapot", 1000, 1000, 1000, 10),
}
METRIC = "Mrays/s (all bounces) on lamp.scene; ms/frame"
PASS_SPP = 20  # MAX_RAYS_PER_PIXEL_PER_PASS, raytracing.cu:109


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=3)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--workload", default="lamp_4k", choices=sorted(WORKLOADS))
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--sort", type=int, default=0, help="1 = reorder live rays between bounces")
    p.add_argument("--nodes", default=None, choices=["f32", "q16", "w4"], help="node format of the tracer (default: the library's)")
    p.add_argument("--pass-group", type=int, default=0, help="passes traced together as one wave; 0 = auto, 1 = one pass per wave")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-ref-gpu", action="store_true", help="skip timing the reference's sm_100a GPU binary")
    p.add_argument("--no-extras", action="store_true", help="skip the extra keys (device BVH build, spatial-split tree)")
    return p.parse_args()


def config_for(args):
    """The `config` object of the JSON line: the same keys and values in both arms (the driver compares them)."""
    scene, w, h, spp, bounces = WORKLOADS[args.workload]
    note = {"lamp": " (lamp-subset geometry: 48 of the 50 meshes are in the checkout)", "teapot": " (synthetic env map: the asset is not in the checkout)",
            "glass_teapot": " (synthetic env map: the asset is not in the checkout)"}.get(scene, "")
    return {"workload": f"{scene}.scene {w}x{h} {spp} spp {bounces} bounces{note}",
            "passes": (spp + PASS_SPP - 1) // PASS_SPP, "pass_spp": PASS_SPP, "sort": int(bool(args.sort)),
            "parallelism": f"per-pixel sample split x{args.gpus} + NCCL reduce",
            "l2": "inputs larger than L2: ray queues are %.1f GB per pass; the packed scene is L2-resident by nature" % (w * h * PASS_SPP * 40 * 2 / 1e9)}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu_index, self.rows, self.proc, self.first = str(gpu_index), [], None, 0

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            f = [x.strip() for x in line.split(",")]
            if len(f) >= 8 and f[0] == self.gpu_index:
                self.rows.append(f)

    def mark(self):
        """Start of the timed region: rows sampled before this point (warm-up) are not reported."""
        self.first = max(0, len(self.rows) - 1)

    def stop(self):
        if self.proc:
            self.proc.terminate()
        self.rows = self.rows[self.first:]
        sm = sorted(float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(r[4 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None,
                "sm_max_mhz": max((float(r[2]) for r in self.rows if r[2].replace(".", "").isdigit()), default=None),
                "reasons": reasons, "samples": len(self.rows)}


def visible_gpu_index(local_rank):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        parts = [v for v in vis.split(",") if v.strip() != ""]
        if local_rank < len(parts) and parts[local_rank].strip().isdigit():
            return int(parts[local_rank])
    return local_rank


# ---- reference arm / cpu_baseline: the reference's own CPU path (oracle/_ref/raytracing_ref cpu no_gpu) -------------
def _scene_file_with_image(scene, w, h, spp, bounces):
    src = open(os.path.join(SCENES, scene + ".scene")).read()
    exposure = re.search(r"(?m)^image\s+\S+\s+\S+\s+\S+\s+\S+\s+(\S+)", src)
    line = f"image {w} {h} {spp} {bounces} {exposure.group(1) if exposure else 1}"
    f = tempfile.NamedTemporaryFile("w", suffix=".scene", delete=False)
    f.write(re.sub(r"(?m)^image .*$", line, src))
    f.close()
    return f.name


def host_threads():
    """Threads the reference's OpenMP loops get: every core this process may run on."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def ref_env():
    """torch.distributed.run exports OMP_NUM_THREADS=1 to its workers; the reference's CPU path is an OpenMP program and is
    given every host core explicitly."""
    env = dict(os.environ)
    env["OMP_NUM_THREADS"] = str(host_threads())
    env.pop("OMP_PROC_BIND", None)
    return env


def live_per_primary(scene, w, h, bounces):
    """Live ray-bounces per primary ray of the reference's CPU path, counted by the reference's OWN functions (load_scene,
    generate_initial_rays, process_ray behind oracle/_ref/libref_shim.so, CPU seeds) on a 1/8-scale image of the same
    camera. The reference binary prints no counts; the ratio converts its `CPU Took` time into the metric's unit."""
    import ctypes as C

    import numpy as np

    os.environ["OMP_NUM_THREADS"] = str(host_threads())  # before libgomp starts (the shim's loops are OpenMP too)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle as po

    ref = po.RefScene(scene + ".scene", cwd=SCENES)
    sw, sh = max(16, w // 8), max(16, h // 8)
    ref.set_image(sw, sh, 20, bounces, 1.0)
    img = np.zeros(sw * sh * 3, np.float32)
    live = C.c_longlong()
    po.shim().ref_render(ref.handle, 0, img, C.byref(live))
    return live.value / float(sw * sh * 20)


def run_reference_cpu(scene, w, h, spp, bounces):
    """One run of the reference program's CPU path; returns seconds from its own `CPU Took` line."""
    path = _scene_file_with_image(scene, w, h, spp, bounces)
    try:
        out = subprocess.run([REF_BIN, path, "cpu", "no_gpu"], cwd=SCENES, capture_output=True, text=True, timeout=1800, env=ref_env()).stdout
    finally:
        os.unlink(path)
    m = re.search(r"CPU Took ([0-9.eE+-]+)s", out)
    if not m:
        raise RuntimeError("reference CPU run printed no `CPU Took` line: " + out[-400:])
    return float(m.group(1))


def size_cpu_sample(scene, w, h, bounces, target_seconds=12.0):
    """Pick the sample: full resolution, spp chosen so one run takes about target_seconds (1..20 spp = at most one pass)."""
    cw, ch = max(64, w // 4), max(64, h // 4)
    t = run_reference_cpu(scene, cw, ch, 4, bounces)
    primaries_per_s = cw * ch * 4 / max(t, 1e-6)
    spp = int(max(1, min(PASS_SPP, target_seconds * primaries_per_s / (w * h))))
    return spp


def reference_arm(args, rank, world, emit):
    if rank != 0:
        return 0
    scene, w, h, spp, bounces = WORKLOADS[args.workload]
    if not os.path.exists(REF_BIN):
        emit({"impl": "reference", "unavailable": "oracle/_ref/raytracing_ref was not built (needs /root/reference at build time)"})
        return 0
    ratio = live_per_primary(scene, w, h, bounces)
    threads = host_threads()
    s_spp = size_cpu_sample(scene, w, h, bounces, 10.0)
    times = []
    for i in range(args.warmup + args.steps):
        t = run_reference_cpu(scene, w, h, s_spp, bounces)
        if i >= args.warmup:
            times.append(t)
    sec = sum(times) / len(times)
    live = ratio * w * h * s_spp
    value = live / sec * 1e-6
    sample = (f"{w}x{h} x {s_spp} spp (one pass of the workload's {spp} spp), `raytracing_ref <scene> cpu no_gpu`, OMP_NUM_THREADS={threads} "
              f"(nproc {os.cpu_count()}), {ratio:.3f} live ray-bounces per primary ray counted by the reference's own process_ray")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": config_for(args),
            "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": threads, "nproc": os.cpu_count(), "kind": "reference", "sample": sample},
            "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)
    return 0


# ---- the B200 arm ------------------------------------------------------------------------------------------------------
def agreeing_runs(measure, least=3, most=5, tol=0.02):
    """Samples of a timer with one-sided noise: at least `least`, then more until the two smallest agree to `tol` (or `most` runs).
    The smallest is the estimate; a single lucky-looking outlier downwards does not exist for such a timer."""
    runs, tries = [], 0
    while len(runs) < most and tries < most + 2:
        tries += 1
        t = measure()
        if t:
            runs.append(t)
        if len(runs) >= least:
            lo = sorted(runs)
            if lo[1] <= lo[0] * (1.0 + tol):
                break
    return runs


def slope_fit(xs, ys):
    """Least-squares line through (xs, ys): slope, intercept, residuals, standard error of the slope."""
    n = len(xs)
    mx, my = sum(xs) / n, sum(ys) / n
    sxx = sum((x - mx) ** 2 for x in xs)
    slope = sum((x - mx) * (y - my) for x, y in zip(xs, ys)) / sxx
    icpt = my - slope * mx
    res = [y - (icpt + slope * x) for x, y in zip(xs, ys)]
    se = (sum(r * r for r in res) / max(1, n - 2) / sxx) ** 0.5
    return slope, icpt, res, se


def robust_line(xs, ys):
    """Theil-Sen line: median of the pairwise slopes, median intercept. One bad point in five does not move it."""
    sl = sorted((ys[j] - ys[i]) / (xs[j] - xs[i]) for i in range(len(xs)) for j in range(i + 1, len(xs)))
    slope = sl[len(sl) // 2] if len(sl) % 2 else 0.5 * (sl[len(sl) // 2 - 1] + sl[len(sl) // 2])
    ic = sorted(y - slope * x for x, y in zip(xs, ys))
    icpt = ic[len(ic) // 2] if len(ic) % 2 else 0.5 * (ic[len(ic) // 2 - 1] + ic[len(ic) // 2])
    return slope, icpt


def fit_with_remeasure(xs, ys, measure_again, rounds=2):
    """slope_fit for a timer whose noise is one-sided (an allocation or an upload taking longer, never shorter): a point that sits
    ABOVE the robust (Theil-Sen) line by more than 3 x the median absolute residual to that line and 2 % of its value is
    measured again (`measure_again(x)` returns further samples, the smallest is kept), at most `rounds` times. The reported
    fit is the least-squares one of the final points. Modifies ys; returns the fit + the re-measured xs."""
    remeasured = []
    for _ in range(rounds):
        slope, icpt = robust_line(xs, ys)
        res = [y - (icpt + slope * x) for x, y in zip(xs, ys)]
        k = max(range(len(xs)), key=lambda j: res[j])
        med = sorted(abs(r) for r in res)[len(xs) // 2]
        if res[k] <= max(3.0 * med, 0.02 * ys[k]):
            break
        ys[k] = min([ys[k]] + list(measure_again(xs[k])))
        remeasured.append(xs[k])
    return (*slope_fit(xs, ys), remeasured)


def ref_gpu_timing(scene, w, h, bounces, lpp):
    """The reference's sm_100a GPU build on the same box (north_star's comparison target). Its `GPU Took` timer
    (raytracing.cu:172,279-281) includes ~13 cudaMalloc / cudaFree, the pageable scene upload and the download, which do not
    depend on the sample count; the steady-state rate is therefore the least-squares SLOPE of `GPU Took` over five sample
    counts (whole passes of 20), every point the best of three runs after one discarded warm-up run per mode. Reported with
    the standard error of the slope (`spread`) and the largest residual, so that a noisy box shows. The timer's hiccups only
    ever ADD time: a point that stands above the robust line through the five is run three more times (fit_with_remeasure)."""
    # sample counts far enough apart that the slope stands clear of the timer's ~20 ms run-to-run jitter (spheres at 1000x1000: 0.2 ms per spp)
    pts = (200, 600, 1000, 1400, 1800) if w * h <= 1_000_000 else (20, 40, 60, 80, 100)
    paths = {spp: _scene_file_with_image(scene, w, h, spp, bounces) for spp in pts}

    def took(spp, flags):
        o = subprocess.run([REF_BIN, paths[spp], *flags], cwd=SCENES, capture_output=True, text=True, timeout=1800).stdout
        m = re.search(r"GPU Took ([0-9.eE+-]+)s", o)
        return float(m.group(1)) if m else None

    out = {"config": f"{scene}.scene {w}x{h}, `GPU Took` at {list(pts)} spp, per point the smallest of 3-5 runs (until the two smallest agree to 2 %), least-squares slope = steady state"}
    try:
        for flags in ((), ("no_sort",)):
            mode = "no_sort" if flags else "sort"
            took(pts[0], flags)  # warm-up, discarded
            ys, n_runs = [], []
            for spp in pts:
                runs = agreeing_runs(lambda: took(spp, flags))
                ys.append(min(runs) if runs else None)
                n_runs.append(len(runs))
            if any(y is None for y in ys):
                out[mode] = {"unavailable": "no `GPU Took` line"}
                continue
            slope, icpt, res, se, remeasured = fit_with_remeasure(pts, ys, lambda spp: [t for t in (took(spp, flags) for _ in range(3)) if t])
            rate = lpp * w * h / max(slope, 1e-12) * 1e-6
            out[mode] = {"value": rate, "unit": "Mrays/s", "spread": rate * se / max(slope, 1e-12), "spread_rel": se / max(slope, 1e-12),
                         "s_per_spp": slope, "fixed_cost_s": icpt, "max_residual_s": max(abs(r) for r in res), "gpu_took_s": dict(zip(pts, ys)),
                         "runs_per_point": n_runs, "remeasured_spp": remeasured}
    finally:
        for pth in paths.values():
            os.unlink(pth)
    return out


def newest_ncu_summary(build=None, kernels=None):
    """The ncu summary (profiles/r*_traffic.json, written by tools/ncu_summary.py) that measured THIS library's kernels: the
    newest one whose `kernel_id` (sha1 of the SASS of trace_rays_kernel / shade_kernel, core.kernel_id()) or `build_id` (hash of
    the sources) equals the running library's; else the one with the highest name (= the latest capture; bench.py then
    reports it as stale and leaves the derived keys out)."""
    found = []
    for f in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_traffic.json"))):
        try:
            d = json.load(open(f))
        except (OSError, ValueError):
            continue
        if "trace_rays_kernel" not in d or "build_id" not in d:
            continue
        d["_file"] = os.path.relpath(f, ROOT)
        found.append(d)
    f32 = [d for d in found if "nodes=f32" in d.get("workload", "")]  # the default node format's captures
    same = [d for d in f32 if d.get("build_id") == build or (kernels is not None and d.get("kernel_id") == kernels)]
    return (same or f32 or found or [None])[-1]


def main():
    args = parse_args()
    # Libraries (NCCL prints its version banner) may write to stdout; the contract is ONE JSON line there. Keep the real
    # stdout aside and point fd 1 at stderr for the duration of the run.
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)

    def emit(obj):
        real_stdout.write(json.dumps(obj) + "\n")
        real_stdout.flush()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return reference_arm(args, rank, world, emit)

    import numpy as np
    import torch

    import cuda_raytracer_b200 as rt
    from cuda_raytracer_b200.core import build_id, kernel_id
    from cuda_raytracer_b200.multi_gpu import DeviceFramebuffer as _DevBuf, sample_range

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the render core has no CPU fallback (use --impl reference for the CPU path)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    scene_name, w, h, spp, bounces = WORKLOADS[args.workload]
    scene = rt.load_scene(scene_name + ".scene", cwd=SCENES).set_image(w, h, spp, bounces)
    total_passes = (spp + PASS_SPP - 1) // PASS_SPP
    total_rays = w * h * spp
    first, count = sample_range(spp, rank, world)
    share = dict(sample_first=first, sample_count=count) if world > 1 else {}  # one GPU: the whole frame, no range
    sort = bool(args.sort)

    r = rt.Renderer(local_rank, nodes=args.nodes)
    r.upload(scene)
    n_floats = w * h * 3
    host_out = torch.empty(n_floats, dtype=torch.float32).pin_memory().numpy()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def frame(timing=False, keep=None):
        """One step, inputs resident: this rank's share of the rays, then the reduce of the accumulation buffer to rank 0."""
        if count > 0 or world == 1:
            ptr = r.render_device(sort=sort, timing=timing, pass_group=args.pass_group, **share)
            st = r.last_stats
            dev_s, live, launches, tr_ms, sh_ms, so_ms = st.loop_seconds, st.live_ray_bounces, st.kernel_launches, st.trace_ms, st.shade_ms, st.sort_ms
        else:  # more ranks than rays: contribute zeros
            ptr = r.render_device(sort=sort, sample_first=spp, sample_count=1)
            dev_s = live = launches = tr_ms = sh_ms = so_ms = 0
        red_s = 0.0
        if dist is not None:
            fb = torch.as_tensor(_DevBuf(ptr, n_floats), device=f"cuda:{local_rank}")
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            dist.reduce(fb, dst=0, op=dist.ReduceOp.SUM)
            e1.record()
            e1.synchronize()  # the core's next render writes this buffer on its own stream: the reduce must be over
            red_s = e0.elapsed_time(e1) * 1e-3
            if keep is not None and rank == 0:
                keep.append(fb.clone())
        return dev_s + red_s, live, launches, tr_ms, so_ms, red_s, sh_ms

    e2e_upload = []

    def frame_e2e():
        """The reference-facing call with host buffers: upload (H2D of the scene), render, D2H of the sums."""
        t0 = time.perf_counter()
        r.upload(scene)
        e2e_upload.append(time.perf_counter() - t0)
        if dist is None:
            r.render(sort=sort, out=host_out.reshape(h, w, 3), pass_group=args.pass_group)
        else:
            ptr = r.render_device(sort=sort, pass_group=args.pass_group, **(share if count > 0 else dict(sample_first=spp, sample_count=1)))
            fb = torch.as_tensor(_DevBuf(ptr, n_floats), device=f"cuda:{local_rank}")
            dist.reduce(fb, dst=0, op=dist.ReduceOp.SUM)
            if rank == 0:
                torch.from_numpy(host_out).copy_(fb, non_blocking=False)
            torch.cuda.synchronize()
        return time.perf_counter() - t0

    # stats-mode render of one pass: N_int / N_tri per live ray-bounce for the roofline's algorithmic bytes
    r.render_device(sort=sort, pass_first=0, pass_count=1, collect_stats=True)
    s1 = r.last_stats
    n_int = s1.internal_visits / max(1, s1.live_ray_bounces)
    n_tri = s1.triangle_tests / max(1, s1.live_ray_bounces)
    bytes_per_unit = 104.0 + 64.0 * n_int + 48.0 * n_tri  # the whole bounce (ray state r/w + key + BVH records)
    # the dominant kernel (trace_rays_kernel) alone: ray read 32 B + hit record 8 B + BVH records
    trace_bytes_per_unit = 40.0 + 64.0 * n_int + 48.0 * n_tri

    sampler = ClockSampler(visible_gpu_index(local_rank))
    sampler.start()  # nvidia-smi needs ~0.5 s to deliver its first row: start it ahead of the warm-up
    first_frame = []
    for i in range(args.warmup):
        frame(keep=first_frame if i == 0 else None)
    barrier()
    sampler.mark()
    wall0 = time.perf_counter()
    per_step = [frame() for _ in range(args.steps)]
    barrier()
    wall = time.perf_counter() - wall0
    clocks = sampler.stop()
    # Attribution steps, right after the timed region: the same frames with a CUDA event pair around every kernel. The
    # product path runs two waves at a time on two streams (render_impl's pipes), where a kernel's event interval would
    # include the other stream's work; with per-kernel events the core traces one wave at a time, so the durations are
    # each kernel's own. `value` comes from the timed region above, the roofline's kernel durations from these steps.
    attr_step = [frame(timing=True) for _ in range(args.steps)]
    barrier()

    dev_s = sum(p[0] for p in per_step)
    live = sum(p[1] for p in per_step)
    launches = sum(p[2] for p in per_step)
    red_s = sum(p[5] for p in per_step)
    attr_s = sum(p[0] for p in attr_step)
    tr_ms = sum(p[3] for p in attr_step)
    so_ms = sum(p[4] for p in attr_step)
    sh_ms = sum(p[6] for p in attr_step)

    # e2e: warm once, then time K steps (wall clock around the public call, barrier + sync on both sides)
    frame_e2e()
    frame_e2e()  # two warm frames: at N = 8 the first e2e frame after the attribution steps ran 34 ms against 19-21 for the rest
    barrier()
    e2e_t = [frame_e2e() for _ in range(args.steps)]
    barrier()
    e2e_s = sum(e2e_t)

    # N > 1: the reduced frame of the first step against a one-GPU render of the same frame on rank 0 (same pass seeds, so the
    # images agree up to float summation order) and the summed live-ray counters against that render's
    image_check = None
    if dist is not None:
        live_first = torch.tensor([float(per_step[0][1])], dtype=torch.float64, device="cuda")
        dist.all_reduce(live_first, op=dist.ReduceOp.SUM)
        if rank == 0 and first_frame:
            ptr = r.render_device(sort=sort, pass_group=args.pass_group)
            one = torch.as_tensor(_DevBuf(ptr, n_floats), device=f"cuda:{local_rank}")
            got = first_frame[0]
            rel_rmse = float(torch.sqrt(torch.mean((got - one) ** 2)) / torch.sqrt(torch.mean(one ** 2)))
            image_check = {"rel_rmse_vs_1gpu": rel_rmse, "live_ray_bounces_n_gpu": int(live_first.item()),
                           "live_ray_bounces_1gpu": int(r.last_stats.live_ray_bounces),
                           "ok": bool(rel_rmse <= 1e-4 and int(live_first.item()) == int(r.last_stats.live_ray_bounces))}
        dist.barrier()

    per_rank_ms = None
    if dist is not None:
        # every rank's own pass loop and reduce interval (ms per step): the reduce interval of a rank that finishes its
        # share early is mostly the wait for the slowest rank, so the spread of the loops is the number to read
        mine = torch.tensor([(dev_s - red_s) / args.steps * 1e3, red_s / args.steps * 1e3], dtype=torch.float64, device="cuda")
        every = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(every, mine)
        per_rank_ms = {"pass_loop": [round(float(v[0]), 3) for v in every], "reduce_interval": [round(float(v[1]), 3) for v in every]}
        t = torch.tensor([dev_s, e2e_s, tr_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_s, e2e_s, tr_ms_max = (float(v) for v in t)
        t = torch.tensor([float(live), float(launches)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        live, launches = (float(v) for v in t)
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return 0

    a = scene.desc()
    scene_bytes = (a.sphere_count * 16 + a.triangle_count * 48 + (a.sphere_count + a.triangle_count) * 2 + a.material_count * 48 +
                   a.bvh_node_count * 32 + a.env_width * a.env_height * 12)
    value = live / dev_s * 1e-6
    e2e_value = live / e2e_s * 1e-6
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    l2_peak = r.probe_copy_gbs(16 << 20, 200)  # 2 x 16 MB: resident in the 126 MB L2
    # dominant kernel = trace_rays_kernel: algorithmic bytes of rank 0's launches / their summed CUDA-event durations
    live_rank0 = sum(p[1] for p in attr_step)
    trace_s, shade_s = max(1e-9, tr_ms * 1e-3), max(1e-9, sh_ms * 1e-3)
    achieved = live_rank0 * trace_bytes_per_unit / trace_s * 1e-9
    n_trace_launches = max(1, sum(p[2] for p in attr_step) // 2) if not sort else None  # a bounce = one trace + one shade launch
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "kernel": "trace_rays_kernel", "bytes_per_ray_bounce": trace_bytes_per_unit, "n_int": n_int, "n_tri": n_tri,
                "whole_bounce_bytes_per_ray_bounce": bytes_per_unit,
                "whole_step_achieved": sum(p[1] for p in per_step) * bytes_per_unit / max(1e-9, sum(p[0] for p in per_step)) * 1e-9,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)",
                "kernel_share_of_step": tr_ms * 1e-3 / max(1e-9, attr_s),
                "frac_note": "SURVEY §8(d)'s contract figure: ALGORITHMIC bytes (ray state + every visited BVH record and tested triangle) over the "
                             "HBM copy peak. The scene is L1/L2-resident, so this fraction exceeds 1 and bounds nothing; the fractions of what each "
                             "memory level actually carries and the limiter are the keys below",
                "l2_peak_gbs": l2_peak, "l2_peak_source": "rt_probe_copy_gbs: 2 x 16 MB device copy, read + write, measured in this run",
                "timing_note": "kernel durations: CUDA event pairs around every launch in %d attribution step(s) run right after the timed region "
                               "(one wave at a time, %.1f ms per step); the timed region itself runs two waves at a time on two streams, where "
                               "per-kernel intervals would overlap" % (args.steps, attr_s / args.steps * 1e3)}
    ncu = newest_ncu_summary(build_id(), kernel_id())
    if ncu is not None:
        # the capture counts for this run if it measured the same machine code: same sources, or the same SASS of the two kernels
        # (a host-only change gives a new source hash and leaves the kernels alone)
        same_kernels = kernel_id() is not None and ncu.get("kernel_id") == kernel_id()
        same_build = ncu.get("build_id") == build_id()
        roofline["ncu_source"] = {"file": ncu["_file"], "raw_csv": ncu.get("raw_csv"), "build_id": ncu.get("build_id"), "this_build_id": build_id(),
                                  "kernel_id": ncu.get("kernel_id"), "this_kernel_id": kernel_id(), "same_kernels": same_kernels,
                                  "same_build": same_build, "valid_for_this_run": bool(same_build or same_kernels),
                                  "validity_rule": "same sources (build_id) or byte-identical SASS of trace_rays_kernel* / shade_kernel* (kernel_id): "
                                                   "a host-only change moves build_id and leaves the captured kernels alone",
                                  "workload": ncu.get("workload")}
        tk, sk = ncu.get("trace_rays_kernel", {}), ncu.get("shade_kernel", {})
        if (same_build or same_kernels) and args.workload == "lamp_4k" and not sort and tk:
            # per live ray-bounce figures of the capture x this run's ray-bounces / this run's kernel time
            roofline["traffic"] = tk["dram_bytes_per_live_ray_bounce"] * live_rank0 / n_trace_launches
            roofline["traffic_note"] = "DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) per average trace launch"
            roofline["dram_frac"] = tk["dram_bytes_per_live_ray_bounce"] * live_rank0 / trace_s * 1e-9 / peak
            roofline["l2_frac"] = tk["lts_bytes_per_live_ray_bounce"] * live_rank0 / trace_s * 1e-9 / l2_peak
            for k in ("l1_wavefront_pct", "issue_active", "lanes_per_inst", "l1_hit_pct", "warp_inst_per_live_ray_bounce"):
                if k in tk:
                    roofline[k] = tk[k]
            roofline["limiter"] = ncu.get("limiter")
            if sk:
                roofline["shade_kernel"] = {"share_of_step": sh_ms * 1e-3 / max(1e-9, attr_s),
                                            "dram_frac": sk["dram_bytes_per_live_ray_bounce"] * live_rank0 / shade_s * 1e-9 / peak,
                                            "l2_frac": sk["lts_bytes_per_live_ray_bounce"] * live_rank0 / shade_s * 1e-9 / l2_peak,
                                            **{k: sk[k] for k in ("l1_wavefront_pct", "issue_active", "lanes_per_inst") if k in sk}}
        elif not (same_build or same_kernels):
            roofline["traffic_note"] = "the newest ncu capture under profiles/ is of other kernels: traffic and the per-level fractions are left out"
    line = {
        "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dev_s / args.steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": config_for(args), "run": {"pass_group": args.pass_group or "auto (70 % of free memory)", "nodes": r.node_mode, "build_id": build_id(), "kernel_id": kernel_id()},
        "e2e": {"value": e2e_value, "unit": "Mrays/s", "ms_per_step": e2e_s / args.steps * 1e3,
                "h2d_bytes_per_step": int(scene_bytes), "d2h_bytes_per_step": int(n_floats * 4),
                "ms_each": [round(t * 1e3, 2) for t in e2e_t], "upload_ms_each": [round(t * 1e3, 2) for t in e2e_upload[-len(e2e_t):]]},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roofline,
        "breakdown_ms_per_step": {"trace_kernels": tr_ms / args.steps, "shade_kernels": sh_ms / args.steps, "reorder": so_ms / args.steps, "nccl_reduce": red_s / args.steps * 1e3,
                                  "wall": wall / args.steps * 1e3},
    }
    if image_check is not None:
        line["multi_gpu_image_check"] = image_check
    if per_rank_ms is not None:
        line["per_rank_ms_per_step"] = per_rank_ms
    extras = world == 1 and not args.no_extras
    if extras:  # the same tree built on the device (csrc/device/bvh_build.cuh) vs the host builder
        dev_scene = rt.load_scene(scene_name + ".scene", use_bvh="device", cwd=SCENES).set_image(w, h, spp, bounces)
        r.upload(dev_scene)
        r.upload(dev_scene)  # second upload: buffers exist, steady state
        line["bvh_build_ms"] = {"host_binned_sah": scene.bvh_build_ms, "device_binned_sah": r.bvh_device_ms,
                                "triangles": int(a.triangle_count), "nodes": int(a.bvh_node_count)}
        r.upload(scene)
    if extras:  # the same workload on the host tree WITH spatial splits (rt_scene_load(use_bvh = 3)): not the headline, whose
        # tree is the reference's own; same hits and image, fewer visited records (profiles/r1m_sbvh.md)
        t0 = time.perf_counter()
        sb_scene = rt.load_scene(scene_name + ".scene", use_bvh="sbvh", cwd=SCENES).set_image(w, h, spp, bounces)
        sb_load_s = time.perf_counter() - t0
        r.upload(sb_scene)
        r.render_device(sort=sort, pass_group=args.pass_group)  # warm-up
        sb_s, sb_live = 0.0, 0
        for _ in range(args.steps):
            r.render_device(sort=sort, pass_group=args.pass_group)
            sb_s += r.last_stats.loop_seconds
            sb_live += r.last_stats.live_ray_bounces
        line["sbvh_tree"] = {"value": sb_live / max(sb_s, 1e-9) * 1e-6, "unit": "Mrays/s", "ms_per_step": sb_s / args.steps * 1e3,
                             "triangle_references": int(sb_scene.desc().triangle_count), "triangles": int(a.triangle_count),
                             "nodes": int(sb_scene.desc().bvh_node_count), "load_and_build_ms": sb_load_s * 1e3,
                             "note": "rt_scene_load(use_bvh = 3): host tree with spatial splits instead of the reference's tree; same core, same image"}
        r.upload(scene)
    if world == 1 and not args.no_cpu_baseline and os.path.exists(REF_BIN):
        ratio = live_per_primary(scene_name, w, h, bounces)
        s_spp = size_cpu_sample(scene_name, w, h, bounces, 12.0)
        sec = run_reference_cpu(scene_name, w, h, s_spp, bounces)
        line["cpu_baseline"] = {"value": ratio * w * h * s_spp / sec * 1e-6, "unit": "Mrays/s", "cores": host_threads(), "nproc": os.cpu_count(),
                                "kind": "reference",
                                "sample": f"{w}x{h} x {s_spp} spp of the workload, raytracing_ref cpu no_gpu, OMP_NUM_THREADS={host_threads()}, {sec:.2f} s"}
    if world == 1 and not args.no_ref_gpu and os.path.exists(REF_BIN):
        lpp = live / args.steps / (w * h * spp)
        line["reference_gpu_sm100a"] = ref_gpu_timing(scene_name, w, h, bounces, lpp)
    emit(line)
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
