#!/usr/bin/env python
"""bench.py — headline benchmark of the SyNeRFgine render hot path on B200 (BASELINE.json: Mrays/s and ms/frame,
1080p NeRF + mesh + shadows, at 1/2/4/8 B200).

    python bench.py --gpus 1 --steps K --warmup W [--config c2]     # this repo's CUDA path (libsnb.so) through the C ABI
    python bench.py --impl reference ...                             # the reference's own kernels on the same GPU (oracle/_ref)
    torchrun --nproc-per-node N bench.py --gpus N ...                # one frame split into N row bands, NCCL point-to-point assembly

A "step" is one complete frame (snb_frame: mesh layer, NeRF march, both shadow directions, overlay) of the configuration
(synerfgine_b200/workloads.py: c1, c2 = default = the configuration the metric is quoted on, c3, c4, c5-8k, c5-batch; for c5-batch a
step is the whole 64-frame batch). value = primary rays / device time with everything resident in HBM; e2e = the same through
snb_frame_host with pinned HOST output buffers and the per-frame scene upload inside the timed region. One JSON line on stdout.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BYTES_PER_SAMPLE = 512          # hash gather: 16 levels x 8 corners x 2 features x 2 B
BYTES_PER_RAY = 64 + 20 + 12    # queue record read (ray + its SH-4 block) + rgba/depth write + position write
FLOP_PER_SAMPLE = 20480         # both MLPs, padded
METRIC = "Mrays/s, 1080p NeRF+mesh+shadows"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c2", help="c1 | c2 | c3 | c4 | c5-8k | c5-batch (synerfgine_b200/workloads.py)")
    ap.add_argument("--width", type=int, default=0, help="override the configuration's resolution")
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--march", default="", help="ws | lane (default: the library's)")
    ap.add_argument("--alt-schedule", default="", help="auto | fixed (default: the reference's count-dependent batch lengths)")
    ap.add_argument("--batch-assign", default="contiguous", choices=["contiguous", "balanced", "strided"],
                    help="c5-batch: which frames of the path a rank renders (balanced = contiguous chunks sized by measured per-frame cost)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-cuda-baseline", action="store_true")
    ap.add_argument("--cpu-sample", default="240x135")
    return ap.parse_args()


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return d.get("hbm_gbs", 6650.0), d.get("bf16_tflops_sustained", 1400.0), "measured"
    return 6650.0, 1400.0, "fallback"


def sectors_per_sample():
    """L2 sectors one evaluated sample costs (ncu: lts__t_sectors_srcunit_tex.sum / samples of the captured launch), profiles/fused_traffic.json."""
    p = os.path.join(ROOT, "profiles", "fused_traffic.json")
    try:
        with open(p) as f:
            return float(json.load(f)["l2_sectors_per_sample"]["c2"])
    except (OSError, KeyError, ValueError):
        return 128.0 * 0.25                            # 128 gathers per sample, a quarter of them reaching the L2 as their own sector


def l2_peaks():
    """L2 gather bandwidth measured on the box by profiles/l2_gather_bench.py (random 4-byte reads, 32-byte sectors), if committed."""
    p = os.path.join(ROOT, "profiles", "r02_l2_gather_peak.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f)
    return None


def claim_stdout():
    """stdout carries exactly ONE JSON line: anything a library prints there (NCCL's banner, tinylogger) goes to stderr."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)

    def emit(line):
        os.write(real, (line + "\n").encode())
    return emit


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True); self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)                      # at least one sample even for a very short timed region
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            pass
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


def config_of(args):
    from synerfgine_b200 import workloads
    cfg = dict(workloads.CONFIGS[args.config])
    if args.width and args.height:
        cfg["width"], cfg["height"] = args.width, args.height
    return cfg


# ---- CPU oracle legs (cpu_baseline) -----------------------------------------------------------------------------------------
_W = {}


def _oracle_worker_init(cfg):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    from synerfgine_b200 import scenes, workloads
    from synerfgine_b200.scene import Scene as HostScene
    _W["O"] = O; O.load()
    _W["desc"] = workloads.model(cfg["model"]).desc()
    _W["model"] = workloads.model(cfg["model"])
    _W["scene"] = HostScene(scenes.scene_json(cfg["scene"]), scenes.mesh_dir()) if cfg.get("scene") else None
    _W["cfg"] = cfg


def _oracle_frame(args):
    w, h, spp = args
    from synerfgine_b200 import workloads
    O, sc, cfg = _W["O"], _W["scene"], _W["cfg"]
    cam = sc.camera_matrix() if sc is not None else workloads.camera_matrix(zoom=cfg.get("zoom", 1.0))
    v = workloads.make_view(w, h, cam=cam, spp=spp)
    t0 = time.perf_counter()
    if sc is not None:
        O.frame(_W["desc"], v, sc, sc.rt, 1999, 1)
    else:
        O.render_nerf(_W["desc"], v, 0)
    return time.perf_counter() - t0


def cpu_oracle_rate(cfg, w, h, procs, steps, warmup):
    """Mrays/s of the oracle frame loop on `procs` host processes; each step = every process renders one w x h frame."""
    import multiprocessing as mp
    ctx = mp.get_context("fork")
    with ctx.Pool(procs, initializer=_oracle_worker_init, initargs=(cfg,)) as pool:
        for i in range(warmup):
            pool.map(_oracle_frame, [(w, h, i)] * procs)
        t0 = time.perf_counter()
        for i in range(steps):
            pool.map(_oracle_frame, [(w, h, warmup + i)] * procs)
        dt = time.perf_counter() - t0
    return procs * steps * w * h / dt / 1e6, dt / steps * 1e3


# ---- the reference's own CUDA kernels on this GPU (oracle/_ref/libref_dev.so, test infrastructure) ----------------------------------
def reference_cuda_frames(cfg, steps, warmup):
    """ms per frame of the reference's kernels on the same configuration: sng::raytrace + NerfTracer::trace_alt (kernel_grid,
    kernel_mlp_fused, composite_kernel_nerf_alt, compact_kernel_nerf, with the reference's own host loop and its two host syncs per
    iteration) + extract_from_payload + write_normals_to_buffer + shade_with_shadow + overlay_nerf, i.e. Engine::frame minus GL
    (oracle/ref_dev_shim.cu). Returns None when the library is not available here."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refdev_lib as R
    ok, why = R.available()
    if not ok:
        return None, why
    from synerfgine_b200 import scenes, workloads
    from synerfgine_b200.scene import Scene as HostScene
    m = workloads.model(cfg["model"])
    W, H = cfg["width"], cfg["height"]
    rm = R.Model(m); eng = R.Engine(W, H)
    stage = np.zeros(4)
    times = []
    if cfg.get("scene"):
        sc = HostScene(scenes.scene_json(cfg["scene"]), scenes.mesh_dir())
        rs = R.Scene(sc)
        v = workloads.make_view(W, H, cam=sc.camera_matrix())
        # frames > 1: the camera follows the scene's keyframe path and the animated objects advance every frame, exactly as our arm's
        # Engine.advance_camera_path / update_world_objects do (CamPath::update, Engine::update_world_objects); one pass over the path is timed
        path = cfg["frames"] > 1 and sc.n_keyframes >= 2 and sc.move_on_start
        total = cfg["frames"]; between = max(1, total // max(sc.n_keyframes - 1, 1)); cur = 0
        if path:
            steps = total
        for i in range(warmup + steps):
            if path:
                cur += 1
                if cur // between + 1 >= sc.n_keyframes:
                    cur = 0
                view, at, zoom = sc.camera_at(cur, total)
                v = workloads.make_view(W, H, cam=workloads.camera_matrix(view, at, zoom))
                if sc.animations:
                    sc.advance(); rs.close(); rs = R.Scene(sc)          # the reference re-uploads the world every frame too (engine.cu:80-127); not timed
            eng.reset(reseed=False)                                   # accumulation reset before every frame, like our arm
            ms = eng.frame(rm, rs, v, sc.rt)
            if i >= warmup:
                stage += np.asarray(ms) / steps; times.append(ms[3])
        rs.close()
        names = ["mesh", "nerf", "overlay", "frame"]
    else:
        v = workloads.make_view(W, H, cam=workloads.camera_matrix(zoom=cfg.get("zoom", 1.0)))
        for i in range(warmup + steps):
            ms = eng.render_nerf(rm, v, alt=False)
            if i >= warmup:
                stage[3] += ms / steps; times.append(ms)
        names = ["mesh", "nerf", "overlay", "frame"]; stage[1] = stage[3]
    eng.close(); rm.close()
    mean = float(np.mean(times))
    return {"ms_per_frame": mean, "ms_per_frame_p95": float(np.percentile(times, 95)), "value": W * H / (mean * 1e-3) / 1e6, "unit": "Mrays/s", "stage_ms": {n: round(float(x), 4) for n, x in zip(names, stage)},
            "frames": steps, "kind": "reference", "same_config": True,
            "what": "the reference's own kernels and NerfTracer host loop compiled for sm_100a from /root/reference (oracle/_ref/libref_dev.so), same model, scene, camera and resolution"}, ""


def run_reference(args, rank, world):
    """Reference arm. The reference is a CUDA program without a CPU path: its own kernels (tiny-cuda-nn, NerfTracer, sng::raytrace,
    shade_with_shadow, overlay_nerf) run on this GPU through oracle/_ref/libref_dev.so — Engine::frame minus GL, on the same
    configuration. When that library cannot be loaded the CPU oracle port of the same frame loop is timed on all host cores on a
    bounded sample instead. Rank 0 only (the reference has no way to split a frame)."""
    if rank != 0:
        return
    emit = claim_stdout()
    cfg = config_of(args)
    try:
        res, why = reference_cuda_frames(cfg, args.steps, max(1, args.warmup))
    except Exception as e:  # noqa: BLE001  (e.g. the reference's wavefront buffers not fitting at 8K)
        res, why = None, f"{type(e).__name__}: {e}"
    line = {"impl": "reference", "metric": METRIC, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "data": "synthetic", "config": {"workload": cfg["workload"]}}
    if res is not None:
        line.update({"value": res["value"], "ms_per_step": res["ms_per_frame"], "dtype": "f32 march (fast-math) / f16 MLP (fp16 accumulate)",
                     "cuda_baseline": res,
                     "cpu_baseline": {"value": res["value"], "unit": "Mrays/s", "cores": 0, "kind": "reference",
                                      "sample": "the reference ships no CPU path: its own CUDA kernels on 1 GPU of this box, full frames of the same workload"},
                     "e2e": {"value": res["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0})
        line["config"]["reference_stage_ms"] = res["stage_ms"]
    else:
        procs = os.cpu_count() or 1
        budget_s = 150.0 / max(1, args.steps + args.warmup)
        rays = int(max(32 * 18, min(240 * 135, budget_s / 0.6e-3)))
        h = max(18, int((rays / (16 / 9)) ** 0.5)); w = h * 16 // 9
        mrays, ms = cpu_oracle_rate(cfg, w, h, procs, args.steps, args.warmup)
        sample = f"{procs} processes x one {w}x{h} frame of the same scene/camera per step (CPU oracle port; reference CUDA library unavailable: {why})"
        line.update({"value": mrays, "ms_per_step": ms, "dtype": "f32 march / f16 MLP (fp32 accumulate)", "config": {"workload": cfg["workload"], "sample": sample},
                     "cpu_baseline": {"value": mrays, "unit": "Mrays/s", "cores": procs, "kind": "port", "sample": sample},
                     "e2e": {"value": mrays, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0})
    emit(json.dumps(line))


def run_ours(args, rank, local_rank, world):
    emit = claim_stdout()
    import torch
    import torch.distributed as dist
    from synerfgine_b200 import Engine, Testbed, scenes, tiles, workloads
    from synerfgine_b200._lib import Tile

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    cfg = config_of(args)
    W, H, K, Wu = cfg["width"], cfg["height"], args.steps, max(3, args.warmup)
    frames_per_step = cfg["frames"] if cfg.get("batch") else 1
    split_frame = world > 1 and not cfg.get("batch")        # one frame -> N row bands; a frame batch shards whole frames instead
    if split_frame:
        Wu = max(Wu, 20)                                  # the band edges settle during warm-up (tiles.StripBalancer); reported as `warmup`
    stream = torch.cuda.Stream()
    tb = Testbed(local_rank, stream=stream)
    lib, ctx = tb.lib, tb.ctx
    if args.march:
        assert lib.snb_set_option(ctx, b"march", args.march.encode()) == 0, lib.snb_last_error(ctx)
    model = workloads.model(cfg["model"])
    tb.set_model(model)                                   # weights / bitfield replicated once per GPU
    stock = bool(cfg.get("stock"))
    eng = None
    halo, align = 0, 1
    if not stock:
        eng = Engine()
        eng.set_virtual_world_json(scenes.scene_json(cfg["scene"]), scenes.mesh_dir())
        align = max(1, eng.rt.syn_px_scale)
        halo = tiles.halo_for(eng.rt.nerf_shadow_kernel, bool(eng.rt.show_shadow_on_nerf), align)
        if cfg["frames"] > 1:
            eng.path_frames = cfg["frames"]               # the scene's keyframe orbit resampled to 24 / 64 frames
    else:
        tb.camera_matrix = np.asarray(workloads.camera_matrix(zoom=cfg.get("zoom", 1.0)), np.float32).reshape(4, 3)
    bal = tiles.StripBalancer(H, world if split_frame else 1, halo, align)       # row bands: rows of a band are contiguous in the frame
    band = bal.strip(rank if split_frame else 0)
    band_tile = lambda s: Tile(0, s.hx0, W, s.render_width)
    if eng is not None:
        eng.init(W, H, tb, tile=band_tile(band) if split_frame else None)
        if cfg.get("zoom") and not eng.is_playing:
            tb.set_camera_from_view(eng.scene.view_dir, eng.scene.at_pos, cfg["zoom"])
    if args.alt_schedule:
        assert lib.snb_set_option(ctx, b"alt_schedule", args.alt_schedule.encode()) == 0
    elif split_frame and not stock:
        assert lib.snb_set_option(ctx, b"alt_schedule", b"external") == 0          # n_alive_j is a whole-frame quantity: summed over the ranks below
    n_rays_step = W * H * frames_per_step
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")   # > 126 MB L2
    # rank 0's assembled frame, double-buffered: the D2H copy of frame i (e2e) runs on a copy stream while frame i + 1 is assembled into the other
    full_bufs = [(torch.empty((H, W, 4), device="cuda"), torch.empty((H, W), device="cuda")) for _ in range(2)] if (split_frame and rank == 0) else None
    full_rgba, full_depth = full_bufs[0] if full_bufs else (None, None)
    asm = {"slot": 0, "copy_stream": torch.cuda.Stream() if (split_frame and rank == 0) else None, "done": [None, None], "pins": None}
    ev_a = torch.cuda.Event(enable_timing=True); ev_b = torch.cuda.Event(enable_timing=True)
    row_cost = torch.zeros(H, device="cuda", dtype=torch.float32) if split_frame else None
    stock_fb = torch.empty((H, W, 4), device="cuda") if stock else None
    stock_depth = torch.empty((H, W), device="cuda") if stock else None
    sched_checks = {"n": 0, "changed": 0}
    pipe = {"on": False, "slot": 0, "bufs": None}
    # frame batch: whole frames per rank. Frames of a camera path differ in cost (C4: p95 / mean = 1.3), so equal contiguous chunks leave the
    # rank with the expensive stretch of the path as the step time. Strided assignment (frame f -> rank f mod N) evens the mix but makes every
    # frame a camera jump for the batch-schedule predictor (more march passes: measured 906 against 1011 Mrays/s at N = 8). "balanced" keeps
    # chunks contiguous and sizes them by the per-frame cost measured during warm-up (one all_reduce of a 64-float vector, control plane):
    # measured 991. Equal contiguous chunks are the default.
    my_frames = None
    if cfg.get("batch"):
        my_frames = set(range(rank, cfg["frames"], world)) if args.batch_assign == "strided" else set(tiles.frame_batch_for(rank, world, cfg["frames"]))
    batch_bounds = None

    def sync_alt_schedule():
        """external schedule: the counts of all bands are summed and every rank sets the sequence they imply (snb.h)."""
        st = (C.c_uint32 * 132)()
        assert lib.snb_get_alt_status(ctx, st) == 0
        cnt = torch.tensor(list(st[4:68]), dtype=torch.int64, device="cuda")
        dist.all_reduce(cnt)
        target = int(st[3])
        want = [min(8, max(1, target // int(c))) if int(c) > 0 else 8 for c in cnt.tolist()]
        used = list(st[68:132])
        sched_checks["n"] += 1
        if [u if u else want[0] for u in used] != want:
            sched_checks["changed"] += 1                   # the frame just rendered used a sequence the whole-frame counts contradict
        arr = (C.c_uint8 * 64)(*want)
        assert lib.snb_set_alt_schedule(ctx, arr, 64) == 0

    def render_one(host=None):
        """One frame (or this rank's band of it). host = (pinned rgba, pinned depth): through snb_frame_host / a D2H copy."""
        if stock:
            v = tb.make_view(W, H, 0)
            rc = lib.snb_render_nerf(ctx, C.byref(v), None, 1, stock_fb.data_ptr(), stock_depth.data_ptr())
            assert rc == 0, lib.snb_last_error(ctx)
            if host is not None:
                host[0].copy_(stock_fb, non_blocking=True); host[1].copy_(stock_depth, non_blocking=True)
            return stock_fb, stock_depth
        if cfg["frames"] > 1:
            eng.advance_camera_path()
            if eng.enable_animations:
                eng.update_world_objects()
        else:
            lib.snb_reset_accumulation(ctx)               # accumulation reset before each timed frame
        if not split_frame:
            if host is not None:
                sc = eng.scene
                lib.snb_update_scene(ctx, sc.objects, sc.n_objects, sc.materials, sc.n_materials, sc.lights, sc.n_lights)   # per-frame H2D of the scene PODs
                v = tb.make_view(W, H, eng.frame_index); eng.frame_index += 1
                if pipe["on"]:
                    # streamed frames (what the reference's main loop does: frame after frame): the D2H copy of this frame runs on the
                    # context's copy stream while the next frame renders; two pinned buffer pairs, a slot is reused only after its copy landed
                    slot = pipe["slot"]; pipe["slot"] ^= 1
                    assert lib.snb_frame_host_wait(ctx, slot) == 0
                    hb = pipe["bufs"][slot]
                    rc = lib.snb_frame_host_async(ctx, C.byref(v), None, C.byref(eng.rt), hb[0].data_ptr(), hb[1].data_ptr(), slot)
                else:
                    rc = lib.snb_frame_host(ctx, C.byref(v), None, C.byref(eng.rt), host[0].data_ptr(), host[1].data_ptr())
                assert rc == 0, lib.snb_last_error(ctx)
                return None
            return eng.frame(sync=False)
        s = bal.strip(rank)
        if host is not None:
            sc = eng.scene
            lib.snb_update_scene(ctx, sc.objects, sc.n_objects, sc.materials, sc.n_materials, sc.lights, sc.n_lights)
        ev_a.record(stream)
        rgba, depth = eng.frame(sync=False, tile=band_tile(s))
        ev_b.record(stream)
        own = slice(s.x0 - s.hx0, s.x1 - s.hx0)
        if host is not None and rank == 0:
            # e2e: streamed like the single-GPU form — assemble into buffer `slot`, copy it to pinned host memory on the copy stream
            slot = asm["slot"]; asm["slot"] ^= 1
            if asm["done"][slot] is not None:
                stream.wait_event(asm["done"][slot])          # the slot's previous copy has left the device buffer
                asm["done"][slot].synchronize()               # ... and its pinned host buffers may be overwritten
            fr, fd = full_bufs[slot]
            tiles.assemble_rows_on_root([rgba[own], depth[own]], bal.bounds, rank, world, dist, [fr, fd])
            ready = torch.cuda.Event(); ready.record(stream)
            with torch.cuda.stream(asm["copy_stream"]):
                asm["copy_stream"].wait_event(ready)
                asm["pins"][slot][0].copy_(fr, non_blocking=True); asm["pins"][slot][1].copy_(fd, non_blocking=True)
                asm["done"][slot] = torch.cuda.Event(); asm["done"][slot].record(asm["copy_stream"])
        else:
            tiles.assemble_rows_on_root([rgba[own], depth[own]], bal.bounds, rank, world, dist, [full_rgba, full_depth])   # rgba + depth, 20 B/px, over NVLink
        if host is None and bal.due():
            if not args.alt_schedule:
                sync_alt_schedule()
            hits = (rgba[own][..., 3] > 0).sum(1, dtype=torch.float32)
            wgt = hits + 0.05 * torch.clamp(hits.max(), min=1.0)
            ev_b.synchronize()
            row_cost.zero_()
            row_cost[s.x0:s.x1] = wgt * (ev_a.elapsed_time(ev_b) / float(wgt.sum()))
            dist.all_reduce(row_cost)
            bal.observe(row_cost.cpu().numpy().astype(np.float64))
        bal.tick()
        return rgba, depth

    def one_step(host=None, cost=None):
        if not cfg.get("batch"):
            return render_one(host)
        for f in range(cfg["frames"]):                    # every rank walks the whole path on the host; it renders its own frames
            if f in my_frames:
                if cost is not None:
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(stream)
                render_one(host)
                if cost is not None:
                    e1.record(stream); cost.append((f, e0, e1))
            else:
                eng.skip_frames(1)                         # host-side position only (no camera matrix, no array refresh): ~2 µs
        return None

    with torch.cuda.stream(stream):
        for _ in range(Wu):
            one_step()
        if cfg.get("batch") and world > 1 and args.batch_assign == "balanced":
            cost = []
            one_step(cost=cost)
            stream.synchronize()
            vec = torch.zeros(cfg["frames"], device="cuda")
            for f, e0, e1 in cost:
                vec[f] = e0.elapsed_time(e1)
            dist.all_reduce(vec)
            batch_bounds = tiles.balanced_bounds(vec.cpu().numpy().astype(np.float64), world, 1, 1)
            my_frames = set(range(batch_bounds[rank], batch_bounds[rank + 1]))
            for _ in range(2):                             # settle on the new chunks (their first frames are the camera jumps now)
                one_step()
        stream.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        launches0 = tb.launch_count()
        sampler = ClockSampler(local_rank); sampler.start()
        step_ms, march_ms, stage_acc, samples, rays, march_stats, alt, pass_log = [], [], {}, 0, 0, {}, None, []
        reused = 0
        for _ in range(K):
            flush.zero_()                                  # L2 flush between timed iterations (outside the event pair)
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            one_step()
            e1.record(stream)
            e1.synchronize()
            step_ms.append(e0.elapsed_time(e1))
            st = tb.stage_ms(); march_ms.append(st["nerf_march"])
            for k_, v_ in st.items():
                stage_acc[k_] = stage_acc.get(k_, 0.0) + v_ / K
            c = tb.counters(); samples += c["samples"]; rays += c["rays"]; reused += c.get("reused", 0)
            ast_ = (C.c_uint32 * 132)(); lib.snb_get_alt_status(ctx, ast_); pass_log.append((int(ast_[0]), int(ast_[1])))
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        clocks = sampler.stop()
        launches = tb.launch_count() - launches0
        march_stats = tb.march_stats()
        march_stats["ws_pipeline"] = {k_: round(float(v_), 4) for k_, v_ in tb.ws_profile().items()}
        ast = (C.c_uint32 * 132)(); lib.snb_get_alt_status(ctx, ast)
        alt = {"passes_last_frame": int(ast[0]), "converged": int(ast[1]), "batch_lengths": [int(x) for x in ast[68:84]], "alive_rays_per_batch": [int(x) for x in ast[4:20]],
               "external_checks": sched_checks,
               "march_passes_per_frame_mean": float(np.mean([p for p, _ in pass_log])) if pass_log else None, "frames_converged": (int(sum(c_ for _, c_ in pass_log)) if not (split_frame and not args.alt_schedule) else None), "frames": len(pass_log),
               "note": ("row bands: every band marches with the WHOLE frame's batch lengths; the ranks sum their counts on the balancer's cadence and 'changed' counts the checks that found the sequence stale (0 = every checked frame exact)" if split_frame else None)}
        total_ms = torch.tensor([sum(step_ms)], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)     # device time, max over ranks
        total_ms = float(total_ms.item())

        # ---- end to end: host buffers; scene upload + frame (+ assembly) + D2H inside the timed region -----------------------------
        pin_rgba = pin_depth = None
        h2d = d2h = 0
        if eng is not None:
            sc = eng.scene
            h2d = sc.n_objects * 72 + sc.n_materials * C.sizeof(type(sc.materials[0])) + sc.n_lights * C.sizeof(type(sc.lights[0])) + 96
        else:
            h2d = 96
        if rank == 0 or not split_frame:
            pin_rgba = torch.empty((H, W, 4), dtype=torch.float32).pin_memory(); pin_depth = torch.empty((H, W), dtype=torch.float32).pin_memory()
            d2h = (pin_rgba.numel() + pin_depth.numel()) * 4 * (len(my_frames) if my_frames is not None else 1)
        host = (pin_rgba, pin_depth)
        if split_frame and rank != 0:
            host = (None, None)
        if split_frame and rank == 0:
            asm["pins"] = [(pin_rgba, pin_depth), (torch.empty((H, W, 4), dtype=torch.float32).pin_memory(), torch.empty((H, W), dtype=torch.float32).pin_memory())]
        e2e_sync_s = None
        if eng is not None and not split_frame:
            # first the strictly synchronous call (snb_frame_host returns with the host buffers filled), reported as e2e.synchronous ...
            for _ in range(2):
                one_step(host)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(K):
                one_step(host)
            torch.cuda.synchronize()
            e2e_sync_s = time.perf_counter() - t0
            # ... then the streamed form, which is the e2e figure: every frame's copy still completes inside the timed region
            pipe["on"] = True
            pipe["bufs"] = [(pin_rgba, pin_depth), (torch.empty((H, W, 4), dtype=torch.float32).pin_memory(), torch.empty((H, W), dtype=torch.float32).pin_memory())]
        for _ in range(2):
            one_step(host)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(K):
            one_step(host)
        if pipe["on"]:
            assert lib.snb_frame_host_wait(ctx, 0) == 0 and lib.snb_frame_host_wait(ctx, 1) == 0
        torch.cuda.synchronize()
        e2e_s = torch.tensor([time.perf_counter() - t0], device="cuda", dtype=torch.float64)
        pipe["on"] = False
        if world > 1:
            dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
        e2e_s = float(e2e_s.item())

        # ---- the march kernel without the mesh layer running beside it (same frames, overlap off): explains `roofline` ----
        march_alone_ms = None
        if world == 1 and eng is not None and cfg["frames"] == 1:
            lib.snb_set_option(ctx, b"overlap", b"0")
            alone = []
            for i in range(2 + 5):
                flush.zero_()
                one_step(); stream.synchronize()
                if i >= 2:
                    alone.append(tb.stage_ms()["nerf_march"])
            lib.snb_set_option(ctx, b"overlap", b"1")
            march_alone_ms = float(np.mean(alone))

        # ---- present stage, timed on its own: FXAA blit of the final frame into an RGB8 window image ----
        present = None
        if world == 1 and eng is not None and cfg["frames"] == 1 and W * H <= 3840 * 2160:
            out8 = torch.empty((H, W, 3), dtype=torch.uint8, device="cuda")
            final_rgba = eng.frame(sync=True)[0]
            ms = []
            for i in range(3 + 10):
                flush.zero_()
                p0 = torch.cuda.Event(enable_timing=True); p1 = torch.cuda.Event(enable_timing=True)
                p0.record(stream)
                rc = lib.snb_present(ctx, final_rgba.data_ptr(), W, H, W, H, None, None, None, out8.data_ptr())
                p1.record(stream); p1.synchronize()
                assert rc == 0, lib.snb_last_error(ctx)
                if i >= 3:
                    ms.append(p0.elapsed_time(p1))
            pm = float(np.mean(ms))
            present = {"kernel": "k_present", "ms": round(pm, 4), "bound": "hbm", "algorithmic_bytes": W * H * 19,
                       "achieved_gbs": round(W * H * 19 / (pm * 1e-3) / 1e9, 1), "note": "16 B in + 3 B out per pixel, L2 flushed before each launch"}
    tb_march = (args.march or os.environ.get("SNB_MARCH") or "lane")

    if rank == 0:
        hbm, tflops, src = measured_peaks()
        if present:
            present["frac"] = round(present["achieved_gbs"] / hbm, 4)
        mm = float(np.mean(march_ms)) if march_ms else 0.0
        kernel = "k_nerf_march_ws" if tb_march == "ws" else "k_nerf_march_fused"
        samples_per_launch = samples / max(1, K * frames_per_step) * (world if cfg.get("batch") else 1)
        rays_per_launch = rays / max(1, K * frames_per_step) * (world if cfg.get("batch") else 1)
        algo_bytes = samples / K * BYTES_PER_SAMPLE + rays / K * BYTES_PER_RAY
        if cfg.get("batch"):                              # counters and the march time are those of the LAST frame of a step
            algo_bytes = (c["samples"] * BYTES_PER_SAMPLE + c["rays"] * BYTES_PER_RAY)
        ach = algo_bytes / max(mm, 1e-9) / 1e6            # bytes / ms -> GB/s
        traffic = None
        tp = os.path.join(ROOT, "profiles", "fused_traffic.json")
        if os.path.exists(tp):
            with open(tp) as f:
                traffic = json.load(f).get(kernel, {}).get(args.config)
        l2 = l2_peaks()
        roof = {"kernel": kernel, "bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s", "frac": ach / hbm, "traffic": traffic,
                "peak_source": f"{src} (MEASURED_PEAKS.json hbm_gbs)", "launch_ms": mm,
                "launch_ms_note": "CUDA events around the march on the context stream (all passes of the batch-schedule iteration; later passes return at once when the first verified); the mesh layer runs concurrently on the second stream",
                "alone": ({"launch_ms": march_alone_ms, "achieved": algo_bytes / march_alone_ms / 1e6, "frac": algo_bytes / march_alone_ms / 1e6 / hbm,
                           "note": "same kernel, same frames, mesh layer not overlapped (option overlap=0)"} if march_alone_ms else None),
                "note": "algorithmic bytes = samples x 512 B hash gather + rays x 96 B. With the 2^19 table (24 MB) the gathers are L2 hits by design and DRAM traffic << algorithmic bytes: see the l2 view; with the 2^22 table (c4, 268 MB > L2) ncu still shows 96 % L2 hits (the touched slice fits) and the march time of a frame covers 2-3 passes of the batch-schedule iteration",
                "tensor": {"achieved_tflops": (samples / K if not cfg.get("batch") else c["samples"]) * FLOP_PER_SAMPLE / max(mm, 1e-9) / 1e9, "peak_tflops": tflops,
                           "note": "both MLPs, padded FLOP, against bf16_tflops_sustained"}}
        if l2:
            spp = sectors_per_sample()                 # 32 B sectors the L2 serves per evaluated sample, counted by ncu (the rest are L1 hits)
            l2_ach = samples_per_launch * spp * 32 / max(mm, 1e-9) / 1e6
            roof["l2"] = {"achieved": l2_ach, "peak": l2.get("l2_gather_sector_gbs"), "unit": "GB/s", "frac": (l2_ach / l2["l2_gather_sector_gbs"]) if l2.get("l2_gather_sector_gbs") else None,
                          "sectors_per_sample": spp,
                          "peak_source": "profiles/r02_l2_gather_peak.json: random 4-byte reads from a 24 MB table, counted as 32-byte sectors; sectors per sample from the ncu capture (profiles/fused_traffic.json)"}
            roof["hash_grid"] = {"achieved_l2_gbs": samples_per_launch * BYTES_PER_SAMPLE / max(mm, 1e-9) / 1e6,
                                 "sectors_gbs": samples_per_launch * 16 * 8 * 32 / max(mm, 1e-9) / 1e6, "note": "feature bytes gathered / march time; as 32-byte sectors (4 useful bytes each)"}
        nframes = K * frames_per_step
        out = {
            "metric": METRIC, "value": n_rays_step * K / (total_ms * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": world, "steps": K, "warmup": Wu,
            "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32 march / f16 MLP (fp32 accumulate)", "data": "synthetic",
            "config": {"workload": cfg["workload"], "name": args.config, "resolution": [W, H], "frames_per_step": frames_per_step, "march": tb_march,
                       "parallelism": (f"{world} row bands, halo {halo}px, edges balanced from measured frame times: {bal.bounds} ({bal.changes} moves); rgba + depth assembled on rank 0 (NCCL p2p, 20 B/px)" if split_frame
                                       else (f"{cfg['frames']} frames sharded over {world} ranks ({args.batch_assign}{': chunk edges ' + str(batch_bounds) if batch_bounds else ''}), whole frames, no halo" if cfg.get("batch") else "single GPU")),
                       "l2": "256 MB buffer written between timed iterations", "samples_per_frame": samples_per_launch, "samples_reused_per_frame": reused / max(1, K * frames_per_step) * (world if cfg.get("batch") else 1),
                       "samples_note": "samples_per_frame = network evaluations (hash gather + both MLPs) the march performs; samples_reused = samples the reference evaluates a second time after its compositor's rewind and this kernel composites from the first evaluation (option dup_reuse); their sum is the reference's sample count", "rays_marched_per_frame": rays_per_launch,
                       "march_stats_last_frame": march_stats, "alt_schedule": alt, "ms_per_frame": total_ms / nframes,
                       "stage_ms": {k_: round(v_, 4) for k_, v_ in stage_acc.items()},
                       "step_ms_min_med_p95_max": [round(float(np.min(step_ms)), 4), round(float(np.median(step_ms)), 4), round(float(np.percentile(step_ms, 95)), 4), round(float(np.max(step_ms)), 4)],
                       "present_stage": present},
            "e2e": {"value": n_rays_step * K / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": h2d * frames_per_step, "d2h_bytes_per_step": d2h, "ms_per_step": e2e_s / K * 1e3,
                    "synchronous": ({"value": n_rays_step * K / e2e_sync_s / 1e6, "ms_per_step": e2e_sync_s / K * 1e3, "note": "snb_frame_host: returns with the host buffers filled, no overlap between frames"} if e2e_sync_s else None),
                    "note": ("scene upload + snb_frame_host_async into two pinned host buffer pairs: a frame's device-to-host copy overlaps the next frame's render; all copies complete inside the timed region" if (eng is not None) else "stock render_nerf + D2H into pinned host buffers") if not split_frame else "scene upload + bands + assembly on rank 0 + D2H of the assembled rgba + depth frame into pinned host memory (double-buffered: the copy overlaps the next frame; all copies complete inside the timed region)"},
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": roof,
        }
        if world == 1 and not args.no_cuda_baseline:
            try:
                res, why = reference_cuda_frames(cfg, 5, 2)
            except Exception as e:  # noqa: BLE001
                res, why = None, f"{type(e).__name__}: {e}"
            out["cuda_baseline"] = res if res is not None else {"unavailable": why}
            if res is not None:
                out["cuda_baseline"]["speedup_device"] = out["value"] / res["value"]
        if world == 1 and not args.no_cpu_baseline:
            cw, ch = [int(x) for x in args.cpu_sample.split("x")]
            mrays, ms = cpu_oracle_rate(cfg, cw, ch, 1, 1, 0)
            out["cpu_baseline"] = {"value": mrays, "unit": "Mrays/s", "cores": 1, "kind": "port", "sample": f"one {cw}x{ch} frame of the same scene/camera (oracle, scalar C++)", "ms": ms}
        emit(json.dumps(out))
    tb.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0")); local_rank = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit("launch with torchrun --nproc-per-node N for --gpus N")
    run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
