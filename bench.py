#!/usr/bin/env python
"""bench.py — headline benchmark of the SyNeRFgine render hot path on B200 (BASELINE.json: Mrays/s and ms/frame,
1080p NeRF + mesh + shadows, at 1/2/4/8 B200).

    python bench.py --gpus 1 --steps K --warmup W            # this repo's CUDA path (libsnb.so) through the C ABI
    python bench.py --impl reference ...                      # the reference algorithm on the host cores (CPU oracle)
    torchrun --nproc-per-node N bench.py --gpus N ...         # one frame split into N column strips, NCCL gather

A "step" is one complete frame (snb_frame: mesh layer, NeRF march, both shadow directions, overlay) of the C2 workload:
1920x1080, base.json-architecture NeRF (16 levels x 2 features, 2^19 table, 64-wide MLPs; synthetic seeded weights),
armadillo lambertian + 1 point light, nerf_shadow_samples 4, syn_shadow_samples 2. value = primary rays / device time
with everything resident in HBM; e2e = the same through snb_frame_host with pinned HOST output buffers and the per-frame
scene upload inside the timed region. One JSON line on stdout (rank 0).
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
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOAD = "C2: 1920x1080 fox-architecture NeRF (16x2 hash grid, T=2^19, 64-wide MLPs) + armadillo.obj lambertian + 1 point light, nerf_shadow_samples=4, syn_shadow_samples=2"
BYTES_PER_SAMPLE = 512          # hash gather: 16 levels x 8 corners x 2 features x 2 B
BYTES_PER_RAY = 64 + 20 + 12    # queue record read (ray + its SH-4 block) + rgba/depth write + position write
FLOP_PER_SAMPLE = 20480         # both MLPs, padded


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--scene", default="c2")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-sample", default="240x135")
    return ap.parse_args()


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return d.get("hbm_gbs", 6650.0), d.get("bf16_tflops_sustained", 1400.0), "measured"
    return 6650.0, 1400.0, "fallback"


def claim_stdout():
    """stdout carries exactly ONE JSON line: anything a library prints there (NCCL's version banner) goes to stderr."""
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


# ---- CPU oracle legs (cpu_baseline and --impl reference) -----------------------------------------------------------------
_W = {}


def _oracle_worker_init(scene_name):
    import common
    import oracle_lib as O
    from synerfgine_b200 import scenes
    from synerfgine_b200.scene import Scene as HostScene
    _W["O"] = O; O.load()
    _W["model"] = common.model("fox16x2")
    _W["desc"] = _W["model"].desc()
    _W["scene"] = HostScene(scenes.scene_json(scene_name), scenes.mesh_dir())


def _oracle_frame(args):
    w, h, spp = args
    import common
    O, sc = _W["O"], _W["scene"]
    v = common.make_view(w, h, cam=sc.camera_matrix(), spp=spp)
    t0 = time.perf_counter()
    O.frame(_W["desc"], v, sc, sc.rt, 1999, 1)
    return time.perf_counter() - t0


def cpu_oracle_rate(scene_name, w, h, procs, steps, warmup):
    """Mrays/s of the oracle frame loop on `procs` host processes; each step = every process renders one w x h frame."""
    import multiprocessing as mp
    ctx = mp.get_context("fork")
    with ctx.Pool(procs, initializer=_oracle_worker_init, initargs=(scene_name,)) as pool:
        for i in range(warmup):
            pool.map(_oracle_frame, [(w, h, i)] * procs)
        t0 = time.perf_counter()
        for i in range(steps):
            pool.map(_oracle_frame, [(w, h, warmup + i)] * procs)
        dt = time.perf_counter() - t0
    return procs * steps * w * h / dt / 1e6, dt / steps * 1e3


def run_reference(args, rank, world):
    """Reference arm: the reference's algorithm for this path on the host cores. The reference ships no CPU path and its
    CUDA application cannot be built here (DESIGN.md §Reference build), so this is the oracle port of the same frame loop,
    on every host core, on a bounded sample of the workload (same scene and camera at reduced resolution)."""
    if rank != 0:
        return
    procs = os.cpu_count() or 1
    budget_s = 150.0 / max(1, args.steps + args.warmup)          # whole run within a few minutes
    rays = int(max(32 * 18, min(240 * 135, budget_s / 0.6e-3)))  # ~0.6 ms of CPU per ray for C2 (measured)
    h = max(18, int((rays / (16 / 9)) ** 0.5)); w = h * 16 // 9
    mrays, ms = cpu_oracle_rate(args.scene, w, h, procs, args.steps, args.warmup)
    sample = f"{procs} processes x one {w}x{h} frame of the same scene/camera per step"
    print(json.dumps({
        "impl": "reference", "metric": "Mrays/s, 1080p NeRF+mesh+shadows", "value": mrays, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32 march / f16 MLP (fp32 accumulate)",
        "data": "synthetic", "config": {"workload": WORKLOAD, "sample": sample},
        "cpu_baseline": {"value": mrays, "unit": "Mrays/s", "cores": procs, "kind": "port", "sample": sample},
        "e2e": {"value": mrays, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def run_ours(args, rank, local_rank, world):
    emit = claim_stdout()
    import torch
    import torch.distributed as dist
    import common
    from synerfgine_b200 import Engine, Testbed, scenes, tiles
    from synerfgine_b200._lib import Tile

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    W, H, K, Wu = args.width, args.height, args.steps, max(3, args.warmup)
    if world > 1:
        Wu = max(Wu, 20)                                  # the strip edges settle during warm-up (tiles.StripBalancer); reported as `warmup`
    stream = torch.cuda.Stream()
    tb = Testbed(local_rank, stream=stream)
    model = common.model("fox16x2")
    tb.set_model(model)                                   # weights / bitfield replicated once per GPU
    eng = Engine()
    eng.set_virtual_world_json(scenes.scene_json(args.scene), scenes.mesh_dir())
    halo = tiles.halo_for(eng.rt.nerf_shadow_kernel, bool(eng.rt.show_shadow_on_nerf))
    align = max(1, eng.rt.syn_px_scale)
    bal = tiles.StripBalancer(W, world, halo, align)       # starts from equal strips, moves the edges from measured frame times
    strip = bal.strip(rank)
    eng.init(W, H, tb, tile=Tile(strip.hx0, 0, strip.render_width, H) if world > 1 else None)
    lib, ctx = tb.lib, tb.ctx
    n_rays_total = W * H
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")   # > 126 MB L2
    full_frame = torch.empty((H, W, 4), device="cuda") if (world > 1 and rank == 0) else None
    ev_a = torch.cuda.Event(enable_timing=True); ev_b = torch.cuda.Event(enable_timing=True)
    col_cost = torch.zeros(W, device="cuda", dtype=torch.float32) if world > 1 else None

    def one_frame():
        lib.snb_reset_accumulation(ctx)                   # accumulation reset before each timed frame
        if world == 1:
            return eng.frame(sync=False)[0]
        s = bal.strip(rank)
        ev_a.record(stream)
        rgba, depth = eng.frame(sync=False, tile=Tile(s.hx0, 0, s.render_width, H))
        ev_b.record(stream)
        mine = rgba[:, s.crop].contiguous()
        tiles.assemble_on_root(mine, bal.bounds, rank, world, dist, full_frame)   # framebuffer assembly over NVLink (NCCL p2p), once per frame
        if bal.due():                                     # strip edges follow the measured cost: one 7.7 KB all_reduce
            hits = (mine[..., 3] > 0).sum(0, dtype=torch.float32)
            wgt = hits + 0.05 * torch.clamp(hits.max(), min=1.0)
            ev_b.synchronize()
            col_cost.zero_()
            col_cost[s.x0:s.x1] = wgt * (ev_a.elapsed_time(ev_b) / float(wgt.sum()))
            dist.all_reduce(col_cost)
            t_ms = ev_a.elapsed_time(ev_b)
            moved = bal.observe(col_cost.cpu().numpy().astype(np.float64))
            if os.environ.get("SNB_BENCH_DEBUG"):
                print(f"[balance] frame {bal.frames} rank {rank} strip {s.x0}-{s.x1} render {t_ms:.3f} ms moved={moved} -> {bal.bounds}", file=sys.stderr)
        bal.tick()
        return rgba

    with torch.cuda.stream(stream):
        for _ in range(Wu):
            one_frame()
        stream.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        launches0 = tb.launch_count()
        sampler = ClockSampler(local_rank); sampler.start()
        step_ms, march_ms, stage_acc, samples, rays = [], [], {}, 0, 0
        for _ in range(K):
            flush.zero_()                                  # L2 flush between timed iterations (outside the event pair)
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            one_frame()
            e1.record(stream)
            e1.synchronize()
            step_ms.append(e0.elapsed_time(e1))
            st = tb.stage_ms(); march_ms.append(st["nerf_march"])
            for k_, v_ in st.items():
                stage_acc[k_] = stage_acc.get(k_, 0.0) + v_ / K
            c = tb.counters(); samples += c["samples"]; rays += c["rays"]
            march_stats = tb.march_stats()
            march_stats["ws_pipeline"] = {k_: round(float(v_), 4) for k_, v_ in tb.ws_profile().items()}
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        clocks = sampler.stop()
        launches = tb.launch_count() - launches0
        total_ms = torch.tensor([sum(step_ms)], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)     # device time, max over ranks
        total_ms = float(total_ms.item())

        # ---- end to end: host buffers, scene upload + frame + D2H inside the timed region --------------------------------
        eng_frame_host = eng.frame
        sc = eng.scene
        h2d = sc.n_objects * 72 + sc.n_materials * C.sizeof(type(sc.materials[0])) + sc.n_lights * C.sizeof(type(sc.lights[0])) + 96
        strip = bal.strip(rank)
        tile = Tile(strip.hx0, 0, strip.render_width, H) if world > 1 else None
        tw = strip.render_width if world > 1 else W
        pin_rgba = torch.empty((H, tw, 4), dtype=torch.float32).pin_memory(); pin_depth = torch.empty((H, tw), dtype=torch.float32).pin_memory()
        d2h = pin_rgba.numel() * 4 + pin_depth.numel() * 4

        def one_frame_host():
            lib.snb_update_scene(ctx, sc.objects, sc.n_objects, sc.materials, sc.n_materials, sc.lights, sc.n_lights)   # per-frame H2D of the scene PODs
            v = tb.make_view(W, H, eng.frame_index); eng.frame_index += 1
            rc = lib.snb_frame_host(ctx, C.byref(v), C.byref(tile) if tile is not None else None, C.byref(eng.rt), pin_rgba.data_ptr(), pin_depth.data_ptr())
            assert rc == 0, lib.snb_last_error(ctx)
        for _ in range(2):
            one_frame_host()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(K):
            one_frame_host()
        torch.cuda.synchronize()
        e2e_s = torch.tensor([time.perf_counter() - t0], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
        e2e_s = float(e2e_s.item())

        # ---- the march kernel without the mesh layer running beside it (same frames, overlap off): explains `roofline` ----
        march_alone_ms = None
        if world == 1:
            lib.snb_set_option(ctx, b"overlap", b"0")
            alone = []
            for i in range(2 + 5):
                flush.zero_()
                one_frame(); stream.synchronize()
                if i >= 2:
                    alone.append(tb.stage_ms()["nerf_march"])
            lib.snb_set_option(ctx, b"overlap", b"1")
            march_alone_ms = float(np.mean(alone))

        # ---- present stage, timed on its own: FXAA blit of the final frame into an RGB8 window image ----
        present = None
        if world == 1:
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

    if rank == 0:
        hbm, tflops, src = measured_peaks()
        if present:
            present["frac"] = round(present["achieved_gbs"] / hbm, 4)
        mm = float(np.mean(march_ms))
        algo_bytes = (samples * BYTES_PER_SAMPLE + rays * BYTES_PER_RAY) / K
        ach = algo_bytes / (mm * 1e-3) / 1e9
        traffic = None
        tp = os.path.join(ROOT, "profiles", "fused_traffic.json")
        if os.path.exists(tp):
            with open(tp) as f:
                traffic = json.load(f).get("dram_bytes_per_launch")
        out = {
            "metric": "Mrays/s, 1080p NeRF+mesh+shadows", "value": n_rays_total * K / (total_ms * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": world, "steps": K, "warmup": Wu,
            "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32 march / f16 MLP (fp32 accumulate)", "data": "synthetic",
            "config": {"workload": WORKLOAD, "resolution": [W, H], "scene": args.scene, "parallelism": (f"{world} column strips, halo {halo}px, edges balanced from measured frame times: {bal.bounds} ({bal.changes} moves)" if world > 1 else "single GPU"),
                       "l2": "256 MB buffer written between timed iterations", "samples_per_frame": samples / K, "rays_marched_per_frame": rays / K, "march_stats_last_frame": march_stats,
                       "stage_ms": {k_: round(v_, 4) for k_, v_ in stage_acc.items()},
                       "step_ms_min_med_max": [round(float(np.min(step_ms)), 4), round(float(np.median(step_ms)), 4), round(float(np.max(step_ms)), 4)],
                       "present_stage": present},
            "e2e": {"value": n_rays_total * K / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": e2e_s / K * 1e3},
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"kernel": "k_nerf_march_fused", "bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s", "frac": ach / hbm, "traffic": traffic,
                         "peak_source": f"{src} (MEASURED_PEAKS.json hbm_gbs)", "launch_ms": mm,
                         "launch_ms_note": "CUDA events around the march launch on the context stream; the mesh layer runs concurrently on the second stream",
                         "alone": ({"launch_ms": march_alone_ms, "achieved": algo_bytes / (march_alone_ms * 1e-3) / 1e9, "frac": algo_bytes / (march_alone_ms * 1e-3) / 1e9 / hbm,
                                    "note": "same kernel, same frames, mesh layer not overlapped (option overlap=0)"} if march_alone_ms else None),
                         "note": "algorithmic bytes = samples x 512 B hash gather + rays x 64 B; the 24 MB table is L2-resident by design, so DRAM traffic << algorithmic bytes",
                         "tensor": {"achieved_tflops": samples / K * FLOP_PER_SAMPLE / (mm * 1e-3) / 1e12, "peak_tflops": tflops}},
        }
        if world == 1 and not args.no_cpu_baseline:
            cw, ch = [int(x) for x in args.cpu_sample.split("x")]
            mrays, ms = cpu_oracle_rate(args.scene, cw, ch, 1, 1, 0)
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
