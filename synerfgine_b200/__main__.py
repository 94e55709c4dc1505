"""Headless command line mirroring the reference's `src/main.cu` (flag names and meaning kept; no window, no GL):

    python -m synerfgine_b200 --snapshot scene.ingp --virtual scene.json --width 1920 --height 1080 --frames 24 --out out/frame_%04d.png

`--snapshot` ≙ Testbed::load_snapshot, `--virtual` / `--rt` ≙ Engine::set_virtual_world, `--width` / `--height` ≙ Engine::init,
`--sshadows` / `--nshadows` ≙ Engine::set_syn_samples / set_nerf_samples (main.cu:200-203). The render loop is `while (engine.frame())`
(main.cu:224): every iteration follows the scene's camera path and object animations, here for `--frames` iterations, each presented
through the display pass (≙ main.frag) and written as a PNG. `--synthetic PRESET` stands in for a snapshot when there is none (the
seeded models of synerfgine_b200/workloads.py). GUI-only flags of the reference (--no-gui, --vr, --frag, --train, --scene, network
config) are accepted and refused with a message: training and the windowed paths are outside this library's scope."""
import argparse
import json
import sys
import time


def build_parser():
    ap = argparse.ArgumentParser(prog="python -m synerfgine_b200", description="synerfgine_b200 headless renderer (flags of the reference's src/main.cu)")
    ap.add_argument("--snapshot", "--load_snapshot", dest="snapshot", help="snapshot (.ingp / .msgpack) to load upon startup")
    ap.add_argument("--rt", "--virtual", dest="virtual", help="path to the virtual scene config")
    ap.add_argument("--width", type=int, default=1280, help="resolution width (reference default 1280)")
    ap.add_argument("--height", type=int, default=720, help="resolution height (reference default 720)")
    ap.add_argument("--sshadows", type=int, help="number of samples of light source for shadows on synthetic surfaces")
    ap.add_argument("--nshadows", type=int, help="filter kernel size for shadows on NeRF surfaces")
    ap.add_argument("-v", "--version", action="store_true", help="display the version")
    ap.add_argument("files", nargs="*", help="files to be loaded: a snapshot and / or a virtual scene config (by extension)")
    # headless additions
    ap.add_argument("--synthetic", help="seeded synthetic model preset instead of a snapshot (fox16x2, small16x2, unbounded22, ...)")
    ap.add_argument("--frames", type=int, default=1, help="iterations of the render loop")
    ap.add_argument("--out", help="PNG path pattern, e.g. out/frame_%%04d.png (one file per frame); omitted: nothing is written")
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--option", action="append", default=[], metavar="KEY=VALUE", help="snb_set_option pairs (march=lane, bvh=lbvh, ...)")
    # the reference's GUI / training flags: recognised so that its command lines fail with a reason instead of a parse error
    for flag in ("--no-gui", "--vr", "--train"):
        ap.add_argument(flag, action="store_true", help=argparse.SUPPRESS)
    for flag in ("--frag", "-s", "--scene", "-n", "-c", "--network", "--config"):
        ap.add_argument(flag, help=argparse.SUPPRESS)
    return ap


def main(argv=None):
    args = build_parser().parse_args(argv)
    from . import _lib
    if args.version:
        print("synerfgine_b200 " + _lib.load().snb_version().decode())
        return 0
    for name in ("vr", "train", "frag", "scene", "network", "config"):
        if getattr(args, name, None):
            print(f"--{name}: windowed display, VR and training are outside this library's scope (SURVEY.md 8: render hot path only)", file=sys.stderr)
            return 2
    for f in args.files:                               # Testbed::load_file: by extension
        if f.endswith((".ingp", ".msgpack")):
            args.snapshot = args.snapshot or f
        elif f.endswith(".json"):
            args.virtual = args.virtual or f
        else:
            print(f"{f}: only snapshots (.ingp / .msgpack) and virtual scene configs (.json) can be loaded", file=sys.stderr)
            return 2
    if not args.snapshot and not args.synthetic:
        print("nothing to render: give --snapshot FILE (or --synthetic PRESET)", file=sys.stderr)
        return 2
    from .testbed import Engine, Testbed
    tb = Testbed(args.device)
    try:
        for kv in args.option:
            k, _, v = kv.partition("=")
            tb._check(tb.lib.snb_set_option(tb.ctx, k.encode(), v.encode()), f"option {kv}")
        if args.snapshot:
            tb.load_snapshot(args.snapshot)
        else:
            from . import workloads
            tb.set_model(workloads.model(args.synthetic))
        stats = []
        if args.virtual:
            eng = Engine()
            eng.set_virtual_world(args.virtual)
            w, h = args.width, args.height
            eng.init(w, h, tb)
            if args.sshadows is not None:              # main.cu:200-203 sets both when --sshadows is given
                eng.set_syn_samples(args.sshadows)
                if args.nshadows is not None:
                    eng.set_nerf_samples(args.nshadows)
            elif args.nshadows is not None:
                eng.set_nerf_samples(args.nshadows)
            for f in range(args.frames):
                t0 = time.perf_counter()
                eng.step()                             # camera path, animations, render (Engine::frame)
                ms = (time.perf_counter() - t0) * 1e3
                if args.out:
                    eng.save_image(args.out % f if "%" in args.out else args.out)
                stats.append({"frame": f, "ms": round(ms, 3), **{k: round(v, 3) for k, v in tb.stage_ms().items()}})
        else:
            import numpy as np
            for f in range(args.frames):
                t0 = time.perf_counter()
                img = tb.render(args.width, args.height, spp=1, linear=False)
                ms = (time.perf_counter() - t0) * 1e3
                if args.out:
                    rgb8 = np.ascontiguousarray((np.clip(img[..., :3], 0.0, 1.0) * 255.0 + 0.5).astype(np.uint8))
                    path = args.out % f if "%" in args.out else args.out
                    tb._check(tb.lib.snb_write_png(path.encode(), rgb8.ctypes.data, args.width, args.height), "snb_write_png")
                stats.append({"frame": f, "ms": round(ms, 3)})
        print(json.dumps({"frames": len(stats), "width": args.width, "height": args.height, "per_frame": stats[:64]}))
        return 0
    finally:
        tb.close()


if __name__ == "__main__":
    sys.exit(main())
