"""Every bundled scene of the reference (Scenes/scene1..9.dat, packed fixtures) at 1920x1080 with the
depth / energy cut-off the scene file asks for: ms/frame through the synchronous call (device time
of the frame and wall time per call), rays, depth levels, kernel launches.  One GPU."""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from __graft_entry__ import load_package

ap = argparse.ArgumentParser()
ap.add_argument("--width", type=int, default=1920)
ap.add_argument("--height", type=int, default=1080)
ap.add_argument("--frames", type=int, default=20)
args = ap.parse_args()
rt = load_package()
r = rt.Renderer(0)
W, H = args.width, args.height
for i in range(1, 10):
    ps = rt.PackedScene.load(os.path.join(ROOT, "tests", "golden", "packed", f"scene{i}.npz"))
    r.set_scene(ps)
    fr = rt.make_frame(ps.frame, W, H)                       # depth and min energy from the scene file
    for _ in range(3):
        r.render(fr, want_f32=False, want_rgba8=False)
    dev, t0 = 0.0, time.perf_counter()
    for _ in range(args.frames):
        _, _, st = r.render(fr, want_f32=False, want_rgba8=False)
        dev += st.ms_render
    wall = (time.perf_counter() - t0) * 1e3 / args.frames
    _, seq = r.render_sequence([fr] * args.frames)
    T, Q, P, L = ps.counts
    row = {"scene": i, "tris": T, "quadrics": Q, "planes": P, "lights": L, "max_bounces": fr.max_bounces,
           "levels": st.depth_levels, "launches": st.kernel_launches, "rays": st.rays,
           "ms_device": dev / args.frames, "ms_wall_per_call": wall, "ms_sequence": seq[0].ms_render,
           "Mrays_s_device": st.rays / (dev / args.frames) / 1e3}
    print(json.dumps(row), flush=True)
