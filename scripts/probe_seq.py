"""One GPU: ms/frame of a pipelined scene1 sequence at the image sizes the weak-scaling bench uses
(1920 x 1080*N).  Separates "the taller image costs more per row" from multi-GPU overheads."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from __graft_entry__ import load_package
rt = load_package()
r = rt.Renderer(0)
ps = rt.PackedScene.load(os.path.join(ROOT, "tests/golden/packed/scene1.npz"))
r.set_scene(ps)
for n in (1, 2, 4, 8):
    W, H = 1920, 1080 * n
    fr = rt.make_frame(ps.frame, W, H, max_bounces=2)
    r.render_sequence([fr] * 20)
    best = 1e9
    for _ in range(3):
        _, st = r.render_sequence([fr] * 50)
        best = min(best, st[0].ms_render)
    rays = st[0].rays_primary + st[0].rays_shadow
    print(json.dumps({"H": H, "ms_per_frame": best, "ms_per_1080_rows": best / n, "rays": rays, "Grays_s": rays / best / 1e6}), flush=True)
