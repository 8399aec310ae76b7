import numpy as np, sys
sys.path.insert(0,'.')
from computeraytracer_b200 import scenes
from computeraytracer_b200.renderer import Renderer
from tests import orc
ps = scenes.mesh_scene(64, 64, n_lon=64, n_lat=32)
r = Renderer(0); r.set_scene(ps)
rays = scenes.camera_rays(ps, 20000, seed=5)
gi, gt = r.trace_rays(rays)
ci, ct, _ = orc.load().scene(ps).trace_rays(rays)
bad = np.nonzero(gi != ci)[0]
print("mismatch", len(bad), "of", len(gi), r.scene_stats())
for k in bad[:12]:
    print(k, "gpu", gi[k], gt[k], "cpu", ci[k], ct[k], "o", rays[k,0:3], "d", rays[k,4:7])
