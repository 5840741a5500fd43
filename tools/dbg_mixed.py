import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import forma_b200
from oracle import binding as ob
from test_oracle_analytic import plane, sphere, tri
rng = np.random.default_rng(5)
objs = [sphere((0, 0, -3), 1.0, mat=0, col=(.8, .3, .3), id=10),
        plane((0, 0, -2.5), (0, 0, -1), (0.6, 0.6, 5), mat=1, col=(.9, .9, .9), id=11),
        tri((-3, -1, -4), (3, -1, -4), (0, 3, -4), id=12),
        sphere((0, 0, -3), 1.0, mat=1, col=(.1, .9, .1), fuzz=0.3, id=13),
        plane((5, 0, -2), (0, 0, -1), (0.5, 0.5, 5), mat=0, col=(.2, .2, .9), id=14),
        tri((-3, -1, -4), (3, -1, -4), (0, 3, -4), id=15),
        sphere((1.5, 0.5, -2), 0.4, mat=2, col=(.9, .9, .9), id=16),
        plane((0.45, -1.2, 3), (0, 0.2, -1), (0.15, 9, 9), mat=0, col=(.5, .5, .1), id=17),
        sphere((-1.5, 0.5, -2), 0.4, mat=3, col=(.1, .1, .1), id=18)]
for _ in range(6):
    c = rng.uniform(-2, 2, 3); c[2] -= 4
    objs.append(sphere(c, rng.uniform(.2, .6), mat=int(rng.integers(0, 5)), col=rng.uniform(.1, .9, 3), fuzz=rng.uniform(0, 1), id=100 + _))
objs = np.stack(objs)
w, h = 96, 64
m = forma_b200.create_model(w, h, seed=77)
m.scene.load_objects(objs)
cam = ob.camera_new(w, h)
ids_g, t_g = m.primary_hits(); ids_o, t_o = ob.primary_hits(objs, cam, w, h)
print("ids equal", np.array_equal(ids_g, ids_o), "t equal", np.array_equal(t_g.view(np.uint32), t_o.view(np.uint32)))
print("gpu ids", np.unique(ids_g, return_counts=True)); print("ora ids", np.unique(ids_o, return_counts=True))
bad = np.argwhere(ids_g != ids_o)
for y, x in bad[:8]: print(y, x, ids_g[y, x], ids_o[y, x], t_g[y, x], t_o[y, x])
for spp in (1, 8):
    rgb_o, lin_o, rays_o = ob.render(objs, cam, w, h, spp, seed=77)
    rgb_g, lin_g = m.render(spp, want_linear=True)
    d = lin_g[..., :3].view(np.uint32) != lin_o.view(np.uint32)
    print("spp", spp, "rays", m.stats().rays, rays_o, "mismatching px", int(d.any(axis=2).sum()))
    for y, x in np.argwhere(d.any(axis=2))[:8]: print("  ", y, x, ids_o[y, x], lin_g[y, x, :3], lin_o[y, x])
