import sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, 'tests'); sys.path.insert(0, 'scenes')
import numpy as np, bench
import cs378hgraphics_raytracer_b200 as rt
import oracle_binding as ob
wl = bench.WORKLOADS['c4a']
path, faces = bench.scene_files(wl)
arena = rt.load_ray_file(path, faces)
c = rt.Context(0); c.upload(arena)
p = rt.Params.default(**wl['params'])
w = h = 2048
a, sa = c.render(p, w, h)
b, sb = c.render(p, w, h)
print('run-to-run differing px', int((a != b).any(axis=-1).sum()), 'camera rays', sa.camera_rays, sb.camera_rays)
out = np.zeros_like(a); cam = 0
for part in range(4):
    _, st = c.render(p, w, h, part, 4, out=out); cam += st.camera_rays
df = (out.astype(int) - a.astype(int))
bad = np.argwhere(np.abs(df).max(axis=-1) > 0)
print('partition differing px', len(bad), 'max abs', np.abs(df).max(), 'camera', cam, sa.camera_rays)
print(bad[:10], [(tuple(x), out[x[0], x[1]].tolist(), a[x[0], x[1]].tolist()) for x in bad[:5]])
print('tile coords of bad', sorted(set((int(y) // 32, int(x) // 32) for y, x in bad))[:20])
