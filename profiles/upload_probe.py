import sys, time, ctypes, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, 'tests'); sys.path.insert(0, 'scenes')
import numpy as np, torch, bench
import cs378hgraphics_raytracer_b200 as rt
wl = bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else 'c5']
a = bench.load_arena(wl); ab = a.bytes()
pinned = torch.empty(len(ab), dtype=torch.uint8, pin_memory=True); pinned.numpy()[...] = np.frombuffer(ab, dtype=np.uint8)
ctx = rt.Context(0)
def up(): ctx._check(rt.lib().rt_upload_scene(ctx._h, ctypes.c_void_p(pinned.data_ptr()), pinned.numel()))
up(); up()
t = time.perf_counter()
for _ in range(10): up()
print('upload ms', (time.perf_counter() - t) * 100, 'MB', len(ab) / 1e6)
d = torch.empty(len(ab), dtype=torch.uint8, device='cuda')
torch.cuda.synchronize(); t = time.perf_counter()
for _ in range(10): d.copy_(pinned, non_blocking=True)
torch.cuda.synchronize(); print('plain H2D ms', (time.perf_counter() - t) * 100)
t = time.perf_counter()
for _ in range(10): rt.lib().rt_arena_validate(ctypes.c_void_p(pinned.data_ptr()), pinned.numel())
print('full validate ms', (time.perf_counter() - t) * 100)
