#!/usr/bin/env python
"""Per-CTA timeline of refine_f16_tiled_kernel (512x384 benchmark frames): staging time, tile life,
spread of the warps' finish times, levels run from global memory, a map of tile lives over the image.
Needs a library built with -DM3R_RT_PROF (not the product build):
  cd lightweight-mast3r-slam_b200 && nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo \
     -Xcompiler -fPIC -shared -DM3R_RT_PROF -o ../tools/_var_prof.so csrc/*.cu
  M3R_B200_LIB=$PWD/tools/_var_prof.so python tools/refine_prof.py"""
import ctypes, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("lightweight-mast3r-slam_b200", "oracle", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
sys.path.insert(0, ROOT)
import torch
import bench
import mast3r_slam_backends as be
from mast3r_slam_backends import _lib

variant = int(os.environ.get("VARIANT", 0))
sets = bench.make_match_sets("cuda", 8)
for s in sets:
    p, _ = be.iter_proj(s["rays"], s["pts"], s["p_init"], 10, 1e-8, 1e-6)
    s["p1"] = p.long()
N = 4096 * 16
buf = (ctypes.c_ulonglong * N)()
for i in range(9):
    s = sets[i % 8]
    be.refine_matches(s["D11"], s["D21"], s["p1"], 3, 5, _variant=variant)
    if i == 7:
        _lib.lib.m3r_debug_refine_profile(buf, N)  # reset before the measured call
_lib.lib.m3r_debug_refine_profile(buf, N)
a = np.frombuffer(buf, dtype=np.uint64).reshape(4096, 16).astype(np.int64)
a = a[a[:, 1] > 0]
t0 = a[:, 1].min()
start, box = a[:, 1] - t0, a[:, 2] - t0
wend = a[:, 3:11] - t0
end = wend.max(1)
print(f"tiles {len(a)}, span {end.max() / 1e3:.1f} us")
print(f"staging (start -> box ready): mean {np.mean(box - start) / 1e3:.2f} us, p90 {np.percentile(box - start, 90) / 1e3:.2f}")
print(f"tile life (start -> last warp): mean {np.mean(end - start) / 1e3:.1f} us, min {np.min(end - start) / 1e3:.1f}, max {np.max(end - start) / 1e3:.1f}")
print(f"warp finish spread per tile (last - first): mean {np.mean(wend.max(1) - wend.min(1)) / 1e3:.1f} us; "
      f"last - mean: {np.mean(wend.max(1) - wend.mean(1)) / 1e3:.1f} us")
print(f"fallback warp-levels: total {a[:, 15].sum()} over {np.count_nonzero(a[:, 15])} tiles")
# per-SM: busy intervals
sm = a[:, 0]
idle = []
for k in np.unique(sm):
    m = sm == k
    idle.append((k, m.sum(), end[m].max()))
idle = np.array(idle)
print(f"per SM: tiles min {idle[:, 1].min()} max {idle[:, 1].max()}; last finish min {idle[:, 2].min() / 1e3:.1f} "
      f"mean {idle[:, 2].mean() / 1e3:.1f} max {idle[:, 2].max() / 1e3:.1f} us")
# tile life by how many tiles were co-resident: first-wave vs last-wave
order = np.argsort(start)
print("start times (us) of tiles, deciles:", np.round(np.percentile(start, [0, 10, 30, 50, 70, 90, 100]) / 1e3, 1))
print("tile life (us) by start decile:", [round(float(np.mean((end - start)[order[i * len(a) // 10:(i + 1) * len(a) // 10]])) / 1e3, 1) for i in range(10)])
fb = a[:, 15] > 0
if fb.any():
    print(f"tile life with fallback {np.mean((end - start)[fb]) / 1e3:.1f} us vs without {np.mean((end - start)[~fb]) / 1e3:.1f} us")
# slowest tiles: index -> (tile row, strip)
ids = np.nonzero(np.frombuffer(buf, dtype=np.uint64).reshape(4096, 16)[:, 1] > 0)[0]
life = end - start
top = np.argsort(-life)[:16]
print("slowest tiles (row, strip, life us, fallback warp-levels):",
      [(int(ids[k] // 32), int(ids[k] % 32), round(float(life[k]) / 1e3, 1), int(a[k, 15])) for k in top])
grid = np.zeros((24, 32))
for k in range(len(a)):
    grid[ids[k] // 32, ids[k] % 32] = life[k] / 1e3
np.set_printoptions(linewidth=250, precision=0, suppress=True)
print(grid)
