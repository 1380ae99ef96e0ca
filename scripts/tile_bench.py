"""The reference front-end's calling pattern (satellite: 128x128 tiles, 8 spp per call, Config.hpp defaults) through
terra_render on the C2 scene, 1920x1080: calls/s and Mpaths/s. Run with TERRA_B200_GRAPH_PATHS=0 for the
kernel-by-kernel path, default for CUDA-graph replay.  python scripts/tile_bench.py [tile] [spp] [passes] [bounces]"""
import ctypes as C
import json
import os
import sys
import time

sys.path.insert(0, ".")
from terra_b200 import api, scenes

tile = int(sys.argv[1]) if len(sys.argv) > 1 else 128
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 8
passes = int(sys.argv[3]) if len(sys.argv) > 3 else 3
bounces = int(sys.argv[4]) if len(sys.argv) > 4 else 4
lib = api.TerraLib()
s, cam = scenes.sphere_field(lib, spp=spp, bounces=bounces, ggx=True)
w, h = 1920, 1080
fb = api.Framebuffer()
assert lib.terra_framebuffer_create(C.byref(fb), w, h)
tiles = [(x, y, min(tile, w - x), min(tile, h - y)) for y in range(0, h, tile) for x in range(0, w, tile)]


def frame():
    for (x, y, tw, th) in tiles:
        lib.terra_render(C.byref(cam), s.h, C.byref(fb), x, y, tw, th)


frame(); frame()                       # warm-up: buffers sized, graphs captured
lib.terra_b200_reset_stats(s.h)
t0 = time.time()
for _ in range(passes):
    frame()
dt = time.time() - t0
st = s.stats()
print(json.dumps({"tile": tile, "spp": spp, "bounces": bounces, "graph_paths": os.environ.get("TERRA_B200_GRAPH_PATHS", "default"),
                  "calls_per_s": passes * len(tiles) / dt, "us_per_call": dt / (passes * len(tiles)) * 1e6,
                  "mpaths_per_s": passes * w * h * spp / dt / 1e6, "graph_replays": int(st.graph_replays),
                  "kernel_launches": int(st.kernel_launches), "render_ms_device": st.render_ms}))
