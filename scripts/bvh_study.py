"""Dump a config's triangles + camera for scripts/bvh_proto.cpp and run it (CPU study tool)."""
import os, subprocess, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ray_tracer_b200 import configs

name = sys.argv[1] if len(sys.argv) > 1 else "c3_teapot"
step = sys.argv[2] if len(sys.argv) > 2 else "8"
scene, info = configs.build(name)
cam, sp, vs, ix, ms = scene.as_tuple()
tris = []
for m in ms:
    idx = ix[int(m["first"]):int(m["first"]) + 3 * int(m["triangles"])].astype(np.int64) + int(m["offset"])
    p = vs["pos"][idx].astype(np.float32) + m["pos"].astype(np.float32)
    tris.append(p.reshape(-1, 9))
tris = np.concatenate(tris).astype(np.float32)
c = np.concatenate([cam["origin"][0][:3] if cam["origin"].ndim > 1 else cam["origin"][:3]] if False else
                   [np.ravel(cam["origin"])[:3], np.ravel(cam["lower_left_corner"])[:3], np.ravel(cam["horizontal"])[:3], np.ravel(cam["vertical"])[:3]]).astype(np.float32)
path = "/tmp/%s.bin" % name
with open(path, "wb") as f:
    f.write(tris.tobytes()); f.write(c.tobytes())
print(name, len(tris), "triangles")
W, H = info["params"]["width"], info["params"]["height"]
subprocess.check_call(["/tmp/bvh_proto", path, str(W), str(H), step])
