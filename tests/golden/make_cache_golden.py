"""Golden fixtures for the bvhcache file naming and format (Renderer.cc:176-211), written by the REFERENCE'S OWN code
(oracle/_ref/ref_cpu --cache-key / --bvh-out: Scene::hash, Platform::computeHash, BuildParams::computeHash, CudaBVH::serialize):

    python tests/golden/make_cache_golden.py      (needs /root/reference; run in the build container)

Writes tests/golden/cache_keys.json (the key of each fixture scene) and tests/golden/bvhcache/<key>.dat for the room scene --
the file a reference build would leave in its bvhcache directory for that mesh."""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as O      # noqa: E402

keys = {}
os.makedirs(os.path.join(HERE, "bvhcache"), exist_ok=True)
for name in ("room", "blob"):
    obj = os.path.join(HERE, name + ".obj")
    line = [l for l in O.run_ref_cpu("--obj", obj, "--cache-key").splitlines() if l.startswith("{")][0]
    keys[name] = json.loads(line)
    if name == "room":
        O.run_ref_cpu("--obj", obj, "--bvh-out", os.path.join(HERE, "bvhcache", keys[name]["cache_key"] + ".dat"))
with open(os.path.join(HERE, "cache_keys.json"), "w") as f:
    json.dump(keys, f, indent=1, sort_keys=True)
print(keys)
