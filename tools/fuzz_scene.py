"""Randomised check of the scene loader (ovis_scene_load, host only) on DAMAGED worlds: the committed fixture with
properties dropped, retyped or given odd values, references to objects that do not exist, entities listed as their own
descendants, components on two entities, duplicated ids, and plain byte damage. Every file must be either loaded or
refused with an error - never crash or hang (the batches run in child processes under a time limit) - and a file the
loader accepts must be accepted by the Python restatement of the reference's loader too, with the same arrays, and the
other way round.
Runs on the CPU.

    python tools/fuzz_scene.py [--cases 600] [--seed 1]"""
import argparse
import copy
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
SCENE = os.path.join(ROOT, "tests", "golden", "scene")
FOREST = os.path.join(SCENE, "worlds", "forest.object")
KEYS = ("parent", "local_trs", "local_aabb", "flags", "entity_node", "lights", "shading", "cameras")


def odd_value(rng):
    return [None, True, False, 0, -1, 1e30, -0.0, "x", "", [], {}, [1], [1, 2], [1, 2, 3, 4, 5], {"objectID": 99999}, {"asset": "game/models/none"},
            {"asset": "engine/x"}, {"asset": 3}, [[1, 2, 3]], ["a", "b", "c"], 1e-40, 2**40][int(rng.integers(0, 22))]


def damage(rng, objects):
    objects = copy.deepcopy(objects)
    for _ in range(int(rng.integers(1, 5))):
        o = objects[int(rng.integers(0, len(objects)))]
        k = int(rng.integers(0, 12))
        props = o.get("objectProperties") if isinstance(o.get("objectProperties"), dict) else None
        if k == 0 and props:
            props.pop(str(rng.choice(list(props.keys()))), None)
        elif k == 1 and props:
            props[str(rng.choice(list(props.keys())))] = odd_value(rng)
        elif k == 2:
            o[str(rng.choice(list(o.keys())))] = odd_value(rng)
        elif k == 3 and isinstance(o.get("children"), list) and o["children"]:
            o["children"][int(rng.integers(0, len(o["children"])))] = {"objectID": int(rng.choice([o.get("objectID", 0), 1, 99999, -3]))}
        elif k == 4 and isinstance(o.get("components"), list) and o["components"]:
            o["components"][int(rng.integers(0, len(o["components"])))] = {"objectID": int(rng.choice([1, 99999, int(rng.integers(0, len(objects)))]))}
        elif k == 5:
            o["objectID"] = int(rng.choice([0, 1, int(rng.integers(0, len(objects))), -1, 2**31, 2**33]))
        elif k == 6:
            o["objectClass"] = str(rng.choice(["Entity", "World", "MeshRenderer", "Camera", "PointLight", "SpotLight", "Nonsense", ""]))
        elif k == 7:
            objects.append(copy.deepcopy(o))                       # the same id twice
        elif k == 8:
            objects.remove(o)
        elif k == 9 and props is not None:
            props["active"] = odd_value(rng)
        elif k == 10 and isinstance(o.get("children"), list):
            o["children"] = o["children"] + o["children"][:2]    # a child listed twice
        elif k == 11 and props is not None:
            for name in ("position", "orientation", "scale", "range", "cutoff", "fov", "zNear", "zFar", "intensity", "colour", "attenuation"):
                if name in props and rng.random() < 0.5:
                    props[name] = odd_value(rng)
    return objects


def child(paths):
    from orion_b200 import lib
    from oracle import pyoracle, scene_import
    port = pyoracle.load("port")
    for path in paths:
        ours = theirs = None
        try:
            ours = lib.load_scene(path, SCENE, 640, 360)
        except lib.OvisError as e:
            ours = None
        try:
            theirs = scene_import.import_world(path, SCENE, 640, 360, oracle=port)
        except Exception:
            theirs = None
        if ours is None:
            verdict = "refused" if theirs is None else "ONLY_RESTATEMENT_LOADS"
        elif theirs is None:
            verdict = "ONLY_LOADER_LOADS"
        else:
            same = all(np.asarray(ours[k]).tobytes() == np.asarray(theirs[k]).tobytes() for k in KEYS) and \
                int(ours["shadow_map_resolution"]) == int(theirs["shadow_map_resolution"])
            verdict = "equal" if same else "DIFFERENT"
        print(os.path.basename(path), verdict, flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=600)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--child", nargs="*", default=None)
    ap.add_argument("--keep", default=None, help="directory to keep the files that crashed or differed in")
    a = ap.parse_args()
    if a.child is not None:
        child(a.child)
        return
    rng = np.random.default_rng(a.seed)
    objects = json.load(open(FOREST))
    text = open(FOREST, "rb").read()
    counts = {"refused": 0, "equal": 0, "DIFFERENT": 0, "ONLY_LOADER_LOADS": 0, "ONLY_RESTATEMENT_LOADS": 0}
    problems = []
    with tempfile.TemporaryDirectory() as tmp:
        for b in range(0, a.cases, 25):
            paths = []
            for i in range(b, min(a.cases, b + 25)):
                p = os.path.join(tmp, f"w{i}.object")
                if rng.random() < 0.8:
                    open(p, "w").write(json.dumps(damage(rng, objects)))
                else:
                    raw = bytearray(text)
                    for _ in range(int(rng.integers(1, 4))):
                        j = int(rng.integers(0, len(raw)))
                        if rng.random() < 0.5:
                            raw[j] = int(rng.integers(0, 256))
                        else:
                            del raw[j:j + int(rng.integers(1, 40))]
                    open(p, "wb").write(bytes(raw))
                paths.append(p)
            try:
                r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child"] + paths, capture_output=True, text=True, timeout=300)
                rc, out, err = r.returncode, r.stdout, r.stderr
            except subprocess.TimeoutExpired as e:
                rc, out, err = "timeout", (e.stdout or b"").decode() if isinstance(e.stdout, bytes) else (e.stdout or ""), "timed out"
            seen = {}
            for line in out.splitlines():
                name, verdict = line.split()
                seen[name] = verdict
                counts[verdict] += 1
            for p in paths:
                v = seen.get(os.path.basename(p))
                if v is None or v in ("DIFFERENT", "ONLY_LOADER_LOADS", "ONLY_RESTATEMENT_LOADS"):
                    problems.append({"file": os.path.basename(p), "what": v or f"the process ended ({rc})", "stderr": err[-600:] if v is None else ""})
                    if a.keep:
                        os.makedirs(a.keep, exist_ok=True)
                        open(os.path.join(a.keep, os.path.basename(p)), "wb").write(open(p, "rb").read())
                    if v is None:
                        break                                      # (the files after a crash were not reached)
            if len(problems) >= 5:
                break
    print(json.dumps({"files": sum(counts.values()), **counts, "problems": problems}, indent=1))
    sys.exit(1 if problems else 0)


if __name__ == "__main__":
    main()
