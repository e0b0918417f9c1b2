"""Randomised check of the OBJ reader (ovis_obj_submesh_aabbs / ovis_obj_mesh, host only) on damaged meshes: the fixture
models with tokens replaced (signs, huge and negative indices, empty fields, hex, exponents, trailing junk), lines
dropped, duplicated or swapped, and plain byte damage. Every file is read or refused - never a crash (child processes) -
and the reader and the Python restatement of the reference's OBJLoader accept the same files with the same boxes.

    python tools/fuzz_obj.py [--cases 2000] [--seed 1]"""
import argparse
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
MODELS = os.path.join(ROOT, "tests", "golden", "scene", "models")

TOKENS = ["", "0", "-1", "+2", "1", "65535", "65536", "65537", "70000", "4294967297", "99999999999999999999999", "1.5", "1e2", "0x10", "1a", "a",
          "-0", "nan", "inf", "-inf", "1e40", "1e-50", ".5", "5.", "--1", "1/2", "//", "1//1", "/1/1", "1/1/", "1/1/1/1", "3/2/1", "\t1"]


def damage(rng, text: str) -> bytes:
    lines = text.split("\n")
    for _ in range(int(rng.integers(1, 5))):
        k = int(rng.integers(0, 7))
        i = int(rng.integers(0, len(lines)))
        tok = lines[i].split(" ")
        if k == 0 and len(tok) > 1:
            tok[int(rng.integers(1, len(tok)))] = str(rng.choice(TOKENS))
            lines[i] = " ".join(tok)
        elif k == 1 and len(tok) > 1 and "/" in tok[-1]:
            sub = tok[-1].split("/")
            sub[int(rng.integers(0, len(sub)))] = str(rng.choice(TOKENS[:14]))
            tok[-1] = "/".join(sub)
            lines[i] = " ".join(tok)
        elif k == 2:
            del lines[i]
        elif k == 3:
            lines.insert(i, lines[int(rng.integers(0, len(lines)))])
        elif k == 4:
            lines[i] = str(rng.choice(["g", "g a b", "usemtl", "usemtl x", "usemtl x y", "f", "f 1/1/1", "f 1/1/1 2/1/1", "v", "v 1 2", "vt 1",
                                       "vn 1 2", "o", "s off", "# c", "f 1/1/1 2/1/1 3/1/1 4/1/1 5/1/1", "v 1 2 3 4 5", "\r", "v 1 2 3\r"]))
        elif k == 5:
            lines[i] = lines[i] + str(rng.choice([" ", "  ", " x", "\r", " 7"]))
        else:
            tok.insert(int(rng.integers(0, len(tok) + 1)), "")       # a doubled space
            lines[i] = " ".join(tok)
    raw = bytearray("\n".join(lines).encode("latin-1"))
    if rng.random() < 0.2 and raw:
        for _ in range(int(rng.integers(1, 4))):
            j = int(rng.integers(0, len(raw)))
            if rng.random() < 0.5:
                raw[j] = int(rng.integers(0, 256))
            else:
                del raw[j:j + int(rng.integers(1, 30))]
    return bytes(raw)


def child(paths):
    from orion_b200 import lib
    from oracle import scene_import
    for path in paths:
        ours = theirs = None
        try:
            ours = lib.obj_submesh_aabbs(path)
        except lib.OvisError:
            ours = None
        try:
            theirs = scene_import.obj_submesh_aabbs(path)
        except Exception:
            theirs = None
        try:                                                    # the vertex-data form: must not crash; boxes of what it returns
            pos, nrm, tc, idx, first = lib.obj_mesh(path)
            mesh_ok = True
            if ours is not None and len(idx):
                for s in range(len(first) - 1):
                    p = pos[idx[first[s]:first[s + 1]]]
                    if len(p) and not (np.array_equal(p.min(axis=0), ours[s, :3]) and np.array_equal(p.max(axis=0), ours[s, 3:])):
                        if not np.isnan(p).any():
                            mesh_ok = False
        except lib.OvisError:
            mesh_ok = True
        if ours is None:
            verdict = "refused" if theirs is None else "ONLY_RESTATEMENT_READS"
        elif theirs is None:
            verdict = "ONLY_READER_READS"
        else:
            same = ours.shape == theirs.shape and np.array_equal(ours.view(np.uint32), theirs.view(np.uint32))
            verdict = "equal" if same else "DIFFERENT"
        if not mesh_ok:
            verdict = "MESH_AND_BOXES_DISAGREE"
        print(os.path.basename(path), verdict, flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=2000)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--child", nargs="*", default=None)
    ap.add_argument("--keep", default=None)
    a = ap.parse_args()
    if a.child is not None:
        child(a.child)
        return
    rng = np.random.default_rng(a.seed)
    sources = [open(os.path.join(MODELS, f), "rb").read().decode("latin-1") for f in sorted(os.listdir(MODELS)) if f.endswith(".obj")]
    counts = {}
    problems = []
    with tempfile.TemporaryDirectory() as tmp:
        for b in range(0, a.cases, 100):
            paths = []
            for i in range(b, min(a.cases, b + 100)):
                p = os.path.join(tmp, f"m{i}.obj")
                open(p, "wb").write(damage(rng, sources[int(rng.integers(0, len(sources)))]))
                paths.append(p)
            try:
                r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child"] + paths, capture_output=True, text=True, timeout=300)
                rc, out, err = r.returncode, r.stdout, r.stderr
            except subprocess.TimeoutExpired:
                rc, out, err = "timeout", "", "timed out"
            seen = {}
            for line in out.splitlines():
                name, verdict = line.split()
                seen[name] = verdict
                counts[verdict] = counts.get(verdict, 0) + 1
            for p in paths:
                v = seen.get(os.path.basename(p))
                if v is None or v not in ("refused", "equal"):
                    problems.append({"file": os.path.basename(p), "what": v or f"the process ended ({rc})", "stderr": err[-500:] if v is None else ""})
                    if a.keep:
                        os.makedirs(a.keep, exist_ok=True)
                        open(os.path.join(a.keep, os.path.basename(p)), "wb").write(open(p, "rb").read())
                    if v is None:
                        break
            if len(problems) >= 6:
                break
    print(json.dumps({"files": sum(counts.values()), **counts, "problems": problems}, indent=1))
    sys.exit(1 if problems else 0)


if __name__ == "__main__":
    main()
