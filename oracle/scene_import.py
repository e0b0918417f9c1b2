"""TEST INFRASTRUCTURE - CPU restatement of the reference's scene loading for the visibility path (SURVEY.md section
8(f) row 3). Only tests/ may import this; the product (orion_b200/csrc/ovis_scene.cu) never does.

What it follows:
  * serialised worlds: engine/src/json_serialiser.cc (a JSON array of objects with objectClass / objectID /
    objectProperties), engine/src/world.cc:56-71 (World -> root), engine/src/entity.cc:106-148 (properties unless root,
    then components, then children), activation order engine/src/entity.cc:414-421, defaults engine/src/entity.cc:36,
    engine/src/component.cc:29, graphics/src/renderer.cc:32, graphics/src/light.cc:35-74, graphics/include/camera.h:67;
  * OBJ sub-mesh boxes: engine/src/loaders/obj_loader.cc:94-279 and mesh_loader.cc:187-199.

The reference's own loader cannot be linked here (it needs the objgen-generated reflection tables and rapidjson), so this
is a restatement: "parity unpinned by reference code" for the parsing itself. It is pinned instead against (1) the values
SURVEY.md section 8(d) transcribes from apps/cubes/assets/worlds/main.object (orion_b200/scenes.cubes_scene) and (2) the
reference-compiled oracle downstream: world transforms, bounds and culls of the imported scene go through oracle/_ref.
All arithmetic after parsing (world TRS, quat * vec3) is done by the C oracle, not here.
"""
from __future__ import annotations

import ctypes as C
import json
import os

import numpy as np

from . import pyoracle

_libc = C.CDLL(None)
_libc.strtof.restype = C.c_float
_libc.strtof.argtypes = [C.c_char_p, C.POINTER(C.c_char_p)]

FLT_MAX = np.finfo(np.float32).max
CASTS_SHADOW = 1
LIGHT_TYPES = {"AmbientLight": 0, "DirectionalLight": 1, "PointLight": 2, "SpotLight": 3}


def _strtof(token: str):
    """strtof + the reference's 'consumed the whole token' check (obj_loader.cc:170-174)."""
    raw = token.encode()
    end = C.c_char_p()
    buf = C.create_string_buffer(raw)
    value = _libc.strtof(buf, C.byref(end))
    consumed = C.cast(end, C.c_void_p).value - C.addressof(buf)
    if consumed != len(raw):
        raise ValueError(f"expected float value: {token!r}")
    return np.float32(value)


def _tokenize(s: str, delims: str, trim_empty=True):
    out, last = [], 0
    while last <= len(s):
        pos = min([p for p in (s.find(d, last) for d in delims) if p >= 0], default=len(s))
        if pos != last or not trim_empty:
            out.append(s[last:pos])
        last = pos + 1
    return out


def obj_submesh_aabbs(path: str) -> np.ndarray:
    """(n_submeshes, 6) float32: min.xyz, max.xyz of the positions each sub-mesh's faces reference."""
    positions, n_tex, n_norm = [], 0, 0
    material, current, boxes = "default", None, []
    with open(path, "rb") as f:
        text = f.read().decode("latin-1")
    for line in text.split("\n"):
        tok = _tokenize(line, " \r")
        if not tok:
            continue
        if tok[0] in ("v", "vt", "vn"):
            want = 2 if tok[0] == "vt" else 3
            if len(tok) < want + 1:
                raise ValueError("too few values")
            vals = [_strtof(t) for t in tok[1:1 + want]]
            if tok[0] == "v":
                positions.append(vals)
            elif tok[0] == "vt":
                n_tex += 1
            else:
                n_norm += 1
        elif tok[0] == "f":
            nv = len(tok) - 1
            if nv not in (3, 4):
                raise ValueError("expected 3 or 4 vertices")
            if current is None:
                boxes.append(np.array([FLT_MAX] * 3 + [-FLT_MAX] * 3, dtype=np.float32))
                current = len(boxes) - 1
            for t in tok[1:]:
                sub = _tokenize(t, "/", trim_empty=False)
                if len(sub) != 3:
                    raise ValueError("expected v/vt/vn")
                idx = [((int(x) & 0xffff) - 1) & 0xffff for x in sub]      # uint16 arithmetic, obj_loader.cc:220-227
                if idx[0] >= len(positions) or idx[1] > n_tex or idx[2] > n_norm:
                    raise ValueError("invalid index")
                p = positions[idx[0]]
                b = boxes[current]
                for k in range(3):                                          # glm::min(minimum, p), glm::max(maximum, p)
                    b[k] = b[k] if b[k] < p[k] else p[k]
                    b[3 + k] = b[3 + k] if b[3 + k] > p[k] else p[k]
        elif tok[0] == "usemtl":
            if len(tok) != 2:
                raise ValueError("expected single material name")
            if tok[1] != material:
                material, current = tok[1], None
        elif tok[0] == "g":
            if len(tok) != 2:
                raise ValueError("expected single group name")
            current = None
    return np.array(boxes, dtype=np.float32).reshape(-1, 6)


class SceneError(ValueError):
    """A file the reference would refuse or assert on."""


# ---- JSONSerialiser::read's type rules (engine/src/json_serialiser.cc:486-626): a value of the wrong JSON type leaves the
# property at its default. Python's json keeps what rapidjson's flags say: `50` is an int, `50.0` / `5e1` a float.
def _is_number(x):
    return isinstance(x, (int, float)) and not isinstance(x, bool)


def _is_double(x):                                              # rapidjson IsDouble(): fraction / exponent, or beyond 64 bits
    return isinstance(x, float) or (_is_number(x) and not (-2 ** 63 <= x <= 2 ** 64 - 1))


def _is_uint(x):                                                # rapidjson IsUint()
    return isinstance(x, int) and not isinstance(x, bool) and 0 <= x <= 0xFFFFFFFF


def _get_bool(props, name, default):
    v = props.get(name)
    return v if isinstance(v, bool) else default


def _get_float(props, name, default):                           # IsFloat() (document.h:971-976), then GetFloat()
    v = props.get(name)
    if v is not None and _is_double(v) and -3.4028234e38 <= float(v) <= 3.4028234e38:
        return np.float32(float(v))
    return np.float32(default)


def _get_floats(props, name, n, default):                       # vec3 / quat: an array of exactly n numbers of any kind
    v = props.get(name)
    if isinstance(v, list) and len(v) == n and all(_is_number(x) for x in v):
        return [np.float32(float(x)) for x in v]
    return [np.float32(x) for x in default]


def import_world(world_path: str, game_dir: str, vp_width: int, vp_height: int, oracle=None) -> dict:
    """The render-world content of a serialised world, in activation order."""
    oracle = oracle or pyoracle.load("best")
    with open(world_path, "rb") as f:
        raw = f.read()
    for mark in (0xEF, 0xBB, 0xBF):                             # rapidjson's stream skips these, each on its own
        if raw[:1] == bytes([mark]):
            raw = raw[1:]
    raw = raw.split(b"\0")[0]                                   # ... and reads a NUL as the end of the text
    def first_wins(pairs):                                      # rapidjson's FindMember returns the FIRST member of that name
        d = {}
        for k, v in pairs:
            d.setdefault(k, v)
        return d

    # (rapidjson does not validate the encoding of strings)
    objects = json.loads(raw.decode("utf-8", "surrogateescape"), object_pairs_hook=first_wins)
    if not isinstance(objects, list):
        raise SceneError("not an array of objects")

    def scene_object(ref):
        """findObject (json_serialiser.cc:232-244): {"objectID": uint} names a POSITION in the document."""
        if not isinstance(ref, dict):
            return None, None
        if not ref:
            raise SceneError("null reference where the reference asserts")
        i = ref.get("objectID")
        if not _is_uint(i) or i >= len(objects):
            return None, None
        o = objects[i]
        if not isinstance(o, dict) or not isinstance(o.get("objectClass"), str):
            return None, None
        return i, o

    if not objects or not isinstance(objects[0], dict) or objects[0].get("objectClass") != "World":
        raise SceneError("the first object is not a World")
    root_ref = objects[0].get("root")
    if not isinstance(root_ref, dict) or not root_ref:
        raise SceneError("the World has no root")
    root_id, root = scene_object(root_ref)
    if root is None or root["objectClass"] != "Entity":
        raise SceneError("the World has no root entity")

    parent, local, ent_aabb, ent_flags, ent_node = [], [], [], [], []
    pending_lights, pending_cameras = [], []
    meshes = {}
    shadow_res = [0]
    taken = {root_id}

    def props_of(o):
        p = o.get("objectProperties")
        return p if isinstance(p, dict) else {}

    def visit(e, parent_node, depth):
        if depth > 4096:
            raise SceneError("entity tree too deep")
        props = props_of(e)
        trs = [np.float32(x) for x in (0, 0, 0, 1, 0, 0, 0, 1, 1, 1)]
        if parent_node >= 0:                                    # the root's properties are not read (entity.cc:127-130)
            if not _get_bool(props, "active", False):
                return
            trs = _get_floats(props, "position", 3, (0, 0, 0)) + _get_floats(props, "orientation", 4, (1, 0, 0, 0)) + \
                _get_floats(props, "scale", 3, (1, 1, 1))
        node = len(parent)
        parent.append(parent_node)
        local.append(trs)
        # `while (serialiser.pop(x))`: the list ends at the first element that cannot be read (entity.cc:132-147)
        comps = e.get("components")
        for ref in comps if isinstance(comps, list) else []:
            i, c = scene_object(ref)
            if c is None or c["objectClass"] in ("Entity", "World"):
                break
            if i in taken:
                raise SceneError("a component is attached twice")
            taken.add(i)
            cp = props_of(c)
            if not _get_bool(cp, "active", False):
                continue
            cls = c["objectClass"]
            if cls in ("MeshRenderer", "Skybox"):
                flags = CASTS_SHADOW if _get_bool(cp, "castsShadow", True) else 0
                if cls == "Skybox":
                    boxes = np.array([[-FLT_MAX] * 3 + [FLT_MAX] * 3], dtype=np.float32)
                else:
                    mesh = cp.get("mesh")
                    asset = mesh.get("asset") if isinstance(mesh, dict) else None
                    if not isinstance(asset, str) or not asset.startswith("game/"):
                        raise SceneError("MeshRenderer without a usable mesh asset")
                    if asset not in meshes:
                        path = os.path.join(game_dir, asset[5:] + ".obj")
                        if not os.path.exists(path):
                            raise SceneError("missing mesh")
                        meshes[asset] = obj_submesh_aabbs(path)
                        if len(meshes[asset]) == 0:
                            raise SceneError("mesh without faces")
                    boxes = meshes[asset]
                for b in boxes:
                    ent_aabb.append(b)
                    ent_flags.append(flags)
                    ent_node.append(node)
            elif cls in LIGHT_TYPES:
                pending_lights.append((node, cls, cp))
            elif cls == "Camera":
                pending_cameras.append((node, cp))
                pipe_ref = cp.get("renderPipeline")
                pipe = scene_object(pipe_ref)[1] if isinstance(pipe_ref, dict) and pipe_ref else None
                if pipe is not None and not shadow_res[0]:
                    r = props_of(pipe).get("shadowMapResolution")       # PROPERTY() uint16_t: IsUint(), then truncated
                    if _is_uint(r):
                        shadow_res[0] = r & 0xFFFF
        kids = e.get("children")
        for ref in kids if isinstance(kids, list) else []:
            i, k = scene_object(ref)
            if k is None or k["objectClass"] != "Entity":
                break
            if i in taken:
                raise SceneError("an entity is attached twice")
            taken.add(i)
            visit(k, node, depth + 1)

    import sys
    sys.setrecursionlimit(max(sys.getrecursionlimit(), 20000))
    visit(root, -1, 0)
    parent = np.array(parent, dtype=np.int32)
    local = np.array(local, dtype=np.float32).reshape(-1, 10)
    world_trs = oracle.world_trs(parent, local)

    lights = np.zeros(len(pending_lights), dtype=pyoracle.LIGHT_DTYPE)
    shading = np.zeros((len(pending_lights), 4), dtype=np.float32)
    for i, (node, cls, cp) in enumerate(pending_lights):
        l = lights[i]
        t = LIGHT_TYPES[cls]
        l["type"] = t
        l["flags"] = CASTS_SHADOW if _get_bool(cp, "castsShadows", False) else 0
        l["colour"] = _get_floats(cp, "colour", 3, (1, 1, 1))
        l["intensity"] = _get_float(cp, "intensity", 0.8)
        att = [0, 0, 0]
        if t == 2:
            l["range"] = _get_float(cp, "range", 100.0)
            att = [1.0, 0.045, 0.0075]
        elif t == 3:
            l["range"] = _get_float(cp, "range", 50.0)
            att = [1.0, 0.09, 0.032]
            l["cutoff"] = _get_float(cp, "cutoff", 20.0)
            if l["cutoff"] > 45.0:
                raise SceneError("spot light cutoff must be <= 45 degrees")          # render_light.cc:104
        shading[i, :3] = _get_floats(cp, "attenuation", 3, att)
        shading[i, 3] = _get_float(cp, "shadowBiasConstant", -0.001)
        l["position"] = world_trs[node, 0:3]
        l["direction"] = oracle.quat_rotate(world_trs[node, 3:7], [0, 0, -1])
    cameras = np.zeros(len(pending_cameras), dtype=pyoracle.VIEW_DTYPE)
    for i, (node, cp) in enumerate(pending_cameras):
        v = cameras[i]
        v["position"] = world_trs[node, 0:3]
        v["orientation"] = world_trs[node, 3:7]
        v["hfov_deg"] = _get_float(cp, "fov", 75.0)
        v["znear"] = _get_float(cp, "zNear", 0.1)
        v["zfar"] = _get_float(cp, "zFar", 1000.0)
        v["vp_width"], v["vp_height"] = vp_width, vp_height
    return {"parent": parent, "local_trs": local, "world_trs": world_trs,
            "local_aabb": np.array(ent_aabb, dtype=np.float32).reshape(-1, 6),
            "flags": np.array(ent_flags, dtype=np.uint32), "entity_node": np.array(ent_node, dtype=np.uint32),
            "lights": lights, "shading": shading, "cameras": cameras, "shadow_map_resolution": shadow_res[0]}
