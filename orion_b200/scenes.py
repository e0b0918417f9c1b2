"""Seeded synthetic scenes for the BASELINE.json configs (host-side numpy; no oracle, no CUDA).

Every generator returns plain numpy arrays in the layouts of include/orion_vis.h:
TRS = (n,10) position xyz / orientation wxyz / scale xyz; AABB = (n,6) min xyz / max xyz; flags uint32
(bit0 = casts shadow, engine/runtime/render/include/render_entity.h:50-53); views and lights as the
structured dtypes below (byte-identical to OvisView / OvisLight).

Distributions follow SURVEY.md section 8(d); the cubes scene is transcribed from
apps/cubes/assets/worlds/main.object (reference), camera aspect from apps/cubes/src/cubes_game.cc:58-59.
"""
from __future__ import annotations

import numpy as np

LIGHT_AMBIENT, LIGHT_DIRECTIONAL, LIGHT_POINT, LIGHT_SPOT = 0, 1, 2, 3
CASTS_SHADOW = 1

LIGHT_DTYPE = np.dtype([("type", "<i4"), ("flags", "<u4"), ("position", "<f4", 3), ("direction", "<f4", 3),
                        ("colour", "<f4", 3), ("intensity", "<f4"), ("range", "<f4"), ("cutoff", "<f4")])
VIEW_DTYPE = np.dtype([("position", "<f4", 3), ("orientation", "<f4", 4), ("hfov_deg", "<f4"),
                       ("znear", "<f4"), ("zfar", "<f4"), ("vp_width", "<i4"), ("vp_height", "<i4")])

IDENTITY_TRS = np.array([0, 0, 0, 1, 0, 0, 0, 1, 1, 1], dtype=np.float32)


def make_view(position, orientation, hfov_deg, znear, zfar, vp_width, vp_height) -> np.ndarray:
    v = np.zeros((), dtype=VIEW_DTYPE)
    v["position"] = position
    v["orientation"] = orientation
    v["hfov_deg"], v["znear"], v["zfar"] = hfov_deg, znear, zfar
    v["vp_width"], v["vp_height"] = vp_width, vp_height
    return v


def make_light(type, position=(0, 0, 0), direction=(0, 0, -1), colour=(1, 1, 1), intensity=1.0, range=0.0,
               cutoff=0.0, flags=0) -> np.ndarray:
    l = np.zeros((), dtype=LIGHT_DTYPE)
    l["type"], l["flags"] = type, flags
    l["position"], l["direction"], l["colour"] = position, direction, colour
    l["intensity"], l["range"], l["cutoff"] = intensity, range, cutoff
    return l


def quat_normalize_f32(q: np.ndarray) -> np.ndarray:
    q = np.asarray(q, dtype=np.float32)
    n = np.sqrt(np.sum(q * q, axis=-1, keepdims=True, dtype=np.float32), dtype=np.float32)
    return (q / n).astype(np.float32)


def quat_yaw_pitch(yaw_deg: float, pitch_deg: float) -> np.ndarray:
    """Camera orientation: yaw about +Y then pitch about local +X (w x y z), normalised in fp32."""
    hy, hp = np.radians(yaw_deg) / 2, np.radians(pitch_deg) / 2
    qy = np.array([np.cos(hy), 0, np.sin(hy), 0])
    qp = np.array([np.cos(hp), np.sin(hp), 0, 0])
    w1, x1, y1, z1 = qy
    w2, x2, y2, z2 = qp
    q = np.array([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                  w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2])
    return quat_normalize_f32(q)


def quat_rotate(q, v) -> np.ndarray:
    """q * v (w x y z) in float64 - only used to derive scene constants, never for parity maths."""
    w, u = q[0], np.asarray(q[1:4], dtype=np.float64)
    uv = np.cross(u, v)
    uuv = np.cross(u, uv)
    return (np.asarray(v, dtype=np.float64) + 2.0 * (w * uv + uuv)).astype(np.float32)


def random_unit_quats(rng: np.random.Generator, n: int) -> np.ndarray:
    q = rng.standard_normal((n, 4)).astype(np.float32)
    return quat_normalize_f32(q)


def default_camera(hfov=75.0, znear=0.1, zfar=1000.0, width=1920, height=1080) -> np.ndarray:
    """Config-2 camera: origin, yaw 30, pitch -10, hfov 75, 16:9, near 0.1, far 1000."""
    return make_view((0, 0, 0), quat_yaw_pitch(30.0, -10.0), hfov, znear, zfar, width, height)


def static_boxes(n: int, seed: int, extent: float):
    """Config 2/4/5 entities: centres U[-extent,extent]^3, half-extents log-U[0.25,8] per axis,
    80 % casters. Returns (aabb (n,6) f32, flags (n,) u32). Generated in chunks so 100M fits."""
    rng = np.random.default_rng(seed)
    aabb = np.empty((n, 6), dtype=np.float32)
    flags = np.empty(n, dtype=np.uint32)
    chunk = 1 << 22
    lo, hi = np.log(0.25), np.log(8.0)
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        c = rng.uniform(-extent, extent, size=(e - s, 3)).astype(np.float32)
        h = np.exp(rng.uniform(lo, hi, size=(e - s, 3))).astype(np.float32)
        aabb[s:e, 0:3] = c - h
        aabb[s:e, 3:6] = c + h
        flags[s:e] = (rng.random(e - s) < 0.8).astype(np.uint32)
    return aabb, flags


def config2(n: int = 1_000_000, seed: int = 1):
    """synthetic 1M static entities, 1 camera frustum."""
    aabb, flags = static_boxes(n, seed, 1000.0)
    return {"aabb": aabb, "flags": flags, "view": default_camera()}


def config5(n: int = 100_000_000, seed: int = 4):
    """100M entities in [-4000,4000]^3 (sharded by contiguous index range across GPUs)."""
    aabb, flags = static_boxes(n, seed, 4000.0)
    return {"aabb": aabb, "flags": flags, "view": default_camera()}


def cascade_views(camera: np.ndarray, splits=((0.1, 10.0), (10.0, 50.0), (50.0, 200.0), (200.0, 1000.0))):
    """Config-4 "cascades": the reference has none (render_light.cc:328-330 is a TODO), so they are
    synthesised as extra RenderViews with the camera transform and sliced near/far."""
    out = np.zeros(len(splits), dtype=VIEW_DTYPE)
    for i, (zn, zf) in enumerate(splits):
        out[i] = camera
        out[i]["znear"], out[i]["zfar"] = zn, zf
    return out


def point_lights(n: int, seed: int, extent: float, rmin: float, rmax: float, shadows=True) -> np.ndarray:
    rng = np.random.default_rng(seed)
    l = np.zeros(n, dtype=LIGHT_DTYPE)
    l["type"] = LIGHT_POINT
    l["flags"] = CASTS_SHADOW if shadows else 0
    l["position"] = rng.uniform(-extent, extent, size=(n, 3)).astype(np.float32)
    l["direction"] = (0, 0, -1)
    l["colour"] = (1, 1, 1)
    l["intensity"] = 1.0
    l["range"] = rng.uniform(rmin, rmax, size=n).astype(np.float32)
    return l


def config4(n: int = 10_000_000, seed: int = 3, n_lights: int = 64):
    aabb, flags = static_boxes(n, seed, 2000.0)
    cam = default_camera()
    return {"aabb": aabb, "flags": flags, "view": cam, "cascades": cascade_views(cam),
            "lights": point_lights(n_lights, seed + 1000, 2000.0, 50.0, 200.0)}


def config3(n_groups: int = 10_000, objects_per_group: int = 99, n_parts: int = 9_000_000, seed: int = 2,
            n_lights: int = 4096):
    """10M dynamic entities with a depth-3 transform hierarchy + 4096 shadow-casting point lights.

    Nodes are depth-sorted (parent index < child index): node 0 = root (identity), then groups
    (depth 1), objects (depth 2), parts (depth 3). Every depth-2/3 node carries one render entity.
    Returns parent (int32, -1 for the root), local TRS, node->entity binding, local AABBs, flags.
    """
    rng = np.random.default_rng(seed)
    n_obj = n_groups * objects_per_group
    n_nodes = 1 + n_groups + n_obj + n_parts
    parent = np.empty(n_nodes, dtype=np.int32)
    trs = np.empty((n_nodes, 10), dtype=np.float32)
    parent[0] = -1
    trs[0] = IDENTITY_TRS
    g0, o0, p0 = 1, 1 + n_groups, 1 + n_groups + n_obj
    # depth 1: groups
    parent[g0:o0] = 0
    trs[g0:o0, 0:3] = rng.uniform(-2000, 2000, size=(n_groups, 3))
    trs[g0:o0, 3:7] = random_unit_quats(rng, n_groups)
    trs[g0:o0, 7:10] = rng.uniform(0.5, 2.0, size=(n_groups, 1)).astype(np.float32)
    # depth 2: objects
    parent[o0:p0] = g0 + np.repeat(np.arange(n_groups, dtype=np.int32), objects_per_group)
    trs[o0:p0, 0:3] = rng.uniform(-50, 50, size=(n_obj, 3))
    trs[o0:p0, 3:7] = random_unit_quats(rng, n_obj)
    trs[o0:p0, 7:10] = rng.uniform(0.5, 2.0, size=(n_obj, 3))
    # depth 3: parts, spread evenly over objects (parents ascending keeps subtrees contiguous-ish)
    parent[p0:] = o0 + (np.arange(n_parts, dtype=np.int64) * n_obj // max(n_parts, 1)).astype(np.int32)
    trs[p0:, 0:3] = rng.uniform(-5, 5, size=(n_parts, 3))
    trs[p0:, 3:7] = random_unit_quats(rng, n_parts)
    trs[p0:, 7:10] = rng.uniform(0.25, 1.5, size=(n_parts, 3))
    n_ent = n_obj + n_parts
    entity_node = np.arange(o0, n_nodes, dtype=np.int32)          # entity e is bound to node o0 + e
    h = np.exp(rng.uniform(np.log(0.1), np.log(2.0), size=(n_ent, 3))).astype(np.float32)
    local_aabb = np.concatenate([-h, h], axis=1).astype(np.float32)
    flags = (rng.random(n_ent) < 0.8).astype(np.uint32)
    lights = point_lights(n_lights, seed + 1000, 2000.0, 10.0, 60.0)
    return {"parent": parent, "local_trs": trs, "entity_node": entity_node, "local_aabb": local_aabb,
            "flags": flags, "view": default_camera(), "lights": lights,
            "group_range": (g0, o0), "n_nodes": n_nodes, "n_entities": n_ent}


def move_groups(scene: dict, frame: int) -> np.ndarray:
    """New local TRS for every depth-1 node for the given frame (config 3: everything below is dirty)."""
    g0, o0 = scene["group_range"]
    rng = np.random.default_rng(10_000 + frame)
    n = o0 - g0
    t = scene["local_trs"][g0:o0].copy()
    t[:, 0:3] += rng.uniform(-1, 1, size=(n, 3)).astype(np.float32)
    t[:, 3:7] = random_unit_quats(rng, n)
    return t


def cubes_scene():
    """apps/cubes/assets/worlds/main.object rebuilt headlessly.

    Render entities in insertion order: skybox, floor, cube_0, cube_1. Camera = player (0,1,0) + child
    (0,1,0); fov 90, near 0.25, far 200, 1440x900. Lights: ambient, spot light_0, point light_1..3, all
    casting shadows -> 1 + 1 + 3*6 = 20 views per frame.
    """
    fmax = np.finfo(np.float32).max
    cube = (2.977269, 2.973644, 2.977269)
    local_aabb = np.array([
        [-fmax, -fmax, -fmax, fmax, fmax, fmax],               # graphics/src/skybox.cc:50
        [-0.5, -0.5, 0.0, 0.5, 0.5, 0.0],                      # floor.obj unit quad
        [-cube[0], -cube[1], -cube[2], cube[0], cube[1], cube[2]],
        [-cube[0], -cube[1], -cube[2], cube[0], cube[1], cube[2]],
    ], dtype=np.float32)
    trs = np.array([
        IDENTITY_TRS,
        [0, 0, 0, 0.70710677, -0.70710677, 0, 0, 100, 100, 1],
        [0, 4, -7, 0.9098436832427979, 0.06645228713750839, 0.3768696188926697, -0.16042999923229218, 0.2, 0.2, 0.2],
        [0.4, 7, -7, 0.9098436832427979, -0.06645228713750839, 0.3768696188926697, 0.16042999923229218, 0.2, 0.2, 0.2],
    ], dtype=np.float32)
    flags = np.array([0, 0, 1, 1], dtype=np.uint32)
    view = make_view((0, 2, 0), (1, 0, 0, 0), 90.0, 0.25, 200.0, 1440, 900)
    lights = np.zeros(5, dtype=LIGHT_DTYPE)
    lights[0] = make_light(LIGHT_AMBIENT, intensity=0.003)
    # light_0: direction = worldOrientation * (0,0,-1) (graphics/src/light.cc:104); the light normalises it
    spot_dir = quat_rotate(np.array([0.7071067690849304, -0.5521576404571533, 0.44172608852386475, 0.0]),
                           np.array([0.0, 0.0, -1.0]))
    lights[1] = make_light(LIGHT_SPOT, position=(2, 3, -7), direction=spot_dir, intensity=0.6, range=20.0,
                           cutoff=45.0, flags=CASTS_SHADOW)
    for i, (p, c) in enumerate([((-2, 3, -3.5), (0, 0, 1)), ((2, 3, -3.5), (0, 1, 0)), ((0, 3, -9), (1, 0, 0))]):
        lights[2 + i] = make_light(LIGHT_POINT, position=p, colour=c, range=50.0, flags=CASTS_SHADOW)
    return {"trs": trs, "local_aabb": local_aabb, "flags": flags, "view": view, "lights": lights}
