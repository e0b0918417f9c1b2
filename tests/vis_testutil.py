"""Shared helpers for the parity tests."""
import numpy as np

from orion_b200 import scenes


def same(a, b) -> bool:
    """Value-exact equality for float arrays: identical values, NaNs in identical places."""
    a, b = np.asarray(a), np.asarray(b)
    return a.shape == b.shape and bool(np.array_equal(a, b, equal_nan=True))


def random_views(rng, n, extent=100.0):
    out = np.zeros(n, dtype=scenes.VIEW_DTYPE)
    sizes = [(1920, 1080), (1440, 900), (512, 512), (800, 600), (1000, 333)]
    for i in range(n):
        w, h = sizes[rng.integers(len(sizes))]
        zn = float(rng.uniform(0.05, 2.0))
        out[i] = scenes.make_view(rng.uniform(-extent, extent, 3), scenes.random_unit_quats(rng, 1)[0],
                                  float(rng.uniform(20, 120)), zn, zn + float(rng.uniform(5, 2000)), w, h)
    return out


def random_trs(rng, n, pos=100.0):
    t = np.empty((n, 10), dtype=np.float32)
    t[:, 0:3] = rng.uniform(-pos, pos, (n, 3))
    t[:, 3:7] = scenes.random_unit_quats(rng, n)
    t[:, 7:10] = rng.uniform(-3, 3, (n, 3))          # negative scale included
    return t


def random_boxes(rng, n, extent=100.0, hmax=8.0):
    c = rng.uniform(-extent, extent, (n, 3)).astype(np.float32)
    h = rng.uniform(0, hmax, (n, 3)).astype(np.float32)
    return np.concatenate([c - h, c + h], axis=1).astype(np.float32)


def edge_boxes():
    """Edge-case pack: skybox, zero-size, inverted, NaN, Inf, signed zeros."""
    fmax = np.finfo(np.float32).max
    inf, nan = np.float32(np.inf), np.float32(np.nan)
    return np.array([
        [-fmax, -fmax, -fmax, fmax, fmax, fmax],
        [0, 0, 0, 0, 0, 0],
        [1, 2, -30, 1, 2, -30],
        [5, 5, -5, -5, -5, -50],
        [nan, 0, 0, 1, 1, 1],
        [0, 0, 0, nan, nan, nan],
        [-inf, -inf, -inf, inf, inf, inf],
        [inf, inf, inf, inf, inf, inf],
        [-0.0, -0.0, -0.0, 0.0, 0.0, 0.0],
        [-1e-30, -1e-30, -1e-30, 1e-30, 1e-30, 1e-30],
        [1e30, 1e30, 1e30, 2e30, 2e30, 2e30],
    ], dtype=np.float32)


def random_lights(rng, n, extent=100.0):
    l = np.zeros(n, dtype=scenes.LIGHT_DTYPE)
    l["type"] = rng.integers(0, 4, n)
    l["flags"] = rng.integers(0, 2, n)
    l["position"] = rng.uniform(-extent, extent, (n, 3))
    d = rng.standard_normal((n, 3)).astype(np.float32)
    axis = rng.random(n) < 0.2                        # exact axis directions hit quatRotateBetween's special branch
    d[axis] = np.array([[0, 0, 1], [0, 0, -1], [0, 1, 0], [1, 0, 0], [0, -1, 0]], dtype=np.float32)[rng.integers(0, 5, axis.sum())]
    l["direction"] = d
    l["colour"] = rng.uniform(0, 1, (n, 3))
    black = rng.random(n) < 0.1
    l["colour"][black] = 0
    l["intensity"] = np.where(rng.random(n) < 0.1, 0.0, rng.uniform(0.01, 2.0, n))
    l["range"] = rng.uniform(1, 80, n)
    l["cutoff"] = rng.uniform(1, 45, n)
    return l


def random_pass_counts(rng, n):
    """Packed shader-pass counts (deferred | basic << 8 | shadow caster << 16); some entities have no pass of a type."""
    deferred = np.where(rng.random(n) < 0.6, rng.integers(1, 4, n), 0)
    basic = np.where(rng.random(n) < 0.8, rng.integers(1, 3, n), 0)
    shadow = np.where(rng.random(n) < 0.7, rng.integers(1, 3, n), 0)
    return (deferred | (basic << 8) | (shadow << 16)).astype(np.uint32)


def gpu_draw_lists(world, mode, visible=None, cap=None):
    """Run ovis_build_draw_lists_device on a host list (or, with visible=None, on the world's last cull result);
    returns (draws0, draws1, (count0, count1)) with the lists cut to min(count, cap)."""
    import torch
    n_slots = world.counts()[0]
    cap = cap if cap is not None else 8 * max(n_slots, 1)
    d0 = torch.zeros((cap + 1, 2), dtype=torch.int32, device="cuda")
    d1 = torch.zeros((cap + 1, 2), dtype=torch.int32, device="cuda")
    cnt = torch.full((2,), -1, dtype=torch.int32, device="cuda")
    if visible is None:
        world.build_draw_lists_device(mode, d0.data_ptr(), cap, d1.data_ptr(), cap, d_draw_counts=cnt.data_ptr())
    else:
        vis = torch.from_numpy(np.ascontiguousarray(visible, dtype=np.uint32).view(np.int32)).cuda()
        vcnt = torch.tensor([len(visible)], dtype=torch.int32, device="cuda")
        world.build_draw_lists_device(mode, d0.data_ptr(), cap, d1.data_ptr(), cap, vis.data_ptr() if len(visible) else d0.data_ptr(),
                                      vcnt.data_ptr(), len(visible), cnt.data_ptr())
    torch.cuda.synchronize()
    c = cnt.cpu().numpy()
    assert not d0[cap].any() and not d1[cap].any()           # nothing written past the capacity
    return (d0[:min(int(c[0]), cap)].cpu().numpy().view(np.uint32), d1[:min(int(c[1]), cap)].cpu().numpy().view(np.uint32),
            (int(c[0]), int(c[1])))


def gpu_entity_uniforms(world, visible=None, cap=None):
    import torch
    n_slots = world.counts()[0]
    m = n_slots if visible is None else len(visible)
    cap = cap if cap is not None else max(m, 1)
    out = torch.full((cap + 1, 20), -7.0, dtype=torch.float32, device="cuda")
    if visible is None:
        world.pack_entity_uniforms_device(out.data_ptr(), cap)
    else:
        vis = torch.from_numpy(np.ascontiguousarray(visible, dtype=np.uint32).view(np.int32)).cuda()
        vcnt = torch.tensor([len(visible)], dtype=torch.int32, device="cuda")
        world.pack_entity_uniforms_device(out.data_ptr(), cap, vis.data_ptr() if len(visible) else out.data_ptr(), vcnt.data_ptr(),
                                          len(visible))
    torch.cuda.synchronize()
    assert bool((out[cap] == -7.0).all())
    return out[:cap].cpu().numpy()


def gpu_light_uniforms(world, visible=None, cap=None):
    """ovis_pack_light_uniforms_device -> (cap, 208) uint8 blocks (rows past the list are untouched 0x55 filler)."""
    import torch
    n_lights = world.counts()[2]
    cap = cap if cap is not None else max(n_lights if visible is None else len(visible), 1)
    out = torch.full((cap + 1, 208), 0x55, dtype=torch.uint8, device="cuda")
    if visible is None:
        world.pack_light_uniforms_device(out.data_ptr(), cap)
    else:
        vis = torch.from_numpy(np.ascontiguousarray(visible, dtype=np.uint32).view(np.int32)).cuda()
        vcnt = torch.tensor([len(visible)], dtype=torch.int32, device="cuda")
        world.pack_light_uniforms_device(out.data_ptr(), cap, vis.data_ptr() if len(visible) else out.data_ptr(), vcnt.data_ptr(),
                                         len(visible))
    torch.cuda.synchronize()
    assert bool((out[cap] == 0x55).all())
    return out[:cap].cpu().numpy()


def gpu_view_uniforms(world, views, viewport_pos=None):
    import torch
    views = np.ascontiguousarray(views, dtype=scenes.VIEW_DTYPE).reshape(-1)
    out = torch.full((len(views) + 1, 288), 0x55, dtype=torch.uint8, device="cuda")
    world.pack_view_uniforms_device(views, out.data_ptr(), viewport_pos)
    torch.cuda.synchronize()
    assert bool((out[len(views)] == 0x55).all())
    return out[:len(views)].cpu().numpy()


def random_gbuffer(rng, width, height, sky_fraction=0.1, linear_z=None):
    """Headless G-buffer with every pixel independent: encoded unit normals, colours, shininess, depth in (0, 1); a
    fraction of the pixels is left as cleared sky (depth 1, zero colour)."""
    n = rng.standard_normal((height, width, 3)).astype(np.float32)
    n /= np.linalg.norm(n, axis=2, keepdims=True).astype(np.float32)
    a = np.ones((height, width, 4), dtype=np.float32)
    a[..., :3] = (n + np.float32(1.0)) / np.float32(2.0)                 # lit_deferred_frag.glsl: (normalize(n) + 1) / 2
    b = np.ones((height, width, 4), dtype=np.float32)
    b[..., :3] = rng.uniform(0, 1, (height, width, 3))
    c = np.empty((height, width, 4), dtype=np.float32)
    c[..., :3] = rng.uniform(0, 1, (height, width, 3))
    c[..., 3] = rng.uniform(1, 64, (height, width))
    depth = rng.uniform(0.5, 0.9999, (height, width)).astype(np.float32)
    if linear_z is not None:                                             # (znear, zfar, lo, hi): view distance uniform in [lo, hi]
        zn, zf, lo, hi = linear_z
        dist = rng.uniform(lo, hi, (height, width))
        depth = (zf / (zf - zn) * (1.0 - zn / dist)).astype(np.float32)  # perspectiveRH_ZO depth of that distance
    sky = rng.random((height, width)) < sky_fraction
    depth[sky] = 1.0
    b[sky, :3] = 0
    c[sky] = 0
    return a, b, c, depth


def lights_for_shading(rng, n, view, extent=30.0):
    """Lights of all four types scattered in front of a camera (so that many pixels are in range), with shading parameters."""
    lights = random_lights(rng, n, extent=extent)
    lights["intensity"] = rng.uniform(0.2, 2.0, n)
    lights["colour"] = rng.uniform(0.1, 1.0, (n, 3))
    lights["range"] = rng.uniform(20.0, 200.0, n)
    lights["flags"] = rng.integers(0, 2, n)
    # a spot light pointing exactly along +-Y has a NaN shadowSpace in the reference too (quatLookAt with up = Y is
    # degenerate, render_light.cc:270); one such light would turn every pixel into NaN and hide everything else
    vertical = (lights["direction"][:, 0] == 0) & (lights["direction"][:, 2] == 0)
    lights["direction"][vertical, 0] = 0.3
    lights["direction"][vertical, 2] = 0.1
    att = np.stack([rng.uniform(0.5, 1.0, n), rng.uniform(0.0, 0.1, n), rng.uniform(0.0, 0.01, n)], axis=1).astype(np.float32)
    bias = rng.uniform(-0.005, 0.0, n).astype(np.float32)
    return lights, att, bias


def ground_gbuffer(rng, oracle, view, width, height):
    """A spatially coherent headless G-buffer: what the camera sees of the ground plane y = 0 (normal +Y, mottled
    colours), sky above the horizon. Depth is derived in float64 and stored as float32; the shader reconstructs its own
    positions from it."""
    mats, _ = oracle.view_matrices(view)
    mats = np.asarray(mats, dtype=np.float64).reshape(4, 4, 4)
    vp, inv = mats[2].T, mats[3].T                                      # stored column-major
    px, py = np.meshgrid(np.arange(width, dtype=np.float64), np.arange(height, dtype=np.float64))
    nx, ny = px / width * 2.0 - 1.0, py / height * 2.0 - 1.0

    def unproject(z):
        h = np.stack([nx, ny, np.full_like(nx, z), np.ones_like(nx)], axis=-1) @ inv.T
        return h[..., :3] / h[..., 3:4]

    p0, p1 = unproject(0.0), unproject(1.0)
    dy = p1[..., 1] - p0[..., 1]
    with np.errstate(divide="ignore", invalid="ignore"):
        t = -p0[..., 1] / dy
    hit = (dy < 0) & (t > 0) & (t <= 1)
    pos = p0 + np.where(hit, t, 1.0)[..., None] * (p1 - p0)
    clip = np.concatenate([pos, np.ones_like(pos[..., :1])], axis=-1) @ vp.T
    depth = np.where(hit, clip[..., 2] / clip[..., 3], 1.0).astype(np.float32)
    a = np.ones((height, width, 4), dtype=np.float32)
    a[..., :3] = (np.array([0.0, 1.0, 0.0], dtype=np.float32) + np.float32(1.0)) / np.float32(2.0)
    b = np.ones((height, width, 4), dtype=np.float32)
    b[..., :3] = rng.uniform(0.3, 0.9, (height, width, 3))
    c = np.empty((height, width, 4), dtype=np.float32)
    c[..., :3] = rng.uniform(0.1, 0.5, (height, width, 3))
    c[..., 3] = rng.uniform(8, 64, (height, width))
    b[~hit, :3] = 0
    c[~hit] = 0
    return a, b, c, depth
