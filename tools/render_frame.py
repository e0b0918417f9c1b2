#!/usr/bin/env python
"""A whole frame of the reference's deferred pipeline, headless, on the GPU: scene import -> transform update -> cull
(entities + lights) -> per-light shadow-caster lists -> shadow-map passes -> G-buffer pass -> deferred light pass ->
tonemap + gamma -> image file (binary PPM, top row first).

    python tools/render_frame.py <world.object> <game asset dir> out.ppm [width height]

Without arguments a built-in scene (cubes on a floor, one spot and two point lights) is drawn, so the tool also runs where
the reference tree is absent. Materials are flat colours (the rasteriser implements the shader variants without TEXTURED /
NORMALMAP; the demo's texture files are partly not in the reference repository anyway)."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orion_b200 import lib, scenes  # noqa: E402


def cube_mesh():
    faces = [((1, 0, 0), (0, 1, 0), (0, 0, 1)), ((-1, 0, 0), (0, 0, 1), (0, 1, 0)), ((0, 1, 0), (0, 0, 1), (1, 0, 0)),
             ((0, -1, 0), (1, 0, 0), (0, 0, 1)), ((0, 0, 1), (1, 0, 0), (0, 1, 0)), ((0, 0, -1), (0, 1, 0), (1, 0, 0))]
    pos, nrm, idx = [], [], []
    for n, u, v in faces:
        n, u, v = np.array(n, np.float32), np.array(u, np.float32), np.array(v, np.float32)
        base = len(pos)
        for su, sv in ((-1, -1), (1, -1), (1, 1), (-1, 1)):
            pos.append(0.5 * (n + su * u + sv * v))
            nrm.append(n)
        idx += [base, base + 1, base + 2, base, base + 2, base + 3]
    return np.array(pos, np.float32), np.array(nrm, np.float32), np.array(idx, np.uint32)


def builtin_scene(w, r, width, height, textured=False):
    rng = np.random.default_rng(7)
    n = 60
    trs = np.zeros((n + 1, 10), dtype=np.float32)
    trs[:n, 0] = rng.uniform(-12, 12, n)
    trs[:n, 1] = rng.uniform(0.5, 5, n)
    trs[:n, 2] = rng.uniform(-25, -3, n)
    trs[:n, 3:7] = scenes.random_unit_quats(rng, n)
    trs[:n, 7:10] = rng.uniform(0.6, 2.0, (n, 1))
    trs[n] = (0, 0, -12, 1, 0, 0, 0, 60, 0.2, 60)                      # the floor: a flattened cube
    aabb = np.tile(np.array([[-0.5, -0.5, -0.5, 0.5, 0.5, 0.5]], np.float32), (n + 1, 1))
    e0 = w.add_entities(aabb, np.ones(n + 1, dtype=np.uint32))
    ids = np.arange(e0, e0 + n + 1, dtype=np.uint32)
    w.set_transforms(ids, trs)
    mesh = r.add_mesh(*cube_mesh())
    mats = [r.add_material((0.85, 0.25, 0.2), (0.5, 0.5, 0.5), 16.0, lib.MATERIAL_SPECULAR),
            r.add_material((0.2, 0.6, 0.85), (0.5, 0.5, 0.5), 32.0, lib.MATERIAL_SPECULAR),
            r.add_material((0.7, 0.7, 0.7), (0.2, 0.2, 0.2), 64.0, lib.MATERIAL_SPECULAR)]
    which = rng.integers(0, 2, n + 1).astype(np.uint32)
    which[n] = 2
    meshes = np.full(n + 1, mesh, dtype=np.uint32)
    if textured:
        # the floor as a textured, normal-mapped slab: its own mesh with texture coordinates and generated tangents,
        # a procedural tile texture (sRGB) and a normal map of shallow bumps
        pos, nrm, idx = cube_mesh()
        uv = np.tile(np.array([[0, 0], [24, 0], [24, 24], [0, 24]], np.float32), (6, 1))
        floor_mesh = r.add_mesh(pos, nrm, idx, uv)
        r.set_mesh_tangents(floor_mesh, lib.mesh_tangents(pos, nrm, uv, idx))
        yy, xx = np.mgrid[0:64, 0:64]
        tile = np.where(((xx // 32) + (yy // 32)) % 2 == 0, 200, 120).astype(np.uint8)
        grout = (xx % 32 < 2) | (yy % 32 < 2)
        rgba = np.stack([np.where(grout, 60, tile), np.where(grout, 60, tile), np.where(grout, 64, (tile * 0.95).astype(np.uint8)), np.full_like(tile, 255)], axis=2)
        bump = np.stack([128 + 40 * np.sin(xx / 64 * 2 * np.pi * 2), 128 + 40 * np.sin(yy / 64 * 2 * np.pi * 2), np.full((64, 64), 235.0),
                         np.full((64, 64), 255.0)], axis=2).astype(np.uint8)
        t_d, t_n = r.add_texture(rgba, srgb=True, wrap=True), r.add_texture(bump, srgb=False, wrap=True)
        mats.append(r.add_material((1, 1, 1), (0.3, 0.3, 0.3), 64.0, lib.MATERIAL_SPECULAR | lib.MATERIAL_TEXTURED | lib.MATERIAL_NORMALMAP, t_d, t_n))
        which[n] = 3
        meshes[n] = floor_mesh
    r.bind_entities(ids, meshes, np.array(mats, np.uint32)[which])
    lights = np.zeros(4, dtype=scenes.LIGHT_DTYPE)
    lights[0] = scenes.make_light(scenes.LIGHT_AMBIENT, colour=(1, 1, 1), intensity=0.08)
    lights[1] = scenes.make_light(scenes.LIGHT_SPOT, position=(0, 14, -2), direction=(0, -1, -0.6), colour=(1, 0.95, 0.8), intensity=2.5,
                                  range=60.0, cutoff=40.0, flags=1)
    lights[2] = scenes.make_light(scenes.LIGHT_POINT, position=(-7, 3, -10), colour=(0.4, 0.6, 1.0), intensity=2.0, range=25.0, flags=1)
    lights[3] = scenes.make_light(scenes.LIGHT_POINT, position=(8, 2.5, -16), colour=(1.0, 0.5, 0.3), intensity=2.0, range=25.0, flags=1)
    w.add_lights(lights)
    view = scenes.make_view((0, 4, 4), scenes.quat_yaw_pitch(0.0, -12.0), 80.0, 0.1, 200.0, width, height)
    builtin_scene.description = dict(trs=trs, mesh=cube_mesh(), material_of=which.copy(),
                                     materials=[((0.85, 0.25, 0.2), (0.5, 0.5, 0.5), 16.0, 1), ((0.2, 0.6, 0.85), (0.5, 0.5, 0.5), 32.0, 1),
                                                ((0.7, 0.7, 0.7), (0.2, 0.2, 0.2), 64.0, 1)])
    return view, 512


def imported_scene(w, r, world_path, asset_dir, width, height):
    info, cams = w.import_scene(world_path, asset_dir, width, height)
    parsed = lib.load_scene(world_path, asset_dir, width, height)
    import json
    cache, palette, textures = {}, {}, {}
    colours = [((0.75, 0.75, 0.78), 64.0), ((0.9, 0.85, 0.8), 16.0), ((0.3, 0.7, 0.4), 32.0)]

    def texture(asset):
        """game/textures/<name> -> a texture id, if <name>.tga is there (the demo's cube texture is not in the repository)."""
        if asset not in textures:
            base = os.path.join(asset_dir, asset[len("game/"):])
            textures[asset] = None
            if os.path.exists(base + ".tga"):
                srgb, wrap = True, False                              # TextureLoader defaults (sRGB on, clamp)
                if os.path.exists(base + ".loader"):
                    props = json.load(open(base + ".loader"))[0].get("objectProperties", {})
                    srgb, wrap = props.get("sRGB", True), props.get("addressMode", "kClamp") == "kWrap"
                textures[asset] = r.add_texture(lib.tga_load(base + ".tga"), srgb, wrap)
        return textures[asset]

    def material_for(path, fallback):
        """materials/<mesh name>.object beside the mesh (the demo names them alike): shader keywords and parameters."""
        name = os.path.splitext(os.path.basename(path))[0]
        mpath = os.path.join(asset_dir, "materials", name + ".object")
        c, sh = fallback
        if not os.path.exists(mpath):
            return r.add_material(c, (0.5, 0.5, 0.5), sh, lib.MATERIAL_SPECULAR), False
        m = json.load(open(mpath))[0]
        shader, params = m["shader"]["asset"], m.get("parameters", {})
        flags = lib.MATERIAL_SPECULAR if "specular" in shader else 0
        d_tex = texture(params["diffuseTexture"]["asset"]) if "diffuseTexture" in params else None
        n_tex = texture(params["normalTexture"]["asset"]) if "normal" in shader and "normalTexture" in params else None
        flags |= (lib.MATERIAL_TEXTURED if d_tex is not None else 0) | (lib.MATERIAL_NORMALMAP if n_tex is not None else 0)
        return r.add_material(params.get("diffuseColour", c), params.get("specularColour", (0.5, 0.5, 0.5)), params.get("shininess", sh), flags,
                              d_tex or 0, n_tex or 0), n_tex is not None

    for e, (path, sub) in enumerate(parsed["entity_mesh"]):
        if not path:
            continue                                                     # the skybox has no mesh to rasterise
        if (path, sub) not in cache:
            pos, nrm, tc, idx, first = lib.obj_mesh(path)
            palette[(path, sub)], wants_tangents = material_for(path, colours[len(palette) % len(colours)])
            cache[(path, sub)] = r.add_mesh(pos, nrm, idx[first[sub]:first[sub + 1]], tc)
            if wants_tangents:                                           # OBJLoader generateTangents (models/<name>.loader)
                r.set_mesh_tangents(cache[(path, sub)], lib.mesh_tangents(pos, nrm, tc, idx))
        r.bind_entities([info.first_entity + e], [cache[(path, sub)]], [palette[(path, sub)]])
    return cams[0], (info.shadow_map_resolution or 512)


def render(world_path=None, asset_dir=None, out_path=None, width=960, height=540, timing=False, keep_state=None, shadow_size=None,
           textured=False):
    """keep_state: a dict that receives the frame's intermediate results (for tests): the G-buffer tensors, the HDR target,
    the shadow maps by light, the visible lists."""
    dev = "cuda"
    w = lib.World(0)
    r = lib.Raster(w)
    if world_path:
        view, default_shadow = imported_scene(w, r, world_path, asset_dir, width, height)
        shadow_size = shadow_size or default_shadow
    else:
        view, default_shadow = builtin_scene(w, r, width, height, textured=textured)
        shadow_size = shadow_size or default_shadow
    n_ent, _, n_lights = w.counts()
    d_vis = torch.zeros(n_ent + 1, dtype=torch.int32, device=dev)
    a, b, c = (torch.zeros((height, width, 4), dtype=torch.float32, device=dev) for _ in range(3))
    depth = torch.zeros((height, width), dtype=torch.float32, device=dev)
    hdr = torch.zeros((height, width, 4), dtype=torch.float32, device=dev)
    ldr = torch.zeros_like(hdr)
    d_view = torch.zeros(288, dtype=torch.uint8, device=dev)
    lights = [w.light(i)[0] for i in range(n_lights)]

    def frame():
        w.update_transforms()
        planes = lib.view_planes(view)
        _, visible_lights = w.cull(view)                                # the light list stays in the world for the light pass
        w.cull_planes_device(planes, d_vis.data_ptr() + 4, d_vis.data_ptr())
        casters = np.array([i for i in visible_lights if lights[i]["flags"] & 1 and lights[i]["type"] in (scenes.LIGHT_POINT, scenes.LIGHT_SPOT)],
                           dtype=np.uint32)
        d_maps = torch.zeros(max(n_lights, 1), dtype=torch.int64, device=dev)
        keep = []
        if len(casters):
            ll, lv, off, _, d_ent = w.cull_shadow_views_device(casters, shadow_size)
            views = np.stack([lib.light_shadow_views(lights[ll[k]], shadow_size)[lv[k]] for k in range(len(ll))])
            per_light = {}
            for i in casters:
                faces = 6 if lights[i]["type"] == scenes.LIGHT_POINT else 1
                per_light[int(i)] = torch.zeros((faces, shadow_size, shadow_size), dtype=torch.float32, device=dev)
            keep = list(per_light.values())
            ptrs = [per_light[int(ll[k])][lv[k]].data_ptr() for k in range(len(ll))]
            r.shadow_maps(views, off, d_ent, shadow_size, ptrs)
            table = np.zeros(max(n_lights, 1), dtype=np.int64)
            for i, t in per_light.items():
                table[i] = t.data_ptr()
            d_maps.copy_(torch.from_numpy(table))
        r.gbuffer(view, d_vis.data_ptr() + 4, d_vis.data_ptr(), width, height, a.data_ptr(), b.data_ptr(), c.data_ptr(), depth.data_ptr())
        w.pack_view_uniforms_device(view, d_view.data_ptr())
        w.shade_deferred_device(width, height, a.data_ptr(), b.data_ptr(), c.data_ptr(), depth.data_ptr(), d_view.data_ptr(), hdr.data_ptr(),
                                d_shadow_maps=d_maps.data_ptr() if len(casters) else 0, shadow_size=shadow_size)
        w.post_effects_device(width * height, hdr.data_ptr(), ldr.data_ptr(), lib.POST_TONEMAP | lib.POST_GAMMA)
        torch.cuda.synchronize()
        if keep_state is not None:
            vis = d_vis.cpu().numpy().view(np.uint32)
            keep_state.update(view=view, lights=lights, visible_lights=np.array(visible_lights), visible_entities=vis[1:1 + vis[0]].copy(),
                              a=a.cpu().numpy(), b=b.cpu().numpy(), c=c.cpu().numpy(), depth=depth.cpu().numpy(), hdr=hdr.cpu().numpy(),
                              ldr=ldr.cpu().numpy(), shadow_size=shadow_size, casters=casters,
                              shadow_maps={i: t.cpu().numpy() for i, t in per_light.items()} if len(casters) else {},
                              shadow_lists=(ll, lv, off, w.read_device(d_ent, int(off[-1]))) if len(casters) else None,
                              world=w)
        return keep

    frame()
    ms = None
    if timing:
        t0 = time.perf_counter()
        for _ in range(10):
            frame()
        ms = (time.perf_counter() - t0) / 10 * 1e3
    img = ldr.cpu().numpy()[::-1, :, :3]                                 # row 0 of the buffers is the bottom row
    covered = float((depth.cpu().numpy() < 1.0).mean())
    if out_path:
        with open(out_path, "wb") as f:
            f.write(b"P6\n%d %d\n255\n" % (width, height))
            f.write((np.clip(img, 0.0, 1.0) * 255.0 + 0.5).astype(np.uint8).tobytes())
    r.close()
    if keep_state is None:
        w.close()
    return img, covered, ms


if __name__ == "__main__":
    args = sys.argv[1:]
    if len(args) >= 3:
        wh = (int(args[3]), int(args[4])) if len(args) >= 5 else (960, 540)
        img, covered, ms = render(args[0], args[1], args[2], *wh, timing=True)
    else:
        out = args[0] if args else os.path.join(ROOT, "gpurun_out", "frame.ppm")
        os.makedirs(os.path.dirname(out), exist_ok=True)
        img, covered, ms = render(None, None, out, timing=True, textured=True)
    print(f"covered {covered:.3f} of the pixels, mean colour {img.reshape(-1, 3).mean(0)}, {ms:.2f} ms per frame (host loop, synchronous)")
