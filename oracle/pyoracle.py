"""ctypes front-end for the two CPU oracles (TEST INFRASTRUCTURE ONLY).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs import this module; nothing
under ``orion_b200/`` does.  ``load("port")`` is the plain-C restatement (``oracle/ovis_oracle.c``),
``load("reference")`` is the reference's own sources compiled in place (``oracle/_ref``); both export
the flat ``orc_*`` API declared in ``oracle/ovis_oracle.h``, the reference build additionally exports
the object-world ``oref_world_*`` API (tier A: SimpleRenderWorld/RenderEntity/RenderLight as shipped).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "libovis_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "liborion_ref.so")

LIGHT_AMBIENT, LIGHT_DIRECTIONAL, LIGHT_POINT, LIGHT_SPOT = 0, 1, 2, 3
CASTS_SHADOW = 1
CULL_LIGHTS = 1


class OrcLight(C.Structure):
    _fields_ = [("type", C.c_int32), ("flags", C.c_uint32), ("position", C.c_float * 3),
                ("direction", C.c_float * 3), ("colour", C.c_float * 3), ("intensity", C.c_float),
                ("range", C.c_float), ("cutoff", C.c_float)]


class OrcView(C.Structure):
    _fields_ = [("position", C.c_float * 3), ("orientation", C.c_float * 4), ("hfov_deg", C.c_float),
                ("znear", C.c_float), ("zfar", C.c_float), ("vp_width", C.c_int32), ("vp_height", C.c_int32)]


LIGHT_DTYPE = np.dtype([("type", "<i4"), ("flags", "<u4"), ("position", "<f4", 3), ("direction", "<f4", 3),
                        ("colour", "<f4", 3), ("intensity", "<f4"), ("range", "<f4"), ("cutoff", "<f4")])
VIEW_DTYPE = np.dtype([("position", "<f4", 3), ("orientation", "<f4", 4), ("hfov_deg", "<f4"),
                       ("znear", "<f4"), ("zfar", "<f4"), ("vp_width", "<i4"), ("vp_height", "<i4")])
assert LIGHT_DTYPE.itemsize == C.sizeof(OrcLight) == 56
assert VIEW_DTYPE.itemsize == C.sizeof(OrcView) == 48


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


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _f32(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float32)
    if shape is not None:
        a = a.reshape(shape)
    return a


def build(reference_root: str = "/root/reference") -> None:
    """Compile the C restatement (always) and, when the reference tree is present, oracle/_ref."""
    subprocess.run(["make", "-s", "-C", HERE, "libovis_oracle.so"], check=True)
    if os.path.isdir(os.path.join(reference_root, "engine", "runtime")):
        subprocess.run(["make", "-s", "-C", HERE, f"REF={reference_root}", "ref"], check=True)
        if os.path.exists(os.path.join(os.path.dirname(HERE), "orion_b200", "liborion_vis.so")):
            subprocess.run(["make", "-s", "-C", HERE, f"REF={reference_root}", "adapter_check"], check=True)


def available(kind: str) -> bool:
    return os.path.exists(PORT_SO if kind == "port" else REF_SO)


class Oracle:
    def __init__(self, path: str):
        self.lib = C.CDLL(path)
        L = self.lib
        L.orc_name.restype = C.c_char_p
        L.orc_cull_aabb.restype = C.c_int64
        L.orc_cull_aabb_mt.restype = C.c_int64
        L.orc_light_cull.restype = C.c_int32
        L.orc_light_shadow_views.restype = C.c_int32
        L.orc_shadow_lists.restype = C.c_int32
        self.name = L.orc_name().decode()
        self.has_world = hasattr(L, "oref_world_create")
        if self.has_world:
            L.oref_world_create.restype = C.c_void_p
            L.oref_world_add_entities.restype = C.c_int64
            L.oref_world_add_lights.restype = C.c_int32
            L.oref_world_cull.restype = C.c_int64
            L.oref_world_shadow_lists.restype = C.c_int32

    # ---- views
    def view_planes(self, view: np.ndarray) -> np.ndarray:
        view = np.ascontiguousarray(view, dtype=VIEW_DTYPE)
        out = np.zeros(24, dtype=np.float32)
        self.lib.orc_view_planes(_p(view), _p(out))
        return out.reshape(6, 4)

    def view_matrices(self, view: np.ndarray):
        view = np.ascontiguousarray(view, dtype=VIEW_DTYPE)
        mats = np.zeros(64, dtype=np.float32)
        corners = np.zeros(24, dtype=np.float32)
        self.lib.orc_view_matrices(_p(view), _p(mats), _p(corners))
        return mats.reshape(4, 16), corners.reshape(8, 3)

    # ---- transforms
    def world_trs(self, parent, local_trs) -> np.ndarray:
        parent = np.ascontiguousarray(parent, dtype=np.int32)
        local_trs = _f32(local_trs, (-1, 10))
        out = np.zeros_like(local_trs)
        self.lib.orc_world_trs(C.c_int64(len(parent)), _p(parent), _p(local_trs), _p(out))
        return out

    def trs_matrix(self, trs) -> np.ndarray:
        trs = _f32(trs, (-1, 10))
        out = np.zeros((len(trs), 16), dtype=np.float32)
        self.lib.orc_trs_matrix(C.c_int64(len(trs)), _p(trs), _p(out))
        return out

    def world_aabb(self, world_trs, local_aabb) -> np.ndarray:
        world_trs = _f32(world_trs, (-1, 10))
        local_aabb = _f32(local_aabb, (-1, 6))
        out = np.zeros_like(local_aabb)
        self.lib.orc_world_aabb(C.c_int64(len(world_trs)), _p(world_trs), _p(local_aabb), _p(out))
        return out

    # ---- cull
    def cull_aabb(self, aabb, planes, flags=None, require=0, threads=1) -> np.ndarray:
        aabb = _f32(aabb, (-1, 6))
        planes = _f32(planes, (24,))
        n = len(aabb)
        if flags is not None:
            flags = np.ascontiguousarray(flags, dtype=np.uint32)
        out = np.empty(max(n, 1), dtype=np.uint32)
        if threads > 1:
            c = self.lib.orc_cull_aabb_mt(C.c_int64(n), _p(aabb), _p(planes), _p(flags), C.c_uint32(require),
                                          _p(out), C.c_int(threads))
        else:
            c = self.lib.orc_cull_aabb(C.c_int64(n), _p(aabb), _p(planes), _p(flags), C.c_uint32(require), _p(out))
        return out[:c].copy()

    # ---- lights
    def light_spot_aabb(self, light: np.ndarray) -> np.ndarray:
        light = np.ascontiguousarray(light, dtype=LIGHT_DTYPE)
        out = np.zeros(6, dtype=np.float32)
        self.lib.orc_light_spot_aabb(_p(light), _p(out))
        return out

    def light_cull(self, lights: np.ndarray, planes) -> np.ndarray:
        lights = np.ascontiguousarray(lights, dtype=LIGHT_DTYPE).reshape(-1)
        planes = _f32(planes, (24,))
        out = np.empty(max(len(lights), 1), dtype=np.uint32)
        c = self.lib.orc_light_cull(C.c_int32(len(lights)), _p(lights), _p(planes), _p(out))
        return out[:c].copy()

    def light_shadow_views(self, light: np.ndarray) -> np.ndarray:
        light = np.ascontiguousarray(light, dtype=LIGHT_DTYPE)
        out = np.zeros(6, dtype=VIEW_DTYPE)
        c = self.lib.orc_light_shadow_views(_p(light), _p(out))
        return out[:c].copy()

    def shadow_lists(self, aabb, flags, lights, main_planes):
        aabb = _f32(aabb, (-1, 6))
        n = len(aabb)
        flags = np.ascontiguousarray(flags, dtype=np.uint32) if flags is not None else None
        lights = np.ascontiguousarray(lights, dtype=LIGHT_DTYPE).reshape(-1)
        main_planes = _f32(main_planes, (24,))
        cap_lists = 6 * max(len(lights), 1)
        cap_ent = cap_lists * max(n, 1)
        ll = np.zeros(cap_lists, dtype=np.int32)
        lv = np.zeros(cap_lists, dtype=np.int32)
        off = np.zeros(cap_lists + 1, dtype=np.int64)
        ent = np.zeros(cap_ent, dtype=np.uint32)
        k = self.lib.orc_shadow_lists(C.c_int64(n), _p(aabb), _p(flags), C.c_int32(len(lights)), _p(lights),
                                      _p(main_planes), C.c_int32(cap_lists), _p(ll), _p(lv), _p(off),
                                      C.c_int64(cap_ent), _p(ent))
        assert k >= 0
        return ll[:k].copy(), lv[:k].copy(), off[:k + 1].copy(), ent[:off[k]].copy()

    # ---- draw lists / uniforms (section 8(f))
    def draw_lists(self, visible, pass_counts, flags, mode=0):
        visible = np.ascontiguousarray(visible, dtype=np.uint32)
        pass_counts = np.ascontiguousarray(pass_counts, dtype=np.uint32)
        flags = np.ascontiguousarray(flags, dtype=np.uint32)
        pc = pass_counts[visible] if len(visible) else np.zeros(0, np.uint32)
        cap = int(((pc & 0xff) + ((pc >> 8) & 0xff) + ((pc >> 16) & 0xff)).sum(dtype=np.int64)) + 1
        d0 = np.zeros((cap, 2), dtype=np.uint32)
        d1 = np.zeros((cap, 2), dtype=np.uint32)
        counts = np.zeros(2, dtype=np.int64)
        self.lib.orc_draw_lists(C.c_int64(len(visible)), _p(visible), _p(pass_counts), _p(flags), C.c_int32(mode),
                                _p(d0), _p(d1), _p(counts))
        return d0[:counts[0]].copy(), d1[:counts[1]].copy()

    def entity_uniforms(self, world_trs) -> np.ndarray:
        world_trs = _f32(world_trs, (-1, 10))
        out = np.zeros((len(world_trs), 20), dtype=np.float32)
        self.lib.orc_entity_uniforms(C.c_int64(len(world_trs)), _p(world_trs), _p(out))
        return out

    def quat_rotate(self, q, v) -> np.ndarray:
        q, v = _f32(q, (4,)), _f32(v, (3,))
        out = np.zeros(3, dtype=np.float32)
        self.lib.orc_quat_rotate(_p(q), _p(v), _p(out))
        return out

    def uniform_sizes(self):
        out = np.zeros(3, dtype=np.int32)
        self.lib.orc_uniform_sizes(_p(out))
        return tuple(int(x) for x in out)

    def entity_uniform_blocks(self, world_trs) -> np.ndarray:
        """The 80-byte blocks as they sit in the entities' uniform buffers (reference tier: through RenderEntity)."""
        world_trs = _f32(world_trs, (-1, 10))
        out = np.zeros((len(world_trs), 20), dtype=np.float32)
        self.lib.orc_entity_uniform_blocks(C.c_int64(len(world_trs)), _p(world_trs), _p(out))
        return out

    def view_uniforms(self, view, viewport_pos=(0, 0)) -> np.ndarray:
        view = np.ascontiguousarray(view, dtype=VIEW_DTYPE).reshape(1)
        pos = np.asarray(viewport_pos, dtype=np.int32)
        out = np.zeros(288, dtype=np.uint8)
        self.lib.orc_view_uniforms(_p(view), _p(pos), _p(out))
        return out

    def light_uniforms(self, light, attenuation=(1.0, 0.045, 0.0075), shadow_bias=-0.001) -> np.ndarray:
        light = np.ascontiguousarray(light, dtype=LIGHT_DTYPE).reshape(1)
        att = np.asarray(attenuation, dtype=np.float32)
        out = np.zeros(208, dtype=np.uint8)
        self.lib.orc_light_uniforms(_p(light), _p(att), C.c_float(shadow_bias), _p(out))
        return out

    # ---- headless rasteriser (ovis_raster.c)
    def raster(self, matrix, mesh, material, vertices, indices, vp, width, height, depth_only=False):
        """G-buffer pass (or a depth-only pass) over instances drawn in order -> (a, b, c, depth), rows bottom-up."""
        matrix = _f32(matrix, (-1, 16))
        n = len(matrix)
        mesh = np.ascontiguousarray(mesh, dtype=np.int64).reshape(n, 3)
        material = _f32(material, (n, 8))
        vertices = _f32(vertices, (-1, 8))
        indices = np.ascontiguousarray(indices, dtype=np.uint32)
        vp = _f32(vp, (16,))
        depth = np.zeros((height, width), dtype=np.float32)
        if depth_only:
            a = b = c = None
        else:
            a, b, c = (np.zeros((height, width, 4), dtype=np.float32) for _ in range(3))
        self.lib.orc_raster(C.c_int64(n), _p(matrix), _p(mesh), _p(material), _p(vertices), _p(indices), _p(vp), C.c_int32(width),
                            C.c_int32(height), _p(a), _p(b), _p(c), _p(depth))
        return a, b, c, depth

    # ---- textured materials
    class _Tex(C.Structure):
        _fields_ = [("rgba", C.c_void_p), ("width", C.c_int32), ("height", C.c_int32), ("srgb", C.c_int32), ("wrap", C.c_int32)]

    def _textures(self, textures):
        """textures: list of (rgba uint8 [h, w, 4], srgb, wrap) -> (ctypes array, keep-alive list)."""
        arr = (self._Tex * max(len(textures), 1))()
        keep = []
        for i, (rgba, srgb, wrap) in enumerate(textures):
            rgba = np.ascontiguousarray(rgba, dtype=np.uint8)
            keep.append(rgba)
            arr[i] = self._Tex(rgba.ctypes.data, rgba.shape[1], rgba.shape[0], int(bool(srgb)), int(bool(wrap)))
        return arr, keep

    def texture_sample(self, texture, u, v):
        arr, keep = self._textures([texture])
        out = np.zeros(4, dtype=np.float32)
        self.lib.orc_texture_sample(C.byref(arr[0]), C.c_float(u), C.c_float(v), _p(out))
        return out

    def raster_textured(self, matrix, mesh, material, vertices, tangents, indices, textures, vp, width, height):
        """orc_raster_textured -> (a, b, c, depth); material [n, 10], tangents [vertices, 4]."""
        matrix = _f32(matrix, (-1, 16))
        n = len(matrix)
        mesh = np.ascontiguousarray(mesh, dtype=np.int64).reshape(n, 3)
        material = _f32(material, (n, 10))
        vertices = _f32(vertices, (-1, 8))
        tangents = _f32(tangents, (len(vertices), 4))
        indices = np.ascontiguousarray(indices, dtype=np.uint32)
        vp = _f32(vp, (16,))
        arr, keep = self._textures(textures)
        depth = np.zeros((height, width), dtype=np.float32)
        a, b, c = (np.zeros((height, width, 4), dtype=np.float32) for _ in range(3))
        self.lib.orc_raster_textured(C.c_int64(n), _p(matrix), _p(mesh), _p(material), _p(vertices), _p(tangents), _p(indices), arr, _p(vp),
                                     C.c_int32(width), C.c_int32(height), _p(a), _p(b), _p(c), _p(depth))
        return a, b, c, depth

    def raster_fragment_textured(self, normal, uv, tangent, bitangent, material, textures, glsl=False):
        arr, keep = self._textures(textures)
        a, b, c = (np.zeros(4, dtype=np.float32) for _ in range(3))
        fn = self.lib.orc_glsl_lit_fragment_textured if glsl else self.lib.orc_raster_fragment_textured
        fn(_p(_f32(normal, (3,))), _p(_f32(uv, (2,))), _p(_f32(tangent, (3,))), _p(_f32(bitangent, (3,))), _p(_f32(material, (10,))), arr,
           _p(a), _p(b), _p(c))
        return a, b, c

    def raster_bitangent(self, n, t, w):
        out = np.zeros(3, dtype=np.float32)
        self.lib.orc_raster_bitangent(_p(_f32(n, (3,))), _p(_f32(t, (3,))), C.c_float(w), _p(out))
        return out

    def glsl_vertex_normalmap(self, transform, view_projection, position, normal, tangent4):
        """lit_vtx.glsl with NORMALMAP from the reference's source -> (world normal, tangent, bitangent)."""
        n, t, b = (np.zeros(3, dtype=np.float32) for _ in range(3))
        self.lib.orc_glsl_lit_vertex_normalmap(_p(_f32(transform, (16,))), _p(_f32(view_projection, (16,))), _p(_f32(position, (3,))),
                                               _p(_f32(normal, (3,))), _p(_f32(tangent4, (4,))), _p(n), _p(t), _p(b))
        return n, t, b

    def raster_basic(self, matrix, mesh, material, vertices, indices, vp, rgba, depth):
        """The Basic (unlit) pass over rgba [h, w, 4] and depth [h, w] as they stand; returns the updated copies."""
        matrix = _f32(matrix, (-1, 16))
        n = len(matrix)
        mesh = np.ascontiguousarray(mesh, dtype=np.int64).reshape(n, 3)
        material = _f32(material, (n, 8))
        vertices = _f32(vertices, (-1, 8))
        indices = np.ascontiguousarray(indices, dtype=np.uint32)
        vp = _f32(vp, (16,))
        h, w = depth.shape
        rgba, depth = _f32(rgba, (h, w, 4)).copy(), _f32(depth, (h, w)).copy()
        self.lib.orc_raster_basic(C.c_int64(n), _p(matrix), _p(mesh), _p(material), _p(vertices), _p(indices), _p(vp), C.c_int32(w),
                                  C.c_int32(h), _p(rgba), _p(depth))
        return rgba, depth

    def glsl_unlit(self, transform, view_projection, position, diffuse):
        """unlit_vtx.glsl + unlit_frag.glsl from the reference's source (reference tier only) -> (clip[4], colour[4])."""
        clip, colour = np.zeros(4, dtype=np.float32), np.zeros(4, dtype=np.float32)
        self.lib.orc_glsl_unlit(_p(_f32(transform, (16,))), _p(_f32(view_projection, (16,))), _p(_f32(position, (3,))),
                                _p(_f32(diffuse, (3,))), _p(clip), _p(colour))
        return clip, colour

    def raster_vertex(self, transform, view_projection, position, normal, glsl=False, shadow=False):
        """lit_vtx.glsl / shadow_vtx.glsl for one vertex -> (clip[4], world normal[3]); glsl=True runs the reference's shader
        source compiled against GLM (reference tier only)."""
        t, vp = _f32(transform, (16,)), _f32(view_projection, (16,))
        pos, nrm = _f32(position, (3,)), _f32(normal, (3,))
        clip, out_n = np.zeros(4, dtype=np.float32), np.zeros(3, dtype=np.float32)
        if glsl and shadow:
            self.lib.orc_glsl_shadow_vertex(_p(t), _p(vp), _p(pos), _p(clip))
        elif glsl:
            self.lib.orc_glsl_lit_vertex(_p(t), _p(vp), _p(pos), _p(nrm), _p(clip), _p(out_n))
        else:
            self.lib.orc_raster_vertex(_p(t), _p(vp), _p(pos), _p(nrm), _p(clip), _p(out_n))
        return clip, out_n

    def raster_fragment(self, normal, material, glsl=False):
        """lit_deferred_frag.glsl (no TEXTURED / NORMALMAP) for one fragment -> (bufferA, bufferB, bufferC)."""
        n, m = _f32(normal, (3,)), _f32(material, (8,))
        a, b, c = (np.zeros(4, dtype=np.float32) for _ in range(3))
        (self.lib.orc_glsl_lit_fragment if glsl else self.lib.orc_raster_fragment)(_p(n), _p(m), _p(a), _p(b), _p(c))
        return a, b, c

    # ---- deferred shading and post effects (section 8(f) row 4)
    def shade_deferred(self, buffer_a, buffer_b, buffer_c, depth, view_uniforms, light_uniforms, light_types,
                       shadow_maps=None, shadow_size=0, glsl=False):
        """Sum of the lights' contributions per pixel -> (h, w, 4) float32. glsl=True runs the reference's shader source
        compiled against GLM (reference tier only) instead of the C restatement."""
        h, w = depth.shape
        a, b, c = (_f32(x, (h, w, 4)) for x in (buffer_a, buffer_b, buffer_c))
        depth = _f32(depth, (h, w))
        vu = np.ascontiguousarray(view_uniforms, dtype=np.uint8).reshape(288)
        lu = np.ascontiguousarray(light_uniforms).view(np.float32).reshape(-1, 52)
        types = np.ascontiguousarray(light_types, dtype=np.int32).reshape(-1)
        assert len(types) == len(lu)
        maps = None
        keep = []
        if shadow_maps is not None:
            maps = (C.c_void_p * len(lu))()
            for i, m in enumerate(shadow_maps):
                if m is not None:
                    m = np.ascontiguousarray(m, dtype=np.float32)
                    keep.append(m)
                    maps[i] = m.ctypes.data
        out = np.zeros((h, w, 4), dtype=np.float32)
        fn = self.lib.orc_glsl_shade_deferred if glsl else self.lib.orc_shade_deferred
        fn(C.c_int32(w), C.c_int32(h), _p(a), _p(b), _p(c), _p(depth), _p(vu), C.c_int32(len(lu)), _p(lu), _p(types), maps,
           C.c_int32(shadow_size), _p(out))
        return out

    def tonemap(self, rgba, exposure, white_point, glsl=False):
        rgba = _f32(rgba, (-1, 4))
        out = np.zeros_like(rgba)
        fn = self.lib.orc_glsl_tonemap if glsl else self.lib.orc_tonemap
        fn(C.c_int64(len(rgba)), _p(rgba), C.c_float(exposure), C.c_float(white_point), _p(out))
        return out

    def gamma(self, rgba, gamma, glsl=False):
        rgba = _f32(rgba, (-1, 4))
        out = np.zeros_like(rgba)
        fn = self.lib.orc_glsl_gamma if glsl else self.lib.orc_gamma
        fn(C.c_int64(len(rgba)), _p(rgba), C.c_float(gamma), _p(out))
        return out

    # ---- tier A object world (reference build only)
    def world(self) -> "RefWorld":
        assert self.has_world, "object world exists only in the reference-compiled oracle"
        return RefWorld(self)


class RefWorld:
    """SimpleRenderWorld + RenderEntity + RenderLight + RenderView exactly as the reference ships them."""

    def __init__(self, oracle: Oracle):
        self.o = oracle
        self.lib = oracle.lib
        self.h = C.c_void_p(self.lib.oref_world_create())
        self.n_entities = 0
        self.n_lights = 0

    def close(self):
        if self.h:
            self.lib.oref_world_destroy(self.h)
            self.h = None

    def abandon(self):
        """Leak the world on purpose: destroying n entities one by one is O(n^2) in the reference."""
        if self.h:
            self.lib.oref_world_abandon(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def add_entities(self, trs, local_aabb, flags=None) -> int:
        trs = _f32(trs, (-1, 10))
        local_aabb = _f32(local_aabb, (-1, 6))
        flags = np.ascontiguousarray(flags, dtype=np.uint32) if flags is not None else None
        first = self.lib.oref_world_add_entities(self.h, C.c_int64(len(trs)), _p(trs), _p(local_aabb), _p(flags))
        self.n_entities += len(trs)
        return first

    def set_transforms(self, idx, trs):
        idx = np.ascontiguousarray(idx, dtype=np.uint32)
        trs = _f32(trs, (-1, 10))
        self.lib.oref_world_set_transforms(self.h, C.c_int64(len(idx)), _p(idx), _p(trs))

    def set_local_aabb(self, idx, box):
        box = _f32(box, (6,))
        self.lib.oref_world_set_local_aabb(self.h, C.c_uint32(idx), _p(box))

    def set_flags(self, idx, flags):
        self.lib.oref_world_set_flags(self.h, C.c_uint32(idx), C.c_uint32(flags))

    def remove_entity(self, idx):
        self.lib.oref_world_remove_entity(self.h, C.c_uint32(idx))

    def bounds(self, first=0, n=None) -> np.ndarray:
        n = self.n_entities - first if n is None else n
        out = np.zeros((n, 6), dtype=np.float32)
        self.lib.oref_world_get_bounds(self.h, C.c_int64(first), C.c_int64(n), _p(out))
        return out

    def matrices(self, first=0, n=None) -> np.ndarray:
        n = self.n_entities - first if n is None else n
        out = np.zeros((n, 16), dtype=np.float32)
        self.lib.oref_world_get_matrices(self.h, C.c_int64(first), C.c_int64(n), _p(out))
        return out

    def add_lights(self, lights) -> int:
        lights = np.ascontiguousarray(lights, dtype=LIGHT_DTYPE).reshape(-1)
        first = self.lib.oref_world_add_lights(self.h, C.c_int32(len(lights)), _p(lights))
        self.n_lights += len(lights)
        return first

    def update_light(self, idx, light):
        light = np.ascontiguousarray(light, dtype=LIGHT_DTYPE)
        self.lib.oref_world_update_light(self.h, C.c_uint32(idx), _p(light))

    def remove_light(self, idx):
        self.lib.oref_world_remove_light(self.h, C.c_uint32(idx))

    def cull(self, view, flags=CULL_LIGHTS):
        ents = np.empty(max(self.n_entities, 1), dtype=np.uint32)
        lights = np.empty(max(self.n_lights, 1), dtype=np.uint32)
        nl = C.c_int32(0)
        ne = self.lib.oref_world_cull(self.h, _p(view), C.c_uint32(flags), _p(ents), _p(lights), C.byref(nl))
        return ents[:ne].copy(), lights[:nl.value].copy()

    def shadow_lists(self, main_view, shadow_res=512):
        cap_lists = 6 * max(self.n_lights, 1)
        cap_ent = cap_lists * max(self.n_entities, 1)
        ll = np.zeros(cap_lists, dtype=np.int32)
        lv = np.zeros(cap_lists, dtype=np.int32)
        off = np.zeros(cap_lists + 1, dtype=np.int64)
        ent = np.zeros(cap_ent, dtype=np.uint32)
        k = self.lib.oref_world_shadow_lists(self.h, _p(main_view), C.c_int32(shadow_res), C.c_int32(cap_lists),
                                             _p(ll), _p(lv), _p(off), C.c_int64(cap_ent), _p(ent))
        assert k >= 0
        return ll[:k].copy(), lv[:k].copy(), off[:k + 1].copy(), ent[:off[k]].copy()

    def time_cull(self, view, flags=0, iters=5):
        ms = np.zeros(iters, dtype=np.float64)
        vis = C.c_int64(0)
        self.lib.oref_world_time_cull(self.h, _p(view), C.c_uint32(flags), C.c_int32(iters), _p(ms), C.byref(vis))
        return ms, vis.value


_cache: dict[str, Oracle] = {}


def load(kind: str = "port") -> Oracle:
    """kind = "port" | "reference" | "best" (reference when built, else port)."""
    if kind == "best":
        kind = "reference" if available("reference") else "port"
    if kind not in _cache:
        path = PORT_SO if kind == "port" else REF_SO
        if not os.path.exists(path):
            if kind == "port":
                build()
            else:
                raise FileNotFoundError(f"{path} is not built (needs /root/reference; run make -C oracle ref)")
        _cache[kind] = Oracle(path)
    return _cache[kind]
