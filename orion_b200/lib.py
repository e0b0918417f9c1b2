"""ctypes binding of liborion_vis.so (include/orion_vis.h) - what a cgo/JNI/ctypes user of the C ABI writes.

The library is the product: if it is missing, or no CUDA device is present, calls fail loudly
(``OvisError``); there is no Python or CPU fallback anywhere in this package.
"""
from __future__ import annotations

import ctypes as C
import os
import re
import subprocess

import numpy as np

from . import scenes

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB_PATH = os.path.join(HERE, "liborion_vis.so")
HEADER = os.path.join(ROOT, "include", "orion_vis.h")

OVIS_OK, OVIS_ERR_INVALID, OVIS_ERR_CUDA, OVIS_ERR_CAPACITY, OVIS_ERR_NOMEM = 0, 1, 2, 3, 4
CASTS_SHADOW = 1
CULL_LIGHTS = 1
DRAWS_MAIN, DRAWS_SHADOW = 0, 1
NO_PARENT = -1

LIGHT_DTYPE = scenes.LIGHT_DTYPE
VIEW_DTYPE = scenes.VIEW_DTYPE


class OvisError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"liborion_vis error {code}: {message}")
        self.code = code


def build(verbose: bool = False) -> str:
    """Compile the CUDA library in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    subprocess.run(["make", "-j8", "-C", os.path.join(HERE, "csrc")] + ([] if verbose else ["-s"]), check=True)
    return LIB_PATH


def declared_symbols() -> list[str]:
    """Every entry point include/orion_vis.h declares."""
    text = open(HEADER).read()
    return sorted(set(re.findall(r"\b(ovis_[a-z0-9_]+)\s*\(", text)))


_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OvisError(OVIS_ERR_CUDA, f"{LIB_PATH} is not built; run `make -C orion_b200/csrc` "
                                           "(there is no CPU fallback)")
        lib = C.CDLL(LIB_PATH)
        lib.ovis_last_error.restype = C.c_char_p
        lib.ovis_version.restype = C.c_char_p
        _lib = lib
    return _lib


def _check(rc: int):
    if rc != OVIS_OK:
        raise OvisError(rc, load().ovis_last_error().decode("utf-8", "replace"))


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _f32(a, shape):
    return np.ascontiguousarray(a, dtype=np.float32).reshape(shape)


def _u32(a):
    return np.ascontiguousarray(a, dtype=np.uint32).reshape(-1)


# ---- host-side set-up (no device needed) -----------------------------------------------------------

def view_planes(view: np.ndarray) -> np.ndarray:
    view = np.ascontiguousarray(view, dtype=VIEW_DTYPE)
    out = np.zeros(24, dtype=np.float32)
    _check(load().ovis_view_planes(_p(view), _p(out)))
    return out.reshape(6, 4)


def view_matrices(view: np.ndarray):
    view = np.ascontiguousarray(view, dtype=VIEW_DTYPE)
    mats = np.zeros(64, dtype=np.float32)
    corners = np.zeros(24, dtype=np.float32)
    _check(load().ovis_view_matrices(_p(view), _p(mats), _p(corners)))
    return mats.reshape(4, 16), corners.reshape(8, 3)


def light_spot_aabb(light: np.ndarray) -> np.ndarray:
    light = np.ascontiguousarray(light, dtype=LIGHT_DTYPE)
    out = np.zeros(6, dtype=np.float32)
    _check(load().ovis_light_spot_aabb(_p(light), _p(out)))
    return out


def view_uniforms(view: np.ndarray, viewport_pos=(0, 0)) -> np.ndarray:
    """ViewUniforms block (288 bytes) of a view."""
    view = np.ascontiguousarray(view, dtype=VIEW_DTYPE).reshape(1)
    pos = np.asarray(viewport_pos, dtype=np.int32)
    out = np.zeros(288, dtype=np.uint8)
    _check(load().ovis_view_uniforms(_p(view), _p(pos), _p(out)))
    return out


def light_uniforms(light: np.ndarray, attenuation=(1.0, 0.045, 0.0075), shadow_bias=-0.001) -> np.ndarray:
    """LightUniforms block (208 bytes) of a light."""
    light = np.ascontiguousarray(light, dtype=LIGHT_DTYPE).reshape(1)
    att = np.asarray(attenuation, dtype=np.float32)
    out = np.zeros(208, dtype=np.uint8)
    _check(load().ovis_light_uniforms(_p(light), _p(att), C.c_float(shadow_bias), _p(out)))
    return out


def obj_submesh_aabbs(path: str) -> np.ndarray:
    """Local AABB of every sub-mesh of a Wavefront .obj (host-side parsing; no device needed)."""
    cap = 64
    while True:
        out = np.zeros((cap, 6), dtype=np.float32)
        n = C.c_int32()
        rc = load().ovis_obj_submesh_aabbs(os.fsencode(path), C.c_int32(cap), _p(out), C.byref(n))
        if rc == OVIS_ERR_CAPACITY:
            cap = n.value
            continue
        _check(rc)
        return out[:n.value].copy()


class GBuffer(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("d_buffer_a", C.c_void_p), ("d_buffer_b", C.c_void_p),
                ("d_buffer_c", C.c_void_p), ("d_depth", C.c_void_p)]


POST_TONEMAP, POST_GAMMA = 1, 2


class SceneInfo(C.Structure):
    _fields_ = [("first_node", C.c_uint32), ("first_entity", C.c_uint32), ("first_light", C.c_uint32),
                ("n_nodes", C.c_int32), ("n_entities", C.c_int64), ("n_lights", C.c_int32), ("n_cameras", C.c_int32),
                ("shadow_map_resolution", C.c_int32)]


def load_scene(world_path: str, game_asset_dir: str, vp_width: int, vp_height: int) -> dict:
    """ovis_scene_load + ovis_scene_arrays: the parsed render-world content of a serialised world (host only)."""
    lib = load()
    h = C.c_void_p()
    _check(lib.ovis_scene_load(os.fsencode(world_path), os.fsencode(game_asset_dir), C.c_int32(vp_width), C.c_int32(vp_height),
                               C.byref(h)))
    try:
        info = SceneInfo()
        _check(lib.ovis_scene_info(h, C.byref(info)))
        out = {"parent": np.zeros(info.n_nodes, dtype=np.int32), "local_trs": np.zeros((info.n_nodes, 10), dtype=np.float32),
               "local_aabb": np.zeros((info.n_entities, 6), dtype=np.float32), "flags": np.zeros(info.n_entities, dtype=np.uint32),
               "entity_node": np.zeros(info.n_entities, dtype=np.uint32), "lights": np.zeros(info.n_lights, dtype=LIGHT_DTYPE),
               "shading": np.zeros((info.n_lights, 4), dtype=np.float32), "cameras": np.zeros(info.n_cameras, dtype=VIEW_DTYPE),
               "shadow_map_resolution": info.shadow_map_resolution}
        _check(lib.ovis_scene_arrays(h, _p(out["parent"]), _p(out["local_trs"]), _p(out["local_aabb"]), _p(out["flags"]),
                                     _p(out["entity_node"]), _p(out["lights"]), _p(out["shading"]), _p(out["cameras"])))
        meshes = []                                  # per render entity: (.obj path or "", sub-mesh index or -1)
        buf = C.create_string_buffer(4096)
        sub = C.c_int32()
        for e in range(info.n_entities):
            _check(lib.ovis_scene_entity_mesh(h, C.c_int64(e), buf, C.c_int32(len(buf)), C.byref(sub)))
            meshes.append((buf.value.decode(), sub.value))
        out["entity_mesh"] = meshes
        return out
    finally:
        lib.ovis_scene_free(h)


def light_shadow_views(light: np.ndarray, resolution: int = 512) -> np.ndarray:
    light = np.ascontiguousarray(light, dtype=LIGHT_DTYPE)
    out = np.zeros(6, dtype=VIEW_DTYPE)
    n = C.c_int32(0)
    _check(load().ovis_light_shadow_views(_p(light), C.c_int32(resolution), _p(out), C.byref(n)))
    return out[:n.value].copy()


NOALLOC_MAX_ENTITIES = 1 << 20


def cull_planes_noalloc(device: int, stream: int, d_world_aabb: int, stride: int, n_entities: int, planes,
                        d_out: int, capacity: int, d_count: int, id_base: int = 0):
    """ovis_cull_planes_noalloc: the single-view cull with library-internal fixed scratch (no allocation).

    All pointers are raw device addresses owned by the caller; asynchronous on `stream`.
    """
    planes = _f32(planes, (24,))
    _check(load().ovis_cull_planes_noalloc(C.c_int(device), C.c_void_p(stream), C.c_void_p(d_world_aabb),
                                           C.c_int64(stride), C.c_int64(n_entities), _p(planes), C.c_uint32(id_base),
                                           C.c_void_p(d_out), C.c_int64(capacity), C.c_void_p(d_count)))


class PinnedBuffer:
    """Page-locked host memory from ovis_host_alloc, viewed as a numpy array."""

    def __init__(self, count: int, dtype=np.uint32):
        self.ptr = C.c_void_p()
        self.dtype = np.dtype(dtype)
        nbytes = max(int(count), 1) * self.dtype.itemsize
        _check(load().ovis_host_alloc(C.c_int64(nbytes), C.byref(self.ptr)))
        buf = (C.c_char * nbytes).from_address(self.ptr.value)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=max(int(count), 1))

    def free(self):
        if self.ptr:
            self.array = None
            load().ovis_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def pin_thread_near_device(device: int) -> int:
    """ovis_host_pin_thread_near_device: keep the calling thread (and those it starts later) on the CPUs next to `device`;
    returns the number of CPUs pinned to (0: nothing changed)."""
    n = C.c_int32()
    _check(load().ovis_host_pin_thread_near_device(C.c_int32(device), C.byref(n)))
    return n.value


def host_register(array: np.ndarray):
    """Page-lock (portable, mapped) memory the caller owns, e.g. a shared-memory segment: ovis_host_register."""
    _check(load().ovis_host_register(C.c_void_p(array.ctypes.data), C.c_int64(array.nbytes)))


def host_unregister(array: np.ndarray):
    _check(load().ovis_host_unregister(C.c_void_p(array.ctypes.data)))


class World:
    """Device-resident render world: the RenderWorld of engine/runtime/render/include/render_world.h:58-98."""

    def __init__(self, device: int = 0):
        self.lib = load()
        self.h = C.c_void_p()
        _check(self.lib.ovis_world_create(C.c_int(device), C.byref(self.h)))
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            self.lib.ovis_world_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- bookkeeping
    def reserve(self, n_entities=0, n_nodes=0):
        _check(self.lib.ovis_world_reserve(self.h, C.c_int64(n_entities), C.c_int64(n_nodes)))

    def set_stream(self, cuda_stream: int):
        _check(self.lib.ovis_world_set_stream(self.h, C.c_void_p(cuda_stream)))

    def set_id_base(self, id_base: int):
        _check(self.lib.ovis_world_set_id_base(self.h, C.c_uint32(id_base)))

    def counts(self):
        ne, nn, nl = C.c_int64(), C.c_int64(), C.c_int32()
        _check(self.lib.ovis_world_counts(self.h, C.byref(ne), C.byref(nn), C.byref(nl)))
        return ne.value, nn.value, nl.value

    def set_timing(self, enabled: bool):
        _check(self.lib.ovis_world_set_timing(self.h, C.c_int(1 if enabled else 0)))

    def read_timing(self, capacity: int = 65536):
        """Per-cull device times (ms) of the test and gather kernels recorded since set_timing(True)."""
        test = np.zeros(capacity, dtype=np.float32)
        gather = np.zeros(capacity, dtype=np.float32)
        n = C.c_int32()
        _check(self.lib.ovis_world_read_timing(self.h, C.c_int32(capacity), _p(test), _p(gather), C.byref(n)))
        k = min(n.value, capacity)
        return test[:k].copy(), gather[:k].copy()

    def launch_count(self) -> int:
        n = C.c_int64()
        _check(self.lib.ovis_world_launch_count(self.h, C.byref(n)))
        return n.value

    # ---- entities
    def add_entities(self, local_aabb, flags=None) -> int:
        local_aabb = _f32(local_aabb, (-1, 6))
        flags = _u32(flags) if flags is not None else None
        first = C.c_uint32()
        _check(self.lib.ovis_entity_add(self.h, C.c_int64(len(local_aabb)), _p(local_aabb), _p(flags), C.byref(first)))
        return first.value

    def remove_entities(self, ids):
        ids = _u32(ids)
        _check(self.lib.ovis_entity_remove(self.h, C.c_int64(len(ids)), _p(ids)))

    def set_local_aabbs(self, ids, local_aabb):
        ids = _u32(ids)
        local_aabb = _f32(local_aabb, (-1, 6))
        _check(self.lib.ovis_entity_set_local_aabbs(self.h, C.c_int64(len(ids)), _p(ids), _p(local_aabb)))

    def set_flags(self, ids, flags):
        ids, flags = _u32(ids), _u32(flags)
        _check(self.lib.ovis_entity_set_flags(self.h, C.c_int64(len(ids)), _p(ids), _p(flags)))

    def set_transforms(self, ids, world_trs):
        ids = _u32(ids)
        world_trs = _f32(world_trs, (-1, 10))
        _check(self.lib.ovis_entity_set_transforms(self.h, C.c_int64(len(ids)), _p(ids), _p(world_trs)))

    def set_world_aabbs(self, first_id, world_aabb):
        world_aabb = _f32(world_aabb, (-1, 6))
        _check(self.lib.ovis_entity_set_world_aabbs(self.h, C.c_uint32(first_id), C.c_int64(len(world_aabb)), _p(world_aabb)))

    def set_world_aabbs_by_id(self, ids, world_aabb):
        ids = _u32(ids)
        world_aabb = _f32(world_aabb, (-1, 6))
        _check(self.lib.ovis_entity_set_world_aabbs_by_id(self.h, C.c_int64(len(ids)), _p(ids), _p(world_aabb)))

    # ---- transform tree
    def add_nodes(self, parent, local_trs) -> int:
        parent = np.ascontiguousarray(parent, dtype=np.int32).reshape(-1)
        local_trs = _f32(local_trs, (-1, 10))
        first = C.c_uint32()
        _check(self.lib.ovis_node_add(self.h, C.c_int64(len(parent)), _p(parent), _p(local_trs), C.byref(first)))
        return first.value

    def set_node_local_trs(self, ids, local_trs):
        ids = _u32(ids)
        local_trs = _f32(local_trs, (-1, 10))
        _check(self.lib.ovis_node_set_local_trs(self.h, C.c_int64(len(ids)), _p(ids), _p(local_trs)))

    def bind_entities(self, node_ids, entity_ids):
        node_ids, entity_ids = _u32(node_ids), _u32(entity_ids)
        _check(self.lib.ovis_node_bind_entities(self.h, C.c_int64(len(node_ids)), _p(node_ids), _p(entity_ids)))

    def update_transforms(self):
        _check(self.lib.ovis_update_transforms(self.h))

    # ---- lights
    def add_lights(self, lights) -> int:
        lights = np.ascontiguousarray(lights, dtype=LIGHT_DTYPE).reshape(-1)
        first = C.c_uint32()
        _check(self.lib.ovis_light_add(self.h, C.c_int32(len(lights)), _p(lights), C.byref(first)))
        return first.value

    def update_light(self, light_id, light):
        light = np.ascontiguousarray(light, dtype=LIGHT_DTYPE)
        _check(self.lib.ovis_light_update(self.h, C.c_uint32(light_id), _p(light)))

    def remove_light(self, light_id):
        _check(self.lib.ovis_light_remove(self.h, C.c_uint32(light_id)))

    # ---- culling
    def cull(self, view, flags=CULL_LIGHTS, out_entities=None, out_lights=None):
        """RenderWorld::cull(view, results, flags) -> (entity ids, light ids), host arrays."""
        view = np.ascontiguousarray(view, dtype=VIEW_DTYPE)
        ne_slots, _, nl_slots = self.counts()
        ents = out_entities if out_entities is not None else np.empty(max(ne_slots, 1), dtype=np.uint32)
        lights = out_lights if out_lights is not None else np.empty(max(nl_slots, 1), dtype=np.uint32)
        ne, nl = C.c_int64(), C.c_int32()
        _check(self.lib.ovis_cull(self.h, _p(view), C.c_uint32(flags), _p(ents), C.c_int64(len(ents)), C.byref(ne),
                                  _p(lights), C.c_int32(len(lights)), C.byref(nl)))
        return ents[:ne.value], lights[:nl.value]

    def bind_cull(self, view, out_entities: np.ndarray, flags=0, out_lights: np.ndarray | None = None):
        """ovis_cull with its arguments marshalled once, for a frame loop: returns step() -> (n_entities, n_lights).
        `view` (a 1-element VIEW_DTYPE array) and the output arrays are referenced, not copied: update the view in place
        between frames and read the ids from out_entities[:n_entities]."""
        view = np.ascontiguousarray(view, dtype=VIEW_DTYPE).reshape(1)
        ne, nl = C.c_int64(), C.c_int32()
        fn, h = self.lib.ovis_cull, self.h
        args = (h, _p(view), C.c_uint32(flags), _p(out_entities), C.c_int64(len(out_entities)), C.byref(ne),
                _p(out_lights), C.c_int32(len(out_lights) if out_lights is not None else 0), C.byref(nl))
        keep = (view, out_entities, out_lights)

        def step():
            rc = fn(*args)
            if rc != OVIS_OK:
                _check(rc)
            return ne.value, nl.value
        step.keep = keep
        return step

    def bind_shadow_lists(self, main_view, cap_entities: int, shadow_res=512, pinned=False):
        """ovis_build_shadow_lists with its arguments marshalled once: step() -> (n_lists, n_entities); the lists are in
        step.list_light / list_view / offsets / entities (referenced arrays, valid until the next step). pinned: the entity
        buffer is page-locked (the copy of the ids is then asynchronous)."""
        main_view = np.ascontiguousarray(main_view, dtype=VIEW_DTYPE).reshape(1)
        _, _, nl_slots = self.counts()
        cap_lists = 6 * max(nl_slots, 1)
        ll, lv = np.zeros(cap_lists, dtype=np.int32), np.zeros(cap_lists, dtype=np.int32)
        off = np.zeros(cap_lists + 1, dtype=np.int64)
        hold = PinnedBuffer(max(int(cap_entities), 1)) if pinned else None
        ent = hold.array if pinned else np.zeros(max(int(cap_entities), 1), dtype=np.uint32)
        nl, ne = C.c_int32(), C.c_int64()
        fn = self.lib.ovis_build_shadow_lists
        args = (self.h, _p(main_view), C.c_int32(shadow_res), C.c_int32(cap_lists), _p(ll), _p(lv), _p(off), C.byref(nl),
                C.c_int64(len(ent)), _p(ent), C.byref(ne))

        def step():
            rc = fn(*args)
            if rc != OVIS_OK:
                _check(rc)
            return nl.value, ne.value
        step.keep = (main_view, hold)
        step.list_light, step.list_view, step.offsets, step.entities = ll, lv, off, ent
        return step

    def cull_planes(self, planes, require=0, out=None):
        planes = _f32(planes, (24,))
        ne_slots, _, _ = self.counts()
        ents = out if out is not None else np.empty(max(ne_slots, 1), dtype=np.uint32)
        ne = C.c_int64()
        _check(self.lib.ovis_cull_planes(self.h, _p(planes), C.c_uint32(require), _p(ents), C.c_int64(len(ents)), C.byref(ne)))
        return ents[:ne.value]

    def cull_planes_device(self, planes, d_out_ptr: int, d_count_ptr: int, require=0):
        """Asynchronous; results stay on the device (raw device pointers, e.g. torch tensor .data_ptr())."""
        planes = _f32(planes, (24,))
        _check(self.lib.ovis_cull_planes_device(self.h, _p(planes), C.c_uint32(require), C.c_void_p(d_out_ptr),
                                                C.c_void_p(d_count_ptr)))

    def cull_planes_device_multi(self, planes, d_out_ptrs, d_count_ptrs, capacity: int, require=0):
        """Cull with the result stored to several device buffers (local + peers' segments). Asynchronous."""
        planes = _f32(planes, (24,))
        n = len(d_out_ptrs)
        outs = (C.c_void_p * n)(*d_out_ptrs)
        counts = (C.c_void_p * n)(*d_count_ptrs)
        _check(self.lib.ovis_cull_planes_device_multi(self.h, _p(planes), C.c_uint32(require), C.c_int64(capacity),
                                                      C.c_int32(n), outs, counts))

    def cull_planes_device_exchange(self, planes, d_out_ptrs, d_count_ptrs, capacity, wait_flags_ptr, n_wait, wait_value,
                                    signal_ptrs, signal_value, require=0):
        """Cull + fused exchange with one-way ready flags (see ovis_cull_planes_device_exchange). Asynchronous."""
        planes = _f32(planes, (24,))
        n = len(d_out_ptrs)
        outs = (C.c_void_p * n)(*d_out_ptrs)
        counts = (C.c_void_p * n)(*d_count_ptrs)
        sig = (C.c_void_p * len(signal_ptrs))(*signal_ptrs)
        _check(self.lib.ovis_cull_planes_device_exchange(
            self.h, _p(planes), C.c_uint32(require), C.c_int64(capacity), C.c_int32(n), outs, counts,
            C.c_int32(n_wait), C.c_void_p(wait_flags_ptr), C.c_uint32(wait_value & 0xffffffff),
            C.c_int32(len(signal_ptrs)), sig, C.c_uint32(signal_value & 0xffffffff)))

    @staticmethod
    def exchange_plan(d_out_ptrs, d_count_ptrs, signal_ptrs):
        """Pre-built ctypes argument arrays for cull_exchange_planned (keeps per-frame host work minimal)."""
        n = len(d_out_ptrs)
        return ((C.c_void_p * n)(*d_out_ptrs), (C.c_void_p * n)(*d_count_ptrs), C.c_int32(n),
                (C.c_void_p * len(signal_ptrs))(*signal_ptrs), C.c_int32(len(signal_ptrs)))

    def cull_exchange_planned(self, planes_f32, plan, capacity, wait_flags_ptr, n_wait, wait_value, signal_value, require=0):
        outs, counts, n, sig, n_sig = plan
        _check(self.lib.ovis_cull_planes_device_exchange(
            self.h, planes_f32.ctypes.data_as(C.c_void_p), C.c_uint32(require), C.c_int64(capacity), n, outs, counts,
            C.c_int32(n_wait), C.c_void_p(wait_flags_ptr), C.c_uint32(wait_value & 0xffffffff), n_sig, sig,
            C.c_uint32(signal_value & 0xffffffff)))

    def wait_flags(self, d_flags_ptr, n, value):
        _check(self.lib.ovis_wait_flags(self.h, C.c_void_p(d_flags_ptr), C.c_int32(n), C.c_uint32(value & 0xffffffff)))

    def read_segments_host(self, d_segments_ptr: int, segment_words: int, n_segments: int, host_out: np.ndarray,
                           d_flags_ptr: int = 0, value: int = 0):
        """Wait for the segments' ready flags and deliver [count | ids] blocks into host_out (synchronous)."""
        assert host_out.dtype == np.uint32 and host_out.size >= segment_words * n_segments
        _check(self.lib.ovis_read_segments_host(self.h, C.c_void_p(d_flags_ptr or None), C.c_uint32(value),
                                                C.c_void_p(d_segments_ptr), C.c_int64(segment_words), C.c_int32(n_segments),
                                                _p(host_out)))

    def exchange_error(self) -> int:
        e = C.c_int32()
        _check(self.lib.ovis_world_exchange_error(self.h, C.byref(e)))
        return e.value

    def cull_views_device(self, planes, require=None):
        """Many views, CSR result on the device. Returns (host offsets, d_offsets ptr, d_entities ptr)."""
        planes = _f32(planes, (-1, 24))
        nv = len(planes)
        require = _u32(require) if require is not None else None
        d_off, d_ent = C.c_void_p(), C.c_void_p()
        host = np.zeros(nv + 1, dtype=np.int64)
        _check(self.lib.ovis_cull_views_device(self.h, C.c_int32(nv), _p(planes), _p(require), C.byref(d_off),
                                               C.byref(d_ent), _p(host)))
        return host, d_off.value, d_ent.value

    def bind_cull_views_device(self, planes, require=None):
        """ovis_cull_views_device with its arguments marshalled once: step() -> None; step.offsets (host, n_views + 1),
        step.d_offsets / step.d_entities after the first step."""
        planes = _f32(planes, (-1, 24))
        nv = len(planes)
        require = _u32(require) if require is not None else None
        d_off, d_ent = C.c_void_p(), C.c_void_p()
        host = np.zeros(nv + 1, dtype=np.int64)
        fn = self.lib.ovis_cull_views_device
        args = (self.h, C.c_int32(nv), _p(planes), _p(require), C.byref(d_off), C.byref(d_ent), _p(host))

        def step():
            rc = fn(*args)
            if rc != OVIS_OK:
                _check(rc)
        step.keep = (planes, require)
        step.offsets, step.d_offsets, step.d_entities = host, d_off, d_ent
        return step

    def bind_cull_shadow_views_device(self, light_ids, shadow_res=512):
        """ovis_cull_shadow_views_device with its arguments marshalled once: step() -> number of lists; step.list_light /
        list_view / offsets are referenced host arrays, step.d_offsets / step.d_entities the device CSR."""
        light_ids = _u32(light_ids)
        cap = 6 * max(len(light_ids), 1)
        ll, lv = np.zeros(cap, dtype=np.int32), np.zeros(cap, dtype=np.int32)
        off = np.zeros(cap + 1, dtype=np.int64)
        n = C.c_int32()
        d_off, d_ent = C.c_void_p(), C.c_void_p()
        fn = self.lib.ovis_cull_shadow_views_device
        args = (self.h, C.c_int32(len(light_ids)), _p(light_ids), C.c_int32(shadow_res), C.c_int32(cap), _p(ll), _p(lv), C.byref(n),
                C.byref(d_off), C.byref(d_ent), _p(off))

        def step():
            rc = fn(*args)
            if rc != OVIS_OK:
                _check(rc)
            return n.value
        step.keep = (light_ids,)
        step.list_light, step.list_view, step.offsets, step.d_offsets, step.d_entities = ll, lv, off, d_off, d_ent
        return step

    def cull_shadow_views_device(self, light_ids, shadow_res=512):
        """Shadow-view culls of the given lights -> (list_light, list_view, host offsets, d_offsets, d_entities)."""
        light_ids = _u32(light_ids)
        cap = 6 * max(len(light_ids), 1)
        ll = np.zeros(cap, dtype=np.int32)
        lv = np.zeros(cap, dtype=np.int32)
        off = np.zeros(cap + 1, dtype=np.int64)
        n = C.c_int32()
        d_off, d_ent = C.c_void_p(), C.c_void_p()
        _check(self.lib.ovis_cull_shadow_views_device(self.h, C.c_int32(len(light_ids)), _p(light_ids), C.c_int32(shadow_res),
                                                      C.c_int32(cap), _p(ll), _p(lv), C.byref(n), C.byref(d_off),
                                                      C.byref(d_ent), _p(off)))
        k = n.value
        return ll[:k].copy(), lv[:k].copy(), off[:k + 1].copy(), d_off.value, d_ent.value

    def shadow_lists(self, main_view, shadow_res=512):
        """prepareLights' per-shadow-view caster lists -> (list_light, list_view, offsets, entities)."""
        main_view = np.ascontiguousarray(main_view, dtype=VIEW_DTYPE)
        ne_slots, _, nl_slots = self.counts()
        cap_lists = 6 * max(nl_slots, 1)
        cap_ent = max(ne_slots, 1) * 2
        while True:
            ll = np.zeros(cap_lists, dtype=np.int32)
            lv = np.zeros(cap_lists, dtype=np.int32)
            off = np.zeros(cap_lists + 1, dtype=np.int64)
            ent = np.zeros(cap_ent, dtype=np.uint32)
            nl, ne = C.c_int32(), C.c_int64()
            rc = self.lib.ovis_build_shadow_lists(self.h, _p(main_view), C.c_int32(shadow_res), C.c_int32(cap_lists),
                                                  _p(ll), _p(lv), _p(off), C.byref(nl), C.c_int64(cap_ent), _p(ent),
                                                  C.byref(ne))
            if rc == OVIS_ERR_CAPACITY:
                cap_lists, cap_ent = max(cap_lists, nl.value), max(cap_ent, ne.value)
                continue
            _check(rc)
            k = nl.value
            return ll[:k].copy(), lv[:k].copy(), off[:k + 1].copy(), ent[:ne.value].copy()

    # ---- consumers of the visible lists
    def set_pass_counts(self, ids, deferred, basic, shadow_caster):
        """Shader-pass counts of the entities' materials (numPasses per pass type, each 0..255)."""
        ids = _u32(ids)
        packed = (_u32(deferred) & 0xff) | ((_u32(basic) & 0xff) << 8) | ((_u32(shadow_caster) & 0xff) << 16)
        packed = np.ascontiguousarray(np.broadcast_to(packed, ids.shape), dtype=np.uint32)
        _check(self.lib.ovis_entity_set_pass_counts(self.h, C.c_int64(len(ids)), _p(ids), _p(packed)))

    def build_draw_lists_device(self, mode, d_draws0: int, cap0: int, d_draws1: int = 0, cap1: int = 0, d_visible: int = 0,
                                d_visible_count: int = 0, max_visible: int = -1, d_draw_counts: int = 0):
        """Asynchronous. Pointers are raw device addresses; d_visible = 0 uses the last cull's entity list."""
        _check(self.lib.ovis_build_draw_lists_device(self.h, C.c_void_p(d_visible or None), C.c_void_p(d_visible_count or None),
                                                     C.c_int64(max_visible), C.c_int32(mode), C.c_void_p(d_draws0 or None),
                                                     C.c_int64(cap0), C.c_void_p(d_draws1 or None), C.c_int64(cap1),
                                                     C.c_void_p(d_draw_counts or None)))

    def draw_counts(self):
        out = np.zeros(2, dtype=np.uint32)
        _check(self.lib.ovis_get_draw_counts(self.h, _p(out)))
        return int(out[0]), int(out[1])

    def pack_entity_uniforms_device(self, d_uniforms: int, capacity: int, d_visible: int = 0, d_visible_count: int = 0,
                                    max_visible: int = -1):
        """Asynchronous. 20 floats per visible entity into d_uniforms (16-byte aligned device address)."""
        _check(self.lib.ovis_pack_entity_uniforms_device(self.h, C.c_void_p(d_visible or None),
                                                         C.c_void_p(d_visible_count or None), C.c_int64(max_visible),
                                                         C.c_void_p(d_uniforms or None), C.c_int64(capacity)))

    def import_scene(self, world_path: str, game_asset_dir: str, vp_width: int, vp_height: int, max_cameras: int = 16):
        """ovis_scene_import -> (SceneInfo, cameras as a VIEW_DTYPE array)."""
        info = SceneInfo()
        cams = np.zeros(max_cameras, dtype=VIEW_DTYPE)
        _check(self.lib.ovis_scene_import(self.h, os.fsencode(world_path), os.fsencode(game_asset_dir), C.c_int32(vp_width),
                                          C.c_int32(vp_height), C.c_int32(max_cameras), _p(cams), C.byref(info)))
        return info, cams[:info.n_cameras].copy()

    def light(self, light_id: int):
        """(OvisLight record, attenuation[3], shadow bias) of a light."""
        l = np.zeros(1, dtype=LIGHT_DTYPE)
        att = np.zeros(3, dtype=np.float32)
        bias = C.c_float()
        _check(self.lib.ovis_light_get(self.h, C.c_uint32(light_id), _p(l), _p(att), C.byref(bias)))
        return l[0], att, np.float32(bias.value)

    def set_light_shading(self, light_id, attenuation, shadow_bias):
        att = np.asarray(attenuation, dtype=np.float32)
        _check(self.lib.ovis_light_set_shading(self.h, C.c_uint32(light_id), _p(att), C.c_float(shadow_bias)))

    def pack_light_uniforms_device(self, d_uniforms: int, capacity: int, d_visible: int = 0, d_visible_count: int = 0,
                                   max_visible: int = -1):
        """Asynchronous. 208 bytes per visible light into d_uniforms; d_visible = 0 uses the last cull's light list."""
        _check(self.lib.ovis_pack_light_uniforms_device(self.h, C.c_void_p(d_visible or None),
                                                        C.c_void_p(d_visible_count or None), C.c_int64(max_visible),
                                                        C.c_void_p(d_uniforms or None), C.c_int64(capacity)))

    def pack_view_uniforms_device(self, views, d_uniforms: int, viewport_pos=None):
        """Asynchronous. 288 bytes per view into d_uniforms."""
        views = np.ascontiguousarray(views, dtype=VIEW_DTYPE).reshape(-1)
        pos = np.ascontiguousarray(viewport_pos, dtype=np.int32).reshape(-1) if viewport_pos is not None else None
        assert pos is None or len(pos) == 2 * len(views)
        _check(self.lib.ovis_pack_view_uniforms_device(self.h, C.c_int32(len(views)), _p(views), _p(pos), C.c_void_p(d_uniforms)))

    # ---- deferred shading over headless buffers
    def shade_deferred_device(self, width, height, d_buffer_a, d_buffer_b, d_buffer_c, d_depth, d_view_uniforms, d_out_rgba,
                              d_visible=0, d_visible_count=0, max_visible=-1, d_shadow_maps=0, shadow_size=0):
        """Asynchronous. All pointers are raw device addresses; d_visible = 0 uses the last cull's light list."""
        g = GBuffer(width, height, d_buffer_a, d_buffer_b, d_buffer_c, d_depth)
        _check(self.lib.ovis_shade_deferred_device(self.h, C.byref(g), C.c_void_p(d_view_uniforms), C.c_void_p(d_visible or None),
                                                   C.c_void_p(d_visible_count or None), C.c_int64(max_visible),
                                                   C.c_void_p(d_shadow_maps or None), C.c_int32(shadow_size),
                                                   C.c_void_p(d_out_rgba)))

    def post_effects_device(self, n_pixels, d_rgba, d_out_rgba, effects, exposure=1.0, white_point=11.2, gamma=2.2):
        _check(self.lib.ovis_post_effects_device(self.h, C.c_int64(n_pixels), C.c_void_p(d_rgba), C.c_uint32(effects),
                                                 C.c_float(exposure), C.c_float(white_point), C.c_float(gamma),
                                                 C.c_void_p(d_out_rgba)))

    # ---- read-back
    def read_device(self, d_ptr: int, count: int, dtype=np.uint32) -> np.ndarray:
        out = np.empty(max(count, 1), dtype=dtype)
        _check(self.lib.ovis_read_device(self.h, C.c_void_p(d_ptr), C.c_int64(count * out.itemsize), _p(out)))
        return out[:count]

    def world_aabbs(self, first=0, n=None) -> np.ndarray:
        if n is None:
            n = self.counts()[0] - first
        out = np.zeros((n, 6), dtype=np.float32)
        _check(self.lib.ovis_get_world_aabbs(self.h, C.c_uint32(first), C.c_int64(n), _p(out)))
        return out

    def node_world_trs(self, first=0, n=None) -> np.ndarray:
        if n is None:
            n = self.counts()[1] - first
        out = np.zeros((n, 10), dtype=np.float32)
        _check(self.lib.ovis_get_node_world_trs(self.h, C.c_uint32(first), C.c_int64(n), _p(out)))
        return out


class ShardedWorld:
    """ovis_sharded_*: one world over several GPUs driven by one host process (worker thread per device inside the library)."""

    def __init__(self, device_ids):
        self.lib = load()
        ids = np.ascontiguousarray(device_ids, dtype=np.int32)
        self.h = C.c_void_p()
        self.n = len(ids)
        _check(self.lib.ovis_sharded_create(_p(ids), C.c_int32(len(ids)), C.byref(self.h)))

    def close(self):
        if self.h:
            self.lib.ovis_sharded_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def add_entities(self, world_aabb, flags=None):
        aabb = _f32(world_aabb, (-1, 6))
        fl = _u32(flags) if flags is not None else None
        _check(self.lib.ovis_sharded_add_entities(self.h, C.c_int64(len(aabb)), _p(aabb), _p(fl)))

    def shard(self, r: int):
        """(raw OvisWorld handle of shard r, its id base) - for the ordinary entry points."""
        w, base = C.c_void_p(), C.c_int64()
        _check(self.lib.ovis_sharded_shard(self.h, C.c_int32(r), C.byref(w), C.byref(base)))
        return w, base.value

    def remove_entities(self, global_ids):
        """Convenience for tests: route removals to the owning shards (ovis_entity_remove takes shard-local slot ids)."""
        ids = np.sort(_u32(global_ids))
        for r in range(self.n):
            w, base = self.shard(r)
            _, next_base = self.shard(r + 1) if r + 1 < self.n else (None, 1 << 62)
            local = (ids[(ids >= base) & (ids < next_base)] - np.uint32(base)).astype(np.uint32)
            if len(local):
                _check(self.lib.ovis_entity_remove(w, C.c_int64(len(local)), _p(local)))

    def bind_cull(self, view, out_entities: np.ndarray, require=0):
        """Frame function: step() -> (n_entities, per-shard counts); ids land in out_entities[:n]."""
        view = np.ascontiguousarray(view, dtype=VIEW_DTYPE).reshape(1)
        ne = C.c_int64()
        counts = np.zeros(self.n, dtype=np.int64)
        fn = self.lib.ovis_sharded_cull
        args = (self.h, _p(view), C.c_uint32(require), _p(out_entities), C.c_int64(len(out_entities)), C.byref(ne), _p(counts))

        def step():
            rc = fn(*args)
            if rc != OVIS_OK:
                _check(rc)
            return ne.value, counts
        step.keep = (view, out_entities, counts)
        return step

    def cull(self, view, require=0, capacity=None):
        ns, total = C.c_int32(), C.c_int64()
        _check(self.lib.ovis_sharded_count(self.h, C.byref(ns), C.byref(total)))
        out = np.empty(max(int(capacity if capacity is not None else total.value), 1), dtype=np.uint32)
        n, counts = self.bind_cull(view, out, require)()
        return out[:n].copy(), counts.copy()

    def add_lights(self, lights) -> int:
        lights = np.ascontiguousarray(lights, dtype=LIGHT_DTYPE).reshape(-1)
        first = C.c_uint32()
        _check(self.lib.ovis_sharded_add_lights(self.h, C.c_int32(len(lights)), _p(lights), C.byref(first)))
        return first.value

    def cull_views(self, planes, require=None, capacity=1 << 16):
        """ovis_sharded_cull_views -> (offsets, entities): list v = entities[offsets[v]:offsets[v + 1]], global ids."""
        planes = _f32(planes, (-1, 24))
        nv = len(planes)
        req = _u32(require) if require is not None else None
        off = np.zeros(nv + 1, dtype=np.int64)
        while True:
            ent = np.empty(max(int(capacity), 1), dtype=np.uint32)
            ne = C.c_int64()
            rc = self.lib.ovis_sharded_cull_views(self.h, C.c_int32(nv), _p(planes), _p(req), _p(off), C.c_int64(len(ent)), _p(ent),
                                                  C.byref(ne))
            if rc == OVIS_ERR_CAPACITY and ne.value > len(ent):
                capacity = ne.value
                continue
            _check(rc)
            return off, ent[:ne.value]

    def shadow_lists(self, main_view, n_lights: int, shadow_res=512, capacity=1 << 16):
        """ovis_sharded_build_shadow_lists -> (list_light, list_view, offsets, entities)."""
        main_view = np.ascontiguousarray(main_view, dtype=VIEW_DTYPE).reshape(1)
        cap_lists = 6 * max(int(n_lights), 1)
        ll = np.zeros(cap_lists, dtype=np.int32)
        lv = np.zeros(cap_lists, dtype=np.int32)
        off = np.zeros(cap_lists + 1, dtype=np.int64)
        while True:
            ent = np.empty(max(int(capacity), 1), dtype=np.uint32)
            nl, ne = C.c_int32(), C.c_int64()
            rc = self.lib.ovis_sharded_build_shadow_lists(self.h, _p(main_view), C.c_int32(shadow_res), C.c_int32(cap_lists), _p(ll), _p(lv),
                                                          _p(off), C.byref(nl), C.c_int64(len(ent)), _p(ent), C.byref(ne))
            if rc == OVIS_ERR_CAPACITY and ne.value > len(ent):
                capacity = ne.value
                continue
            _check(rc)
            k = nl.value
            return ll[:k].copy(), lv[:k].copy(), off[:k + 1].copy(), ent[:ne.value]

    def allgather_device(self, planes, capacity: int, require=0):
        """NCCL mode: returns (per-device block pointers, per-shard counts)."""
        planes = _f32(planes, (24,))
        blocks = C.POINTER(C.c_void_p)()
        counts = np.zeros(self.n, dtype=np.int64)
        _check(self.lib.ovis_sharded_allgather_device(self.h, _p(planes), C.c_uint32(require), C.c_int64(capacity),
                                                      C.byref(blocks), _p(counts)))
        return [blocks[d] for d in range(self.n)], counts


def _world_extra_methods():
    def compact(self):
        """ovis_world_compact -> remap (old id -> new id, 0xffffffff = removed)."""
        live, slots = C.c_int64(), C.c_int64()
        _check(self.lib.ovis_world_live_counts(self.h, C.byref(live), C.byref(slots)))
        remap = np.empty(max(slots.value, 1), dtype=np.uint32)
        old, new = C.c_int64(), C.c_int64()
        _check(self.lib.ovis_world_compact(self.h, C.c_int64(len(remap)), _p(remap), C.byref(old), C.byref(new)))
        return remap[:old.value], new.value

    def live_counts(self):
        live, slots = C.c_int64(), C.c_int64()
        _check(self.lib.ovis_world_live_counts(self.h, C.byref(live), C.byref(slots)))
        return live.value, slots.value

    def set_node_parent(self, node: int, parent: int):
        _check(self.lib.ovis_node_set_parent(self.h, C.c_uint32(node), C.c_int32(parent)))

    def remove_node(self, node: int):
        a, b = C.c_int64(), C.c_int64()
        _check(self.lib.ovis_node_remove(self.h, C.c_uint32(node), C.byref(a), C.byref(b)))
        return a.value, b.value

    def bind_light_to_node(self, light: int, node: int):
        _check(self.lib.ovis_light_bind_node(self.h, C.c_uint32(light), C.c_int32(node)))

    for f in (compact, live_counts, set_node_parent, remove_node, bind_light_to_node):
        setattr(World, f.__name__, f)


_world_extra_methods()


def _world_shadow_views_host():
    def cull_shadow_views(self, light_ids, require=CASTS_SHADOW, shadow_res=512):
        """ovis_cull_shadow_views: (list_light, list_view, offsets, entities) on the host; require = 0 keeps every entity."""
        light_ids = _u32(light_ids)
        cap_lists = 6 * max(len(light_ids), 1)
        cap_ent = 1 << 16
        while True:
            ll = np.zeros(cap_lists, dtype=np.int32)
            lv = np.zeros(cap_lists, dtype=np.int32)
            off = np.zeros(cap_lists + 1, dtype=np.int64)
            ent = np.zeros(cap_ent, dtype=np.uint32)
            nl, ne = C.c_int32(), C.c_int64()
            rc = self.lib.ovis_cull_shadow_views(self.h, C.c_int32(len(light_ids)), _p(light_ids), C.c_int32(shadow_res),
                                                 C.c_uint32(require), C.c_int32(cap_lists), _p(ll), _p(lv), _p(off), C.byref(nl),
                                                 C.c_int64(cap_ent), _p(ent), C.byref(ne))
            if rc == OVIS_ERR_CAPACITY:
                cap_lists, cap_ent = max(cap_lists, nl.value), max(cap_ent, ne.value)
                continue
            _check(rc)
            return ll[:nl.value], lv[:nl.value], off[:nl.value + 1], ent[:ne.value]

    World.cull_shadow_views = cull_shadow_views


_world_shadow_views_host()


MATERIAL_DTYPE = np.dtype([("diffuse", np.float32, 3), ("specular", np.float32, 3), ("shininess", np.float32), ("flags", np.uint32),
                           ("diffuse_texture", np.uint32), ("normal_texture", np.uint32)])
MATERIAL_SPECULAR, MATERIAL_TEXTURED, MATERIAL_NORMALMAP = 1, 2, 4


class _GBufferTarget(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("a", C.c_void_p), ("b", C.c_void_p), ("c", C.c_void_p),
                ("depth", C.c_void_p)]


class Raster:
    """ovis_raster_*: the headless rasteriser of one world (G-buffer pass and shadow-map passes as CUDA kernels)."""

    def __init__(self, world: "World"):
        self.lib = world.lib
        self.world = world
        self.h = C.c_void_p()
        _check(self.lib.ovis_raster_create(world.h, C.byref(self.h)))

    def close(self):
        if self.h:
            self.lib.ovis_raster_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def add_mesh(self, positions, normals, indices, texcoords=None) -> int:
        positions, normals = _f32(positions, (-1, 3)), _f32(normals, (-1, 3))
        indices = _u32(indices)
        tc = _f32(texcoords, (-1, 2)) if texcoords is not None else None
        out = C.c_uint32()
        _check(self.lib.ovis_raster_add_mesh(self.h, C.c_int64(len(positions)), _p(positions), _p(normals), _p(tc),
                                             C.c_int64(len(indices)), _p(indices), C.byref(out)))
        return out.value

    def set_mesh_tangents(self, mesh: int, tangents):
        t = _f32(tangents, (-1, 4))
        _check(self.lib.ovis_raster_set_mesh_tangents(self.h, C.c_uint32(mesh), C.c_int64(len(t)), _p(t)))

    def add_texture(self, rgba, srgb=True, wrap=True) -> int:
        """rgba: uint8 [height, width, 4], rows bottom-up (as tga_load returns them)."""
        rgba = np.ascontiguousarray(rgba, dtype=np.uint8)
        assert rgba.ndim == 3 and rgba.shape[2] == 4
        out = C.c_uint32()
        _check(self.lib.ovis_raster_add_texture(self.h, C.c_int32(rgba.shape[1]), C.c_int32(rgba.shape[0]), _p(rgba), C.c_int32(int(srgb)),
                                                C.c_int32(int(wrap)), C.byref(out)))
        return out.value

    def add_material(self, diffuse, specular=(0.0, 0.0, 0.0), shininess=0.0, flags=0, diffuse_texture=0, normal_texture=0) -> int:
        m = np.zeros(1, dtype=MATERIAL_DTYPE)
        m["diffuse"], m["specular"], m["shininess"], m["flags"] = diffuse, specular, shininess, flags
        m["diffuse_texture"], m["normal_texture"] = diffuse_texture, normal_texture
        out = C.c_uint32()
        _check(self.lib.ovis_raster_add_material(self.h, _p(m), C.byref(out)))
        return out.value

    def bind_entities(self, entity_ids, mesh_ids, material_ids):
        e, me, ma = _u32(entity_ids), _u32(mesh_ids), _u32(material_ids)
        me = np.ascontiguousarray(np.broadcast_to(me, e.shape), dtype=np.uint32)
        ma = np.ascontiguousarray(np.broadcast_to(ma, e.shape), dtype=np.uint32)
        _check(self.lib.ovis_raster_bind_entities(self.h, C.c_int64(len(e)), _p(e), _p(me), _p(ma)))

    def gbuffer(self, view, d_visible: int, d_count: int, width: int, height: int, d_a: int, d_b: int, d_c: int, d_depth: int,
                max_visible=-1):
        """The G-buffer pass for a device list (raw device pointers, e.g. torch tensors' data_ptr())."""
        view = np.ascontiguousarray(view, dtype=VIEW_DTYPE).reshape(1)
        t = _GBufferTarget(width, height, d_a, d_b, d_c, d_depth)
        _check(self.lib.ovis_raster_gbuffer_device(self.h, _p(view), C.c_void_p(d_visible), C.c_void_p(d_count), C.c_int64(max_visible),
                                                   C.byref(t)))

    def gbuffer_draws(self, view, d_draws: int, d_count: int, width: int, height: int, d_a: int, d_b: int, d_c: int, d_depth: int,
                      max_draws=-1):
        """The G-buffer pass fed by a draw list ((entity, pass) pairs of ovis_build_draw_lists_device)."""
        view = np.ascontiguousarray(view, dtype=VIEW_DTYPE).reshape(1)
        t = _GBufferTarget(width, height, d_a, d_b, d_c, d_depth)
        _check(self.lib.ovis_raster_gbuffer_draws_device(self.h, _p(view), C.c_void_p(d_draws), C.c_void_p(d_count), C.c_int64(max_draws),
                                                         C.byref(t)))

    def basic(self, view, d_ids: int, d_count: int, width: int, height: int, d_rgba: int, d_depth: int, max_count=-1, draws=False):
        """The Basic (unlit) pass into the colour target and depth buffer as they stand; draws=True: d_ids is a draw list."""
        view = np.ascontiguousarray(view, dtype=VIEW_DTYPE).reshape(1)
        fn = self.lib.ovis_raster_basic_draws_device if draws else self.lib.ovis_raster_basic_device
        _check(fn(self.h, _p(view), C.c_void_p(d_ids), C.c_void_p(d_count), C.c_int64(max_count), C.c_int32(width), C.c_int32(height),
                  C.c_void_p(d_rgba), C.c_void_p(d_depth)))

    def shadow_maps(self, views, host_offsets, d_entities: int, shadow_size: int, d_maps):
        views = np.ascontiguousarray(views, dtype=VIEW_DTYPE).reshape(-1)
        off = np.ascontiguousarray(host_offsets, dtype=np.int64)
        maps = (C.c_void_p * len(views))(*[int(p) for p in d_maps])
        _check(self.lib.ovis_raster_shadow_maps_device(self.h, C.c_int32(len(views)), _p(views), _p(off), C.c_void_p(d_entities),
                                                       C.c_int32(shadow_size), maps))


def obj_mesh(path: str):
    """ovis_obj_mesh: (positions [n, 3], normals [n, 3], texcoords [n, 2], indices, first_index per sub-mesh + end)."""
    L = load()
    nv, ni, ns = C.c_int64(), C.c_int64(), C.c_int32()
    rc = L.ovis_obj_mesh(path.encode(), C.c_int64(0), None, None, None, C.c_int64(0), None, C.c_int32(0), None, C.byref(nv), C.byref(ni),
                         C.byref(ns))
    if rc not in (OVIS_OK, OVIS_ERR_CAPACITY):
        _check(rc)
    pos = np.zeros((nv.value, 3), dtype=np.float32)
    nrm = np.zeros((nv.value, 3), dtype=np.float32)
    tc = np.zeros((nv.value, 2), dtype=np.float32)
    idx = np.zeros(ni.value, dtype=np.uint32)
    first = np.zeros(ns.value + 1, dtype=np.int64)
    _check(L.ovis_obj_mesh(path.encode(), C.c_int64(nv.value), _p(pos), _p(nrm), _p(tc), C.c_int64(ni.value), _p(idx), C.c_int32(ns.value),
                           _p(first), C.byref(nv), C.byref(ni), C.byref(ns)))
    return pos, nrm, tc, idx, first


def mesh_tangents(positions, normals, texcoords, indices) -> np.ndarray:
    """ovis_mesh_tangents: [n, 4] tangents (xyz + handedness) as the reference's mesh loader generates them."""
    pos, nrm, tc, idx = _f32(positions, (-1, 3)), _f32(normals, (-1, 3)), _f32(texcoords, (-1, 2)), _u32(indices)
    out = np.zeros((len(pos), 4), dtype=np.float32)
    _check(load().ovis_mesh_tangents(C.c_int64(len(pos)), _p(pos), _p(nrm), _p(tc), C.c_int64(len(idx)), _p(idx), _p(out)))
    return out


def tga_load(path: str) -> np.ndarray:
    """ovis_tga_load: [height, width, 4] uint8 R, G, B, A, rows in file order."""
    L = load()
    w, h = C.c_int32(), C.c_int32()
    rc = L.ovis_tga_load(os.fsencode(path), C.c_int64(0), None, C.byref(w), C.byref(h))
    if rc not in (OVIS_OK, OVIS_ERR_CAPACITY):
        _check(rc)
    out = np.zeros((h.value, w.value, 4), dtype=np.uint8)
    _check(L.ovis_tga_load(os.fsencode(path), C.c_int64(out.nbytes), _p(out), C.byref(w), C.byref(h)))
    return out
