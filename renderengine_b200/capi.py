"""ctypes mirror of include/reb.h.  Loads the in-tree libreb.so and fails loudly if it is missing —
there is no Python or CPU fallback for any render call."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("REB_LIB") or os.path.join(HERE, "libreb.so")   # REB_LIB: A/B-testing of kernel build variants

REB_OUT_COLOR_SRGB8, REB_OUT_COLOR_F32, REB_OUT_DEPTH_F32, REB_OUT_TRIID_U32 = 1, 2, 4, 8
REB_COLOR_SRGB8, REB_COLOR_F32, REB_DEPTH_F32, REB_TRIID_U32, REB_SHADOW_F32 = 0, 1, 2, 3, 4
REB_MS_DEPTH_F32, REB_MS_TRIID_U32 = 5, 6

SYMBOLS = [
    "reb_abi_version", "reb_create", "reb_destroy", "reb_last_error", "reb_upload_geometry", "reb_upload_transforms",
    "reb_upload_materials", "reb_upload_lights", "reb_register_texture", "reb_shadow_create", "reb_shadow_set_transform",
    "reb_set_view", "reb_render_shadow", "reb_render_forward", "reb_get_output", "reb_readback", "reb_sync", "reb_flush", "reb_readback_async", "reb_set_shading_model", "reb_set_skybox", "reb_set_sample_count", "reb_set_shadow_partition", "reb_render_forward2",
    "reb_set_partition", "reb_shadow_mark_valid", "reb_shadow_wait_event", "reb_get_counters", "reb_get_launch_count", "reb_set_profiling",
    "reb_get_kernel_times", "reb_shadow_depth_static", "reb_set_shadow_split",
]
REB_STATIC_MAX_VERTICES, REB_STATIC_MAX_TRIANGLES = 262144, 524288


class RebParams(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("device", C.c_int32), ("width", C.c_uint32), ("height", C.c_uint32),
                ("max_descriptors", C.c_uint32), ("outputs", C.c_uint32), ("rank", C.c_uint32), ("n_ranks", C.c_uint32),
                ("band_h", C.c_uint32), ("reserved", C.c_uint32), ("stream", C.c_void_p)]


class RebTextureDesc(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("format", C.c_uint32), ("srgb", C.c_uint32)]


class RebError(RuntimeError):
    pass


_lib = None


def load():
    """dlopen libreb.so (building nothing here: run `python -m renderengine_b200.build` or __graft_entry__.build())."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RebError(f"{LIB_PATH} is missing: build it with `python -m renderengine_b200.build` (no CPU fallback exists)")
    lib = C.CDLL(LIB_PATH)
    vp, u32, u64 = C.c_void_p, C.c_uint32, C.c_uint64
    lib.reb_abi_version.restype = C.c_int
    lib.reb_last_error.restype = C.c_char_p
    lib.reb_last_error.argtypes = [vp]
    lib.reb_create.argtypes = [C.POINTER(RebParams), C.POINTER(vp)]
    lib.reb_destroy.argtypes = [vp]
    lib.reb_upload_geometry.argtypes = [vp, vp, u64, vp, u64]
    lib.reb_upload_transforms.argtypes = [vp, vp, u32]
    lib.reb_upload_materials.argtypes = [vp, vp, u32, u32]
    lib.reb_upload_lights.argtypes = [vp, vp, u32, u32]
    lib.reb_register_texture.argtypes = [vp, C.POINTER(RebTextureDesc), vp, C.POINTER(u32)]
    lib.reb_shadow_create.argtypes = [vp, u32, C.POINTER(u32)]
    lib.reb_shadow_set_transform.argtypes = [vp, u32, vp, vp]
    lib.reb_set_view.argtypes = [vp, vp]
    lib.reb_render_shadow.argtypes = [vp, u32, vp, u32]
    lib.reb_render_forward.argtypes = [vp, vp, u32]
    lib.reb_render_forward2.argtypes = [vp, vp, u32, vp, u32]
    lib.reb_get_output.argtypes = [vp, C.c_int, C.POINTER(vp), C.POINTER(u64)]
    lib.reb_readback.argtypes = [vp, C.c_int, vp]
    lib.reb_sync.argtypes = [vp]
    lib.reb_flush.argtypes = [vp]
    lib.reb_set_shading_model.argtypes = [vp, u32]
    lib.reb_set_skybox.argtypes = [vp, vp, u32, u32]
    lib.reb_set_sample_count.argtypes = [vp, u32]
    lib.reb_set_shadow_partition.argtypes = [vp, u32, u32, u32]
    lib.reb_set_shadow_split.argtypes = [vp, u32, u32]
    lib.reb_readback_async.argtypes = [vp, C.c_int, vp]
    lib.reb_set_partition.argtypes = [vp, u32, u32, u32]
    lib.reb_shadow_mark_valid.argtypes = [vp, u32]
    lib.reb_shadow_wait_event.argtypes = [vp, u32, vp]
    lib.reb_get_counters.argtypes = [vp, vp]
    lib.reb_get_launch_count.argtypes = [vp, C.POINTER(u64)]
    lib.reb_set_profiling.argtypes = [vp, C.c_int]
    lib.reb_get_kernel_times.argtypes = [vp, C.c_int, vp]
    lib.reb_shadow_depth_static.argtypes = [vp, u32, u32, vp, u32, vp, vp, vp, u32, vp, vp, C.POINTER(C.c_char_p)]
    _lib = lib
    return lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def shadow_depth_static(d_vertices, vertex_stride, n_vertices, d_indices, n_indices, model16, view16, proj16, size, d_depth_out, stream=None):
    """reb_shadow_depth_static (include/reb.h): the allocation-free shadow depth pass of one draw.  d_* are DEVICE addresses (ints)
    the caller owns; the matrices are 16 column-major floats each.  Asynchronous on `stream` (a cudaStream_t as int, None = default)."""
    lib = load()
    m = [np.ascontiguousarray(np.asarray(x, dtype=np.float32).reshape(16)) for x in (model16, view16, proj16)]
    err = C.c_char_p()
    rc = lib.reb_shadow_depth_static(C.c_void_p(d_vertices), vertex_stride, n_vertices, C.c_void_p(d_indices), n_indices,
                                     _ptr(m[0]), _ptr(m[1]), _ptr(m[2]), size, C.c_void_p(d_depth_out), C.c_void_p(stream), C.byref(err))
    if rc != 0:
        e = RebError(f"libreb error {rc}: {(err.value or b'').decode()}")
        e.code = rc
        raise e


_KIND_INFO = {REB_COLOR_SRGB8: (np.uint8, 4), REB_COLOR_F32: (np.float32, 4), REB_DEPTH_F32: (np.float32, 1),
              REB_TRIID_U32: (np.uint32, 1)}


class Context:
    """One raster context per GPU (the reference's Renderer + RenderScene systems, headless)."""

    def __init__(self, width, height, device=0, outputs=0, rank=0, n_ranks=1, band_h=64, stream=None, max_descriptors=0):
        self.lib = load()
        self.width, self.height = int(width), int(height)
        p = RebParams(C.sizeof(RebParams), device, width, height, max_descriptors, outputs, rank, n_ranks, band_h, 0, stream)
        h = C.c_void_p()
        rc = self.lib.reb_create(C.byref(p), C.byref(h))
        if rc != 0:
            raise RebError(f"reb_create failed ({rc}): {self.lib.reb_last_error(None).decode()}")
        self.h = h
        self.shadow_size = 0
        self.shadow_id = None

    def _ck(self, rc):
        if rc != 0:
            raise RebError(f"libreb error {rc}: {self.lib.reb_last_error(self.h).decode()}")

    def close(self):
        if getattr(self, "h", None):
            self.lib.reb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # --- uploads (RenderScene::UploadToGPU, RenderScene.cpp:80-92) ----------------------------
    def upload_geometry(self, verts, indices):
        v = np.ascontiguousarray(verts, np.float32); i = np.ascontiguousarray(indices, np.uint32)
        assert v.ndim == 2 and v.shape[1] == 8
        self._ck(self.lib.reb_upload_geometry(self.h, _ptr(v), v.shape[0], _ptr(i), i.size))

    def upload_transforms(self, transforms):
        t = np.ascontiguousarray(transforms, np.float32).reshape(-1, 16)
        self._ck(self.lib.reb_upload_transforms(self.h, _ptr(t), t.shape[0]))

    def upload_materials(self, records):
        m = np.ascontiguousarray(records, np.uint8)
        self._ck(self.lib.reb_upload_materials(self.h, _ptr(m), m.shape[1], m.shape[0]))

    def upload_lights(self, records):
        l = np.ascontiguousarray(records, np.uint8)
        self._ck(self.lib.reb_upload_lights(self.h, _ptr(l), l.shape[1], l.shape[0]))

    def shadow_create(self, size):
        sid = C.c_uint32()
        self._ck(self.lib.reb_shadow_create(self.h, size, C.byref(sid)))
        self.shadow_size, self.shadow_id = size, sid.value
        return sid.value

    def shadow_set_transform(self, view16, proj16):
        v = np.ascontiguousarray(view16, np.float32); p = np.ascontiguousarray(proj16, np.float32)
        self._ck(self.lib.reb_shadow_set_transform(self.h, self.shadow_id, _ptr(v), _ptr(p)))

    def set_view(self, view160):
        v = np.ascontiguousarray(view160, np.uint8)
        assert v.size == 160
        self._ck(self.lib.reb_set_view(self.h, _ptr(v)))

    def register_texture(self, texels, srgb=True):
        """texels: [H, W, 4] uint16 (R16G16B16A16_UNORM mip 0). Returns the TextureHandle (running index)."""
        t = np.ascontiguousarray(texels, np.uint16)
        assert t.ndim == 3 and t.shape[2] == 4
        desc = RebTextureDesc(t.shape[1], t.shape[0], 0, 1 if srgb else 0)
        h = C.c_uint32()
        self._ck(self.lib.reb_register_texture(self.h, C.byref(desc), _ptr(t), C.byref(h)))
        return h.value

    def upload_scene(self, scene):
        self.set_shading_model(scene.get("shading_model", 0))
        if scene.get("skybox") is not None:
            self.set_skybox(scene["skybox"])
        self.upload_geometry(scene["verts"], scene["indices"])
        self.upload_transforms(scene["transforms"])
        self.upload_materials(scene["materials"])
        self.upload_lights(scene["lights"])
        for k, tx in enumerate(scene.get("textures", [])):
            assert self.register_texture(tx["texels"], tx.get("srgb", 1)) == k
        if scene.get("shadow_size", 0):
            self.shadow_create(scene["shadow_size"])
            self.shadow_set_transform(scene["shadow_view"], scene["shadow_proj"])
        self.set_view(scene["view"])

    # --- the hot path ---------------------------------------------------------------------------
    def render_shadow(self, draws):
        d = np.ascontiguousarray(draws, np.uint32).reshape(-1, 4)
        self._ck(self.lib.reb_render_shadow(self.h, self.shadow_id, _ptr(d), d.shape[0]))

    def render_forward(self, draws, translucent=None):
        """Base pass: the opaque list and, optionally, the translucent list drawn after it (Appendix B.16)."""
        d = np.ascontiguousarray(draws, np.uint32).reshape(-1, 4)
        if translucent is None or len(translucent) == 0:
            self._ck(self.lib.reb_render_forward(self.h, _ptr(d), d.shape[0]))
            return
        t = np.ascontiguousarray(translucent, np.uint32).reshape(-1, 4)
        self._ck(self.lib.reb_render_forward2(self.h, _ptr(d), d.shape[0], _ptr(t), t.shape[0]))

    def sync(self):
        self._ck(self.lib.reb_sync(self.h))

    def set_shading_model(self, model):
        self._ck(self.lib.reb_set_shading_model(self.h, int(model)))

    def set_sample_count(self, samples):
        """1 (default) or 4 samples per pixel for the base pass (Appendix B.15)."""
        self._ck(self.lib.reb_set_sample_count(self.h, int(samples)))
        self.samples = int(samples)

    def set_skybox(self, faces):
        """faces: (6, h, w, 4) uint16 cube map (+X -X +Y -Y +Z -Z), or None to remove it (Skybox.cpp:58-92)."""
        if faces is None:
            self._ck(self.lib.reb_set_skybox(self.h, None, 0, 0))
            return
        f = np.ascontiguousarray(faces, dtype=np.uint16)
        if f.ndim != 4 or f.shape[0] != 6 or f.shape[3] != 4:
            raise ValueError("skybox faces must have shape (6, h, w, 4)")
        self._ck(self.lib.reb_set_skybox(self.h, _ptr(f), f.shape[2], f.shape[1]))

    def flush(self):
        """Submit the recorded shadow pass now (before consuming the shadow map on the stream, e.g. an NCCL broadcast)."""
        self._ck(self.lib.reb_flush(self.h))

    # --- outputs ----------------------------------------------------------------------------------
    def output_ptr(self, kind):
        p = C.c_void_p(); b = C.c_uint64()
        self._ck(self.lib.reb_get_output(self.h, kind, C.byref(p), C.byref(b)))
        return p.value, b.value

    def readback(self, kind, out=None):
        if kind in (REB_MS_DEPTH_F32, REB_MS_TRIID_U32):
            shape, dt = (getattr(self, "samples", 1), self.height, self.width), (np.float32 if kind == REB_MS_DEPTH_F32 else np.uint32)
        elif kind == REB_SHADOW_F32:
            shape, dt = (self.shadow_size, self.shadow_size), np.float32
        else:
            dt, ch = _KIND_INFO[kind]
            shape = (self.height, self.width, ch) if ch > 1 else (self.height, self.width)
        if out is None:
            out = np.empty(shape, dt)
        self._ck(self.lib.reb_readback(self.h, kind, _ptr(out)))
        return out

    def readback_into(self, kind, host_ptr):
        """D2H into a caller-owned (e.g. pinned) buffer given as an integer address."""
        self._ck(self.lib.reb_readback(self.h, kind, C.c_void_p(host_ptr)))

    def readback_async_into(self, kind, host_ptr):
        """Enqueue the D2H copy into a pinned buffer and return; valid after the next sync() (sRGB8: copy stream + 2 images)."""
        self._ck(self.lib.reb_readback_async(self.h, kind, C.c_void_p(host_ptr)))

    def set_partition(self, rank, n_ranks, band_h=64):
        self._ck(self.lib.reb_set_partition(self.h, rank, n_ranks, band_h))

    def set_shadow_partition(self, rank, n_ranks, band_h):
        """Row split of the shadow pass (slabs of band_h rows, band_h % 32 == 0); fill the other rows, then shadow_mark_valid()."""
        self._ck(self.lib.reb_set_shadow_partition(self.h, rank, n_ranks, band_h))

    def set_shadow_split(self, rank, n_ranks):
        """Sort-last split of the shadow pass: this context rasterises its share of the triangle stream into the whole map; combine the
        ranks' maps with an element-wise minimum (parallel.allreduce_shadow_min)."""
        self._ck(self.lib.reb_set_shadow_split(self.h, rank, n_ranks))

    def shadow_mark_valid(self):
        self._ck(self.lib.reb_shadow_mark_valid(self.h, self.shadow_id))

    def shadow_wait_event(self, cuda_event):
        """The map is being filled on another stream (NCCL broadcast / all-gather): the next forward pass waits for `cuda_event`
        (a raw cudaEvent_t, e.g. torch.cuda.Event.cuda_event) right before its tile stage, so its K0 / K1 overlap the transfer."""
        self._ck(self.lib.reb_shadow_wait_event(self.h, self.shadow_id, C.c_void_p(int(cuda_event))))

    def counters(self):
        c = np.zeros(4, np.uint64)
        self._ck(self.lib.reb_get_counters(self.h, _ptr(c)))
        return c

    KERNEL_SLOTS = ("clear", "vertex", "setup_raster", "bin_scatter", "tile_stage", "unused")

    def set_profiling(self, on):
        self._ck(self.lib.reb_set_profiling(self.h, int(on)))

    def kernel_times_ms(self, pass_kind):
        t = np.zeros(6, np.float32)
        self._ck(self.lib.reb_get_kernel_times(self.h, pass_kind, _ptr(t)))
        return {k: float(x) for k, x in zip(self.KERNEL_SLOTS, t) if k != "unused"}

    def launch_count(self):
        n = C.c_uint64()
        self._ck(self.lib.reb_get_launch_count(self.h, C.byref(n)))
        return n.value
