"""Mirror of rpt's Renderer builder (src/renderer.rs:23-156) and photon drivers
(src/photon.rs:642-655) over the CUDA library.  No CPU path exists here."""
import ctypes as C

import numpy as np

from . import _abi as A
from .world import lower_scene, upload_scene


class RptxError(RuntimeError):
    pass


class Filter:
    """src/buffer.rs:96-108"""

    def __init__(self, radius=0):
        self.radius = int(radius)

    @staticmethod
    def Box(radius):
        return Filter(radius)


class PhotonRenderKind:
    PhotonMap, PhotonPointBeam, PhotonBeamBeam = A.PHOTON_MAP, A.PHOTON_POINT_BEAM, A.PHOTON_BEAM_BEAM


def _dp(a):
    return a.ctypes.data_as(A.DP)


class Context:
    """One CUDA device (rptx_ctx)."""

    def __init__(self, device=0):
        self.lib = A.load()
        self.device = int(device)
        h = C.c_void_p()
        rc = self.lib.rptx_ctx_create(int(device), C.byref(h))
        self.h = h
        if rc:
            msg = self.error()
            self.close()
            raise RptxError(f"rptx_ctx_create failed: {msg}")

    def error(self):
        return (self.lib.rptx_last_error(self.h) or b"").decode()

    # ---- multi-GPU: this rank's context joins an NCCL communicator owned by the library ----
    @staticmethod
    def comm_unique_id():
        """128-byte communicator id (an ncclUniqueId); rank 0 makes it, the host hands it to the other ranks."""
        buf = C.create_string_buffer(A.COMM_ID_BYTES)
        if A.load().rptx_comm_unique_id(buf):
            raise RptxError("rptx_comm_unique_id failed (libnccl.so.2 not loadable?)")
        return buf.raw

    def comm_init(self, comm_id, world, rank):
        self.check(self.lib.rptx_comm_init(self.h, C.c_char_p(comm_id), int(world), int(rank)), "rptx_comm_init")

    def comm_info(self):
        w, r = C.c_int(), C.c_int()
        self.lib.rptx_comm_info(self.h, C.byref(w), C.byref(r))
        return w.value, r.value

    def check(self, rc, what):
        if rc:
            raise RptxError(f"{what}: {self.error()}")

    def trim(self):
        """Give the device memory this context caches between calls (pool blocks, the spare film slab) back to the device."""
        self.check(self.lib.rptx_ctx_trim(self.h), "rptx_ctx_trim")

    def close(self):
        if self.h:
            self.lib.rptx_ctx_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_default_ctx = {}


def default_context(device=0):
    if device not in _default_ctx:
        _default_ctx[device] = Context(device)
    return _default_ctx[device]


class DeviceScene:
    """A Scene lowered and committed to one device (rptx_scene)."""

    def __init__(self, scene, ctx=None):
        self.ctx = ctx or default_context()
        self.lib = self.ctx.lib
        self.low = lower_scene(scene)
        h = C.c_void_p()
        self.ctx.check(self.lib.rptx_scene_create(self.ctx.h, C.byref(h)), "rptx_scene_create")
        self.h = h
        upload_scene(self.low, self)

    # backend protocol used by world.upload_scene
    def add_mesh(self, v, n):
        mid = C.c_uint32()
        v = np.ascontiguousarray(v, dtype=np.float64)
        nptr = None
        if n is not None:
            n = np.ascontiguousarray(n, dtype=np.float64)
            nptr = _dp(n)
        self.ctx.check(self.lib.rptx_scene_add_mesh(self.h, _dp(v), nptr, len(v), C.byref(mid)), "rptx_scene_add_mesh")
        return mid.value

    def add_object(self, sd, md):
        self.ctx.check(self.lib.rptx_scene_add_object(self.h, C.byref(sd), C.byref(md), None), "rptx_scene_add_object")

    def add_light(self, ld):
        self.ctx.check(self.lib.rptx_scene_add_light(self.h, C.byref(ld)), "rptx_scene_add_light")

    def set_medium(self, kind, a, s):
        self.ctx.check(self.lib.rptx_scene_set_medium(self.h, kind, a, s), "rptx_scene_set_medium")

    def set_environment_color(self, rgb):
        arr = (C.c_double * 3)(*rgb)
        self.ctx.check(self.lib.rptx_scene_set_environment_color(self.h, arr), "rptx_scene_set_environment_color")

    def set_environment_hdri(self, w, h, buf):
        buf = np.ascontiguousarray(buf, dtype=np.float64)
        self.ctx.check(self.lib.rptx_scene_set_environment_hdri(self.h, w, h, _dp(buf)), "rptx_scene_set_environment_hdri")

    def commit(self):
        self.ctx.check(self.lib.rptx_scene_commit(self.h), "rptx_scene_commit")

    def intersect_batch(self, o, d, t_min=1e-12):
        """get_closest_hit (src/renderer.rs:416-425) over host ray arrays [n,3]."""
        o = np.ascontiguousarray(o, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(d, dtype=np.float64).reshape(-1, 3)
        n = len(o)
        t = np.empty(n, np.float64)
        obj = np.empty(n, np.uint32)
        prim = np.empty(n, np.uint32)
        inst = np.empty(n, np.uint32)
        nrm = np.empty((n, 3), np.float64)
        rc = self.lib.rptx_intersect_batch(self.h, _dp(o), _dp(d), n, t_min, _dp(t), obj.ctypes.data_as(A.U32P),
                                           prim.ctypes.data_as(A.U32P), inst.ctypes.data_as(A.U32P), _dp(nrm))
        self.ctx.check(rc, "rptx_intersect_batch")
        return t, obj, prim, inst, nrm

    def intersect_batch_static(self, o, d, t_min=1e-12):
        """intersect_batch through rptx_intersect_batch_static: no device allocation (the staging is a fixed static
        device array of the library), at most A.INTERSECT_STATIC_MAX_RAYS rays per call.  Returns the same arrays plus
        the number of rays that hit something.  No counterpart in the reference."""
        o = np.ascontiguousarray(o, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(d, dtype=np.float64).reshape(-1, 3)
        n = len(o)
        t = np.empty(n, np.float64)
        obj = np.empty(n, np.uint32)
        prim = np.empty(n, np.uint32)
        inst = np.empty(n, np.uint32)
        nrm = np.empty((n, 3), np.float64)
        hits = C.c_uint64(0)
        rc = self.lib.rptx_intersect_batch_static(self.h, _dp(o), _dp(d), n, t_min, _dp(t), obj.ctypes.data_as(A.U32P),
                                                  prim.ctypes.data_as(A.U32P), inst.ctypes.data_as(A.U32P), _dp(nrm), C.byref(hits))
        self.ctx.check(rc, "rptx_intersect_batch_static")
        return t, obj, prim, inst, nrm, int(hits.value)

    def close(self):
        if self.h:
            self.lib.rptx_scene_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


ROW = A.PHOTON_ROW_F64


def rows_keys(rows):
    """u64 keys (photon index << 16 | depth) of [n,13] f64 photon rows."""
    return np.ascontiguousarray(rows[:, 12]).view(np.uint64)


class PhotonList:
    """rptx_photons: the reference's PhotonList(surface, volume) (src/photon.rs:141-171) on the device."""

    def __init__(self, ctx, h):
        self.ctx, self.lib, self.h = ctx, ctx.lib, h

    @staticmethod
    def emit(dscene, params, photon_count, kind, first=0, count=None):
        count = int(photon_count) - int(first) if count is None else int(count)
        h = C.c_void_p()
        rc = dscene.lib.rptx_photons_emit(dscene.ctx.h, dscene.h, C.byref(params), int(photon_count), int(first), count, int(kind), C.byref(h))
        dscene.ctx.check(rc, "rptx_photons_emit")
        return PhotonList(dscene.ctx, h)

    @staticmethod
    def empty(ctx):
        h = C.c_void_p()
        ctx.check(ctx.lib.rptx_photons_create(ctx.h, C.byref(h)), "rptx_photons_create")
        return PhotonList(ctx, h)

    def allgather(self):
        """PhotonList::merge across the ranks of the context's communicator (NCCL inside the library)."""
        h = C.c_void_p()
        self.ctx.check(self.lib.rptx_photons_allgather(self.h, C.byref(h)), "rptx_photons_allgather")
        return PhotonList(self.ctx, h)

    def counts(self):
        c = (C.c_uint64 * 2)()
        self.ctx.check(self.lib.rptx_photons_count(self.h, c), "rptx_photons_count")
        return int(c[0]), int(c[1])

    def rows(self, which):
        """Download one list as [n,13] f64 (column 12 holds the u64 key bits)."""
        out = np.empty((self.counts()[which], ROW), np.float64)
        self.ctx.check(self.lib.rptx_photons_download(self.h, which, out.ctypes.data_as(C.c_void_p)), "rptx_photons_download")
        return out

    def append_host(self, which, rows):
        rows = np.ascontiguousarray(rows, np.float64).reshape(-1, ROW)
        self.ctx.check(self.lib.rptx_photons_append_host(self.h, which, rows.ctypes.data_as(C.c_void_p), len(rows)), "rptx_photons_append_host")

    def append_device(self, which, ptr, n):
        self.ctx.check(self.lib.rptx_photons_append_device(self.h, which, C.c_void_p(int(ptr)), int(n)), "rptx_photons_append_device")

    def append_tensor(self, which, rows):
        """Append a [n,13] f64 torch tensor (device rows stay on the device)."""
        rows = rows.contiguous()
        if rows.is_cuda:
            import torch

            torch.cuda.current_stream(rows.device).synchronize()
            self.append_device(which, rows.data_ptr(), rows.shape[0])
        else:
            self.append_host(which, rows.numpy())

    def device_tensor(self, which):
        """torch view [n,13] f64 of one list's rows (for the NCCL all-gather)."""
        from .dist import device_f64_view

        n = self.counts()[which]
        ptr = self.lib.rptx_photons_device_ptr(self.h, which)
        return device_f64_view(ptr, n * ROW, f"cuda:{self.ctx.device}").view(n, ROW)

    def stats(self):
        e, r, ms = C.c_uint64(), C.c_uint64(), C.c_double()
        self.ctx.check(self.lib.rptx_photons_stats(self.h, C.byref(e), C.byref(r), C.byref(ms)), "rptx_photons_stats")
        return dict(emitted=e.value, rays=r.value, emit_ms=ms.value)

    def close(self):
        if self.h:
            self.lib.rptx_photons_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PhotonMap:
    """rptx_photon_map: the reference's PhotonMap enum (src/photon.rs:185-189) on the device."""

    def __init__(self, dscene, photons, kind, sharded=False):
        """sharded: `photons` is this rank's part of lists spread over the context's communicator; the ranks split the
        build and every rank gets the complete map (rptx_photon_map_build_sharded, collective)."""
        self.ctx, self.lib, self.kind = dscene.ctx, dscene.lib, int(kind)
        self.dscene = dscene
        h = C.c_void_p()
        if sharded:
            self.ctx.check(self.lib.rptx_photon_map_build_sharded(self.ctx.h, dscene.h, photons.h, int(kind), C.byref(h)), "rptx_photon_map_build_sharded")
        else:
            self.ctx.check(self.lib.rptx_photon_map_build(self.ctx.h, dscene.h, photons.h, int(kind), C.byref(h)), "rptx_photon_map_build")
        self.h = h

    def stats(self):
        c, ms = (C.c_uint64 * 3)(), C.c_double()
        self.ctx.check(self.lib.rptx_photon_map_stats(self.h, c, C.byref(ms)), "rptx_photon_map_stats")
        return dict(surface=int(c[0]), volume=int(c[1]), bvh_items=int(c[2]), build_ms=ms.value)

    def knn(self, which, q, k):
        """KdTree::nearests for host queries [n,3] -> (keys [n,k] u64, d2 [n,k], counts [n])."""
        q = np.ascontiguousarray(q, np.float64).reshape(-1, 3)
        n = len(q)
        keys = np.empty((n, k), np.uint64)
        d2 = np.empty((n, k), np.float64)
        counts = np.empty(n, np.uint32)
        rc = self.lib.rptx_photon_map_knn(self.h, which, _dp(q), n, k, keys.ctypes.data_as(A.U64P), _dp(d2), counts.ctypes.data_as(A.U32P))
        self.ctx.check(rc, "rptx_photon_map_knn")
        return keys, d2, counts

    def spheres(self):
        n = self.stats()["bvh_items"]
        c, r = np.empty((n, 3), np.float64), np.empty(n, np.float64)
        self.ctx.check(self.lib.rptx_photon_map_spheres(self.h, _dp(c), _dp(r)), "rptx_photon_map_spheres")
        return c, r

    def close(self):
        if self.h:
            self.lib.rptx_photon_map_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Buffer:
    """What iterative_render's callback receives (src/buffer.rs:5-94): image(), variance()."""

    def __init__(self, film):
        self._film = film

    def image(self):
        return self._film.image()

    def variance(self):
        return self._film.variance()

    def mean(self):
        return self._film.mean()


class Film:
    """rptx_film: the device Buffer plus the path pool of one render."""

    def __init__(self, dscene, camera, params, gather_only=False):
        self.dscene, self.ctx, self.lib = dscene, dscene.ctx, dscene.lib
        self.params = params
        self.w, self.h_px = params.width, params.height
        h = C.c_void_p()
        cam = camera.lower()
        begin = self.lib.rptx_photon_gather_begin if gather_only else self.lib.rptx_render_begin
        self.ctx.check(begin(self.ctx.h, dscene.h, C.byref(cam), C.byref(params), C.byref(h)), "rptx_render_begin")
        self.h = h

    def gather(self, photon_map, iterations):
        """This rank's shard of `iterations` photon-map samples per pixel into the batch sums."""
        self.ctx.check(self.lib.rptx_photon_gather_trace(self.h, photon_map.h, int(iterations)), "rptx_photon_gather_trace")

    def sample(self, iterations):
        self.ctx.check(self.lib.rptx_render_sample(self.h, int(iterations)), "rptx_render_sample")

    def sample_trace(self, iterations):
        self.ctx.check(self.lib.rptx_render_sample_trace(self.h, int(iterations)), "rptx_render_sample_trace")

    def sample_allreduce(self):
        """The exchange between sample_trace and sample_commit, over the context's communicator."""
        self.ctx.check(self.lib.rptx_render_sample_allreduce(self.h), "rptx_render_sample_allreduce")

    def sample_commit(self, iterations):
        self.ctx.check(self.lib.rptx_render_sample_commit(self.h, int(iterations)), "rptx_render_sample_commit")

    def batch_sum_ptr(self):
        return self.lib.rptx_film_batch_sum_device_ptr(self.h)

    def batch_sum_tensor(self):
        """The device batch-sum film ([h*w*3] f64) as a torch view, for the NCCL all-reduce."""
        if getattr(self, "_bs", None) is None:
            from .dist import device_f64_view

            self._bs = device_f64_view(self.batch_sum_ptr(), self.h_px * self.w * 3, f"cuda:{self.ctx.device}")
        return self._bs

    def image(self):
        out = np.empty((self.h_px, self.w, 3), np.uint8)
        self.ctx.check(self.lib.rptx_film_image_rgb8(self.h, out.ctypes.data_as(A.U8P)), "rptx_film_image_rgb8")
        return out

    def variance(self):
        v = C.c_double()
        self.ctx.check(self.lib.rptx_film_variance(self.h, C.byref(v)), "rptx_film_variance")
        return v.value

    def mean(self):
        out = np.empty((self.h_px, self.w, 3), np.float64)
        self.ctx.check(self.lib.rptx_film_mean(self.h, _dp(out)), "rptx_film_mean")
        return out

    def stats(self):
        st = A.Stats()
        self.ctx.check(self.lib.rptx_stats_get(self.h, C.byref(st)), "rptx_stats_get")
        return {k: getattr(st, k) for k, _ in A.Stats._fields_}

    def close(self):
        if self.h:
            self.lib.rptx_film_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Renderer:
    """Builder object for rendering a scene (src/renderer.rs:23-134); same defaults (:60-76)."""

    def __init__(self, scene, camera, ctx=None):
        self.scene, self.camera, self._ctx = scene, camera, ctx
        self._p = dict(width=800, height=600, exposure_value=0.0, filter_radius=0, stepsize=0.0, max_bounces=0, num_samples=1,
                       gather_size=50, gather_size_volume=50, watts=100.0, seed=0, shard_count=1, shard_index=0)
        self._dscene = None

    @staticmethod
    def new(scene, camera):
        return Renderer(scene, camera)

    def _set(self, k, v):
        self._p[k] = v
        return self

    def width(self, v): return self._set("width", int(v))
    def height(self, v): return self._set("height", int(v))
    def exposure_value(self, v): return self._set("exposure_value", float(v))
    def stepsize(self, v): return self._set("stepsize", float(v))
    def filter(self, f): return self._set("filter_radius", int(f.radius))
    def max_bounces(self, v): return self._set("max_bounces", int(v))
    def num_samples(self, v): return self._set("num_samples", int(v))
    def gather_size(self, v): return self._set("gather_size", int(v))
    def gather_size_volume(self, v): return self._set("gather_size_volume", int(v))
    def watts(self, v): return self._set("watts", float(v))
    # additions (the reference seeds from OS entropy and has one address space)
    def seed(self, v): return self._set("seed", int(v))
    def shard(self, count, index): self._set("shard_count", int(count)); return self._set("shard_index", int(index))

    def params(self):
        p = A.Params()
        for k, v in self._p.items():
            setattr(p, k, v)
        return p

    def device_scene(self):
        if self._dscene is None:
            self._dscene = DeviceScene(self.scene, self._ctx)
        return self._dscene

    def begin(self):
        return Film(self.device_scene(), self.camera, self.params())

    def render(self):
        """Render the scene by path tracing (src/renderer.rs:137-141) -> uint8 [h,w,3]."""
        film = self.begin()
        film.sample(self._p["num_samples"])
        return film.image()

    def iterative_render(self, callback_interval, callback):
        """src/renderer.rs:144-156: callback(iteration, buffer) after every k samples."""
        film = self.begin()
        iteration, total = 0, self._p["num_samples"]
        while iteration < total:
            steps = min(total - iteration, int(callback_interval))
            film.sample(steps)
            iteration += steps
            callback(iteration, Buffer(film))
        return film

    def get_closest_hit(self, origin, direction):
        """src/renderer.rs:416-425 for one ray -> (t, normal, object index) or None."""
        t, obj, _, _, n = self.device_scene().intersect_batch([origin], [direction])
        return None if obj[0] == 0xFFFFFFFF else (float(t[0]), tuple(n[0]), int(obj[0]))

    def photon_render(self, photon_count, kind, want_mean=False):
        """src/photon.rs:655-720: emit -> build -> gather on one device."""
        ds = self.device_scene()
        p = self.params()
        out = np.empty((p.height, p.width, 3), np.uint8)
        mean = np.empty((p.height, p.width, 3), np.float64) if want_mean else None
        cam = self.camera.lower()
        rc = ds.lib.rptx_photon_render(ds.ctx.h, ds.h, C.byref(cam), C.byref(p), int(photon_count), int(kind),
                                       out.ctypes.data_as(A.U8P), _dp(mean) if want_mean else None)
        ds.ctx.check(rc, "rptx_photon_render")
        return (out, mean) if want_mean else out

    # the three passes one by one (multi-GPU hosts shard pass 1 and pass 3; tests check each alone)
    def emit_photons(self, photon_count, kind, first=0, count=None):
        """PASS 1, src/photon.rs:663-675 -> PhotonList on the device."""
        return PhotonList.emit(self.device_scene(), self.params(), photon_count, kind, first, count)

    def build_photon_map(self, photons, kind, sharded=False):
        """PASS 2, src/photon.rs:693-702.  sharded: `photons` is this rank's part; collective over the library communicator."""
        return PhotonMap(self.device_scene(), photons, kind, sharded=sharded)

    def begin_gather(self):
        """Film of PASS 3 (src/photon.rs:706-716); film.gather(map, n) then film.sample_commit(n)."""
        return Film(self.device_scene(), self.camera, self.params(), gather_only=True)

    def photon_map_render(self, photon_count):
        return self.photon_render(photon_count, PhotonRenderKind.PhotonMap)

    def photon_point_query_beam_render(self, photon_count):
        return self.photon_render(photon_count, PhotonRenderKind.PhotonPointBeam)

    def photon_beam_query_beam_render(self, photon_count):
        return self.photon_render(photon_count, PhotonRenderKind.PhotonBeamBeam)
