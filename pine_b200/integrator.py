"""`PathIntegrator(BlueSampler(spp), depth).render(scene)` on a B200, through the C ABI.

Mirrors the reference's script-level interface (program_context.cpp:76-81, path.h:7,
sampler.cpp:115-121): same names, same argument meaning, same error behaviour
(`max_path_length <= 0` is an error, spp > 256 warns and clamps, spp rounds up to a power of two).
Everything numeric happens in pine_b200/lib/libpine_b200.so; there is no CPU path.
"""
import ctypes as C
import os
import struct
import warnings
import zlib

import numpy as np

from . import _abi as A

TABLES_PATH = os.path.join(A.HERE, "data", "bluesobol_tables.bin")
_tables = None


def sampler_tables():
    """{spp: (sobol u8[65536], scramble u8[131072], rank u8[131072])} from the extracted blob
    (pine_b200/tools/extract_tables.cpp documents the layout and provenance)."""
    global _tables
    if _tables is None:
        if not os.path.exists(TABLES_PATH):
            raise FileNotFoundError(
                f"{TABLES_PATH} missing: `make -C oracle tables` extracts it from the reference's "
                "vendored bluesobol sources (needs /root/reference)")
        raw = np.fromfile(TABLES_PATH, dtype=np.uint8)
        if bytes(raw[:8]) != b"PB2BLUE1":
            raise ValueError("bad sampler table file")
        n = struct.unpack_from("<I", raw, 8)[0]
        off, out = 16, {}
        for _ in range(n):
            spp = struct.unpack_from("<I", raw, off)[0]
            off += 8
            sob = raw[off:off + 65536]; off += 65536
            scr = raw[off:off + 131072]; off += 131072
            rnk = raw[off:off + 131072]; off += 131072
            out[spp] = (np.ascontiguousarray(sob), np.ascontiguousarray(scr),
                        np.ascontiguousarray(rnk))
        _tables = out
    return _tables


class BlueSampler:
    """BlueSobolSampler (PRL name `BlueSampler`), sampler.cpp:115-121."""

    def __init__(self, samples_per_pixel):
        if samples_per_pixel > 256:
            warnings.warn("[BlueSobolSampler]Only support up to 256 samples per pixel")
            samples_per_pixel = 256
        self.samples_per_pixel = A.lib().pb2h_round_spp(int(samples_per_pixel))
        if self.samples_per_pixel <= 0:
            raise ValueError("samples per pixel must be positive")

    def spp(self):
        return self.samples_per_pixel


class SobolSampler:
    """pine::SobolSampler (sampler.h:83-164): Owen-scrambled (0,2)-sequence over a Morton-ordered pixel
    index; any positive spp. Draws are bit-identical to the reference's."""
    kind = A.SAMPLER_SOBOL

    def __init__(self, samples_per_pixel):
        if samples_per_pixel <= 0:
            raise ValueError("samples per pixel must be positive")
        self.samples_per_pixel = int(samples_per_pixel)
        self.seed = 0

    def spp(self):
        return self.samples_per_pixel


class HaltonSampler:
    """pine::HaltonSampler (sampler.h:40-81): scrambled radical inverses over the first 1000 primes, the
    pixel's index from the Chinese remainder theorem on (x mod 128, y mod 128). Draws bit-identical."""
    kind = A.SAMPLER_HALTON

    def __init__(self, samples_per_pixel):
        if samples_per_pixel <= 0:
            raise ValueError("`HaltonSampler` should have positive samples per pixel")  # sampler.cpp:45-46
        self.samples_per_pixel = int(samples_per_pixel)
        self.seed = 0

    def spp(self):
        return self.samples_per_pixel


class UniformSampler:
    """pine::UniformSampler (sampler.h:13-39): independent uniform draws. The reference shares one RNG
    per worker thread; here each (pixel, sample, dimension) has its own draw (include/pine_b200.h)."""
    kind = A.SAMPLER_UNIFORM

    def __init__(self, samples_per_pixel, seed=0):
        if samples_per_pixel <= 0:
            raise ValueError("`UniformSampler` should have positive samples per pixel")
        self.samples_per_pixel = int(samples_per_pixel)
        self.seed = int(seed)

    def spp(self):
        return self.samples_per_pixel


def _sampler_of(spp_or_sampler):
    """(spp, kind, seed) of a sampler object, or of BlueSampler's already-rounded spp given as an int."""
    if isinstance(spp_or_sampler, (int, np.integer)):
        return int(spp_or_sampler), A.SAMPLER_BLUE, 0
    return spp_or_sampler.spp(), getattr(spp_or_sampler, "kind", A.SAMPLER_BLUE), getattr(spp_or_sampler, "seed", 0)


def host_sampler_draws(sampler, size, px, py, sample_index, pattern, n_samples):
    """Draws of a SobolSampler / UniformSampler object evaluated by the library's host build of the device
    draw functions (pb2h_sampler_draws; test hook, no GPU, no film)."""
    spp, kind, seed = _sampler_of(sampler)
    width = sum(1 if c == "1" else 2 for c in pattern)
    d = np.zeros(max(width * n_samples, 1), np.float32)
    A.check(A.lib().pb2h_sampler_draws(kind, spp, seed, size[0], size[1], px, py, sample_index,
                                       pattern.encode(), n_samples, d.ctypes.data))
    return d[:width * n_samples].reshape(n_samples, width)


_nccl_preloaded = False


def _preload_bundled_nccl():
    """pb2_multi binds NCCL with dlopen("libnccl.so.2") at its first multi-GPU use. If that loaded the system
    library and the process imported PyTorch LATER, torch's libtorch_cuda.so would be resolved against it (same
    SONAME) and fail on symbols only its own, newer NCCL has. So when PyTorch's bundled NCCL is installed
    (site-packages/nvidia/nccl/lib) it is loaded first and both sides share it; PB2_NCCL_LIB overrides."""
    global _nccl_preloaded
    if _nccl_preloaded or os.environ.get("PB2_NCCL_LIB"):
        return
    _nccl_preloaded = True
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        for d in (spec.submodule_search_locations if spec else []):
            path = os.path.join(d, "lib", "libnccl.so.2")
            if os.path.exists(path):
                C.CDLL(path, mode=C.RTLD_GLOBAL)
                return
    except (ImportError, OSError, ValueError):
        pass


class Backend:
    """One pb2_context (one GPU). `device=-1` makes a host-only context (BVH builder inspection)."""

    def __init__(self, device=0):
        self.L = A.lib()
        h = C.c_void_p()
        A.check(self.L.pb2_create(device, C.byref(h)))
        self.h = h
        self.device = device
        self._tables_loaded = set()
        self._scene_keep = None
        self.size = None

    def close(self):
        if getattr(self, "h", None):
            self.L.pb2_destroy(self.h)
            self.h = None

    def set_progress_callback(self, fn):
        """`fn(fraction)` is called as the render's waves complete on the device (path.cpp:39's set_progress); it
        runs on a CUDA driver thread while the render call blocks. None clears."""
        self._progress_keep = A.PROGRESS_FN((lambda f, _u: fn(f))) if fn else A.PROGRESS_FN()
        A.check(self.L.pb2_set_progress_callback(self.h, self._progress_keep, None))

    def __del__(self):
        self.close()

    def upload_scene(self, scene):
        d, keep = scene.desc()
        A.check(self.L.pb2_upload_scene(self.h, C.byref(d)))
        self._scene_keep = (scene, keep)
        self.size = scene.size

    def export_scene(self, path):
        """Write the scene this backend BUILT to `path` (/dev/shm/...) for the other processes of the node."""
        A.check(self.L.pb2_export_host_scene(self.h, os.fsencode(path)))

    def import_scene(self, path, size):
        """Adopt a scene another process built and exported (instead of upload_scene); `size` = (w, h) of its film."""
        A.check(self.L.pb2_import_host_scene(self.h, os.fsencode(path)))
        self.size = (int(size[0]), int(size[1]))

    def upload_tables(self, spp):
        if spp in self._tables_loaded:
            return
        if spp not in sampler_tables():  # BlueSobolSampler::spp() is a power of two <= 256 (sampler.cpp:115-121)
            raise A.PineB200Error(-1, f"spp must be BlueSobolSampler::spp(): a power of two <= 256, got {spp}")
        sob, scr, rnk = sampler_tables()[spp]
        A.check(self.L.pb2_upload_sampler_tables(self.h, spp, sob.ctypes.data, scr.ctypes.data,
                                                 rnk.ctypes.data))
        self._tables_loaded.add(spp)

    def _params(self, spp, depth, rows=None, samples=None, accel=0, flags=0, bands=None):
        """`spp`: an int (BlueSampler's spp) or a sampler object (BlueSampler/SobolSampler/UniformSampler).
        `bands` = (band, parts, part): only the interleaved row bands of this part (pb2_render_params)."""
        spp, kind, seed = _sampler_of(spp)
        if kind == A.SAMPLER_BLUE:
            self.upload_tables(spp)
        p = A.RenderParams()
        p.spp, p.max_path_length = spp, depth
        p.sampler, p.sampler_seed = kind, seed
        p.row_begin, p.row_end = rows if rows else (0, self.size[1])
        p.sample_begin, p.sample_end = samples if samples else (0, spp)
        p.accel, p.flags = accel, flags
        if bands:
            p.row_band, p.row_parts, p.row_part = bands
        return p

    def restage(self, stream=None):
        """Copy the built scene host->device again (asynchronously on `stream`)."""
        A.check(self.L.pb2_restage_scene(self.h, stream))

    def restage_part(self, part, n_parts, stream=None):
        """Copy only slice `part` of `n_parts` of every large scene array host->device (small ones whole);
        -> [(device pointer, padded bytes, part bytes)] of the large arrays, for an in-place all-gather
        (multigpu.restage_sharded)."""
        arr = (A.DeviceArray * 64)()
        n = C.c_int(0)
        A.check(self.L.pb2_restage_scene_part(self.h, stream, part, n_parts, arr, 64, C.byref(n)))
        return [(int(arr[i].device_ptr), int(arr[i].bytes), int(arr[i].part_bytes)) for i in range(n.value)]

    def render(self, spp, depth, rows=None, samples=None, accel=0, out=None, flags=0, bands=None):
        """-> film float32 [h, w, 4] in host memory (pre-tonemap). `out` may be any writable
        C-contiguous float32 buffer object of that size (e.g. a pinned torch tensor's numpy view)."""
        w, h = self.size
        film = out if out is not None else np.empty((h, w, 4), np.float32)
        p = self._params(spp, depth, rows, samples, accel, flags, bands)
        A.check(self.L.pb2_render(self.h, C.byref(p), film.ctypes.data))
        return film

    def render_recorded(self, spp, depth, accel=0, samples=None):
        """Test hook: render (one wave) and return (film, record) where record holds every ray the
        render traced: closest rays with (geom, prim, t) and shadow rays with their occlusion."""
        w, h = self.size
        cap = w * h * spp * depth
        rec = A.RayRecord()
        bufs = {"closest_rays": np.zeros((cap, 8), np.float32), "closest_t": np.zeros(cap, np.float32),
                "closest_geom": np.zeros(cap, np.int32), "closest_prim": np.zeros(cap, np.int32),
                "shadow_rays": np.zeros((cap, 8), np.float32), "shadow_occluded": np.zeros(cap, np.uint8)}
        rec.capacity = cap
        for k, v in bufs.items():
            setattr(rec, k, v.ctypes.data)
        A.check(self.L.pb2_set_ray_record(self.h, C.byref(rec)))
        try:
            film = self.render(spp, depth, accel=accel, samples=samples)
        finally:
            A.check(self.L.pb2_set_ray_record(self.h, None))
        nc, ns = int(rec.n_closest), int(rec.n_shadow)
        out = {k: (v[:ns] if k.startswith("shadow") else v[:nc]) for k, v in bufs.items()}
        return film, out

    def render_device(self, spp, depth, rows=None, samples=None, accel=0, stream=None, flags=0, bands=None):
        """-> device pointer (int) of the float4 film owned by the context."""
        p = self._params(spp, depth, rows, samples, accel, flags, bands)
        ptr = C.c_void_p()
        A.check(self.L.pb2_render_device(self.h, C.byref(p), stream, C.byref(ptr)))
        return ptr.value

    def finalize(self, film, flip_y=True):
        w, h = self.size
        f = np.ascontiguousarray(film, np.float32)
        rgba8 = np.empty((h, w, 4), np.uint8)
        mapped = np.empty((h, w, 4), np.float32)
        A.check(self.L.pb2_film_finalize(self.h, f.ctypes.data, int(flip_y), rgba8.ctypes.data,
                                         mapped.ctypes.data))
        return rgba8, mapped

    def finalize_device(self, flip_y=True, tonemapper=-1, want_mapped=False):
        """Film::finalize + quantise on the film the last render left on the device -> RGBA8 [h, w, 4]
        (and the tone-mapped float film when asked): 4 bytes per pixel cross PCIe instead of 16."""
        w, h = self.size
        rgba8 = np.empty((h, w, 4), np.uint8)
        mapped = np.empty((h, w, 4), np.float32) if want_mapped else None
        A.check(self.L.pb2_film_finalize_device(self.h, tonemapper, int(flip_y), rgba8.ctypes.data,
                                                mapped.ctypes.data if want_mapped else None))
        return (rgba8, mapped) if want_mapped else rgba8

    def trace_closest(self, rays, accel=0, surf=False):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        n = len(rays)
        geom, prim = np.empty(n, np.int32), np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        s = np.empty((n, 8), np.float32) if surf else None
        A.check(self.L.pb2_trace_closest(self.h, rays.ctypes.data, n, accel, geom.ctypes.data,
                                         prim.ctypes.data, t.ctypes.data,
                                         s.ctypes.data if surf else None))
        out = {"geom": geom, "prim": prim, "t": t}
        if surf:
            out["surf"] = s
        return out

    def trace_any(self, rays, accel=0):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        occ = np.empty(len(rays), np.uint8)
        A.check(self.L.pb2_trace_any(self.h, rays.ctypes.data, len(rays), accel, occ.ctypes.data))
        return occ

    def test_sampler(self, spp, px, py, sample_index, n_rand, pattern, n_samples):
        self.upload_tables(spp)
        width = sum(1 if c == "1" else 2 for c in pattern)
        r = np.zeros(max(n_rand, 1), np.float32)
        d = np.zeros(max(width * n_samples, 1), np.float32)
        A.check(self.L.pb2_test_sampler(self.h, spp, px, py, sample_index, n_rand, r.ctypes.data,
                                        pattern.encode(), n_samples, d.ctypes.data))
        return r[:n_rand], d[:width * n_samples].reshape(n_samples, width)

    def test_sampler_kind(self, sampler, px, py, sample_index, pattern, n_samples):
        """get1d/get2d draws of `sampler` (an object of this module) for the uploaded scene's film size."""
        spp, kind, seed = _sampler_of(sampler)
        if kind == A.SAMPLER_BLUE:
            self.upload_tables(spp)
        width = sum(1 if c == "1" else 2 for c in pattern)
        d = np.zeros(max(width * n_samples, 1), np.float32)
        A.check(self.L.pb2_test_sampler_kind(self.h, kind, spp, seed, px, py, sample_index, pattern.encode(),
                                             n_samples, d.ctypes.data))
        return d[:width * n_samples].reshape(n_samples, width)

    def eval_node(self, start, ctx8):
        """The uploaded scene's node program starting at `start` (1-based: Material.*_node) at the points
        ctx8 = n x {p.xyz, n.xyz, uv} -> n x 3 (pb2_test_eval_node)."""
        c = np.ascontiguousarray(ctx8, np.float32).reshape(-1, 8)
        out = np.zeros((len(c), 3), np.float32)
        A.check(self.L.pb2_test_eval_node(self.h, int(start), len(c), c.ctypes.data, out.ctypes.data))
        return out

    def primary_rays(self, si=0):
        w, h = self.size
        rays = np.empty((w * h, 8), np.float32)
        A.check(self.L.pb2_test_primary_rays(self.h, si, rays.ctypes.data))
        return rays

    def measure_read_gbs(self, nbytes, iters):
        """Sustained device read bandwidth in GB/s of an `nbytes` buffer read `iters` times (L1 bypassed):
        nbytes well below the 126 MB L2 measures the L2, far above it the HBM."""
        g = C.c_double()
        A.check(self.L.pb2_measure_read_gbs(self.h, nbytes, iters, C.byref(g)))
        return g.value

    def counters(self):
        c = A.Counters()
        A.check(self.L.pb2_get_counters(self.h, C.byref(c)))
        return {k: getattr(c, k) for k, _ in A.Counters._fields_}

    def bvh_tree(self, which=-1):
        n, root, npr = C.c_int32(), C.c_int32(), C.c_int32()
        A.check(self.L.pb2_bvh_info(self.h, which, C.byref(n), C.byref(root), C.byref(npr)))
        aabbs = np.zeros((n.value, 12), np.float32)
        children = np.zeros((n.value, 2), np.int32)
        leaf_range = np.zeros((n.value, 2), np.int32)
        leaf_prims = np.zeros(max(npr.value, 1), np.int32)
        if n.value:
            A.check(self.L.pb2_bvh_nodes(self.h, which, aabbs.ctypes.data, children.ctypes.data,
                                         leaf_range.ctypes.data, leaf_prims.ctypes.data))
        return {"root": root.value, "aabbs": aabbs, "children": children,
                "leaf_range": leaf_range, "leaf_prims": leaf_prims[:npr.value]}


class MultiBackend:
    """One pb2_multi: every GPU of the box (or `devices`) behind one render call, inside the library --
    one host build, replicas staged 1/N through PCIe + in-place ncclAllGather, shards rendered by one host
    thread per GPU, ONE ncclReduce(sum) of the float film (include/pine_b200.h pb2_multi_*)."""

    def __init__(self, devices=None):
        self.L = A.lib()
        _preload_bundled_nccl()
        h = C.c_void_p()
        if devices is None:
            A.check(self.L.pb2_create_multi(None, 0, C.byref(h)))
        else:
            arr = (C.c_int * len(devices))(*devices)
            A.check(self.L.pb2_create_multi(arr, len(devices), C.byref(h)))
        self.h = h
        self.n = self.L.pb2_multi_size(h)
        self._tables_loaded = set()
        self._scene_keep = None
        self.size = None

    def close(self):
        if getattr(self, "h", None):
            self.L.pb2_destroy_multi(self.h)
            self.h = None

    def set_progress_callback(self, fn):
        """`fn(fraction)`: mean over the devices of their completed waves (called from the devices' driver threads)."""
        self._progress_keep = A.PROGRESS_FN((lambda f, _u: fn(f))) if fn else A.PROGRESS_FN()
        A.check(self.L.pb2_multi_set_progress_callback(self.h, self._progress_keep, None))

    def __del__(self):
        self.close()

    def upload_scene(self, scene):
        d, keep = scene.desc()
        A.check(self.L.pb2_multi_upload_scene(self.h, C.byref(d)))
        self._scene_keep = (scene, keep)
        self.size = scene.size

    def restage(self):
        A.check(self.L.pb2_multi_restage_scene(self.h))

    def upload_tables(self, spp):
        if spp in self._tables_loaded:
            return
        if spp not in sampler_tables():
            raise A.PineB200Error(-1, f"spp must be BlueSobolSampler::spp(): a power of two <= 256, got {spp}")
        sob, scr, rnk = sampler_tables()[spp]
        A.check(self.L.pb2_multi_upload_sampler_tables(self.h, spp, sob.ctypes.data, scr.ctypes.data, rnk.ctypes.data))
        self._tables_loaded.add(spp)

    def render(self, spp, depth, mode="samples", accel=0, out=None, flags=0):
        """-> film float32 [h, w, 4] on the host. mode "samples" (statistically equal to one GPU) or "rows"
        (bit-identical to one GPU)."""
        spp, kind, seed = _sampler_of(spp)
        if kind == A.SAMPLER_BLUE:
            self.upload_tables(spp)
        w, h = self.size
        film = out if out is not None else np.empty((h, w, 4), np.float32)
        p = A.RenderParams()
        p.spp, p.max_path_length, p.sampler, p.sampler_seed = spp, depth, kind, seed
        p.row_begin, p.row_end, p.sample_begin, p.sample_end = 0, h, 0, spp
        p.accel, p.flags = accel, flags
        A.check(self.L.pb2_multi_render(self.h, C.byref(p), A.SHARD_ROWS if mode == "rows" else A.SHARD_SAMPLES,
                                        film.ctypes.data))
        return film

    def finalize_device(self, flip_y=True, tonemapper=-1):
        """RGBA8 of the reduced film of the last render (tone-mapped on the first device)."""
        w, h = self.size
        rgba8 = np.empty((h, w, 4), np.uint8)
        A.check(self.L.pb2_multi_film_finalize(self.h, tonemapper, int(flip_y), rgba8.ctypes.data, None))
        return rgba8

    def counters(self):
        c = A.Counters()
        r, s = C.c_double(), C.c_double()
        A.check(self.L.pb2_multi_get_counters(self.h, C.byref(c), C.byref(r), C.byref(s)))
        out = {k: getattr(c, k) for k, _ in A.Counters._fields_}
        out["reduce_ms"], out["stage_ms"] = r.value, s.value
        return out


class PathIntegrator:
    """PathIntegrator(sampler, max_path_length) -- path.cpp:9-14 / program_context.cpp:79-81."""

    def __init__(self, sampler, max_path_length, device=0, accel=0):
        if max_path_length <= 0:
            raise ValueError(
                f"`PathIntegrator` expect `max_path_length` to be positive, get{max_path_length}")
        self.sampler, self.max_path_length = sampler, max_path_length
        self.device, self.accel = device, accel
        self.backend = None

    def render(self, scene):
        """Fills scene.camera.film().pixels (float32 [h, w, 4], pre-tonemap), like path.cpp:26-41."""
        if self.backend is None:
            self.backend = Backend(self.device)
        self.backend.upload_scene(scene)
        film = self.backend.render(self.sampler, self.max_path_length, accel=self.accel)
        scene.camera.film().pixels = film
        scene._backend = self.backend
        scene._rendered_film = film  # the array whose device twin is still in the context (save_film)
        return film


def write_png(path, rgba8):
    h, w, _ = rgba8.shape
    raw = b"".join(b"\x00" + rgba8[y].tobytes() for y in range(h))

    def chunk(tag, data):
        return (struct.pack(">I", len(data)) + tag + data +
                struct.pack(">I", zlib.crc32(tag + data) & 0xffffffff))

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 6, 0, 0, 0)) +
                chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def save_film(film, path, scene=None):
    """Film::save: finalize (tone-map) on the device, gamma/quantise/y-flip, PNG on the host
    (film.cpp:12-27, fileio.cpp:42-76)."""
    backend = getattr(scene, "_backend", None) if scene is not None else None
    if backend is None:
        raise RuntimeError("save needs the scene that was rendered (its device context)")
    if film.pixels is getattr(scene, "_rendered_film", None):
        # the film of the render is still on the device: tone-map it there and fetch 4 bytes per pixel
        rgba8 = backend.finalize_device(flip_y=True)
    else:  # the caller replaced or edited Film::pixels: finalize what it holds now
        rgba8, _ = backend.finalize(film.pixels, flip_y=True)
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    write_png(path, rgba8)
