"""B200 render backend behind the viewer's renderer interface.

``GaussianRenderBase`` is the reference's abstract interface (/root/reference/renderer_ogl.py:75-101);
``B200Renderer`` implements it the way ``OpenGLRenderer`` (renderer_ogl.py:104-179) and ``CUDARenderer``
(renderer_cuda.py:101-258) do, with the same method names and the argument lists the viewer's frame
loop actually uses (main.py:500-520, 824-837, 844-887).  Every method forwards to the C ABI of
``csrc/libgsr.so`` (include/gsr.h); PyTorch only provides device buffers and the CUDA stream.
There is no CPU path: construction raises RuntimeError if the library or a CUDA device is missing.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _gsr, util, util_gau


class GaussianRenderBase:
    # /root/reference/renderer_ogl.py:75-101
    def __init__(self):
        self.gaussians = None

    def update_gaussian_data(self, gaus: util_gau.GaussianData):
        raise NotImplementedError()

    def sort_and_update(self):
        raise NotImplementedError()

    def set_scale_modifier(self, modifier: float):
        raise NotImplementedError()

    def set_render_mod(self, mod: int):
        raise NotImplementedError()

    def update_camera_pose(self, camera: util.Camera):
        raise NotImplementedError()

    def update_camera_intrin(self, camera: util.Camera):
        raise NotImplementedError()

    def draw(self):
        raise NotImplementedError()

    def set_render_reso(self, w, h):
        raise NotImplementedError()


def _f32c(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _fptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


class B200Renderer(GaussianRenderBase):
    """Headless: ``draw()`` writes (and returns) a float32 RGBA (H, W, 4) CUDA tensor, row 0 = top of
    the image — the layout CUDARenderer.draw builds at renderer_cuda.py:226-228 — instead of a GL
    framebuffer.  ``read_pixels_rgb8`` / ``read_pixels_r32f`` stand in for the viewer's glReadPixels
    calls (main.py:852, 870)."""

    def __init__(self, w, h, shaderpath=None, device=0, keys64=False, official_compat=False):
        super().__init__()
        import torch  # device memory + streams only

        if not torch.cuda.is_available():
            raise RuntimeError("B200Renderer needs a CUDA device (no CPU fallback)")
        self._torch = torch
        self._lib = _gsr.load()
        self.device = torch.device("cuda", device)
        hptr = C.c_void_p()
        _gsr.check(self._lib.gsr_create(int(device), C.byref(hptr)))
        self._h = hptr
        self.scale_modifier = 1.0
        self.render_mod = 3
        self.is_depth_image = 0
        self.keys64 = bool(keys64)
        # True: the semantics of the viewer's CUDA backend (renderer_cuda.py, diff_gaussian_rasterization forward)
        # instead of the OpenGL shaders' — see GSR_FLAG_OFFICIAL in include/gsr.h
        self.official_compat = bool(official_compat)
        self.auto_sort = True
        self.debug = False
        self.target_unorm = None  # None: unorm (clamped) target except in render_mod -1 (R32F FBO, main.py:866-870)
        self._view = None
        self._proj = None
        self._hff = None
        self._campos = np.zeros(3, dtype=np.float32)
        self._sort_view = None
        self._frame = None
        self._pinned = {}
        self.w = self.h = 0
        self.set_render_reso(w, h)

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                self._lib.gsr_destroy(self._h)
                self._h = None
        except Exception:
            pass

    # ---- interface ----------------------------------------------------------------------------
    def update_gaussian_data(self, gaus: util_gau.GaussianData):
        # renderer_ogl.py:132-137 / renderer_cuda.py:134-136
        self.gaussians = gaus
        xyz, rot, scale, op, sh = (_f32c(gaus.xyz), _f32c(gaus.rot), _f32c(gaus.scale), _f32c(gaus.opacity),
                                   _f32c(gaus.sh))
        n = len(xyz)
        if rot.shape != (n, 4) or scale.shape != (n, 3) or op.size != n or sh.shape[0] != n:
            raise ValueError("GaussianData arrays have inconsistent shapes")
        vp = lambda a: C.c_void_p(a.ctypes.data)  # noqa: E731
        _gsr.check(self._lib.gsr_set_gaussians(self._h, vp(xyz), vp(rot), vp(scale), vp(op), vp(sh), n,
                                               int(sh.shape[1]), 0))

    def update_gaussian_data_device(self, xyz, rot, scale, opacity, sh):
        """Same, from CUDA tensors already on this device (e.g. after a torch.distributed.broadcast)."""
        ts = [t.contiguous().float() for t in (xyz, rot, scale, opacity, sh)]
        n = ts[0].shape[0]
        _gsr.check(self._lib.gsr_set_gaussians(self._h, *[C.c_void_p(t.data_ptr()) for t in ts], n,
                                               int(ts[4].shape[1]), 1))

    def alloc_gaussians(self, n: int, sh_dim: int):
        """Resident scene buffers for n Gaussians, contents undefined: the receiving side of a scene broadcast
        (pose_shard.broadcast_scene_resident)."""
        _gsr.check(self._lib.gsr_alloc_gaussians(self._h, int(n), int(sh_dim)))
        self.gaussians = None

    def scene_buffers(self):
        """The four resident scene planes as float32 CUDA tensors that ALIAS the library's buffers (no copy):
        xyz (N,3), rot (N,4 wxyz), scale+opacity (N,4), sh (N,sh_dim).  Valid until the scene is replaced."""
        torch = self._torch
        out = []
        for which, cols in ((0, 3), (1, 4), (2, 4), (3, None)):
            ptr, nb = C.c_void_p(), C.c_longlong()
            _gsr.check(self._lib.gsr_scene_buffer(self._h, which, C.byref(ptr), C.byref(nb)))
            n_el = nb.value // 4

            class _Raw:  # CUDA array interface over the raw allocation
                pass
            raw = _Raw()
            raw.__cuda_array_interface__ = {"shape": (max(n_el, 0),), "typestr": "<f4", "data": (ptr.value or 0, False),
                                            "version": 2}
            t = torch.as_tensor(raw, device=self.device) if n_el > 0 else torch.empty(0, device=self.device)
            if cols is None:
                n = out[0].shape[0]
                cols = (n_el // n) if n else 0
            out.append(t.view(-1, cols) if n_el else t.view(0, cols or 1))
        return out

    def load_ply(self, path):
        """``load_ply(path)`` + ``update_gaussian_data`` in one step with the decode on the device
        (/root/reference/util_gau.py:63-125 followed by renderer_ogl.py:132-137): the header is parsed
        here, the vertex rows are mapped from the file and ``gsr_load_ply_rows`` turns them
        into the resident scene (activations bit-identical to the host ``util_gau.load_ply``).
        Returns (bounding_box (2,3) float32, center (3,) float32) like the reference's second and third
        results; ``download_gaussians()`` gives the GaussianData if the caller wants it on the host."""
        torch = self._torch
        with open(path, "rb") as f:
            fmt, count, props = util_gau.read_ply_header(f)
            row_bytes, offs = util_gau.ply_row_layout(props)
            if fmt == "ascii":
                rows = util_gau.read_ply_vertices(path).view(np.uint8)
                big = 0
            else:
                big = 1 if fmt == "binary_big_endian" else 0
                nbytes = count * row_bytes
                body = f.tell()
                f.seek(0, 2)
                if f.tell() - body < nbytes:
                    raise ValueError(f"{path}: truncated PLY body ({f.tell() - body} of {nbytes} bytes)")
                # the body is MAPPED, not read: gsr_load_ply_rows uploads it in 64 MB chunks straight out of the page
                # cache while the previous chunk is decoded.  (Pinning a buffer of the file's size first cost more than
                # the whole decode: 0.35 s per 300 k vertices in round 1.)
                rows = np.memmap(path, dtype=np.uint8, mode="r", offset=body, shape=(nbytes,)) if nbytes else \
                    np.zeros(0, np.uint8)
        bbox, center = self.load_ply_rows(rows, count, row_bytes, offs, big)
        return bbox, center

    def load_ply_rows(self, rows, count, row_bytes, offsets, big_endian=0):
        """Vertex rows (a uint8 NumPy array on the host, or a uint8 CUDA tensor) -> resident scene."""
        lay = _gsr.PlyLayout()
        lay.row_bytes = int(row_bytes)
        lay.big_endian = int(big_endian)
        for i, o in enumerate(offsets):
            lay.offset[i] = int(o)
        bbox = np.zeros((2, 3), dtype=np.float32)
        center = np.zeros(3, dtype=np.float32)
        if isinstance(rows, np.ndarray):
            ptr, on_dev = C.c_void_p(rows.ctypes.data), 0
        else:
            ptr, on_dev = C.c_void_p(rows.data_ptr()), 1
        _gsr.check(self._lib.gsr_load_ply_rows(self._h, ptr, int(count), C.byref(lay), on_dev, _fptr(bbox),
                                               _fptr(center), self._stream()))
        self.gaussians = None
        return bbox, center

    def download_gaussians(self) -> util_gau.GaussianData:
        """The resident scene as GaussianData (what OpenGLRenderer keeps in ``self.gaussians``)."""
        st = self.stats()
        n, shd = int(st.n_gaussians), 48
        buf = C.c_void_p()
        nb = C.c_longlong()
        _gsr.check(self._lib.gsr_scene_buffer(self._h, 3, C.byref(buf), C.byref(nb)))
        if n:
            shd = int(nb.value // (4 * n))
        xyz = np.empty((n, 3), np.float32)
        rot = np.empty((n, 4), np.float32)
        scale = np.empty((n, 3), np.float32)
        op = np.empty((n, 1), np.float32)
        sh = np.empty((n, shd), np.float32)
        vp = lambda a: C.c_void_p(a.ctypes.data)  # noqa: E731
        _gsr.check(self._lib.gsr_get_gaussians(self._h, vp(xyz), vp(rot), vp(scale), vp(op), vp(sh)))
        return util_gau.GaussianData(xyz, rot, scale, op, sh)

    def _resolve_view(self, camera, use_file, pose):
        if use_file:
            return camera.get_view_matrix(True, pose["camera_front"], pose["camera_position"], pose["camera_up"],
                                          pose["camera_view"])
        return camera.get_view_matrix(True)

    def sort_and_update(self, camera: util.Camera = None, use_file=False, pose=None):
        """renderer_ogl.py:139-146.  With ``auto_sort`` True (default; the viewer's "auto sort" box, main.py:993-997)
        draw() depth-sorts for the CURRENT view every frame, as the CUDA backend does (renderer_cuda.py:139-140),
        and this call only records the view for ``gaussian_order()``.  With ``auto_sort`` False the call freezes
        the depth order of this view on the device (gsr_freeze_order) and draw() keeps using it until the next
        sort_and_update, which is what the OpenGL backend does between presses of "sort Gaussians"."""
        if camera is not None:
            self._sort_view = _f32c(self._resolve_view(camera, use_file, pose))
        if not self.auto_sort and self._sort_view is not None:
            _gsr.check(self._lib.gsr_freeze_order(self._h, _fptr(self._sort_view.reshape(16)), self._stream()))
            self._frozen = True
        return

    def set_auto_sort(self, on: bool):
        self.auto_sort = bool(on)
        if self.auto_sort and getattr(self, "_frozen", False):
            _gsr.check(self._lib.gsr_thaw_order(self._h))
            self._frozen = False

    def gaussian_order(self, view=None, to_host=True):
        """The OpenGL backend's `gi` buffer: int32 (N, 1), ascending view-space z (far -> near), stable
        (_sort_gaussian, renderer_ogl.py:10-53), computed by the device radix sort."""
        view = _f32c(self._sort_view if view is None else view).reshape(16)
        n = len(self.gaussians) if self.gaussians is not None else self.stats().n_gaussians
        s = C.c_void_p(self._torch.cuda.current_stream(self.device).cuda_stream)
        if to_host:
            out = np.empty((n, 1), dtype=np.int32)
            _gsr.check(self._lib.gsr_sort_depth(self._h, _fptr(view), C.c_void_p(out.ctypes.data), 0, s))
            return out
        out = self._torch.empty((n, 1), dtype=self._torch.int32, device=self.device)
        _gsr.check(self._lib.gsr_sort_depth(self._h, _fptr(view), C.c_void_p(out.data_ptr()), 1, s))
        return out

    def set_scale_modifier(self, modifier):
        self.scale_modifier = float(modifier)

    def set_render_mod(self, mod: int):
        self.render_mod = int(mod)

    def set_is_depth_image(self, is_depth_image: int):
        # declared but unused by the reference's shaders (gau_frag.glsl:9,28-29)
        self.is_depth_image = int(is_depth_image)

    def set_render_reso(self, w, h):
        self.w, self.h = int(w), int(h)
        _gsr.check(self._lib.gsr_set_resolution(self._h, self.w, self.h))
        self._frame = None

    def update_camera_pose(self, camera: util.Camera, use_file=False, pose=None):
        # renderer_ogl.py:160-168: uniforms view_matrix and cam_pos (= camera.position, which the reference
        # overwrites with the pose's position when a pose file is used)
        view_mat = self._resolve_view(camera, use_file, pose)
        if use_file:
            camera.position = pose["camera_position"]
        self._view = _f32c(view_mat)
        self._campos = _f32c(np.asarray(camera.position).reshape(-1)[:3])

    def update_camera_intrin(self, camera: util.Camera):
        # renderer_ogl.py:170-173
        self._proj = _f32c(camera.get_project_matrix())
        self._hff = [float(v) for v in camera.get_htanfovxy_focal()]

    def _push_state(self):
        if self._view is None or self._proj is None:
            raise RuntimeError("update_camera_pose and update_camera_intrin must be called before draw")
        _gsr.check(self._lib.gsr_set_camera(self._h, _fptr(self._view.reshape(16)), _fptr(self._proj.reshape(16)),
                                            _fptr(self._campos), self._hff[0], self._hff[1], self._hff[2]))
        unorm = (self.render_mod != -1) if self.target_unorm is None else bool(self.target_unorm)
        flags = (_gsr.GSR_FLAG_CLAMP_UNORM if unorm else 0) | (_gsr.GSR_FLAG_KEYS64 if self.keys64 else 0) | \
                (_gsr.GSR_FLAG_DEBUG if self.debug else 0) | (_gsr.GSR_FLAG_OFFICIAL if self.official_compat else 0)
        _gsr.check(self._lib.gsr_set_params(self._h, self.scale_modifier, self.render_mod, flags))

    def _stream(self):
        return C.c_void_p(self._torch.cuda.current_stream(self.device).cuda_stream)

    def draw(self, out=None):
        """renderer_ogl.py:175-179 / renderer_cuda.py:205-228.  Asynchronous on the current torch stream."""
        self._push_state()
        if out is None:
            if self._frame is None:
                self._frame = self._torch.empty((self.h, self.w, 4), dtype=self._torch.float32, device=self.device)
            out = self._frame
        try:
            _gsr.check(self._lib.gsr_render(self._h, C.c_void_p(out.data_ptr()), self._stream()))
        except _gsr.GsrOverflow:
            # an EARLIER asynchronous frame did not fit the instance buffers (its image was incomplete); the
            # library has grown them, this frame goes in again
            _gsr.check(self._lib.gsr_render(self._h, C.c_void_p(out.data_ptr()), self._stream()))
        return out

    # ---- read-back (the viewer's glReadPixels) ---------------------------------------------------------
    def _pinned_buf(self, key, shape, dtype):
        buf = self._pinned.get(key)
        if buf is None or tuple(buf.shape) != tuple(shape):
            buf = self._torch.empty(shape, dtype=dtype, pin_memory=True)
            self._pinned[key] = buf
        return buf

    def read_pixels_rgb8(self):
        """draw + glReadPixels(GL_RGB, GL_UNSIGNED_BYTE) + vertical flip (main.py:852, 912): (H, W, 3) uint8
        NumPy view of a pinned buffer (valid until the next call)."""
        self._push_state()
        buf = self._pinned_buf("rgb8", (self.h, self.w, 3), self._torch.uint8)
        _gsr.check(self._lib.gsr_render_rgb8_host(self._h, C.c_void_p(buf.data_ptr()), self._stream()))
        return buf.numpy()

    def read_pixels_rgb8_async(self):
        """Pipelined capture loop (main.py:793-917 renders and reads back every pose): like read_pixels_rgb8,
        but returns at once; the copy to the host overlaps the next frame.  Returns the (H, W, 3) uint8 pinned
        buffer the frame WILL land in: it is complete after the next-but-one call or after sync_host(), and it
        is not written again before the call after that (three pinned buffers in rotation: the library waits for
        the copy of call k when call k + 2 is made, and call k + 3 is the first to reuse call k's buffer)."""
        self._push_state()
        k = self._async_k = (getattr(self, "_async_k", -1) + 1) % 3
        buf = self._pinned_buf(("rgb8a", k), (self.h, self.w, 3), self._torch.uint8)
        _gsr.check(self._lib.gsr_render_rgb8_host_async(self._h, C.c_void_p(buf.data_ptr()), self._stream()))
        return buf.numpy()

    def capture_stereo(self, pose_left, pose_right, want=("left", "disparity", "right"), share_order=True):
        """One pose of the viewer's stereo capture (main.py:844-917): left RGB8, disparity uint16 (render_mod -1
        at the left pose), right RGB8, packed on the device and copied to pinned host buffers while the next
        pose renders.  ``pose_*`` are the dicts ``colmap.load_camera_positions`` returns (keys camera_view,
        camera_position).  Returns {"left": (H,W,3) u8, "disparity": (H,W) u16, "right": (H,W,3) u8} views of
        the buffers the images WILL land in: complete after the next-but-one call or ``sync_host()``, reused by
        the call after that."""
        vl = _f32c(pose_left["camera_view"]).reshape(16)
        cl = _f32c(np.asarray(pose_left["camera_position"], dtype=np.float32)[:3])
        self._view, self._campos = vl.reshape(4, 4).copy(), cl.copy()  # update_camera_pose(poseLeft), main.py:848
        self._push_state()
        k = self._cap_k = (getattr(self, "_cap_k", -1) + 1) % 3  # three host buffers in rotation, see read_pixels_rgb8_async
        out = {}
        ptr = {"left": None, "disparity": None, "right": None}
        for name, shape, dt in (("left", (self.h, self.w, 3), self._torch.uint8),
                                ("disparity", (self.h, self.w), self._torch.int16),
                                ("right", (self.h, self.w, 3), self._torch.uint8)):
            if name in want:
                buf = self._pinned_buf(("cap", name, k), shape, dt)
                ptr[name] = C.c_void_p(buf.data_ptr())
                a = buf.numpy()
                out[name] = a.view(np.uint16) if name == "disparity" else a
        vr = _f32c(pose_right["camera_view"]).reshape(16) if pose_right is not None else vl
        cr = _f32c(np.asarray(pose_right["camera_position"], dtype=np.float32)[:3]) if pose_right is not None else cl
        _gsr.check(self._lib.gsr_capture_stereo_async(self._h, _fptr(vl), _fptr(cl), _fptr(vr), _fptr(cr), ptr["left"],
                                                      ptr["disparity"], ptr["right"], 1 if share_order else 0,
                                                      self._stream()))
        return out

    def sync_host(self):
        """Waits for every pipelined read-back."""
        _gsr.check(self._lib.gsr_host_sync(self._h))

    def read_pixels_r32f(self):
        """draw + glReadPixels(GL_RED, GL_FLOAT) (main.py:870): (H, W) float32."""
        self._push_state()
        buf = self._pinned_buf("r32f", (self.h, self.w), self._torch.float32)
        _gsr.check(self._lib.gsr_render_r32f_host(self._h, C.c_void_p(buf.data_ptr()), self._stream()))
        return buf.numpy()

    def read_pixels_rgba32f(self):
        self._push_state()
        buf = self._pinned_buf("rgba", (self.h, self.w, 4), self._torch.float32)
        _gsr.check(self._lib.gsr_render_host(self._h, C.c_void_p(buf.data_ptr()), self._stream()))
        return buf.numpy()

    # ---- pose batches (render_all.py / camera_data.csv workload) -------------------------------------------
    def draw_batch(self, views, campos, out=None):
        """views (n,4,4), campos (n,3): renders every pose with the current intrinsics/params.  out: optional
        CUDA tensor (n,H,W,4); without it frames go to a scratch buffer (throughput measurement)."""
        self._push_state()
        views = _f32c(views).reshape(-1, 16)
        campos = _f32c(campos).reshape(-1, 3)
        n = len(views)
        optr = C.c_void_p(out.data_ptr()) if out is not None else None
        try:
            _gsr.check(self._lib.gsr_render_batch(self._h, _fptr(views), _fptr(campos), n, optr, self._stream()))
        except _gsr.GsrOverflow:  # see draw()
            _gsr.check(self._lib.gsr_render_batch(self._h, _fptr(views), _fptr(campos), n, optr, self._stream()))
        return out

    # ---- introspection --------------------------------------------------------------------------------
    def draw_timed(self, out=None):
        self._push_state()
        if out is None:
            if self._frame is None:
                self._frame = self._torch.empty((self.h, self.w, 4), dtype=self._torch.float32, device=self.device)
            out = self._frame
        _gsr.check(self._lib.gsr_render_timed(self._h, C.c_void_p(out.data_ptr()), self._stream()))
        return out, self.stats()

    def stats(self) -> _gsr.Stats:
        st = _gsr.Stats()
        _gsr.check(self._lib.gsr_get_stats(self._h, C.byref(st)))
        return st

    def set_capacity(self, n):
        _gsr.check(self._lib.gsr_set_capacity(self._h, int(n)))

    def launch_count(self) -> int:
        return int(self._lib.gsr_launch_count(self._h))

    def measure_peaks(self):
        """(fp32 TFLOP/s, 1e12 ex2/s, copy GB/s) from micro-benchmarks on this device."""
        a, b, c = C.c_float(), C.c_float(), C.c_float()
        _gsr.check(self._lib.gsr_measure_peaks(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def tune(self, key: int, value: int):
        """Kernel-variant knob for measurement runs (gsr_tune)."""
        _gsr.check(self._lib.gsr_tune(self._h, int(key), int(value)))

    def count_pairs(self):
        """(visited, contributed) (pixel, splat) pairs of the last frame's blend."""
        a, b = C.c_ulonglong(), C.c_ulonglong()
        _gsr.check(self._lib.gsr_count_pairs(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def debug_fetch(self, which, dtype, cols=None):
        nbytes = C.c_longlong()
        _gsr.check(self._lib.gsr_debug_fetch(self._h, which, None, 0, C.byref(nbytes)))
        out = np.empty(max(nbytes.value, 0) // np.dtype(dtype).itemsize, dtype=dtype)
        if nbytes.value:
            _gsr.check(self._lib.gsr_debug_fetch(self._h, which, C.c_void_p(out.ctypes.data), nbytes.value,
                                                 C.byref(nbytes)))
        return out.reshape(-1, cols) if cols else out

    def sort_pairs(self, keys, vals, begin_bit=0, end_bit=None):
        """The radix sort primitive on CUDA tensors (uint32 stored as int32 / uint64 stored as int64, values
        int32): sorts in place, stable, on key bits [begin_bit, end_bit)."""
        torch = self._torch
        kb = keys.element_size() * 8
        if end_bit is None:
            end_bit = kb
        kt, vt = torch.empty_like(keys), torch.empty_like(vals)
        _gsr.check(self._lib.gsr_sort_pairs(self._h, kb, C.c_void_p(keys.data_ptr()), C.c_void_p(kt.data_ptr()),
                                            C.c_void_p(vals.data_ptr()), C.c_void_p(vt.data_ptr()), keys.numel(),
                                            begin_bit, end_bit, self._stream()))
        torch.cuda.current_stream(self.device).synchronize()
        return keys, vals
