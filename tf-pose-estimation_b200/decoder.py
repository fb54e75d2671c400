"""Host side of seam 1: network output -> humans, over the C-ABI.

``PoseDecoder`` owns one ``tfpd_ctx`` (one GPU, one thread at a time).  It replaces what
``TfPoseEstimator.inference`` does after the network forward (tf_pose/estimator.py:536-551):
the resize_area / smoother / NMS graph ops and ``PoseEstimator.estimate_paf``.

PyTorch tensors are used only as device buffers (``data_ptr()``); NumPy arrays go through
``tfpd_decode_host`` which does the host<->device copies itself.
"""
import ctypes as C

import numpy as np

from . import _lib
from .humans import BodyPart, Human


_OBJBUILD = False


def _objbuild():
    """The CPython extension that builds Human/BodyPart objects for a whole batch (built by csrc/Makefile), or None
    when it is not built: object construction is host-side Python either way, the per-frame Python loop
    (``humans_from_table``) gives the same objects."""
    global _OBJBUILD
    if _OBJBUILD is False:
        try:
            from . import _objbuild as mod
            _OBJBUILD = mod if mod.HUMAN_RECORD_BYTES == _lib.HUMAN_DTYPE.itemsize else None
        except ImportError:
            _OBJBUILD = None
    return _OBJBUILD


def upsample_size_for(target_w, target_h, ratio=4.0):
    """[H, W] exactly as inference() computes it (estimator.py:522-523) for
    ``resize_to_default=True``: ``int(target / 8 * ratio)``."""
    return [int(target_h / 8 * ratio), int(target_w / 8 * ratio)]


def humans_from_table(table_row, n_humans, H, W):
    """Build ``Human``/``BodyPart`` objects from one frame's tfpd_human records the way
    PoseEstimator.estimate_paf does from the getters (estimator.py:277-297).

    The records are converted with one ``tolist()`` per field (plain Python ints / floats: a float32 widened to a
    Python float is the same double the reference gets from its swig getter); indexing NumPy scalars one by one
    cost several times the object construction itself."""
    n = int(n_humans)
    if n <= 0:
        return []
    recs = table_row[:n]
    cids, xs, ys = recs["cid"].tolist(), recs["x"].tolist(), recs["y"].tolist()
    pss, scores = recs["part_score"].tolist(), recs["score"].tolist()
    humans = []
    for human_id in range(n):
        cid, x, y, ps = cids[human_id], xs[human_id], ys[human_id], pss[human_id]
        human = Human([])
        parts = human.body_parts
        for part_idx in range(18):
            if cid[part_idx] < 0:
                continue
            parts[part_idx] = BodyPart('%d-%d' % (human_id, part_idx), part_idx,
                                       float(x[part_idx]) / W, float(y[part_idx]) / H, ps[part_idx])
        if parts:
            human.score = scores[human_id]
            humans.append(human)
    return humans


class DecodeResult(object):
    """Per-frame tables of one decode call, already on the host.  ``Human`` objects are built
    lazily: the tables are the transport format (also what the multi-GPU gather moves)."""

    def __init__(self, info, table, H, W):
        self.info = info        # structured [N]: n_humans, n_peaks, n_cands, flags, near_ties, n_rows
        self.table = table      # structured [N, max_humans] (tfpd_human)
        self.H, self.W = H, W

    def __len__(self):
        return len(self.info)

    def humans(self, frame=0):
        return humans_from_table(self.table[frame], self.info["n_humans"][frame], self.H, self.W)

    def all_humans(self):
        """``list[list[Human]]`` of the whole batch, built in one call by the ``_objbuild`` extension (csrc/objbuild.c:
        same objects as ``humans(i)`` frame by frame, an order of magnitude faster)."""
        ob = _objbuild()
        if ob is None:
            return [self.humans(i) for i in range(len(self.info))]
        table = np.ascontiguousarray(self.table)
        counts = np.ascontiguousarray(self.info["n_humans"], np.int32)
        return ob.build_humans(BodyPart, Human, table, counts, int(table.shape[1]), int(self.H), int(self.W))

    def compact(self):
        """(counts[N] int32, records[sum(counts)] tfpd_human) -- what the gather ships."""
        counts = self.info["n_humans"].astype(np.int32)
        recs = np.concatenate([self.table[i, :counts[i]] for i in range(len(counts))]) if len(counts) else \
            np.empty(0, _lib.HUMAN_DTYPE)
        return counts, recs


class PoseDecoder(object):
    def __init__(self, max_batch=1, max_upsample=(368, 656), device=0, mode="asbuilt", **overrides):
        """max_upsample = (H, W) of the largest upsampled map.  ``overrides`` are further
        ``tfpd_config`` fields (thresholds, blur_ksize/blur_sigma, max_peaks, max_cands, ...)."""
        L = _lib.lib()
        cfg = _lib.default_config(device=int(device), max_batch=int(max_batch), max_h=int(max_upsample[0]),
                                  max_w=int(max_upsample[1]),
                                  mode={"asbuilt": _lib.MODE_ASBUILT, "upstream": _lib.MODE_UPSTREAM}[mode],
                                  **overrides)
        self._L = L
        self._ctx = C.c_void_p()
        rc = L.tfpd_create(C.byref(cfg), C.byref(self._ctx))
        if rc != _lib.OK:
            msg = L.tfpd_last_error(None).decode()
            self._ctx = None
            raise _lib.DecodeError(rc, msg)
        self.cfg = cfg
        self.max_batch = int(max_batch)
        self.max_humans = int(cfg.max_humans)
        self.device = int(device)
        self._shape = None      # (N, H, W) of the last call
        self._keep = None       # buffers borrowed by the in-flight call
        self._host_maps = None
        self._fetch_bufs = None # (info, table) handed out by fetch(reuse=True)

    # -- lifetime ------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_ctx", None):
            self._L.tfpd_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, rc):
        if rc != _lib.OK:
            msg = self._L.tfpd_last_error(self._ctx).decode()
            raise (_lib.CapacityError if rc == _lib.ERR_CAPACITY else _lib.DecodeError)(rc, msg)

    # -- seam 1 --------------------------------------------------------------------------
    def decode_async(self, net_out, upsample_size, dense=None, stream=None):
        """Enqueue the decode of a CUDA tensor ``net_out[N,h,w,57]`` (float32, contiguous) on
        ``stream`` (default: torch's current stream).  ``dense`` = (heat_up, paf_up, peaks) CUDA
        tensors to fill ([N,H,W,19], [N,H,W,38], [N,H,W,19]) or None for minimal mode."""
        import torch
        if not (isinstance(net_out, torch.Tensor) and net_out.is_cuda):
            raise TypeError("decode_async wants a CUDA torch.Tensor; use decode_host for NumPy arrays")
        if net_out.dtype != torch.float32 or not net_out.is_contiguous() or net_out.dim() != 4:
            raise ValueError("net_out must be a contiguous float32 [N,h,w,57] tensor")
        N, h, w, c = net_out.shape
        if c != 57:
            raise ValueError("net_out must have 57 channels (19 heat + 38 PAF), got %d" % c)
        if net_out.device.index != self.device:
            raise ValueError("net_out is on cuda:%s, decoder on cuda:%d" % (net_out.device.index, self.device))
        H, W = int(upsample_size[0]), int(upsample_size[1])
        dptr = None
        if dense is not None:
            hu, pu, pk = dense
            for t, ch in ((hu, 19), (pu, 38), (pk, 19)):
                if tuple(t.shape) != (N, H, W, ch) or t.dtype != torch.float32 or not t.is_contiguous() \
                        or not t.is_cuda:
                    raise ValueError("dense outputs must be contiguous float32 CUDA [N,H,W,19|38|19]")
            dptr = _lib.Dense(hu.data_ptr(), pu.data_ptr(), pk.data_ptr())
        if stream is None:
            stream = torch.cuda.current_stream(net_out.device)
        sptr = C.c_void_p(stream.cuda_stream)
        self._keep = (net_out, dense)
        self._check(self._L.tfpd_decode_device(self._ctx, net_out.data_ptr(), N, h, w, H, W, sptr,
                                               C.byref(dptr) if dptr is not None else None))
        self._shape = (N, H, W)

    def fetch(self, reuse=False):
        """Wait for the enqueued decode and return its tables (``DecodeResult``).

        ``reuse=True`` hands out views of two buffers the decoder keeps (valid until the next ``fetch(reuse=True)``
        on this decoder) instead of allocating ~185 B x max_humans per frame on every call -- for steady-state loops
        that consume a result before asking for the next one."""
        if self._shape is None:
            raise _lib.DecodeError(_lib.ERR_STATE, "fetch() before decode")
        N, H, W = self._shape
        if reuse:
            if self._fetch_bufs is None:
                self._fetch_bufs = (np.zeros(self.max_batch, _lib.INFO_DTYPE),
                                    np.zeros((self.max_batch, self.max_humans), _lib.HUMAN_DTYPE))
            info, table = self._fetch_bufs[0][:N], self._fetch_bufs[1][:N]
        else:
            info = np.empty(N, _lib.INFO_DTYPE)
            table = np.empty((N, self.max_humans), _lib.HUMAN_DTYPE)
        rc = self._L.tfpd_fetch_humans(self._ctx, table.ctypes.data, info.ctypes.data)
        self._keep = None
        if rc == _lib.ERR_CAPACITY:
            # one frame overflowing a table must not cost the caller the other frames: the tables are filled,
            # the exception carries them and ``info['flags']`` names the frames and tables that overflowed
            err = _lib.CapacityError(rc, self._L.tfpd_last_error(self._ctx).decode())
            err.result = DecodeResult(info, table, H, W)
            raise err
        self._check(rc)
        return DecodeResult(info, table, H, W)

    def decode(self, net_out, upsample_size, dense=None, stream=None):
        """CUDA tensor in -> DecodeResult (synchronous)."""
        self.decode_async(net_out, upsample_size, dense=dense, stream=stream)
        return self.fetch()

    def decode_host_async(self, net_out, upsample_size, want_dense=False, out=None):
        """Enqueue the decode of a NumPy ``net_out[N,h,w,57]`` (or [h,w,57]) in host memory and return at once;
        ``fetch()`` waits and returns the DecodeResult, ``host_maps()`` the dense arrays.  The arrays are kept
        alive (and must not be written) until ``fetch()``.  ``out`` = preallocated (heat_up, paf_up, peaks)
        float32 C-contiguous arrays to fill (page-locked memory makes the device->host copies asynchronous
        and several times faster)."""
        x = np.ascontiguousarray(net_out, np.float32)
        if x.ndim == 3:
            x = x[None]
        if x.ndim != 4 or x.shape[3] != 57:
            raise ValueError("net_out must be [N,h,w,57]")
        N, h, w, _ = x.shape
        H, W = int(upsample_size[0]), int(upsample_size[1])
        hu = pu = pk = None
        if out is not None:
            want_dense = True
            hu, pu, pk = out
            for a, ch in ((hu, 19), (pu, 38), (pk, 19)):
                if not (isinstance(a, np.ndarray) and a.dtype == np.float32 and a.flags.c_contiguous
                        and a.shape == (N, H, W, ch)):
                    raise ValueError("out arrays must be C-contiguous float32 [N,H,W,19|38|19]")
        elif want_dense:
            hu = np.empty((N, H, W, 19), np.float32)
            pu = np.empty((N, H, W, 38), np.float32)
            pk = np.empty((N, H, W, 19), np.float32)
        self._shape = None
        self._keep = (x, hu, pu, pk)
        self._check(self._L.tfpd_decode_host_async(self._ctx, x.ctypes.data, N, h, w, H, W,
                                                   hu.ctypes.data if want_dense else None,
                                                   pu.ctypes.data if want_dense else None,
                                                   pk.ctypes.data if want_dense else None))
        self._shape = (N, H, W)
        self._host_maps = (hu, pu, pk) if want_dense else None

    def host_maps(self):
        """(heat_up, paf_up, peaks) of the last ``decode_host_async(..., want_dense=True)``; valid after ``fetch()``."""
        return self._host_maps

    def decode_host(self, net_out, upsample_size, want_dense=False, out=None):
        """NumPy ``net_out[N,h,w,57]`` (or [h,w,57]) in host memory -> DecodeResult; with
        ``want_dense`` also returns (heat_up, paf_up, peaks) NumPy arrays, i.e. everything one
        ``sess.run`` + ``estimate_paf`` of the reference yields (estimator.py:536-551).  Synchronous."""
        self.decode_host_async(net_out, upsample_size, want_dense=want_dense, out=out)
        res = self.fetch()
        return (res, self._host_maps) if self._host_maps is not None else res

    # -- draw_humans (estimator.py:391-415) on a headless buffer ------------------------------
    @staticmethod
    def _centers_table(humans_per_image, image_h, image_w):
        """int32 [N, max_h, 18, 2] of (x, y) = int(part.x * w + 0.5), int(part.y * h + 0.5) (estimator.py:404) and
        counts[N]; an absent part has x == INT32_MIN."""
        n = len(humans_per_image)
        max_h = max(1, max((len(hs) for hs in humans_per_image), default=1))
        centers = np.full((n, max_h, 18, 2), np.iinfo(np.int32).min, np.int32)
        counts = np.zeros(n, np.int32)
        for i, hs in enumerate(humans_per_image):
            counts[i] = len(hs)
            for j, human in enumerate(hs):
                for k, bp in human.body_parts.items():
                    if 0 <= k < 18:
                        centers[i, j, k, 0] = int(bp.x * image_w + 0.5)
                        centers[i, j, k, 1] = int(bp.y * image_h + 0.5)
        return centers, counts, max_h

    def draw_humans(self, images, humans_per_image, imgcopy=False):
        """``images``: uint8 NumPy [N,H,W,3] / [H,W,3] (drawn in place unless ``imgcopy``) or a CUDA uint8 tensor
        [N,H,W,3]; ``humans_per_image``: list (per image) of lists of ``Human`` (a flat list for one image).
        Sets the pixels ``TfPoseEstimator.draw_humans`` would, bit for bit.  Returns the drawn array/tensor."""
        is_numpy = isinstance(images, np.ndarray)
        single = (images.ndim == 3)
        if single:
            humans_per_image = [humans_per_image]
        if is_numpy:
            if images.dtype != np.uint8 or images.shape[-1] != 3:
                raise ValueError("images must be uint8 [N,H,W,3] or [H,W,3]")
            src = images[None] if single else images
            buf = np.array(src, dtype=np.uint8, order='C', copy=True) if imgcopy else np.ascontiguousarray(src)
            n, h, w = buf.shape[:3]
            ptr, on_device, stream = buf.ctypes.data, 0, None
        else:
            import torch
            if images.dtype != torch.uint8 or images.shape[-1] != 3 or not images.is_cuda or not images.is_contiguous():
                raise ValueError("CUDA images must be contiguous uint8 [N,H,W,3]")
            buf = images.clone() if imgcopy else images
            if single:
                buf = buf.unsqueeze(0)
            n, h, w = buf.shape[:3]
            ptr, on_device = buf.data_ptr(), 1
            stream = C.c_void_p(torch.cuda.current_stream(images.device).cuda_stream)
        if len(humans_per_image) != n:
            raise ValueError("%d images but humans for %d" % (n, len(humans_per_image)))
        centers, counts, max_h = self._centers_table(humans_per_image, h, w)
        self._check(self._L.tfpd_draw_humans(self._ctx, ptr, on_device, n, h, w, centers.ctypes.data,
                                             counts.ctypes.data, max_h, stream))
        if is_numpy and not imgcopy:
            if not np.shares_memory(buf, images):
                images[...] = buf[0] if single else buf      # non-contiguous input: write the result back
            return images
        return buf[0] if single else buf

    # -- image pre-processing in front of the network (estimator.py:383-388, :417-424) ---------
    def resize_cubic(self, images, target_size=None, fx=None, fy=None):
        """``cv2.resize(img, target_size, interpolation=cv2.INTER_CUBIC)`` -- or, with ``target_size=None``,
        ``cv2.resize(img, dsize=None, fx=fx, fy=fy, ...)`` -- for uint8 images: NumPy [H,W], [H,W,C] or
        [N,H,W,C] (host in, host out) or a contiguous CUDA uint8 tensor [N,H,W,C] (device in, device out,
        asynchronous on the current stream).  ``target_size = (w, h)`` as cv2 and the reference ctor take it.
        The bytes are those of OpenCV's own bicubic code (cv2 with IPP off)."""
        if target_size is None:
            if fx is None or fy is None:
                raise ValueError("give target_size or fx and fy")
            shp = images.shape
            sh_, sw_ = (shp[0], shp[1]) if len(shp) < 4 else (shp[1], shp[2])
            dw, dh = int(round(sw_ * float(fx))), int(round(sh_ * float(fy)))     # cv2: round half to even
            scaled = True
        else:
            dw, dh = int(target_size[0]), int(target_size[1])
            scaled = False

        def call(src_ptr, on_device, n, sh, sw, cn, dst_ptr, stream):
            if scaled:
                return self._L.tfpd_resize_cubic_u8_fxfy(self._ctx, src_ptr, on_device, n, sh, sw, cn, float(fx), float(fy),
                                                         dst_ptr, dh, dw, stream)
            return self._L.tfpd_resize_cubic_u8(self._ctx, src_ptr, on_device, n, sh, sw, cn, dst_ptr, dh, dw, stream)

        if isinstance(images, np.ndarray):
            if images.dtype != np.uint8 or images.ndim not in (2, 3, 4):
                raise ValueError("images must be uint8 [H,W], [H,W,C] or [N,H,W,C]")
            src = np.ascontiguousarray(images)
            shp = src.shape
            if src.ndim == 2:
                src = src[None, :, :, None]
            elif src.ndim == 3:
                src = src[None]
            n, sh, sw, cn = src.shape
            dst = np.empty((n, dh, dw, cn), np.uint8)
            self._check(call(src.ctypes.data, 0, n, sh, sw, cn, dst.ctypes.data, None))
            return dst[0, :, :, 0] if len(shp) == 2 else (dst[0] if len(shp) == 3 else dst)
        import torch
        if images.dtype != torch.uint8 or images.dim() != 4 or not images.is_cuda or not images.is_contiguous():
            raise ValueError("CUDA images must be contiguous uint8 [N,H,W,C]")
        n, sh, sw, cn = (int(v) for v in images.shape)
        dst = torch.empty((n, dh, dw, cn), dtype=torch.uint8, device=images.device)
        stream = C.c_void_p(torch.cuda.current_stream(images.device).cuda_stream)
        self._check(call(images.data_ptr(), 1, n, sh, sw, cn, dst.data_ptr(), stream))
        return dst

    def quantize_img(self, npimg):
        """``TfPoseEstimator._quantize_img`` (estimator.py:383-388) for float32 input: NumPy array (host) or a
        contiguous CUDA float32 tensor (device, asynchronous on the current stream); same shape, uint8."""
        if isinstance(npimg, np.ndarray):
            if npimg.dtype != np.float32:
                raise ValueError("quantize_img takes float32 images")
            src = np.ascontiguousarray(npimg)
            dst = np.empty(src.shape, np.uint8)
            self._check(self._L.tfpd_quantize_img(self._ctx, src.ctypes.data, 0, src.size, dst.ctypes.data, None))
            return dst
        import torch
        if npimg.dtype != torch.float32 or not npimg.is_cuda or not npimg.is_contiguous():
            raise ValueError("CUDA images must be contiguous float32")
        dst = torch.empty(npimg.shape, dtype=torch.uint8, device=npimg.device)
        stream = C.c_void_p(torch.cuda.current_stream(npimg.device).cuda_stream)
        self._check(self._L.tfpd_quantize_img(self._ctx, npimg.data_ptr(), 1, npimg.numel(), dst.data_ptr(), stream))
        return dst

    def draw_decoded(self, images, stream=None):
        """Draw the humans of the last decode on this decoder into the CUDA uint8 tensor ``images[N,H,W,3]``
        straight from the device tables (nothing crosses PCIe).  Asynchronous on ``stream``."""
        import torch
        if not (isinstance(images, torch.Tensor) and images.is_cuda and images.dtype == torch.uint8
                and images.dim() == 4 and images.shape[3] == 3 and images.is_contiguous()):
            raise ValueError("images must be a contiguous CUDA uint8 [N,H,W,3] tensor")
        if stream is None:
            stream = torch.cuda.current_stream(images.device)
        self._check(self._L.tfpd_draw_decoded(self._ctx, images.data_ptr(), int(images.shape[1]), int(images.shape[2]),
                                              C.c_void_p(stream.cuda_stream)))
        return images

    # -- introspection -------------------------------------------------------------------
    def peaks(self, frame=0):
        """dict(part, x, y, score) of int32/float32 arrays: the peak table of a frame."""
        cap = int(self.cfg.max_peaks) * 2
        part, x, y = (np.empty(cap, np.int32) for _ in range(3))
        score = np.empty(cap, np.float32)
        n = C.c_int(0)
        self._check(self._L.tfpd_fetch_peaks(self._ctx, frame, cap, part.ctypes.data, x.ctypes.data,
                                             y.ctypes.data, score.ctypes.data, C.byref(n)))
        k = min(n.value, cap)
        return {"part": part[:k].copy(), "x": x[:k].copy(), "y": y[:k].copy(), "score": score[:k].copy()}

    def connections(self, frame=0):
        """dict(limb, cid1, cid2, score): accepted connections, limb-major, acceptance order."""
        cap = int(self.cfg.max_cands) * 4
        limb, c1, c2 = (np.empty(cap, np.int32) for _ in range(3))
        score = np.empty(cap, np.float32)
        n = C.c_int(0)
        self._check(self._L.tfpd_fetch_connections(self._ctx, frame, cap, limb.ctypes.data, c1.ctypes.data,
                                                   c2.ctypes.data, score.ctypes.data, C.byref(n)))
        k = min(n.value, cap)
        return {"limb": limb[:k].copy(), "cid1": c1[:k].copy(), "cid2": c2[:k].copy(), "score": score[:k].copy()}

    def set_option(self, name, value):
        """Tuning options of include/tfpose_decode.h (``k1_ctas``, ``k1_turnstile``, ``k1_early_out``, ``k1_min_band``, ``tail_priority``, ``graph``, ``fuse_sort``, ``k2_prefilter``, ``std_sort_ties``, ``early_fetch``,
        ``host_chunks``, ``k1_l2``, ``pow2_path``)."""
        self._check(self._L.tfpd_set_option(self._ctx, name.encode(), int(value)))

    def set_profiling(self, on=True):
        self._check(self._L.tfpd_set_profiling(self._ctx, 1 if on else 0))

    def stage_ms(self):
        """[K1 maps, peak sort, K2 limb scoring, K3 assembly] device ms of the last call."""
        out = (C.c_float * 4)()
        self._check(self._L.tfpd_last_stage_ms(self._ctx, C.byref(out)))
        return [float(v) for v in out]

    def k1_cta_times(self, n_ctas=148):
        """[n_ctas,4] uint64: CTA start / blur-role end / copy-role end (ns), SM id of the last K1 launch
        (needs TFPD_K1_TIMES=1 in the environment when the decoder is created)."""
        out = np.zeros((n_ctas, 4), np.uint64)
        self._check(self._L.tfpd_debug_k1_times(self._ctx, out.ctypes.data, n_ctas))
        return out

    def launch_count(self):
        return int(self._L.tfpd_last_launch_count(self._ctx))

    # -- seam 2 (used by the pafprocess module) ------------------------------------------
    def process_paf(self, peaks, heat_mat, paf_mat):
        arrs = []
        for a in (peaks, heat_mat, paf_mat):
            a = np.ascontiguousarray(a, np.float32)   # swig IN_ARRAY3 semantics (pafprocess_wrap.cpp:4935-4968)
            if a.ndim != 3:
                raise ValueError("Array must have 3 dimensions.  Given array has %d dimensions" % a.ndim)
            arrs.append(a)
        p, h, f = arrs
        self._shape = None
        rc = self._L.tfpd_process_paf(self._ctx, p.shape[0], p.shape[1], p.shape[2], p.ctypes.data,
                                      h.shape[0], h.shape[1], h.shape[2], h.ctypes.data,
                                      f.shape[0], f.shape[1], f.shape[2], f.ctypes.data)
        self._check(rc)
        self._shape = (1, p.shape[0], p.shape[1])
        return 0
