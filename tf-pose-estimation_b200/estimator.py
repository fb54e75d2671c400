"""The reference-facing operator interface of the decode path.

* ``PoseEstimator.estimate_paf(peaks, heat_mat, paf_mat)`` -- same name, arguments and result as
  tf_pose/estimator.py:272-299, running on the GPU through the ``pafprocess`` drop-in (seam 2).
* ``TfPosePostProcessor`` -- the part of ``TfPoseEstimator`` after the network forward
  (tf_pose/estimator.py:321-337 graph ops + :536-551): give it the tensor the frozen graph
  calls ``Openpose/concat_stage7`` and get ``inference()``'s return value, with the same
  ``heatMat`` / ``pafMat`` side effects (:545-546).  Seam 1.

The network forward itself (TF1 frozen graphs) is outside this repository's scope.
"""
import numpy as np

from . import pafprocess
from .decoder import PoseDecoder, upsample_size_for
from .humans import BodyPart, Human


def _sliding_windows(width, height, win_w, win_h, overlap):
    """``slidingwindow.generate(data, DimOrder.HeightWidthChannel, win_w, win_h, overlap)`` (estimator.py:439,467;
    the reference vendors the package: tf_pose/slidingwindow/SlidingWindow.py:87-146) as [(x, y, w, h)]: windows
    clipped to the data, step = size - floor(size*overlap), a last window flush with the far edge, x outer / y inner."""
    ww, wh = min(win_w, width), min(win_h, height)
    step_x, step_y = ww - int(ww * overlap), wh - int(wh * overlap)
    xs, ys = list(range(0, width - ww + 1, step_x)), list(range(0, height - wh + 1, step_y))
    if not xs or xs[-1] != width - ww:
        xs.append(width - ww)
    if not ys or ys[-1] != height - wh:
        ys.append(height - wh)
    return [(x, y, ww, wh) for x in xs for y in ys]


class PoseEstimator(object):
    def __init__(self):
        pass

    @staticmethod
    def estimate_paf(peaks, heat_mat, paf_mat):
        pafprocess.process_paf(peaks, heat_mat, paf_mat)
        humans = []
        H, W = heat_mat.shape[0], heat_mat.shape[1]
        for human_id in range(pafprocess.get_num_humans()):
            human = Human([])
            is_added = False
            for part_idx in range(18):
                c_idx = int(pafprocess.get_part_cid(human_id, part_idx))
                if c_idx < 0:
                    continue
                is_added = True
                human.body_parts[part_idx] = BodyPart(
                    '%d-%d' % (human_id, part_idx), part_idx,
                    float(pafprocess.get_part_x(c_idx)) / W,
                    float(pafprocess.get_part_y(c_idx)) / H,
                    pafprocess.get_part_score(c_idx))
            if is_added:
                human.score = pafprocess.get_score(human_id)
                humans.append(human)
        return humans


class TfPosePostProcessor(object):
    """``TfPoseEstimator`` minus the TensorFlow session.

    target_size = (w, h) of the network input like the reference ctor (estimator.py:305-306).
    ``inference_from_output(net_out, upsample_size=4.0)`` mirrors
    ``inference(npimg, resize_to_default=True, upsample_size=...)`` from the point where the
    session returns: it sets ``self.heatMat`` / ``self.pafMat`` (when ``keep_maps``) and returns
    ``list[Human]``.
    """

    def __init__(self, target_size=(432, 368), device=0, mode="asbuilt", keep_maps=True, max_ratio=8.0,
                 reuse_map_buffers=False, **overrides):
        """``reuse_map_buffers``: ``heatMat`` / ``pafMat`` / ``peaks`` of NumPy-input calls live in page-locked
        buffers that the next call overwrites (several times faster than fresh pageable arrays; the reference
        allocates new arrays per call, so this is opt-in)."""
        self.target_size = target_size
        self.keep_maps = keep_maps
        self.reuse_map_buffers = reuse_map_buffers
        self._pinned = {}
        H, W = upsample_size_for(target_size[0], target_size[1], max_ratio)
        # the reference's process_paf has no table limits (std::vector everywhere): same generous capacities as the
        # pafprocess drop-in so that a crowded frame decodes instead of raising CapacityError
        for k, v in (("max_peaks", 4096), ("max_cands", 8192), ("max_rows", 1024), ("max_humans", 256)):
            overrides.setdefault(k, v)
        self._decoder = PoseDecoder(max_batch=1, max_upsample=(H, W), device=device, mode=mode, **overrides)
        self.heatMat = self.pafMat = None
        self.peaks = None
        self.last_info = None

    def inference_from_output(self, net_out, upsample_size=1.0, input_hw=None):
        """net_out: [1,h,w,57] or [h,w,57] NumPy array or CUDA tensor.  ``input_hw`` plays the
        role of ``npimg.shape[:2]`` when ``resize_to_default`` is False (estimator.py:524-525)."""
        if net_out is None:
            raise Exception('The network output is not valid.')
        if input_hw is None:
            size = upsample_size_for(self.target_size[0], self.target_size[1], upsample_size)
        else:
            size = [int(input_hw[0] / 8 * upsample_size), int(input_hw[1] / 8 * upsample_size)]
        is_numpy = isinstance(net_out, np.ndarray)
        if is_numpy:
            bufs = None
            if self.keep_maps and self.reuse_map_buffers:
                n = 1 if net_out.ndim == 3 else net_out.shape[0]
                key = (n, size[0], size[1])
                if key not in self._pinned:
                    import torch
                    self._pinned = {key: tuple(torch.empty((n, size[0], size[1], c), dtype=torch.float32).pin_memory().numpy()
                                               for c in (19, 38, 19))}
                bufs = self._pinned[key]
            out = self._decoder.decode_host(net_out, size, want_dense=self.keep_maps, out=bufs)
            if self.keep_maps:
                res, (hu, pu, pk) = out
                self.peaks, self.heatMat, self.pafMat = pk[0], hu[0], pu[0]
            else:
                res = out
        else:
            import torch
            x = net_out if net_out.dim() == 4 else net_out.unsqueeze(0)
            dense = None
            if self.keep_maps:
                n = x.shape[0]
                dense = (torch.empty((n, size[0], size[1], 19), device=x.device),
                         torch.empty((n, size[0], size[1], 38), device=x.device),
                         torch.empty((n, size[0], size[1], 19), device=x.device))
            res = self._decoder.decode(x.contiguous(), size, dense=dense)
            if self.keep_maps:
                self.heatMat, self.pafMat, self.peaks = dense[0][0], dense[1][0], dense[2][0]
        self.last_info = res.info[0]
        return res.humans(0)

    def draw_humans(self, npimg, humans, imgcopy=False):
        """``TfPoseEstimator.draw_humans(npimg, humans, imgcopy=False)`` (estimator.py:391-415), rasterised by the
        CUDA kernels on a headless buffer: same pixels as the cv2.circle / cv2.line calls of the reference."""
        return self._decoder.draw_humans(npimg, humans, imgcopy=imgcopy)

    def _quantize_img(self, npimg):
        """``TfPoseEstimator._quantize_img`` (estimator.py:383-388; a staticmethod there, it needs the GPU
        context here)."""
        return self._decoder.quantize_img(npimg)

    def _get_scaled_img(self, npimg, scale=None):
        """``TfPoseEstimator._get_scaled_img(npimg, scale)`` (estimator.py:417-500) with the bicubic resizes on the
        GPU: ``None`` (the form ``inference`` uses, :534-535) resizes to ``target_size`` unless the image already has
        it; a float or ``(scale, step)`` resizes by ``base_scale`` and cuts overlapping ``target_size`` windows
        (overlap 0.2 / ``step``); ``(want_x, want_y, scale_ratio)`` resizes and crops one window around a point.
        Returns ``(rois, ratios)`` like the reference.  The window enumeration follows the reference's vendored
        ``slidingwindow.generate`` (pinned by tests/golden/scaled_img.json)."""
        tw, th = self.target_size
        img_h, img_w = npimg.shape[:2]
        if scale is None:
            if npimg.shape[:2] != (th, tw):
                npimg = self._decoder.resize_cubic(npimg, self.target_size)
            return [npimg], [(0.0, 0.0, 1.0, 1.0)]
        base = max(tw / float(img_h), th / float(img_w))          # get_base_scale as written (:418)
        if isinstance(scale, float) or (isinstance(scale, tuple) and len(scale) == 2):
            factor, overlap = (scale, 0.2) if isinstance(scale, float) else scale
            npimg = self._decoder.resize_cubic(npimg, None, fx=base * factor, fy=base * factor)
            o_h, o_w = npimg.shape[:2]
            if o_h < th or o_w < tw:                              # pad bottom/right with zeros up to a full window
                padded = np.zeros((max(th, o_h), max(tw, o_w), 3), dtype=np.uint8)
                padded[:o_h, :o_w, :] = npimg
                npimg = padded
            rois, ratios = [], []
            for (x, y, w, h) in _sliding_windows(npimg.shape[1], npimg.shape[0], tw, th, overlap):
                rois.append(npimg[y:y + h, x:x + w])
                ratios.append((float(x) / o_w, float(y) / o_h, float(w) / o_w, float(h) / o_h))
            return rois, ratios
        if isinstance(scale, tuple) and len(scale) == 3:
            factor = base * scale[2]
            npimg = self._decoder.resize_cubic(npimg, None, fx=factor, fy=factor)
            ratio_w, ratio_h = tw / float(npimg.shape[1]), th / float(npimg.shape[0])
            ratio_x, ratio_y = max(scale[0] - ratio_w / 2., 0.0), max(scale[1] - ratio_h / 2., 0.0)
            if ratio_x + ratio_w > 1.0:
                ratio_x = 1. - ratio_w
            if ratio_y + ratio_h > 1.0:
                ratio_y = 1. - ratio_h
            return [self._crop_roi(npimg, ratio_x, ratio_y)], [(ratio_x, ratio_y, ratio_w, ratio_h)]
        return None                                               # any other ``scale``: the reference falls through

    def _crop_roi(self, npimg, ratio_x, ratio_y):
        """``TfPoseEstimator._crop_roi`` (estimator.py:502-516) as written: a crop that comes out smaller than
        ``target_size`` is centred in a blank image which is then dropped -- the method returns None for it."""
        tw, th = self.target_size
        h, w = npimg.shape[:2]
        x, y = max(int(w * ratio_x - .5), 0), max(int(h * ratio_y - .5), 0)
        cropped = npimg[y:y + th, x:x + tw]
        if cropped.shape[1] < tw or cropped.shape[0] < th:
            return None
        return cropped

    def close(self):
        self._decoder.close()
