"""Tier-1 extractor: the reference's ORBExtractor driver (saiga/vision/features/ORBExtractor.cpp:82-289) executed
with the *real* cv2 4.13 primitives and the reference's *own* compiled Distribute / ComputeAngle /
ComputeDescriptor / ScalePyramid (oracle/_ref). Only usable where cv2 and /root/reference exist; it produces the
golden vectors under tests/golden/ that pin the in-repo oracle (and through it the CUDA path)."""
import ctypes as C
import math

import numpy as np

import oracle_py as O


def _p(a):
    return C.c_void_p(a.ctypes.data)


def iround_f32(v):
    v = np.float32(v)
    return int(np.float32(v + (np.float32(0.5) if v >= 0 else np.float32(-0.5))))


def hybrid_detect(img, nfeatures, scale_factor, nlevels, th, thmin):
    import cv2

    cv2.setNumThreads(1)
    R = O.ref()
    assert R is not None
    scale = np.zeros(nlevels, np.float64)
    feats = np.zeros(nlevels, np.int32)
    R.ref_scale_pyramid(nlevels, C.c_float(scale_factor), nfeatures, _p(scale), _p(feats))
    rows, cols = img.shape
    levels = [np.ascontiguousarray(img)]
    for l in range(1, nlevels):
        s = np.float32(1.0 / scale[l])
        lr, lc = iround_f32(np.float32(rows) * s), iround_f32(np.float32(cols) * s)
        levels.append(cv2.resize(levels[-1], (lc, lr), interpolation=cv2.INTER_LINEAR))
    det = {}
    out_k, out_d, stages = [], [], []
    for l, im in enumerate(levels):
        h, w = im.shape
        minB, maxBX, maxBY = 16, w - 16, h - 16
        width, height = np.float32(maxBX - minB), np.float32(maxBY - minB)
        nCols, nRows = int(width / np.float32(30)), int(height / np.float32(30))
        wCell, hCell = int(math.ceil(width / np.float32(nCols))), int(math.ceil(height / np.float32(nRows)))
        raw = []
        for i in range(nRows):
            iniY = minB + i * hCell
            maxY = iniY + hCell + 6
            if iniY >= maxBY - 3:
                continue
            maxY = min(maxY, maxBY)
            for j in range(nCols):
                iniX = minB + j * wCell
                maxX = iniX + wCell + 6
                if iniX >= maxBX - 6:
                    continue
                maxX = min(maxX, maxBX)
                roi = im[iniY:maxY, iniX:maxX]
                k = det.setdefault(th, cv2.FastFeatureDetector_create(threshold=th, nonmaxSuppression=True)).detect(roi)
                if len(k) == 0:
                    k = det.setdefault(thmin, cv2.FastFeatureDetector_create(threshold=thmin, nonmaxSuppression=True)).detect(roi)
                for kp in k:
                    raw.append((kp.pt[0] + j * wCell, kp.pt[1] + i * hCell, kp.size, kp.angle, kp.response, kp.octave))
        raw = np.array(raw, dtype=O.KP32) if raw else np.zeros(0, O.KP32)
        kept, _ = O.distribute(raw, (minB, minB), (maxBX, maxBY), int(feats[l]), which="ref") if len(raw) else (raw, raw)
        kept = kept.copy()
        kept["x"] += minB
        kept["y"] += minB
        kept["octave"] = l
        kept["size"] = float(int(31 * scale[l]))
        if len(kept):
            xy = np.ascontiguousarray(np.stack([kept["x"], kept["y"]], 1), np.float32)
            ang = np.zeros(len(kept), np.float32)
            R.ref_angles(_p(im), h, w, C.c_size_t(w), _p(xy), len(kept), _p(ang))
            kept["angle"] = ang
            g = cv2.GaussianBlur(im, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
            d = np.zeros((len(kept), 4), np.uint64)
            R.ref_descriptors(_p(g), h, w, C.c_size_t(w), _p(xy), _p(ang), len(kept), _p(d))
        else:
            g = im.copy()
            d = np.zeros((0, 4), np.uint64)
        stages.append(dict(level=im, gauss=g, raw=raw, kept=kept.copy()))
        if l != 0:
            s = np.float32(scale[l])
            kept["x"] = kept["x"] * s
            kept["y"] = kept["y"] * s
        out_k.append(kept)
        out_d.append(d)
    return np.concatenate(out_k), np.concatenate(out_d), stages
