"""FactoryAssociation.greedy -> AssociateDescription<TupleDesc_F64> over the C ABI (bs_assoc_*).

Reference: F/factory/feature/associate/FactoryAssociation.java:51-65, WrapAssociateGreedy
(F/abst/feature/associate/WrapAssociateGreedy.java:58-123), AssociateGreedy (F/alg/feature/associate/AssociateGreedy.java:65-118),
ScoreAssociateEuclideanSq_F64 (F/abst/feature/associate/ScoreAssociateEuclideanSq_F64.java:33-43)."""
import ctypes as C

import numpy as np

from . import _native as N
from .surf import AssociatedIndex

DOUBLE_MAX = float(np.finfo(np.float64).max)      # Double.MAX_VALUE


class MatchScoreType:
    NORM_ERROR = "NORM_ERROR"


class ScoreAssociateEuclideanSq_F64:
    """The only score the GPU path implements (the one FactoryAssociation.defaultScore picks for TupleDesc_F64 / SURF)."""
    def getScoreType(self):
        return MatchScoreType.NORM_ERROR


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _as_matrix(features):
    """FastQueue<TupleDesc_F64> -> [n][dof] float64.  Accepts an ndarray or a sequence of objects with `.value`."""
    if features is None:
        return None
    if isinstance(features, np.ndarray):
        a = features
    else:
        features = list(features)
        if len(features) == 0:
            return np.zeros((0, 0), np.float64)
        a = np.stack([np.asarray(getattr(f, "value", f), np.float64) for f in features])
    if a.ndim == 1:
        a = a[:, None]
    return np.ascontiguousarray(a, np.float64)


class AssociateGreedyGPU:
    def __init__(self, score, maxError, backwardsValidation, device=0):
        if not isinstance(score, ScoreAssociateEuclideanSq_F64):
            raise ValueError("only ScoreAssociateEuclideanSq_F64 is supported by the GPU associator")
        self._score = score
        self._max_fit = float(maxError)
        self._backwards = bool(backwardsValidation)
        self._h = C.c_void_p()
        N.check(N.lib().bs_assoc_create(device, float(maxError), int(self._backwards), C.byref(self._h)))
        self.device = device
        self._src = None; self._dst = None
        self._matches = []; self._unassoc_src = np.zeros(0, np.int32); self._unassoc_dst = np.zeros(0, np.int32)
        self._pairs = None; self._fit = None

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            N.lib().bs_assoc_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:      # interpreter shutdown: the module globals may already be gone
            pass

    def setSource(self, listSrc):
        self._src = _as_matrix(listSrc)

    def setDestination(self, listDst):
        self._dst = _as_matrix(listDst)

    def associate(self):
        if self._src is None:
            raise ValueError("source features not specified")
        if self._dst is None:
            raise ValueError("destination features not specified")
        n, m = len(self._src), len(self._dst)
        dof = max(self._src.shape[1] if n else 0, self._dst.shape[1] if m else 0, 1)
        if (n and self._src.shape[1] != dof) or (m and self._dst.shape[1] != dof):
            raise ValueError("source and destination descriptors differ in length")
        self._pairs = np.full(n, -1, np.int32); self._fit = np.zeros(n, np.float64)
        if n:
            N.check(N.lib().bs_assoc_run(self._h, _p(self._src), n, _p(self._dst) if m else None, m, dof, _p(self._pairs),
                                         _p(self._fit)))
        ms = np.zeros(n, np.int32); md = np.zeros(n, np.int32); sc = np.zeros(n, np.float64)
        us = np.zeros(n, np.int32); ud = np.zeros(m, np.int32)
        nm, nus, nud = C.c_int(0), C.c_int(0), C.c_int(0)
        N.check(N.lib().bs_assoc_matches(_p(self._pairs), _p(self._fit), n, m, _p(ms), _p(md), _p(sc), C.byref(nm), _p(us),
                                         C.byref(nus), _p(ud), C.byref(nud)))
        self._matches = [AssociatedIndex(int(ms[k]), int(md[k]), float(sc[k])) for k in range(nm.value)]
        self._unassoc_src = us[:nus.value].copy()
        self._unassoc_dst = ud[:nud.value].copy()

    def getMatches(self):
        return self._matches

    def getUnassociatedSource(self):
        return self._unassoc_src

    def getUnassociatedDestination(self):
        return self._unassoc_dst

    def getPairs(self):            # AssociateGreedyBase.getPairs
        return self._pairs

    def getFitQuality(self):       # AssociateGreedyBase.getFitQuality
        return self._fit

    def setMaxScoreThreshold(self, score):
        N.check(N.lib().bs_assoc_set_max_fit_error(self._h, float(score)))
        self._max_fit = float(score)

    def maxFitError(self):         # AssociateGreedyBase.maxFitError
        return self._max_fit

    def setMode(self, mode):
        N.check(N.lib().bs_assoc_set_mode(self._h, int(mode)))

    def filterScores(self, src, dst, want_scores=True):
        """Diagnostic (bs_assoc_filter_scores): the tensor-core filter's fp32 distance estimates [n][m], the number of
        candidates per source row, and (tensor-core runs, exact fallbacks) of this handle."""
        src = _as_matrix(src); dst = _as_matrix(dst)
        n, m = len(src), len(dst)
        scores = np.zeros((n, m), np.float32) if want_scores else None
        cnt = np.zeros(n, np.int32); stats = np.zeros(2, np.int64)
        N.check(N.lib().bs_assoc_filter_scores(self._h, _p(src), n, _p(dst), m, src.shape[1], _p(scores) if want_scores else None, _p(cnt),
                                               _p(stats)))
        return scores, cnt, (int(stats[0]), int(stats[1]))

    def getScoreType(self):
        return self._score.getScoreType()

    def uniqueSource(self):
        return True

    def uniqueDestination(self):
        return self._backwards

    def associateFrames(self, det, frameSrc, frameDst, surfBasic=False, stream=None):
        """detect -> associate without leaving the GPU (bs_assoc_frames_device): features of two frames of `det`'s last
        batch.  Returns (pairs int32[n], fit float64[n]) CUDA tensors."""
        import torch
        n = det.countFrame(frameSrc)
        dev = torch.device("cuda", self.device)
        pairs = torch.empty(n, dtype=torch.int32, device=dev); fit = torch.empty(n, dtype=torch.float64, device=dev)
        st = (stream if stream is not None else torch.cuda.current_stream(dev).cuda_stream) or 1
        N.check(N.lib().bs_assoc_frames_device(self._h, det._h, int(frameSrc), int(frameDst), int(bool(surfBasic)),
                                               C.c_void_p(pairs.data_ptr()), C.c_void_p(fit.data_ptr()), C.c_void_p(st)))
        return pairs, fit

    def associateFramesBatch(self, det, framesSrc, framesDst, stream=None):
        """associateFrames for a list of frame pairs of `det`'s last batch in one call (bs_assoc_frames_batch_device: the pairs
        are queued back to back, one host synchronisation at the end).  Returns (pairs int32[total], fit float64[total],
        offsets int64[len+1]): the results of pair k are [offsets[k], offsets[k+1])."""
        import torch
        fs = np.ascontiguousarray(framesSrc, dtype=np.int32); fd = np.ascontiguousarray(framesDst, dtype=np.int32)
        if fs.shape != fd.shape or fs.ndim != 1:
            raise ValueError("framesSrc and framesDst must be two lists of the same length")
        total = int(sum(det.countFrame(int(f)) for f in fs))
        dev = torch.device("cuda", self.device)
        pairs = torch.empty(total, dtype=torch.int32, device=dev); fit = torch.empty(total, dtype=torch.float64, device=dev)
        offsets = np.zeros(len(fs) + 1, np.int64)
        st = (stream if stream is not None else torch.cuda.current_stream(dev).cuda_stream) or 1
        N.check(N.lib().bs_assoc_frames_batch_device(self._h, det._h, len(fs), fs.ctypes.data_as(C.c_void_p), fd.ctypes.data_as(C.c_void_p),
                                                     C.c_void_p(pairs.data_ptr()), C.c_void_p(fit.data_ptr()), total,
                                                     offsets.ctypes.data_as(C.c_void_p), C.c_void_p(st)))
        return pairs, fit, offsets

    def associateRetained(self, src, dst, surfBasic=False, stream=None):
        """AssociateDescription.associate on two retained feature sets (DetectDescribeSurf.retainFrame): previous frame ->
        current frame without leaving the device (bs_assoc_features_device).  Returns (pairs, fit) torch tensors."""
        import torch
        n, _ = src.count()
        dev = torch.device("cuda", self.device)
        pairs = torch.empty(max(n, 1), dtype=torch.int32, device=dev)[:n]; fit = torch.empty(max(n, 1), dtype=torch.float64, device=dev)[:n]
        st = (stream if stream is not None else torch.cuda.current_stream(dev).cuda_stream) or 1
        N.check(N.lib().bs_assoc_features_device(self._h, src._h, dst._h, int(bool(surfBasic)), C.c_void_p(pairs.data_ptr()),
                                                 C.c_void_p(fit.data_ptr()), C.c_void_p(st)))
        return pairs, fit

    # device-resident variant (torch float64 tensors [n,dof], [m,dof]); returns (pairs int32[n], fit float64[n]) tensors
    def associateDevice(self, src, dst, stream=None):
        import torch
        n, dof = src.shape; m = dst.shape[0]
        pairs = torch.empty(n, dtype=torch.int32, device=src.device); fit = torch.empty(n, dtype=torch.float64, device=src.device)
        st = (stream if stream is not None else torch.cuda.current_stream(src.device).cuda_stream) or 1   # 0 -> cudaStreamLegacy
        N.check(N.lib().bs_assoc_run_device(self._h, C.c_void_p(src.data_ptr()), n, C.c_void_p(dst.data_ptr()), m, dof,
                                            C.c_void_p(pairs.data_ptr()), C.c_void_p(fit.data_ptr()), C.c_void_p(st)))
        return pairs, fit


def _as_bright(features):
    """FastQueue<BrightFeature> -> ([n][dof] float64, white uint8[n]).  Accepts (desc, white) arrays, a dict with "desc"/"white"
    (DetectDescribeSurf.readFrame) or a sequence of objects with `.value` and `.white`."""
    if features is None:
        return None, None
    if isinstance(features, dict):
        features = (features["desc"], features["white"])
    if isinstance(features, tuple) and len(features) == 2 and isinstance(features[0], np.ndarray):
        d = _as_matrix(features[0]); w = np.ascontiguousarray(np.asarray(features[1]) != 0, np.uint8)
    else:
        features = list(features)
        d = _as_matrix(features)
        w = np.ascontiguousarray([1 if f.white else 0 for f in features], np.uint8)
    if len(w) != len(d):
        raise ValueError("descriptor and laplacian-sign lists differ in length")
    return d, w


class AssociateSurfBasic:
    """AssociateSurfBasic (F/alg/feature/associate/AssociateSurfBasic.java:36-160): features are split by the sign of the
    Laplacian (BrightFeature.white), the two classes are associated independently by `assoc` (FactoryAssociation.greedy)
    and reported with the indices of the input lists.  Runs on the handle of `assoc` (bs_assoc_surf_run)."""
    def __init__(self, assoc):
        if not isinstance(assoc, AssociateGreedyGPU):
            raise ValueError("AssociateSurfBasic needs the GPU greedy associator (FactoryAssociation.greedy)")
        self.assoc = assoc
        self._src = (None, None); self._dst = (None, None)
        self._matches = []; self.unassociatedSrc = np.zeros(0, np.int32); self._unassoc_dst = np.zeros(0, np.int32)
        self._pairs = None; self._fit = None

    def setSrc(self, src):
        self._src = _as_bright(src)

    def setDst(self, dst):
        self._dst = _as_bright(dst)

    def swapLists(self):
        self._src, self._dst = self._dst, self._src

    def associate(self):
        (src, sw), (dst, dw) = self._src, self._dst
        if src is None:
            raise ValueError("source features not specified")
        if dst is None:
            raise ValueError("destination features not specified")
        n, m = len(src), len(dst)
        dof = max(src.shape[1] if n else 0, dst.shape[1] if m else 0, 1)
        if (n and src.shape[1] != dof) or (m and dst.shape[1] != dof):
            raise ValueError("source and destination descriptors differ in length")
        L = N.lib()
        self._pairs = np.full(n, -1, np.int32); self._fit = np.zeros(n, np.float64)
        if n:
            N.check(L.bs_assoc_surf_run(self.assoc._h, _p(src), _p(sw), n, _p(dst) if m else None, _p(dw) if m else None, m, dof,
                                        _p(self._pairs), _p(self._fit)))
        ms = np.zeros(n, np.int32); md = np.zeros(n, np.int32); sc = np.zeros(n, np.float64)
        us = np.zeros(n, np.int32); ud = np.zeros(m, np.int32)
        nm, nus, nud = C.c_int(0), C.c_int(0), C.c_int(0)
        N.check(L.bs_assoc_surf_matches(_p(self._pairs), _p(self._fit), _p(sw) if n else None, n, m, _p(ms), _p(md), _p(sc), C.byref(nm),
                                        _p(us), C.byref(nus), _p(ud), C.byref(nud)))
        self._matches = [AssociatedIndex(int(ms[k]), int(md[k]), float(sc[k])) for k in range(nm.value)]
        self.unassociatedSrc = us[:nus.value].copy()
        self._unassoc_dst = ud[:nud.value].copy()

    def getMatches(self):
        return self._matches

    def totalDestination(self):
        return 0 if self._dst[0] is None else len(self._dst[0])

    def getUnassociatedSrc(self):
        return self.unassociatedSrc

    def getAssoc(self):
        return self.assoc

    def getPairs(self):
        return self._pairs

    def getFitQuality(self):
        return self._fit

    # device-resident variant: torch tensors desc [n,dof] float64 + white [n] uint8 on both sides
    def associateDevice(self, src, src_white, dst, dst_white, stream=None):
        import torch
        n, dof = src.shape; m = dst.shape[0]
        pairs = torch.empty(n, dtype=torch.int32, device=src.device); fit = torch.empty(n, dtype=torch.float64, device=src.device)
        st = (stream if stream is not None else torch.cuda.current_stream(src.device).cuda_stream) or 1
        N.check(N.lib().bs_assoc_surf_run_device(self.assoc._h, C.c_void_p(src.data_ptr()), C.c_void_p(src_white.data_ptr()), n,
                                                 C.c_void_p(dst.data_ptr()), C.c_void_p(dst_white.data_ptr()), m, dof,
                                                 C.c_void_p(pairs.data_ptr()), C.c_void_p(fit.data_ptr()), C.c_void_p(st)))
        return pairs, fit


class WrapAssociateSurfBasic:
    """AssociateDescription<BrightFeature> around AssociateSurfBasic
    (F/abst/feature/associate/WrapAssociateSurfBasic.java:35-90)."""
    def __init__(self, alg):
        self.alg = alg

    def setSource(self, listSrc):
        self.alg.setSrc(listSrc)

    def setDestination(self, listDst):
        self.alg.setDst(listDst)

    def associate(self):
        self.alg.associate()

    def getMatches(self):
        return self.alg.getMatches()

    def getUnassociatedSource(self):
        return self.alg.getUnassociatedSrc()

    def getUnassociatedDestination(self):
        return self.alg._unassoc_dst

    def setMaxScoreThreshold(self, score):
        self.alg.getAssoc().setMaxScoreThreshold(score)

    def getScoreType(self):
        return self.alg.getAssoc().getScoreType()

    def uniqueSource(self):
        return self.alg.getAssoc().uniqueSource()

    def uniqueDestination(self):
        return self.alg.getAssoc().uniqueDestination()


class FactoryAssociation:
    @staticmethod
    def greedy(score, maxError, backwardsValidation, device=0):
        return AssociateGreedyGPU(score, maxError, backwardsValidation, device)

    @staticmethod
    def defaultScore(tupleType=None):
        return ScoreAssociateEuclideanSq_F64()
