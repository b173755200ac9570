"""Host-side mirror of the reference interfaces for the SURF path, over the C ABI (include/boofsurf.h).

  FactoryDetectDescribe.surfFast / surfStable   F/factory/feature/detdesc/FactoryDetectDescribe.java:118-135, 209-226
  DetectDescribePoint<T,BrightFeature>          F/abst/feature/detdesc/DetectDescribePoint.java:32-46 (+ InterestPointDetector,
                                                FoundPointSO, DescriptorInfo) as implemented by WrapDetectDescribeSurf :75-141
  FactoryAssociation.greedy                     F/factory/feature/associate/FactoryAssociation.java:51-65
  AssociateDescription<D>                       F/abst/feature/associate/AssociateDescription.java:42-61 / Associate.java:46-108
                                                as implemented by WrapAssociateGreedy :58-123

Same method names, argument meaning and error behaviour (IllegalArgumentException -> ValueError).  Images are numpy arrays
(float32 = GrayF32, uint8 = GrayU8, int32 = GrayS32); a row-strided numpy view plays the role of a BoofCV sub-image.
The batched / device-resident entry points (detectBatch, associate on device arrays) are additions for throughput use."""
import ctypes as C

import numpy as np

from . import _native as N
from . import stages as S

SURF_SCALE_TO_RADIUS = 2.0      # BoofDefaults.SURF_SCALE_TO_RADIUS, IP/BoofDefaults.java:38


# ------------------------------------------------------------------------------------------------ configs (Java field names)
class ConfigFastHessian:
    """F/abst/feature/detect/interest/ConfigFastHessian.java:33-70"""
    def __init__(self, detectThreshold=1.0, extractRadius=2, maxFeaturesPerScale=-1, initialSampleSize=1, initialSize=9,
                 numberScalesPerOctave=4, numberOfOctaves=4):
        self.detectThreshold = detectThreshold; self.extractRadius = extractRadius
        self.maxFeaturesPerScale = maxFeaturesPerScale; self.initialSampleSize = initialSampleSize
        self.initialSize = initialSize; self.numberScalesPerOctave = numberScalesPerOctave
        self.numberOfOctaves = numberOfOctaves; self.scaleStepSize = 6

    def _c(self):
        return N.FhCfg(self.detectThreshold, self.extractRadius, self.maxFeaturesPerScale, self.initialSampleSize,
                       self.initialSize, self.numberScalesPerOctave, self.numberOfOctaves, self.scaleStepSize)


class ConfigAverageIntegral:
    """F/abst/feature/orientation/ConfigAverageIntegral.java:34-51"""
    def __init__(self):
        self.objectRadiusToScale = 1.0 / 2.0; self.radius = 6; self.samplePeriod = 1.0; self.sampleWidth = 6
        self.weightSigma = -1.0

    def _c(self):
        return N.OriCfg(0, self.objectRadiusToScale, self.radius, self.samplePeriod, self.sampleWidth, self.weightSigma,
                        np.pi / 3.0)


class ConfigSlidingIntegral:
    """F/abst/feature/orientation/ConfigSlidingIntegral.java:34-54"""
    def __init__(self):
        self.objectRadiusToScale = 1.0 / 2.0; self.samplePeriod = 0.65; self.windowSize = np.pi / 3.0; self.radius = 8
        self.weightSigma = -1.0; self.sampleWidth = 6

    def _c(self):
        return N.OriCfg(1, self.objectRadiusToScale, self.radius, self.samplePeriod, self.sampleWidth, self.weightSigma,
                        self.windowSize)


class ConfigSurfDescribe:
    """F/abst/feature/describe/ConfigSurfDescribe.java:34-79"""
    class Speed:
        def __init__(self):
            self.widthLargeGrid = 4; self.widthSubRegion = 5; self.widthSample = 3; self.useHaar = False
            self.weightSigma = 4.5

        def _c(self):
            return N.DescCfg(0, self.widthLargeGrid, self.widthSubRegion, self.widthSample, int(self.useHaar),
                             self.weightSigma, 2, 2.5, 2.5)

    class Stability:
        def __init__(self):
            self.widthLargeGrid = 4; self.widthSubRegion = 5; self.widthSample = 3; self.useHaar = False
            self.overLap = 2; self.sigmaLargeGrid = 2.5; self.sigmaSubRegion = 2.5

        def _c(self):
            return N.DescCfg(1, self.widthLargeGrid, self.widthSubRegion, self.widthSample, int(self.useHaar), 4.5,
                             self.overLap, self.sigmaLargeGrid, self.sigmaSubRegion)


# ------------------------------------------------------------------------------------------------ value types
class Point2D_F64:
    __slots__ = ("x", "y")

    def __init__(self, x=0.0, y=0.0):
        self.x = x; self.y = y


class BrightFeature:
    """F/struct/feature/BrightFeature.java:28-36 — TupleDesc_F64 + white flag"""
    def __init__(self, numFeatures):
        self.value = np.zeros(numFeatures, np.float64)
        self.white = False

    def size(self):
        return len(self.value)


class AssociatedIndex:
    """F/struct/feature/AssociatedIndex.java:27-53"""
    __slots__ = ("src", "dst", "fitScore")

    def __init__(self, src, dst, fitScore):
        self.src = src; self.dst = dst; self.fitScore = fitScore


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


# ------------------------------------------------------------------------------------------------ DetectDescribePoint
class RetainedFeatures:
    """Handle of a bs_features set (device memory); close() or garbage collection frees it."""

    def __init__(self, handle, device=0):
        self._h = handle; self.device = device

    def count(self):
        n, dof = C.c_int(0), C.c_int(0)
        N.check(N.lib().bs_features_count(self._h, C.byref(n), C.byref(dof)))
        return n.value, dof.value

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            N.lib().bs_features_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class HostFeatureBuffers:
    """Host-side result arrays of detectHostBatch, page-locked through the library (bs_host_register) so that the
    device->host copies run asynchronously.  frame f owns rows frame_offset[f]:frame_offset[f+1] of every array."""

    def __init__(self, batch, cap, dof=64, pinned=True):
        self.cap = int(cap); self.dof = int(dof); self.total = 0; self.batch = 0
        self.frame_offset = np.zeros(batch + 1, np.int32)
        self.xy = np.zeros((self.cap, 2)); self.scale = np.zeros(self.cap); self.angle = np.zeros(self.cap)
        self.desc = np.zeros((self.cap, self.dof)); self.white = np.zeros(self.cap, np.uint8)
        self._pinned = []
        if pinned:
            for a in (self.xy, self.scale, self.angle, self.desc, self.white):
                if a.nbytes:
                    N.check(N.lib().bs_host_register(C.c_void_p(a.ctypes.data), a.nbytes))
                    self._pinned.append(a)

    def ptr(self, name):
        return C.c_void_p(getattr(self, name).ctypes.data)

    def frame(self, f):
        b, e = int(self.frame_offset[f]), int(self.frame_offset[f + 1])
        return dict(xy=self.xy[b:e], x=self.xy[b:e, 0], y=self.xy[b:e, 1], scale=self.scale[b:e], angle=self.angle[b:e],
                    desc=self.desc[b:e], white=self.white[b:e].astype(bool))

    def close(self):
        for a in self._pinned:
            N.lib().bs_host_unregister(C.c_void_p(a.ctypes.data))
        self._pinned = []

    def __del__(self):
        try:
            self.close()
        except Exception:      # interpreter shutdown: the module globals may already be gone
            pass


class DetectDescribeSurf:
    """WrapDetectDescribeSurf over a bs_surf handle.  Not thread safe, like the reference (recycled storage)."""

    def __init__(self, fh, ori, desc, imageType, device=0, numBands=0):
        self._pixel = S.pixel_type(imageType)
        self._dtype = np.dtype(imageType)
        self._h = C.c_void_p()
        self._bands = int(numBands)
        if self._bands > 0:
            N.check(N.lib().bs_surf_create_planar(device, C.byref(fh), C.byref(ori), C.byref(desc), self._pixel, self._bands,
                                                  C.byref(self._h)))
        else:
            N.check(N.lib().bs_surf_create(device, C.byref(fh), C.byref(ori), C.byref(desc), self._pixel, C.byref(self._h)))
        self._dof = N.lib().bs_surf_dof(self._h)
        self.device = device
        self._res = None          # results of frame 0 of the last detect (host copies)
        self._batch = 0

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            N.lib().bs_surf_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:      # interpreter shutdown: the module globals may already be gone
            pass

    def setWorkspace(self, chunkFrames=0, poolFeatures=0):
        N.check(N.lib().bs_surf_set_workspace(self._h, chunkFrames, poolFeatures))

    def retainFrame(self, frame=0, stream=None):
        """Device-resident copy of one frame's features that survives later detect calls (bs_surf_retain): what the
        frame-at-a-time DDA tracker keeps of the previous frame (FactoryPointTracker.java:169-175)."""
        h = C.c_void_p()
        N.check(N.lib().bs_surf_retain(self._h, int(frame), C.c_void_p(stream or 0), C.byref(h)))
        return RetainedFeatures(h, self.device)

    def setWalkScratch(self, maxFeatures=0):
        """surfStable: cap the sliding-window walk scratch (features); the rest is walked inside describe_kernel."""
        N.check(N.lib().bs_surf_set_walk_scratch(self._h, maxFeatures))

    def setWays(self, ways):
        N.check(N.lib().bs_surf_set_ways(self._h, ways))

    # --- InterestPointDetector.detect
    def detect(self, input):
        a = np.asarray(input)
        if a.dtype != self._dtype:
            raise ValueError("Unknown input type: expected %s" % self._dtype)
        a, stride = S._strided(a)
        h, w = a.shape
        N.check(N.lib().bs_surf_detect(self._h, _p(a), 0, stride, w, h))
        self._batch = 1
        self._res = self.readFrame(0)

    # --- batched, device-resident (torch tensor [B,H,W] on this handle's device); asynchronous on the current stream
    def detectBatch(self, frames, stream=None):
        import torch
        if frames.dim() != 3 or frames.stride(2) != 1:
            raise ValueError("frames must be [batch, height, width] with unit column stride")
        if {torch.float32: N.PIXEL_F32, torch.uint8: N.PIXEL_U8, torch.int32: N.PIXEL_S32}.get(frames.dtype) != self._pixel:
            raise ValueError("Unknown input type")
        b, h, w = frames.shape
        st = (stream if stream is not None else torch.cuda.current_stream(frames.device).cuda_stream) or 1   # 0 -> cudaStreamLegacy
        N.check(N.lib().bs_surf_detect_device(self._h, C.c_void_p(frames.data_ptr()), frames.stride(0), frames.stride(1),
                                              w, h, b, C.c_void_p(st)))
        self._batch = b
        self._res = None

    # --- batched, HOST frames in / HOST feature lists out (bs_surf_detect_host_batch): copies overlap compute inside the
    # library.  frames: numpy [B,H,W] or a CPU torch tensor (pinned memory makes the copies asynchronous); out: a
    # HostFeatureBuffers sized for at least the expected number of features.
    def detectHostBatch(self, frames, out):
        if hasattr(frames, "data_ptr"):
            b, h, w = frames.shape
            ptr, fs, rs = frames.data_ptr(), frames.stride(0), frames.stride(1)
            if frames.stride(2) != 1 or frames.is_cuda:
                raise ValueError("frames must be a CPU tensor [batch, height, width] with unit column stride")
            import torch
            code = {torch.float32: N.PIXEL_F32, torch.uint8: N.PIXEL_U8, torch.int32: N.PIXEL_S32}.get(frames.dtype)
        else:
            frames = np.ascontiguousarray(frames)
            b, h, w = frames.shape
            ptr, fs, rs = frames.ctypes.data, h * w, w
            code = {np.dtype(np.float32): N.PIXEL_F32, np.dtype(np.uint8): N.PIXEL_U8, np.dtype(np.int32): N.PIXEL_S32}.get(frames.dtype)
        if code != self._pixel:
            raise ValueError("Unknown input type")
        if out.dof != self._dof or len(out.frame_offset) < b + 1:
            raise ValueError("host buffers do not fit this detector / batch")
        total = C.c_long(0)
        N.check(N.lib().bs_surf_detect_host_batch(self._h, C.c_void_p(ptr), fs, rs, w, h, b, out.cap, out.ptr("frame_offset"), out.ptr("xy"),
                                                  out.ptr("scale"), out.ptr("angle"), out.ptr("desc"), out.ptr("white"), C.byref(total)))
        out.total = total.value; out.batch = b
        self._batch = 0; self._res = None
        return out

    def totalFeatures(self):
        n = C.c_long(0)
        N.check(N.lib().bs_surf_total(self._h, C.byref(n)))
        return n.value

    def countFrame(self, frame):
        n = C.c_int(0)
        N.check(N.lib().bs_surf_count(self._h, frame, C.byref(n)))
        return n.value

    def readFrame(self, frame):
        n = self.countFrame(frame)
        r = dict(xy=np.zeros((n, 2)), scale=np.zeros(n), angle=np.zeros(n), desc=np.zeros((n, self._dof)),
                 white=np.zeros(n, np.uint8), meta=np.zeros((n, 4), np.int32), response=np.zeros(n, np.float32))
        N.check(N.lib().bs_surf_read(self._h, frame, _p(r["xy"]), _p(r["scale"]), _p(r["angle"]), _p(r["desc"]),
                                     _p(r["white"]), _p(r["meta"]), _p(r["response"])))
        r["x"] = r["xy"][:, 0]; r["y"] = r["xy"][:, 1]
        r["white"] = r["white"].astype(bool)
        return r

    STAGES = ("integral", "fast_hessian", "nms_refine", "compaction", "orient_describe")

    def setTiming(self, on=True):
        N.check(N.lib().bs_surf_set_timing(self._h, int(on)))

    def getTiming(self):
        """{stage: (accumulated device ms, chunks timed)} since setTiming(True)"""
        ms = (C.c_double * 5)(); calls = (C.c_long * 5)()
        N.check(N.lib().bs_surf_get_timing(self._h, ms, calls))
        return {k: (ms[i], calls[i]) for i, k in enumerate(self.STAGES)}

    def deviceView(self):
        v = N.DeviceView()
        N.check(N.lib().bs_surf_view(self._h, C.byref(v)))
        return v

    def _need(self):
        if self._res is None:
            if self._batch <= 0:
                raise RuntimeError("detect has not been called")
            self._res = self.readFrame(0)
        return self._res

    # --- FoundPointSO
    def getNumberOfFeatures(self):
        return len(self._need()["scale"])

    def getLocation(self, featureIndex):
        r = self._need()
        return Point2D_F64(float(r["xy"][featureIndex, 0]), float(r["xy"][featureIndex, 1]))

    def getRadius(self, featureIndex):
        return float(self._need()["scale"][featureIndex]) * SURF_SCALE_TO_RADIUS

    def getOrientation(self, featureIndex):
        return float(self._need()["angle"][featureIndex])

    # --- DescriptorInfo / DetectDescribePoint
    def getDescription(self, index):
        r = self._need()
        f = BrightFeature(self._dof)
        f.value = r["desc"][index]          # recycled storage, like the reference
        f.white = bool(r["white"][index])
        return f

    def createDescription(self):
        return BrightFeature(self._dof)

    def getDescriptionType(self):
        return BrightFeature

    def hasScale(self):
        return True

    def hasOrientation(self):
        return True


class ImageType:
    """boofcv.struct.image.ImageType, as far as the SURF factories look at it: family, band type, band count."""
    GRAY, PLANAR, INTERLEAVED = "GRAY", "PLANAR", "INTERLEAVED"

    def __init__(self, family, dataType, numBands):
        self.family = family; self.dataType = dataType; self.numBands = int(numBands)

    @staticmethod
    def pl(numBands, dataType):
        return ImageType(ImageType.PLANAR, dataType, numBands)

    @staticmethod
    def single(dataType):
        return ImageType(ImageType.GRAY, dataType, 1)

    def getFamily(self):
        return self.family

    def getNumBands(self):
        return self.numBands

    def getImageClass(self):
        return self.dataType


class SurfPlanar_to_DetectDescribePoint(DetectDescribeSurf):
    """DetectDescribePoint<Planar<T>,BrightFeature> for colour images (F/abst/feature/detdesc/
    SurfPlanar_to_DetectDescribePoint.java:40-118 around DetectDescribeSurfPlanar / DescribePointSurfPlanar): detection and
    orientation on the band average, descriptor = per-band SURF sums concatenated and normalised once (64*numBands values).
    Input: array [numBands, height, width] (Planar.bands)."""

    def detect(self, input):
        a = np.asarray(input)
        if a.ndim != 3:
            raise ValueError("expected a Planar image [numBands, height, width]")
        if a.shape[0] != self._bands:
            raise ValueError("Unexpected number of bands. Expected %d found %d" % (self._bands, a.shape[0]))
        if a.dtype != self._dtype:
            raise ValueError("Unknown input type: expected %s" % self._dtype)
        planes = [S._strided(a[b]) for b in range(self._bands)]
        stride = planes[0][1]
        if any(p[1] != stride for p in planes):
            planes = [(np.ascontiguousarray(p[0]), p[0].shape[1]) for p in planes]
            stride = planes[0][1]
        h, w = planes[0][0].shape
        ptrs = (C.c_void_p * self._bands)(*[p[0].ctypes.data for p in planes])
        N.check(N.lib().bs_surf_detect_planar(self._h, ptrs, self._bands, 0, stride, w, h))
        self._batch = 1
        self._res = self.readFrame(0)

    def detectBatch(self, frames, stream=None):
        """Device-resident batch: torch tensor [B, numBands, H, W] on this handle's device."""
        import torch
        if frames.dim() != 4 or frames.stride(3) != 1 or frames.shape[1] != self._bands:
            raise ValueError("frames must be [batch, numBands, height, width] with unit column stride")
        if {torch.float32: N.PIXEL_F32, torch.uint8: N.PIXEL_U8, torch.int32: N.PIXEL_S32}.get(frames.dtype) != self._pixel:
            raise ValueError("Unknown input type")
        b, _, h, w = frames.shape
        st = (stream if stream is not None else torch.cuda.current_stream(frames.device).cuda_stream) or 1
        N.check(N.lib().bs_surf_detect_planar_device(self._h, C.c_void_p(frames.data_ptr()), frames.stride(0), frames.stride(1),
                                                     frames.stride(2), w, h, b, C.c_void_p(st)))
        self._batch = b
        self._res = None

    def detectHostBatch(self, frames, out):
        raise NotImplementedError("host batches are implemented for single-band frames only")

    def getRadius(self, featureIndex):
        # DetectDescribeSurfPlanar.getRadius (:131-133) returns the scale itself
        return float(self._need()["scale"][featureIndex])


class FactoryDetectDescribe:
    @staticmethod
    def _color(fh, ori, desc, imageType, device):
        if not isinstance(imageType, ImageType) or imageType.getFamily() != ImageType.PLANAR:
            raise ValueError("Image type not supported")           # FactoryDetectDescribe.java:176, :266
        return SurfPlanar_to_DetectDescribePoint(fh, ori, desc, imageType.getImageClass(), device, imageType.getNumBands())

    @staticmethod
    def surfColorFast(configDetector=None, configDesc=None, configOrientation=None, imageType=None, device=0):
        """FactoryDetectDescribe.surfColorFast (F/factory/feature/detdesc/FactoryDetectDescribe.java:153-178)."""
        fh = (configDetector or ConfigFastHessian())._c()
        desc = (configDesc or ConfigSurfDescribe.Speed())._c()
        ori = (configOrientation or ConfigAverageIntegral())._c()
        return FactoryDetectDescribe._color(fh, ori, desc, imageType, device)

    @staticmethod
    def surfColorStable(configDetector=None, configDescribe=None, configOrientation=None, imageType=None, device=0):
        """FactoryDetectDescribe.surfColorStable (:247-268)."""
        fh = (configDetector or ConfigFastHessian())._c()
        desc = (configDescribe or ConfigSurfDescribe.Stability())._c()
        ori = (configOrientation or ConfigSlidingIntegral())._c()
        return FactoryDetectDescribe._color(fh, ori, desc, imageType, device)

    @staticmethod
    def surfFast(configDetector=None, configDesc=None, configOrientation=None, imageType=np.float32, device=0):
        fh = (configDetector or ConfigFastHessian())._c()
        desc = (configDesc or ConfigSurfDescribe.Speed())._c()
        ori = (configOrientation or ConfigAverageIntegral())._c()
        return DetectDescribeSurf(fh, ori, desc, imageType, device)

    @staticmethod
    def surfStable(configDetector=None, configDescribe=None, configOrientation=None, imageType=np.float32, device=0):
        fh = (configDetector or ConfigFastHessian())._c()
        desc = (configDescribe or ConfigSurfDescribe.Stability())._c()
        ori = (configOrientation or ConfigSlidingIntegral())._c()
        return DetectDescribeSurf(fh, ori, desc, imageType, device)
