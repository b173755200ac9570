/*
 * Drop-in for WrapDetectDescribeSurf (main/boofcv-feature/src/main/java/boofcv/abst/feature/detdesc/WrapDetectDescribeSurf.java:75-141)
 * and WrapDetectDescribeSurf_MT: the whole SURF pipeline (integral image, Fast-Hessian, NMS + refinement, orientation, SURF-64)
 * runs in libboofsurf.so on a B200.  Not compiled in the build image (no JDK) — see BoofSurfNative.
 */
package boofcv.gpu;

import boofcv.BoofDefaults;
import boofcv.abst.feature.describe.ConfigSurfDescribe;
import boofcv.abst.feature.detdesc.DetectDescribePoint;
import boofcv.abst.feature.detect.interest.ConfigFastHessian;
import boofcv.abst.feature.orientation.ConfigAverageIntegral;
import boofcv.abst.feature.orientation.ConfigSlidingIntegral;
import boofcv.struct.feature.BrightFeature;
import boofcv.struct.feature.SurfFeatureQueue;
import boofcv.struct.image.GrayF32;
import boofcv.struct.image.GrayS32;
import boofcv.struct.image.GrayU8;
import boofcv.struct.image.ImageGray;
import georegression.struct.point.Point2D_F64;

import java.lang.foreign.Arena;
import java.lang.foreign.MemorySegment;

import static boofcv.gpu.BoofSurfNative.*;
import static java.lang.foreign.ValueLayout.ADDRESS;
import static java.lang.foreign.ValueLayout.JAVA_BYTE;
import static java.lang.foreign.ValueLayout.JAVA_DOUBLE;
import static java.lang.foreign.ValueLayout.JAVA_FLOAT;
import static java.lang.foreign.ValueLayout.JAVA_INT;

public class DetectDescribeSurfB200<T extends ImageGray<T>> implements DetectDescribePoint<T, BrightFeature>, AutoCloseable {

	private final Arena arena = Arena.ofShared();
	private final MemorySegment handle;
	private final Class<T> imageType;
	private final int pixelType;
	private final int dof;

	// results of the last detect(), in the reference's emission order
	private final SurfFeatureQueue features;
	private double[] xy = new double[0], scale = new double[0], angle = new double[0];
	private int count;

	/** surfFast: ConfigAverageIntegral + ConfigSurfDescribe.Speed (FactoryDetectDescribe.java:118-135). */
	public static <T extends ImageGray<T>> DetectDescribeSurfB200<T>
	surfFast(ConfigFastHessian fh, ConfigSurfDescribe.Speed desc, ConfigAverageIntegral ori, Class<T> imageType) {
		return new DetectDescribeSurfB200<>(fh, desc, ori, null, imageType, 0);
	}

	/** surfStable: ConfigSlidingIntegral + ConfigSurfDescribe.Stability (FactoryDetectDescribe.java:209-226). */
	public static <T extends ImageGray<T>> DetectDescribeSurfB200<T>
	surfStable(ConfigFastHessian fh, ConfigSurfDescribe.Stability desc, ConfigSlidingIntegral ori, Class<T> imageType) {
		return new DetectDescribeSurfB200<>(fh, desc, null, ori, imageType, 1);
	}

	private DetectDescribeSurfB200(ConfigFastHessian fh, ConfigSurfDescribe desc, ConfigAverageIntegral avg, ConfigSlidingIntegral slide,
								   Class<T> imageType, int variant) {
		this.imageType = imageType;
		if (imageType == GrayF32.class) pixelType = PIXEL_F32;
		else if (imageType == GrayU8.class) pixelType = PIXEL_U8;
		else if (imageType == GrayS32.class) pixelType = PIXEL_S32;
		else throw new IllegalArgumentException("Unknown input type: " + imageType.getSimpleName());   // GIntegralImageOps.getIntegralType

		try (Arena a = Arena.ofConfined()) {
			MemorySegment cFh = a.allocate(FH_CFG), cOri = a.allocate(ORI_CFG), cDesc = a.allocate(DESC_CFG);
			FH_DEFAULT.invokeExact(cFh);
			ORI_DEFAULT.invokeExact(cOri, variant);
			DESC_DEFAULT.invokeExact(cDesc, variant);
			if (fh != null) {
				fh.checkValidity();
				cFh.set(JAVA_FLOAT, 0, fh.detectThreshold);
				cFh.set(JAVA_INT, 4, fh.extractRadius);
				cFh.set(JAVA_INT, 8, fh.maxFeaturesPerScale);
				cFh.set(JAVA_INT, 12, fh.initialSampleSize);
				cFh.set(JAVA_INT, 16, fh.initialSize);
				cFh.set(JAVA_INT, 20, fh.numberScalesPerOctave);
				cFh.set(JAVA_INT, 24, fh.numberOfOctaves);
				// scale_step_size (offset 28) keeps its default of 6: FactoryInterestPointAlgs.fastHessian passes the constant
			}
			if (avg != null) {
				avg.checkValidity();
				cOri.set(JAVA_DOUBLE, 8, avg.objectRadiusToScale);
				cOri.set(JAVA_INT, 16, avg.radius);
				cOri.set(JAVA_DOUBLE, 24, avg.samplePeriod);
				cOri.set(JAVA_INT, 32, avg.sampleWidth);
				cOri.set(JAVA_DOUBLE, 40, avg.weightSigma);
			}
			if (slide != null) {
				slide.checkValidity();
				cOri.set(JAVA_DOUBLE, 8, slide.objectRadiusToScale);
				cOri.set(JAVA_INT, 16, slide.radius);
				cOri.set(JAVA_DOUBLE, 24, slide.samplePeriod);
				cOri.set(JAVA_INT, 32, slide.sampleWidth);
				cOri.set(JAVA_DOUBLE, 40, slide.weightSigma);
				cOri.set(JAVA_DOUBLE, 48, slide.windowSize);
			}
			if (desc != null) {
				desc.checkValidity();
				cDesc.set(JAVA_INT, 4, desc.widthLargeGrid);
				cDesc.set(JAVA_INT, 8, desc.widthSubRegion);
				cDesc.set(JAVA_INT, 12, desc.widthSample);
				cDesc.set(JAVA_INT, 16, desc.useHaar ? 1 : 0);
				if (desc instanceof ConfigSurfDescribe.Speed s) {
					cDesc.set(JAVA_DOUBLE, 24, s.weightSigma);
				} else if (desc instanceof ConfigSurfDescribe.Stability s) {
					cDesc.set(JAVA_INT, 32, s.overLap);
					cDesc.set(JAVA_DOUBLE, 40, s.sigmaLargeGrid);
					cDesc.set(JAVA_DOUBLE, 48, s.sigmaSubRegion);
				}
			}
			MemorySegment out = a.allocate(ADDRESS);
			check((int) SURF_CREATE.invokeExact(0, cFh, cOri, cDesc, pixelType, out));
			handle = out.get(ADDRESS, 0);
			dof = (int) SURF_DOF.invokeExact(handle);
		} catch (RuntimeException e) {
			throw e;
		} catch (Throwable t) {
			throw new RuntimeException(t);
		}
		features = new SurfFeatureQueue(dof);
	}

	/** WrapDetectDescribeSurf.detect :75-96. The image's data/startIndex/stride go over as they are (sub-images included). */
	@Override
	public void detect(T input) {
		if (input.getClass() != imageType)
			throw new IllegalArgumentException("Unknown input type");
		try (Arena a = Arena.ofConfined()) {
			MemorySegment px;
			if (input instanceof GrayF32 f) px = a.allocateFrom(JAVA_FLOAT, f.data);
			else if (input instanceof GrayU8 u) px = a.allocateFrom(JAVA_BYTE, u.data);
			else px = a.allocateFrom(JAVA_INT, ((GrayS32) input).data);
			check((int) SURF_DETECT.invokeExact(handle, px, input.startIndex, input.stride, input.width, input.height));
			MemorySegment n = a.allocate(JAVA_INT);
			check((int) SURF_COUNT.invokeExact(handle, 0, n));
			count = n.get(JAVA_INT, 0);
			MemorySegment sXy = a.allocate(JAVA_DOUBLE, 2L * Math.max(count, 1)), sScale = a.allocate(JAVA_DOUBLE, Math.max(count, 1));
			MemorySegment sAngle = a.allocate(JAVA_DOUBLE, Math.max(count, 1)), sDesc = a.allocate(JAVA_DOUBLE, (long) dof * Math.max(count, 1));
			MemorySegment sWhite = a.allocate(JAVA_BYTE, Math.max(count, 1));
			check((int) SURF_READ.invokeExact(handle, 0, sXy, sScale, sAngle, sDesc, sWhite, MemorySegment.NULL, MemorySegment.NULL));
			xy = sXy.toArray(JAVA_DOUBLE); scale = sScale.toArray(JAVA_DOUBLE); angle = sAngle.toArray(JAVA_DOUBLE);
			features.reset();
			for (int i = 0; i < count; i++) {
				BrightFeature f = features.grow();                      // recycled between calls, like the reference's queue
				MemorySegment.copy(sDesc, JAVA_DOUBLE, (long) i * dof * 8, f.value, 0, dof);
				f.white = sWhite.get(JAVA_BYTE, i) != 0;
			}
		} catch (RuntimeException e) {
			throw e;
		} catch (Throwable t) {
			throw new RuntimeException(t);
		}
	}

	@Override public int getNumberOfFeatures() { return count; }
	@Override public Point2D_F64 getLocation(int featureIndex) { return new Point2D_F64(xy[2 * featureIndex], xy[2 * featureIndex + 1]); }
	@Override public double getRadius(int featureIndex) { return scale[featureIndex] * BoofDefaults.SURF_SCALE_TO_RADIUS; }
	@Override public double getOrientation(int featureIndex) { return angle[featureIndex]; }
	@Override public BrightFeature getDescription(int index) { return features.get(index); }
	@Override public boolean hasScale() { return true; }
	@Override public boolean hasOrientation() { return true; }
	@Override public BrightFeature createDescription() { return new BrightFeature(dof); }
	@Override public Class<BrightFeature> getDescriptionType() { return BrightFeature.class; }

	@Override
	public void close() {
		try {
			SURF_DESTROY.invokeExact(handle);
		} catch (Throwable ignored) {
		}
		arena.close();
	}
}
