/*
 * Golden-vector dumper: runs the STOCK reference (FactoryDetectDescribe.surfFast / surfStable, FactoryAssociation.greedy with
 * ScoreAssociateEuclideanSq_F64) on deterministic synthetic images and writes its outputs as raw little-endian files, so that
 * anyone with a JDK and the BoofCV jars can turn the oracle's "pinned by the reference's unit tests" into "pinned by the
 * reference's outputs".  Nothing here touches the GPU.  Not compiled in the build image (no JDK, no jars).
 *
 *   javac -cp "$BOOFCV_JARS" -d out java/boofcv/gpu/DumpSurfGolden.java
 *   java  -cp "out:$BOOFCV_JARS" boofcv.gpu.DumpSurfGolden tests/golden/java
 *   python -m pytest tests/test_golden_java.py          # compares oracle/ (and on a GPU box the CUDA path) with the files
 *
 * Images: the generator of GenericTestsDetectDescribePoint (java.util.Random(234), uniform [0,100), 100 x 120 — SURVEY.md 8c) and
 * a 320 x 240 blob image written next to the results as image_*.f32 so that the Python side reads exactly the pixels Java saw.
 * Files per case <name>: <name>.meta (text: width height count dof), image_<name>.f32, <name>.xy.f64 [n][2], .scale.f64 [n],
 * .angle.f64 [n], .desc.f64 [n][dof], .white.u8 [n]; association: assoc.pairs.i32 [n], assoc.fit.f64 [n] for surfFast features
 * of the noise image (source) against those of the same image shifted by (3, 2) pixels (destination), backward validation on.
 */
package boofcv.gpu;

import boofcv.abst.feature.associate.AssociateDescription;
import boofcv.abst.feature.associate.ScoreAssociateEuclideanSq_F64;
import boofcv.abst.feature.detdesc.DetectDescribePoint;
import boofcv.alg.feature.associate.AssociateGreedy;
import boofcv.concurrency.BoofConcurrency;
import boofcv.factory.feature.detdesc.FactoryDetectDescribe;
import boofcv.struct.feature.BrightFeature;
import boofcv.struct.image.GrayF32;
import org.ddogleg.struct.FastQueue;

import java.io.IOException;
import java.nio.ByteBuffer;
import java.nio.ByteOrder;
import java.nio.file.Files;
import java.nio.file.Path;
import java.util.Random;

public class DumpSurfGolden {

	static GrayF32 noise(int w, int h, long seed) {
		Random rand = new Random(seed);
		GrayF32 img = new GrayF32(w, h);
		for (int i = 0; i < img.data.length; i++) img.data[i] = rand.nextFloat() * 100f;      // GImageMiscOps.fillUniform(img, rand, 0, 100)
		return img;
	}

	/** smooth blobs at several scales: sum of a few low-frequency cosine products, exactly representable steps only in float */
	static GrayF32 blobs(int w, int h) {
		GrayF32 img = new GrayF32(w, h);
		for (int y = 0; y < h; y++)
			for (int x = 0; x < w; x++) {
				double v = 120 + 40 * Math.cos(x * 0.11) * Math.cos(y * 0.07) + 30 * Math.cos(x * 0.031 + 1) * Math.cos(y * 0.043 + 2)
						+ 20 * Math.cos(x * 0.29 + y * 0.17);
				img.data[y * w + x] = (float) v;
			}
		return img;
	}

	static void write(Path p, ByteBuffer b) throws IOException { b.flip(); Files.write(p, java.util.Arrays.copyOf(b.array(), b.limit())); }
	static ByteBuffer le(int bytes) { return ByteBuffer.allocate(Math.max(bytes, 1)).order(ByteOrder.LITTLE_ENDIAN); }

	static FastQueue<BrightFeature> dump(Path dir, String name, GrayF32 img, DetectDescribePoint<GrayF32, BrightFeature> alg) throws IOException {
		alg.detect(img);
		final int n = alg.getNumberOfFeatures(), dof = alg.createDescription().size();
		ByteBuffer px = le(4 * img.data.length);
		for (float v : img.data) px.putFloat(v);
		write(dir.resolve("image_" + name + ".f32"), px);
		ByteBuffer xy = le(16 * n), sc = le(8 * n), an = le(8 * n), de = le(8 * n * dof), wh = le(n);
		FastQueue<BrightFeature> copy = new FastQueue<>(n, BrightFeature.class, false);
		for (int i = 0; i < n; i++) {
			xy.putDouble(alg.getLocation(i).x).putDouble(alg.getLocation(i).y);
			sc.putDouble(alg.getRadius(i) / boofcv.BoofDefaults.SURF_SCALE_TO_RADIUS);         // WrapDetectDescribeSurf.getRadius :126-128 -> the scale
			an.putDouble(alg.getOrientation(i));
			BrightFeature f = alg.getDescription(i);
			for (int k = 0; k < dof; k++) de.putDouble(f.value[k]);
			wh.put((byte) (f.white ? 1 : 0));
			BrightFeature c = new BrightFeature(dof); c.setTo(f); c.white = f.white; copy.add(c);
		}
		write(dir.resolve(name + ".xy.f64"), xy); write(dir.resolve(name + ".scale.f64"), sc); write(dir.resolve(name + ".angle.f64"), an);
		write(dir.resolve(name + ".desc.f64"), de); write(dir.resolve(name + ".white.u8"), wh);
		Files.writeString(dir.resolve(name + ".meta"), img.width + " " + img.height + " " + n + " " + dof + "\n");
		return copy;
	}

	public static void main(String[] args) throws IOException {
		Path dir = Path.of(args.length > 0 ? args[0] : "tests/golden/java");
		Files.createDirectories(dir);
		BoofConcurrency.USE_CONCURRENT = false;            // the single-thread emission order is the one the parity tests pin
		GrayF32 n0 = noise(100, 120, 234), b0 = blobs(320, 240);
		FastQueue<BrightFeature> src = dump(dir, "fast_noise", n0, FactoryDetectDescribe.surfFast(null, null, null, GrayF32.class));
		dump(dir, "stable_noise", n0, FactoryDetectDescribe.surfStable(null, null, null, GrayF32.class));
		dump(dir, "fast_blobs", b0, FactoryDetectDescribe.surfFast(null, null, null, GrayF32.class));
		dump(dir, "stable_blobs", b0, FactoryDetectDescribe.surfStable(null, null, null, GrayF32.class));
		// destination: the noise image shifted by (3, 2)
		GrayF32 shifted = new GrayF32(97, 118);
		for (int y = 0; y < 118; y++) for (int x = 0; x < 97; x++) shifted.data[y * 97 + x] = n0.data[(y + 2) * 100 + x + 3];
		FastQueue<BrightFeature> dst = dump(dir, "fast_noise_shifted", shifted, FactoryDetectDescribe.surfFast(null, null, null, GrayF32.class));
		AssociateGreedy<BrightFeature> greedy = new AssociateGreedy<>(new ScoreAssociateEuclideanSq_F64(), true);
		greedy.setMaxFitError(Double.MAX_VALUE);
		greedy.associate(src, dst);
		ByteBuffer pairs = le(4 * src.size), fit = le(8 * src.size);
		for (int i = 0; i < src.size; i++) { pairs.putInt(greedy.getPairs()[i]); fit.putDouble(greedy.getFitQuality()[i]); }
		write(dir.resolve("assoc.pairs.i32"), pairs); write(dir.resolve("assoc.fit.f64"), fit);
		System.out.println("wrote golden vectors to " + dir.toAbsolutePath());
	}
}
