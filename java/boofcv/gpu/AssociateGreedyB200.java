/*
 * Drop-in for WrapAssociateGreedy around AssociateGreedy + ScoreAssociateEuclideanSq_F64
 * (main/boofcv-feature/src/main/java/boofcv/abst/feature/associate/WrapAssociateGreedy.java:58-123,
 *  alg/feature/associate/AssociateGreedy.java:65-118): what FactoryAssociation.greedy(score, maxError, backwardsValidation)
 * returns for TupleDesc_F64 descriptors when libboofsurf.so is available.  Not compiled in the build image — see BoofSurfNative.
 */
package boofcv.gpu;

import boofcv.abst.feature.associate.AssociateDescription;
import boofcv.struct.feature.AssociatedIndex;
import boofcv.struct.feature.MatchScoreType;
import boofcv.struct.feature.TupleDesc_F64;
import org.ddogleg.struct.FastQueue;
import org.ddogleg.struct.GrowQueue_I32;

import java.lang.foreign.Arena;
import java.lang.foreign.MemorySegment;

import static boofcv.gpu.BoofSurfNative.*;
import static java.lang.foreign.ValueLayout.ADDRESS;
import static java.lang.foreign.ValueLayout.JAVA_DOUBLE;
import static java.lang.foreign.ValueLayout.JAVA_INT;

public class AssociateGreedyB200<D extends TupleDesc_F64> implements AssociateDescription<D>, AutoCloseable {

	private final MemorySegment handle;
	private final boolean backwardsValidation;
	private FastQueue<D> listSrc, listDst;
	private final FastQueue<AssociatedIndex> matches = new FastQueue<>(10, AssociatedIndex.class, true);
	private final GrowQueue_I32 unassocSrc = new GrowQueue_I32(), unassocDst = new GrowQueue_I32();
	private int[] pairs = new int[0];
	private double[] fitQuality = new double[0];

	public AssociateGreedyB200(double maxFitError, boolean backwardsValidation) {
		this.backwardsValidation = backwardsValidation;
		try (Arena a = Arena.ofConfined()) {
			MemorySegment out = a.allocate(ADDRESS);
			check((int) ASSOC_CREATE.invokeExact(0, maxFitError, backwardsValidation ? 1 : 0, out));
			handle = out.get(ADDRESS, 0);
		} catch (RuntimeException e) {
			throw e;
		} catch (Throwable t) {
			throw new RuntimeException(t);
		}
	}

	@Override public void setSource(FastQueue<D> listSrc) { this.listSrc = listSrc; }
	@Override public void setDestination(FastQueue<D> listDst) { this.listDst = listDst; }

	private static <D extends TupleDesc_F64> MemorySegment pack(Arena a, FastQueue<D> list, int dof) {
		MemorySegment s = a.allocate(JAVA_DOUBLE, (long) Math.max(list.size, 1) * dof);
		for (int i = 0; i < list.size; i++)
			MemorySegment.copy(list.data[i].value, 0, s, JAVA_DOUBLE, (long) i * dof * 8, dof);     // TupleDesc_F64.value back to back
		return s;
	}

	/** WrapAssociateGreedy.associate :72-92 — pairs/fit from the GPU, the bookkeeping (matches, unassociated lists) from
	 *  bs_assoc_matches, which restates :85-92 and FindUnassociated.checkDestination :56-72. */
	@Override
	public void associate() {
		if (listSrc == null) throw new IllegalArgumentException("source features not specified");
		if (listDst == null) throw new IllegalArgumentException("destination features not specified");
		final int n = listSrc.size, m = listDst.size;
		final int dof = n > 0 ? listSrc.data[0].size() : (m > 0 ? listDst.data[0].size() : 1);
		matches.reset(); unassocSrc.reset(); unassocDst.reset();
		try (Arena a = Arena.ofConfined()) {
			MemorySegment sPairs = a.allocate(JAVA_INT, Math.max(n, 1)), sFit = a.allocate(JAVA_DOUBLE, Math.max(n, 1));
			if (n > 0)
				check((int) ASSOC_RUN.invokeExact(handle, pack(a, listSrc, dof), n, m > 0 ? pack(a, listDst, dof) : MemorySegment.NULL, m, dof, sPairs, sFit));
			pairs = sPairs.asSlice(0, 4L * n).toArray(JAVA_INT);
			fitQuality = sFit.asSlice(0, 8L * n).toArray(JAVA_DOUBLE);
			MemorySegment mSrc = a.allocate(JAVA_INT, Math.max(n, 1)), mDst = a.allocate(JAVA_INT, Math.max(n, 1)), mScore = a.allocate(JAVA_DOUBLE, Math.max(n, 1));
			MemorySegment uSrc = a.allocate(JAVA_INT, Math.max(n, 1)), uDst = a.allocate(JAVA_INT, Math.max(m, 1));
			MemorySegment nm = a.allocate(JAVA_INT), nus = a.allocate(JAVA_INT), nud = a.allocate(JAVA_INT);
			check((int) ASSOC_MATCHES.invokeExact(sPairs, sFit, n, m, mSrc, mDst, mScore, nm, uSrc, nus, uDst, nud));
			for (int k = 0; k < nm.get(JAVA_INT, 0); k++)
				matches.grow().setAssociation(mSrc.getAtIndex(JAVA_INT, k), mDst.getAtIndex(JAVA_INT, k), mScore.getAtIndex(JAVA_DOUBLE, k));
			for (int k = 0; k < nus.get(JAVA_INT, 0); k++) unassocSrc.add(uSrc.getAtIndex(JAVA_INT, k));
			for (int k = 0; k < nud.get(JAVA_INT, 0); k++) unassocDst.add(uDst.getAtIndex(JAVA_INT, k));
		} catch (RuntimeException e) {
			throw e;
		} catch (Throwable t) {
			throw new RuntimeException(t);
		}
	}

	@Override public FastQueue<AssociatedIndex> getMatches() { return matches; }
	@Override public GrowQueue_I32 getUnassociatedSource() { return unassocSrc; }
	@Override public GrowQueue_I32 getUnassociatedDestination() { return unassocDst; }
	@Override public MatchScoreType getScoreType() { return MatchScoreType.NORM_ERROR; }      // ScoreAssociateEuclideanSq_F64.getScoreType
	@Override public boolean uniqueSource() { return true; }
	@Override public boolean uniqueDestination() { return backwardsValidation; }

	@Override
	public void setMaxScoreThreshold(double score) {
		try {
			check((int) ASSOC_SET_MAX.invokeExact(handle, score));
		} catch (RuntimeException e) {
			throw e;
		} catch (Throwable t) {
			throw new RuntimeException(t);
		}
	}

	/** AssociateGreedyBase.getPairs / getFitQuality: what the parity tests compare. */
	public int[] getPairs() { return pairs; }
	public double[] getFitQuality() { return fitQuality; }

	@Override
	public void close() {
		try {
			ASSOC_DESTROY.invokeExact(handle);
		} catch (Throwable ignored) {
		}
	}
}
