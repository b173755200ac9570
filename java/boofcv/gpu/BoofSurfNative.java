/*
 * Java side of the drop-in boundary: Panama (java.lang.foreign, JDK 22+) bindings of include/boofsurf.h.
 *
 * NOT compiled or run in the build image (no JDK there): these sources are what a BoofCV maintainer adds next to
 * main/boofcv-feature to put libboofsurf.so behind FactoryDetectDescribe.surfFast/surfStable and FactoryAssociation.greedy.
 * Compile with BoofCV (boofcv-feature, boofcv-ip, boofcv-types, ddogleg, georegression) on the class path:
 *     javac --release 22 -cp "$BOOFCV_JARS" -d out java/boofcv/gpu/*.java
 * and run with  -Djava.library.path=<dir of libboofsurf.so> --enable-native-access=ALL-UNNAMED.
 */
package boofcv.gpu;

import java.lang.foreign.Arena;
import java.lang.foreign.FunctionDescriptor;
import java.lang.foreign.Linker;
import java.lang.foreign.MemoryLayout;
import java.lang.foreign.MemorySegment;
import java.lang.foreign.StructLayout;
import java.lang.foreign.SymbolLookup;
import java.lang.invoke.MethodHandle;

import static java.lang.foreign.ValueLayout.ADDRESS;
import static java.lang.foreign.ValueLayout.JAVA_DOUBLE;
import static java.lang.foreign.ValueLayout.JAVA_FLOAT;
import static java.lang.foreign.ValueLayout.JAVA_INT;
import static java.lang.foreign.ValueLayout.JAVA_LONG;

/** One static downcall handle per C-ABI entry point used by the two drop-in classes. */
public final class BoofSurfNative {
	public static final int BS_OK = 0, BS_ERR_ARG = 1;              // include/boofsurf.h status codes (ARG -> IllegalArgumentException)
	public static final int PIXEL_F32 = 0, PIXEL_U8 = 1, PIXEL_S32 = 2;

	static final Linker LINKER = Linker.nativeLinker();
	static final SymbolLookup LIB = SymbolLookup.libraryLookup(System.mapLibraryName("boofsurf"), Arena.global());

	static MethodHandle h(String name, FunctionDescriptor d) {
		return LINKER.downcallHandle(LIB.find(name).orElseThrow(() -> new UnsatisfiedLinkError(name)), d);
	}

	// struct bs_fh_cfg { float detect_threshold; int extract_radius, max_features_per_scale, initial_sample_size, initial_size,
	//                    scales_per_octave, octaves, scale_step_size; }
	public static final StructLayout FH_CFG = MemoryLayout.structLayout(
			JAVA_FLOAT.withName("detect_threshold"), JAVA_INT.withName("extract_radius"), JAVA_INT.withName("max_features_per_scale"),
			JAVA_INT.withName("initial_sample_size"), JAVA_INT.withName("initial_size"), JAVA_INT.withName("scales_per_octave"),
			JAVA_INT.withName("octaves"), JAVA_INT.withName("scale_step_size"));
	// struct bs_ori_cfg { int variant; double object_radius_to_scale; int radius; double sample_period; int sample_width;
	//                     double weight_sigma; double window_size; }  (natural C alignment: 4 bytes of padding after each int)
	public static final StructLayout ORI_CFG = MemoryLayout.structLayout(
			JAVA_INT.withName("variant"), MemoryLayout.paddingLayout(4), JAVA_DOUBLE.withName("object_radius_to_scale"),
			JAVA_INT.withName("radius"), MemoryLayout.paddingLayout(4), JAVA_DOUBLE.withName("sample_period"),
			JAVA_INT.withName("sample_width"), MemoryLayout.paddingLayout(4), JAVA_DOUBLE.withName("weight_sigma"),
			JAVA_DOUBLE.withName("window_size"));
	// struct bs_desc_cfg { int variant, width_large_grid, width_sub_region, width_sample, use_haar; double weight_sigma;
	//                      int overlap; double sigma_large_grid, sigma_sub_region; }
	public static final StructLayout DESC_CFG = MemoryLayout.structLayout(
			JAVA_INT.withName("variant"), JAVA_INT.withName("width_large_grid"), JAVA_INT.withName("width_sub_region"),
			JAVA_INT.withName("width_sample"), JAVA_INT.withName("use_haar"), MemoryLayout.paddingLayout(4),
			JAVA_DOUBLE.withName("weight_sigma"), JAVA_INT.withName("overlap"), MemoryLayout.paddingLayout(4),
			JAVA_DOUBLE.withName("sigma_large_grid"), JAVA_DOUBLE.withName("sigma_sub_region"));

	public static final MethodHandle LAST_ERROR = h("bs_last_error", FunctionDescriptor.of(ADDRESS));
	public static final MethodHandle DEVICE_COUNT = h("bs_device_count", FunctionDescriptor.of(JAVA_INT, ADDRESS));
	public static final MethodHandle FH_DEFAULT = h("bs_fh_cfg_default", FunctionDescriptor.ofVoid(ADDRESS));
	public static final MethodHandle ORI_DEFAULT = h("bs_ori_cfg_default", FunctionDescriptor.ofVoid(ADDRESS, JAVA_INT));
	public static final MethodHandle DESC_DEFAULT = h("bs_desc_cfg_default", FunctionDescriptor.ofVoid(ADDRESS, JAVA_INT));

	// int bs_surf_create(int device, const bs_fh_cfg*, const bs_ori_cfg*, const bs_desc_cfg*, int pixel_type, bs_surf** out)
	public static final MethodHandle SURF_CREATE = h("bs_surf_create",
			FunctionDescriptor.of(JAVA_INT, JAVA_INT, ADDRESS, ADDRESS, ADDRESS, JAVA_INT, ADDRESS));
	public static final MethodHandle SURF_DESTROY = h("bs_surf_destroy", FunctionDescriptor.of(JAVA_INT, ADDRESS));
	// int bs_surf_detect(bs_surf*, const void* pixels, int start_index, int stride, int width, int height)
	public static final MethodHandle SURF_DETECT = h("bs_surf_detect",
			FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, JAVA_INT, JAVA_INT, JAVA_INT));
	public static final MethodHandle SURF_COUNT = h("bs_surf_count", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS));
	// int bs_surf_read(bs_surf*, int frame, double* xy, double* scale, double* angle, double* desc, uint8_t* white, int32_t* meta, float* response)
	public static final MethodHandle SURF_READ = h("bs_surf_read",
			FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS));
	public static final MethodHandle SURF_DOF = h("bs_surf_dof", FunctionDescriptor.of(JAVA_INT, ADDRESS));

	// int bs_assoc_create(int device, double max_fit_error, int backwards_validation, bs_assoc** out)
	public static final MethodHandle ASSOC_CREATE = h("bs_assoc_create", FunctionDescriptor.of(JAVA_INT, JAVA_INT, JAVA_DOUBLE, JAVA_INT, ADDRESS));
	public static final MethodHandle ASSOC_DESTROY = h("bs_assoc_destroy", FunctionDescriptor.of(JAVA_INT, ADDRESS));
	public static final MethodHandle ASSOC_SET_MAX = h("bs_assoc_set_max_fit_error", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_DOUBLE));
	// int bs_assoc_run(bs_assoc*, const double* src, int n, const double* dst, int m, int dof, int32_t* pairs, double* fit)
	public static final MethodHandle ASSOC_RUN = h("bs_assoc_run",
			FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, ADDRESS, JAVA_INT, JAVA_INT, ADDRESS, ADDRESS));
	// int bs_assoc_matches(pairs, fit, n, m, match_src, match_dst, match_score, n_matches, unassoc_src, n_unassoc_src, unassoc_dst, n_unassoc_dst)
	public static final MethodHandle ASSOC_MATCHES = h("bs_assoc_matches",
			FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS));

	private BoofSurfNative() {}

	/** true when the library loads and reports at least one CUDA device: the factories branch on this. */
	public static boolean available() {
		try (Arena a = Arena.ofConfined()) {
			MemorySegment n = a.allocate(JAVA_INT);
			return (int) DEVICE_COUNT.invokeExact(n) == BS_OK && n.get(JAVA_INT, 0) > 0;
		} catch (Throwable t) {
			return false;
		}
	}

	/** Status code -> exception, with the library's message.  BS_ERR_ARG mirrors Configuration.checkValidity / the reference's
	 *  IllegalArgumentExceptions; anything else is a RuntimeException (there is no CPU fallback behind it). */
	public static void check(int status) {
		if (status == BS_OK) return;
		String msg;
		try {
			MemorySegment p = (MemorySegment) LAST_ERROR.invokeExact();
			msg = p.reinterpret(Long.MAX_VALUE).getString(0);
		} catch (Throwable t) {
			msg = "libboofsurf status " + status;
		}
		if (status == BS_ERR_ARG) throw new IllegalArgumentException(msg);
		throw new RuntimeException("libboofsurf: " + msg + " (status " + status + ")");
	}

	static long unused() { return JAVA_LONG.byteSize(); }
}
