/*
 * GpuBVGraph -- a BVGraph whose successor lists are decoded on a B200 through libwgbv (include/wgbv.h).
 *
 * This is the class a maintainer of webgraph-big would add next to BVGraph.  It is SOURCE ONLY in this repository:
 * the build image has no JDK and none of webgraph-big's dependency jars, so it has never been compiled here.  The
 * binding layer is Panama FFM (JDK 22+); no JNI glue is needed.  INTEGRATION.md explains the mapping.
 *
 * It accepts the files BVGraph accepts (graphclass may name the standard or the big BVGraph, as
 * BVGraph.loadInternal does, BVGraph.java:1491-1492) and is selected like any other graph class:
 * -g GpuBVGraph (GraphClassParser.java:42-52), --spec, or ImmutableGraph.load after editing graphclass.
 */
package it.unimi.dsi.big.webgraph;

import static java.lang.foreign.ValueLayout.ADDRESS;
import static java.lang.foreign.ValueLayout.JAVA_INT;
import static java.lang.foreign.ValueLayout.JAVA_LONG;

import java.io.IOException;
import java.io.InputStream;
import java.io.UncheckedIOException;
import java.lang.foreign.Arena;
import java.lang.foreign.FunctionDescriptor;
import java.lang.foreign.Linker;
import java.lang.foreign.MemorySegment;
import java.lang.foreign.SymbolLookup;
import java.lang.invoke.MethodHandle;
import java.util.NoSuchElementException;

import it.unimi.dsi.fastutil.longs.LongBigArrays;
import it.unimi.dsi.logging.ProgressLogger;

public class GpuBVGraph extends ImmutableGraph {
	/** One downcall handle per entry point of include/wgbv.h that this class uses. */
	static final class Wgbv {
		static final Linker L = Linker.nativeLinker();
		static final SymbolLookup LIB = SymbolLookup.libraryLookup(System.getProperty("wgbv.library", "libwgbv.so"), Arena.global());

		static MethodHandle h(final String name, final FunctionDescriptor d) {
			return L.downcallHandle(LIB.find(name).orElseThrow(), d);
		}

		static final MethodHandle OPEN = h("wgbv_open", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, JAVA_INT, JAVA_INT, ADDRESS));
		static final MethodHandle CLOSE = h("wgbv_close", FunctionDescriptor.ofVoid(ADDRESS));
		static final MethodHandle NUM_NODES = h("wgbv_num_nodes", FunctionDescriptor.of(JAVA_LONG, ADDRESS));
		static final MethodHandle NUM_ARCS = h("wgbv_num_arcs", FunctionDescriptor.of(JAVA_LONG, ADDRESS));
		static final MethodHandle STREAM_CREATE = h("wgbv_stream_create", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS));
		static final MethodHandle STREAM_DESTROY = h("wgbv_stream_destroy", FunctionDescriptor.ofVoid(ADDRESS));
		static final MethodHandle OUTDEGREES = h("wgbv_outdegrees", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS));
		static final MethodHandle DECODE_CSR_HOST = h("wgbv_decode_csr_host",
				FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_LONG, JAVA_LONG, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS));
		static final MethodHandle SUCCESSORS_BATCH_HOST = h("wgbv_successors_batch_host",
				FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS));
		static final MethodHandle LAST_ERROR = h("wgbv_last_error", FunctionDescriptor.of(ADDRESS));
		// Transform.symmetrizeOffline / union on device CSRs, and BVGraph.store for a device CSR (include/wgbv.h)
		static final MethodHandle SYMMETRIZE_CSR = h("wgbv_symmetrize_csr",
			FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS));
		static final MethodHandle UNION_CSR = h("wgbv_union_csr",
			FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS));
		static final MethodHandle ENCODE_CSR = h("wgbv_encode_csr",
			FunctionDescriptor.of(JAVA_INT, JAVA_INT, ADDRESS, JAVA_LONG, JAVA_LONG, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS));
		// arc labels (labelling/BitStreamArcLabelledImmutableGraph.java:65-100): an array parallel to the successors
		static final MethodHandle WRITE_OFFSETS = h("wgbv_write_offsets",
				FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, JAVA_LONG, ADDRESS));
		static final MethodHandle LABELS_OPEN_MEM = h("wgbv_labels_open_mem",
				FunctionDescriptor.of(JAVA_INT, JAVA_INT, JAVA_LONG, JAVA_INT, JAVA_INT, ADDRESS, JAVA_LONG, ADDRESS, JAVA_LONG, ADDRESS));
		static final MethodHandle LABELS_DECODE_HOST = h("wgbv_labels_decode_host",
				FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, JAVA_LONG, ADDRESS, ADDRESS, JAVA_LONG));
		static final MethodHandle LABELS_DECODE_LISTS_HOST = h("wgbv_labels_decode_lists_host",
				FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, JAVA_LONG, ADDRESS, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS));
		static final MethodHandle LABELS_CLOSE = h("wgbv_labels_close", FunctionDescriptor.ofVoid(ADDRESS));

		/** Status code to the exception the reference throws at the same place (include/wgbv.h). */
		static void check(final int rc) {
			if (rc == 0) return;
			String msg;
			try {
				msg = ((MemorySegment)LAST_ERROR.invokeExact()).reinterpret(512).getString(0);
			} catch (final Throwable t) {
				msg = "libwgbv error " + rc;
			}
			switch (rc) {
			case -1: case -6: throw new IllegalArgumentException(msg);            // BVGraph.java:823, 863, 1000, 1128
			case -2: throw new IllegalStateException(msg);                        // BVGraph.java:701
			case -3: throw new UncheckedIOException(new IOException(msg));        // BVGraph.java:1491-1497
			case -4: throw new UnsupportedOperationException(msg);                // BVGraph.java:631, 658, 699, 733, 763, 794
			default: throw new RuntimeException(msg);                             // CUDA failure
			}
		}
	}

	/** The device-resident graph shared by all flyweight copies (BVGraph.java:553-578 shares graphMemory and offsets). */
	private static final class Shared {
		final MemorySegment graph;
		final CharSequence basename;
		final long n, m;
		Shared(final MemorySegment graph, final CharSequence basename, final long n, final long m) {
			this.graph = graph; this.basename = basename; this.n = n; this.m = m;
		}
	}

	/** Nodes decoded per bulk call behind nodeIterator(). */
	public static final int CHUNK_NODES = 1 << 20;

	private final Shared shared;
	private final MemorySegment stream;
	private final Arena arena = Arena.ofShared();

	private GpuBVGraph(final Shared shared) {
		this.shared = shared;
		try {
			final MemorySegment out = arena.allocate(ADDRESS);
			Wgbv.check((int)Wgbv.STREAM_CREATE.invokeExact(shared.graph, MemorySegment.NULL, out));
			stream = out.get(ADDRESS, 0);
		} catch (final RuntimeException e) {
			throw e;
		} catch (final Throwable t) {
			throw new RuntimeException(t);
		}
	}

	/* The static factories the reflective loader looks for (ImmutableGraph.java:104-121, 674-713). */
	public static GpuBVGraph load(final CharSequence basename) throws IOException { return load(basename, null); }
	public static GpuBVGraph load(final CharSequence basename, final ProgressLogger pl) throws IOException {
		try (Arena a = Arena.ofConfined()) {
			final MemorySegment out = a.allocate(ADDRESS);
			final int device = Integer.getInteger("wgbv.device", 0);
			Wgbv.check((int)Wgbv.OPEN.invokeExact(a.allocateFrom(basename.toString()), device, 0, 1, out));
			final MemorySegment g = out.get(ADDRESS, 0);
			return new GpuBVGraph(new Shared(g, basename, (long)Wgbv.NUM_NODES.invokeExact(g), (long)Wgbv.NUM_ARCS.invokeExact(g)));
		} catch (final UncheckedIOException e) {
			throw e.getCause();
		} catch (final RuntimeException e) {
			throw e;
		} catch (final Throwable t) {
			throw new RuntimeException(t);
		}
	}
	public static GpuBVGraph loadMapped(final CharSequence basename) throws IOException { return load(basename, null); }
	public static GpuBVGraph loadMapped(final CharSequence basename, final ProgressLogger pl) throws IOException { return load(basename, pl); }
	public static GpuBVGraph loadOffline(final CharSequence basename) throws IOException { return load(basename, null); }
	public static GpuBVGraph loadOffline(final CharSequence basename, final ProgressLogger pl) throws IOException { return load(basename, pl); }
	public static GpuBVGraph loadOnce(final InputStream is) { throw new UnsupportedOperationException("GpuBVGraph needs random access to the bit stream"); }

	@Override public long numNodes() { return shared.n; }
	@Override public long numArcs() { return shared.m; }
	@Override public boolean randomAccess() { return true; }
	@Override public boolean hasCopiableIterators() { return true; }
	@Override public CharSequence basename() { return shared.basename; }
	@Override public GpuBVGraph copy() { return new GpuBVGraph(shared); }

	@Override
	public long outdegree(final long x) {
		if (x < 0 || x >= shared.n) throw new IllegalArgumentException("Node index out of range: " + x);
		try (Arena a = Arena.ofConfined()) {
			final MemorySegment id = a.allocate(JAVA_LONG), d = a.allocate(JAVA_LONG);
			id.set(JAVA_LONG, 0, x);
			Wgbv.check((int)Wgbv.OUTDEGREES.invokeExact(stream, id, 1L, d));
			return d.get(JAVA_LONG, 0);
		} catch (final RuntimeException e) {
			throw e;
		} catch (final Throwable t) {
			throw new RuntimeException(t);
		}
	}

	/** The successors of x in an off-heap segment of outdegree(x) longs (owned by the returned arena-less copy). */
	private long[][] successorsOf(final long x) {
		if (x < 0 || x >= shared.n) throw new IllegalArgumentException("Node index out of range: " + x);
		try (Arena a = Arena.ofConfined()) {
			final MemorySegment id = a.allocate(JAVA_LONG), rp = a.allocate(JAVA_LONG, 2), arcs = a.allocate(JAVA_LONG);
			id.set(JAVA_LONG, 0, x);
			Wgbv.check((int)Wgbv.SUCCESSORS_BATCH_HOST.invokeExact(stream, id, 1L, rp, MemorySegment.NULL, 0L, arcs));
			final long d = arcs.get(JAVA_LONG, 0);
			final MemorySegment col = a.allocate(JAVA_LONG, Math.max(d, 1));
			Wgbv.check((int)Wgbv.SUCCESSORS_BATCH_HOST.invokeExact(stream, id, 1L, rp, col, d, arcs));
			final long[][] big = LongBigArrays.newBigArray(d);
			for (long i = 0; i < d; i++) it.unimi.dsi.fastutil.BigArrays.set(big, i, col.getAtIndex(JAVA_LONG, i));
			return big;
		} catch (final RuntimeException e) {
			throw e;
		} catch (final Throwable t) {
			throw new RuntimeException(t);
		}
	}

	@Override public long[][] successorBigArray(final long x) { return successorsOf(x); }

	@Override
	public LazyLongIterator successors(final long x) {
		final long[][] a = successorsOf(x);
		return LazyLongIterators.wrap(a, it.unimi.dsi.fastutil.BigArrays.length(a));
	}

	/** The bulk call: CSR of [from, to) in two off-heap segments (row pointers, successors) allocated in {@code a}. */
	public MemorySegment[] decodeToCSR(final long from, final long to, final Arena a) {
		try {
			final MemorySegment rp = a.allocate(JAVA_LONG, to - from + 1), arcs = a.allocate(JAVA_LONG);
			Wgbv.check((int)Wgbv.DECODE_CSR_HOST.invokeExact(stream, from, to, rp, MemorySegment.NULL, 0L, arcs));
			final long m = arcs.get(JAVA_LONG, 0);
			final MemorySegment col = a.allocate(JAVA_LONG, Math.max(m, 1));
			Wgbv.check((int)Wgbv.DECODE_CSR_HOST.invokeExact(stream, from, to, rp, col, m, arcs));
			return new MemorySegment[] { rp, col.asSlice(0, m * Long.BYTES) };
		} catch (final RuntimeException e) {
			throw e;
		} catch (final Throwable t) {
			throw new RuntimeException(t);
		}
	}

	@Override
	/** List-valued labels (FixedWidthIntListLabel, FixedWidthLongListLabel) of the arcs of [from, to): returns { listPtr
	 *  (arcs + 1 longs), values }; the list of arc j is values[listPtr[j] .. listPtr[j + 1]). */
	public MemorySegment[] decodeLabelLists(final MemorySegment labels, final long from, final long to, final MemorySegment rowptr, final Arena a) {
		final long arcs = rowptr.getAtIndex(JAVA_LONG, to - from) - rowptr.getAtIndex(JAVA_LONG, 0);
		final MemorySegment listPtr = a.allocate((arcs + 1) * 8, 8), total = a.allocate(8, 8);
		try {
			Wgbv.check((int)Wgbv.LABELS_DECODE_LISTS_HOST.invokeExact(stream, labels, from, to, rowptr, listPtr, MemorySegment.NULL, 0L, total));
			final long count = total.get(JAVA_LONG, 0);
			final MemorySegment values = a.allocate(Math.max(count, 1) * 8, 8);
			Wgbv.check((int)Wgbv.LABELS_DECODE_LISTS_HOST.invokeExact(stream, labels, from, to, rowptr, listPtr, values, count, total));
			return new MemorySegment[] { listPtr, values };
		} catch (final RuntimeException e) {
			throw e;
		} catch (final Throwable t) {
			throw new RuntimeException(t);
		}
	}

	/** Labels of the arcs of [from, to), parallel to the successors that {@link #decodeToCSR} returns for the same range:
	 *  labels = handle from wgbv_labels_open_mem over the .labels / .labeloffsets of a BitStreamArcLabelledImmutableGraph
	 *  whose underlying graph is this one; rowptr = first segment returned by decodeToCSR(from, to, a). */
	public MemorySegment decodeLabels(final MemorySegment labels, final long from, final long to, final MemorySegment rowptr, final Arena a) {
		final long arcs = rowptr.getAtIndex(JAVA_LONG, to - from) - rowptr.getAtIndex(JAVA_LONG, 0);
		final MemorySegment out = a.allocate(Math.max(arcs, 1) * 8, 8);
		try {
			Wgbv.check((int)Wgbv.LABELS_DECODE_HOST.invokeExact(stream, labels, from, to, rowptr, out, arcs));
		} catch (final RuntimeException e) {
			throw e;
		} catch (final Throwable t) {
			throw new RuntimeException(t);
		}
		return out;
	}

	public NodeIterator nodeIterator(final long from) {
		if (from < 0 || from > shared.n) throw new IllegalArgumentException("Node index out of range: " + from);
		return new GpuNodeIterator(this, from, Long.MAX_VALUE);
	}

	/** BVGraphNodeIterator (BVGraph.java:1100-1245) over chunked bulk decodes. */
	private static final class GpuNodeIterator extends NodeIterator {
		private final GpuBVGraph g;
		private final long from, limit;
		private long curr, c0, c1;
		private Arena chunk;
		private MemorySegment rp, col;

		GpuNodeIterator(final GpuBVGraph g, final long from, final long upperBound) {
			this.g = g; this.from = from; this.curr = from - 1; this.c0 = this.c1 = from;
			this.limit = Math.min(upperBound, g.shared.n) - 1;   // hasNextLimit, BVGraph.java:1147
		}
		@Override public boolean hasNext() { return curr < limit; }
		@Override public long nextLong() {
			if (!hasNext()) throw new NoSuchElementException();      // BVGraph.java:1165
			if (++curr >= c1) {
				if (chunk != null) chunk.close();
				chunk = Arena.ofConfined();
				c0 = curr; c1 = Math.min(c0 + CHUNK_NODES, limit + 1);
				final MemorySegment[] csr = g.decodeToCSR(c0, c1, chunk);
				rp = csr[0]; col = csr[1];
			}
			return curr;
		}
		@Override public long outdegree() {
			if (curr == from - 1) throw new IllegalStateException(); // BVGraph.java:1207
			return rp.getAtIndex(JAVA_LONG, curr - c0 + 1) - rp.getAtIndex(JAVA_LONG, curr - c0);
		}
		@Override public long[][] successorBigArray() {
			final long a = rp.getAtIndex(JAVA_LONG, curr - c0), d = outdegree();
			final long[][] big = LongBigArrays.newBigArray(d);
			for (long i = 0; i < d; i++) it.unimi.dsi.fastutil.BigArrays.set(big, i, col.getAtIndex(JAVA_LONG, a + i));
			return big;
		}
		@Override public NodeIterator copy(final long upperBound) { return new GpuNodeIterator(g.copy(), curr + 1, upperBound); } // BVGraph.java:1223-1229
	}

	/** Releases this copy's stream; the device-resident graph goes with the last copy (call from a Cleaner in production). */
	public void close() {
		try {
			Wgbv.STREAM_DESTROY.invokeExact(stream);
		} catch (final Throwable t) {
			throw new RuntimeException(t);
		}
		arena.close();
	}
}
