// Gaussian kernel for sm_100a (Sobel and Canny are in pf_edges.cu).
//
// A "column-march" stencil: a thread owns a few adjacent columns of a
// horizontal strip of the page and walks down the rows, keeping the vertical
// window of the stencil in registers.  Every RGBA8 pixel is read from HBM
// once (row halos come out of L1/L2) and written once: 8 algorithmic bytes
// per pixel.
//
// Reference semantics (upstream src/pillowfight/util.c pf_dbl_matrix_convolution,
// _gaussian.c -- not in the mount):
// true convolution with the kernel centre at size/2, taps outside the page
// skipped.  Skipping a tap and adding +0.0 give the same double, so the
// kernels zero-pad.
#include <cstdlib>

#include "pf_march.cuh"

namespace pfb {

namespace {

constexpr int kBlock = 128;

__device__ __forceinline__ uint32_t load_px(const uint32_t *row, int x, int w)
{
	return (row != nullptr && x >= 0 && x < w) ? __ldg(row + x) : 0u;
}

// ---------------------------------------------------------------------------
// Gaussian
// ---------------------------------------------------------------------------

// uint8 -> double without the conversion pipe: 2^52 + v is exact, minus 2^52.
__device__ __forceinline__ double u8_to_double(uint32_t v)
{
	return __dadd_rn(__hiloint2double(0x43300000, (int)v), -4503599627370496.0);
}

// trunc(v) for 0 <= v < 2^31 with one DADD (round toward zero into 2^52's ulp)
__device__ __forceinline__ int trunc_nonneg(double v)
{
	return __double2loint(__dadd_rz(v, 4503599627370496.0));
}

template <int N>
struct TapsN {
	double t[N];
};

// Strict-order separable convolution in FP64.  Each thread produces PX
// adjacent pixels per row.  N taps, centre N/2.  The horizontal sums of the
// last N rows sit in a register ring; __dmul_rn/__dadd_rn keep the compiler
// from contracting to FMA so that every rounding matches the C reference.
template <int N, int PX>
__global__ void __launch_bounds__(kBlock)
gaussian_march_kernel(const uint32_t *__restrict__ in, uint32_t *__restrict__ out,
		int w, int h, int y_begin, int y_end, int rows_per_strip, const TapsN<N> tp)
{
	pdl_wait();
	constexpr int C = N / 2;             // taps reach x+C ... x-(N-1-C)
	constexpr int L = N - 1 - C;         // left / up reach
	constexpr int NPIX = PX + N - 1;     // input pixels a thread needs per row

	const int x0 = (blockIdx.x * kBlock + threadIdx.x) * PX;
	if (x0 >= w)
		return;
	const size_t page = (size_t)blockIdx.z * w * h;
	in += page;
	out += page;
	const int ys = y_begin + blockIdx.y * rows_per_strip;
	const int ye = min(y_end, ys + rows_per_strip);

	double taps[N];
#pragma unroll
	for (int k = 0; k < N; k++)
		taps[k] = tp.t[k];

	double ring[N][PX * 3];

	// row r of the horizontal pass lands in ring slot (r - (ys - L)) % N
	for (int r0 = ys - L; r0 < ye + C; r0 += N) {
#pragma unroll
		for (int j = 0; j < N; j++) {
			const int r = r0 + j;
			if (r < ye + C) {
				// ---- horizontal pass of row r into ring[j] ----
				const uint32_t *row = (r >= 0 && r < h) ? in + (size_t)r * w : nullptr;
				double px[NPIX][3];
#pragma unroll
				for (int i = 0; i < NPIX; i++) {
					const uint32_t p = load_px(row, x0 - L + i, w);
					px[i][0] = u8_to_double(p & 0xFF);
					px[i][1] = u8_to_double((p >> 8) & 0xFF);
					px[i][2] = u8_to_double((p >> 16) & 0xFF);
				}
#pragma unroll
				for (int o = 0; o < PX; o++) {
#pragma unroll
					for (int c = 0; c < 3; c++) {
						// tap k multiplies pixel x - k + C = index (o + L) - k + C
						double v = __dmul_rn(px[o + N - 1][c], taps[0]);
#pragma unroll
						for (int k = 1; k < N; k++)
							v = __dadd_rn(v, __dmul_rn(px[o + N - 1 - k][c], taps[k]));
						ring[j][o * 3 + c] = v;
					}
				}
				// ---- vertical pass for output row y = r - C ----
				const int y = r - C;
				if (y >= ys) {
					uint32_t res[PX];
#pragma unroll
					for (int o = 0; o < PX; o++) {
						uint32_t rgba = 0xFF000000u;
#pragma unroll
						for (int c = 0; c < 3; c++) {
							// tap k multiplies row y - k + C = r - k: slot (j - k) mod N
							double v = __dmul_rn(ring[j][o * 3 + c], taps[0]);
#pragma unroll
							for (int k = 1; k < N; k++)
								v = __dadd_rn(v, __dmul_rn(
										ring[(j - k + N) % N][o * 3 + c], taps[k]));
							const int q = min(trunc_nonneg(v), 255);
							rgba |= (uint32_t)q << (8 * c);
						}
						res[o] = rgba;
					}
					uint32_t *orow = out + (size_t)y * w;
					if (PX == 2 && x0 + 1 < w && ((w & 1) == 0)) {
						*reinterpret_cast<uint2 *>(orow + x0) = make_uint2(res[0], res[PX - 1]);
					} else {
#pragma unroll
						for (int o = 0; o < PX; o++)
							if (x0 + o < w)
								orow[x0 + o] = res[o];
					}
				}
			}
		}
	}
}

// One pixel by the reference's exact operation order (N*N multiply-adds, no
// FMA): the slow path of the fast kernel below and the body of the generic one.
__device__ __noinline__ uint32_t gaussian_exact_pixel(const uint32_t *__restrict__ in,
		int w, int h, int x, int y, const double *__restrict__ t, int n)
{
	const int c0 = n / 2;
	double acc[3] = { 0.0, 0.0, 0.0 };
	bool first_row = true;
	for (int ky = 0; ky < n; ky++) {
		const int iy = y - ky + c0;
		if (iy < 0 || iy >= h)
			continue;
		const uint32_t *row = in + (size_t)iy * w;
		double hv[3] = { 0.0, 0.0, 0.0 };
		bool first = true;
		for (int kx = 0; kx < n; kx++) {
			const int ix = x - kx + c0;
			if (ix < 0 || ix >= w)
				continue;
			const uint32_t p = __ldg(row + ix);
			const double tk = t[kx];
			for (int c = 0; c < 3; c++) {
				const double m = __dmul_rn(u8_to_double((p >> (8 * c)) & 0xFF), tk);
				hv[c] = first ? m : __dadd_rn(hv[c], m);
			}
			first = false;
		}
		const double tv = t[ky];
		for (int c = 0; c < 3; c++) {
			const double m = __dmul_rn(hv[c], tv);
			acc[c] = first_row ? m : __dadd_rn(acc[c], m);
		}
		first_row = false;
	}
	uint32_t rgba = 0xFF000000u;
	for (int c = 0; c < 3; c++)
		rgba |= (uint32_t)min(trunc_nonneg(acc[c]), 255) << (8 * c);
	return rgba;
}

// The same for a compile-time tap count, with all N*N loads issued before the
// arithmetic (one memory latency instead of N): the fast kernel's slow path.
template <int N>
__device__ __noinline__ uint32_t gaussian_exact_pixel_n(const uint32_t *__restrict__ in,
		int w, int h, int x, int y, const double *__restrict__ t)
{
	constexpr int C = N / 2;
	uint32_t p[N][N];
#pragma unroll
	for (int ky = 0; ky < N; ky++) {
		const int iy = y - ky + C;
#pragma unroll
		for (int kx = 0; kx < N; kx++) {
			const int ix = x - kx + C;
			p[ky][kx] = (iy >= 0 && iy < h && ix >= 0 && ix < w) ? __ldg(in + (size_t)iy * w + ix) : 0u;
		}
	}
	// a tap outside the page is skipped by the reference; adding +0.0 instead
	// gives the same double (all terms are >= +0)
	uint32_t rgba = 0xFF000000u;
#pragma unroll
	for (int c = 0; c < 3; c++) {
		double acc = 0.0;
#pragma unroll
		for (int ky = 0; ky < N; ky++) {
			double hv = 0.0;
#pragma unroll
			for (int kx = 0; kx < N; kx++)
				hv = __dadd_rn(hv, __dmul_rn(u8_to_double((p[ky][kx] >> (8 * c)) & 0xFF), t[kx]));
			acc = __dadd_rn(acc, __dmul_rn(hv, t[ky]));
		}
		rgba |= (uint32_t)min(trunc_nonneg(acc), 255) << (8 * c);
	}
	return rgba;
}

// Any tap count: one thread per pixel.
__global__ void __launch_bounds__(kBlock)
gaussian_generic_kernel(const uint32_t *__restrict__ in, uint32_t *__restrict__ out,
		int w, int h, int y_begin, int y_end, const Taps tp)
{
	pdl_wait();
	const int x = blockIdx.x * kBlock + threadIdx.x;
	if (x >= w)
		return;
	const size_t page = (size_t)blockIdx.z * w * h;
	// grid.y is capped at 65535: taller row ranges wrap around
	for (int y = y_begin + blockIdx.y; y < y_end; y += gridDim.y)
		out[page + (size_t)y * w + x] = gaussian_exact_pixel(in + page, w, h, x, y, tp.t, tp.n);
}

// ---------------------------------------------------------------------------
// Fast kernel for 3 and 5 taps: FP64 with FMA, exactness by a guard band.
//
// The reference rounds after every multiply and every add; the result is then
// truncated to an integer, so its last bits only matter when the value lies
// within rounding distance of an integer (which happens for whole regions of
// flat colour, where every value is "k * sum(taps)^2").  This kernel computes
// the same sums with FMAs (about half the FP64 instructions; error ~1e-13),
// truncates with one round-toward-zero add that also exposes 24 fraction
// bits, and only when a channel lies within 2e-6 of an integer does it ask
// for the reference's own rounding: from a 256-entry table if the pixel's
// whole N x N window holds one colour (flat areas: most of a document page,
// and every pixel of them is such a case), from gaussian_exact_pixel_n
// otherwise (a few pixels per million).  Results are bit-identical to the
// reference and the run time does not depend on the page's content.
// ---------------------------------------------------------------------------

// v + 2^28 rounded toward zero leaves floor(v * 2^24) in the low word for
// 0 <= v < 256: the integer part is the top byte, 24 fraction bits below it.
constexpr double kFracMagic = 268435456.0;         // 2^28
constexpr uint32_t kFracMask = (1u << 24) - 1;
constexpr uint32_t kFracGuard = 32;                // 32 * 2^-24 ~ 2e-6

template <int N>
struct GaussRing {
	double h[N][12];       // horizontal sums of the last N rows: 4 columns x 3 channels
	// bit 24 + o: the N pixels under output column o of that row share one
	// colour; bits 0..23: that colour (windows of two columns overlap, so
	// there is only one), 0 if no column is flat
	uint32_t flat[N];
};

// Taps plus the reference's result for a window of N x N equal values v, in
// its own operation order (computed on the host, launch_gaussian_fast).  On a
// flat area every value is v * sum(taps)^2, within rounding of the integer v,
// so every pixel is "risky": this table answers for it without the N*N loads
// and ~7 N*N strict operations of the exact path.  Flat areas are most of a
// document page.
template <int N>
struct FastTaps {
	double t[N];        // the taps (exact path)
	double th[N];       // * 2^1000: horizontal pass over subnormal pixel values
	double tv[N];       // * 2^34: vertical pass, back to the true scale
	unsigned char flat[256];
};

// Step J of N (s = steps since the top of the strip): the row at the head of
// the N-deep row queue is filtered horizontally into ring slot J, its queue
// slot is refilled with the row N below, and the output row C rows above
// leaves.  The first N-1 steps of a strip only fill the ring.
template <int N, int J>
__device__ __forceinline__ void gauss_step(const uint32_t *__restrict__ in,
		const StripGeom &g, const WarpPlace &wp, int s, const FastTaps<N> &tp,
		const unsigned char *__restrict__ flat_lut,
		RowReader<true> &rd, RowWriter<true> &wr, uint4 (&q)[N], GaussRing<N> &ring)
{
	constexpr int C = N / 2;
	const double (&t)[N] = tp.t, (&th)[N] = tp.th, (&tv)[N] = tp.tv;
	const uint4 cur = q[J];
	q[J] = rd.next(g);
	// pixels x0-C .. x0+3+C: own four plus C from each neighbouring lane
	uint32_t px[4 + 2 * C];
	px[C] = cur.x;
	px[C + 1] = cur.y;
	px[C + 2] = cur.z;
	px[C + 3] = cur.w;
	if (C >= 1) {
		px[C - 1] = __shfl_up_sync(kFull, cur.w, 1);
		px[C + 4] = __shfl_down_sync(kFull, cur.x, 1);
	}
	if (C >= 2) {
		px[C - 2] = __shfl_up_sync(kFull, cur.z, 1);
		px[C + 5] = __shfl_down_sync(kFull, cur.y, 1);
	}
	double v[4 + 2 * C][3];
#pragma unroll
	for (int i = 0; i < 4 + 2 * C; i++) {
#pragma unroll
		for (int c = 0; c < 3; c++) {
			// byte c as bits 8..15 of the high word of a double whose exponent
			// field is 0: the subnormal b * 2^-1034, exactly, by one PRMT
			// instead of a trip through the conversion unit (16 lanes/clk).
			// The horizontal taps carry 2^1000 and the vertical ones 2^34, so
			// the sums are the unscaled ones times a power of two throughout.
			v[i][c] = __hiloint2double((int)__byte_perm(px[i], 0u, 0x4404 + (c << 4)), 0);
		}
	}
	{
		// which of the four N-pixel row segments hold one colour (alpha ignored)
		uint32_t d[3 + 2 * C];
#pragma unroll
		for (int i = 0; i < 3 + 2 * C; i++)
			d[i] = px[i] ^ px[i + 1];
		uint32_t word = 0;
#pragma unroll
		for (int o = 0; o < 4; o++) {
			uint32_t any = 0;
#pragma unroll
			for (int i = 0; i < 2 * C; i++)
				any |= d[o + i];
			if (!(any & 0x00FFFFFFu)) {
				const uint32_t colour = px[o + C] & 0x00FFFFFFu;
				if ((word & 0x00FFFFFFu) != colour)
					word = 0;          // (3 taps: columns 0 and 3 do not overlap)
				word |= colour | (0x01000000u << o);
			}
		}
		ring.flat[J] = word;
	}
	double (&hrow)[12] = ring.h[J];
#pragma unroll
	for (int o = 0; o < 4; o++) {
#pragma unroll
		for (int c = 0; c < 3; c++) {
			// tap k multiplies pixel x - k + C, i.e. px index o + 2C - k
			double a = v[o + 2 * C][c] * th[0];
#pragma unroll
			for (int k = 1; k < N; k++)
				a = fma(v[o + 2 * C - k][c], th[k], a);
			hrow[o * 3 + c] = a;
		}
	}
	if (s < N - 1)
		return;
	const int y = wp.ys + s - (N - 1);
	uint32_t res[4];
	unsigned risky = 0;           // bit o: pixel o has a channel too close to an integer
#pragma unroll
	for (int o = 0; o < 4; o++) {
		uint32_t fx[3];
#pragma unroll
		for (int c = 0; c < 3; c++) {
			// tap k multiplies row y - k + C, which sits in slot (J - k) mod N.
			// The sum is built on top of the 2^28 offset (each partial sum is
			// then rounded to 2^-24, N-1 times: far inside the guard band) and
			// the last FMA rounds toward zero, which is the truncation.
			double a = fma(ring.h[J][o * 3 + c], tv[0], kFracMagic);
#pragma unroll
			for (int k = 1; k < N - 1; k++)
				a = fma(ring.h[(J - k + N) % N][o * 3 + c], tv[k], a);
			if (N > 1)
				a = __fma_rz(ring.h[(J - (N - 1) + N) % N][o * 3 + c], tv[N - 1], a);
			fx[c] = (uint32_t)__double2loint(a);
		}
		// (fraction + guard) mod 2^24 < 2*guard  <=>  within guard of an integer
		const uint32_t nearest = min(min((fx[0] + kFracGuard) & kFracMask,
				(fx[1] + kFracGuard) & kFracMask), (fx[2] + kFracGuard) & kFracMask);
		if (nearest < 2 * kFracGuard)
			risky |= 1u << o;
		// the three top bytes are R, G, B (the sums never reach 256)
		res[o] = __byte_perm(__byte_perm(fx[0], fx[1], 0x4473), fx[2], 0x4710) | 0xFF000000u;
	}
	if (risky) {
		// columns whose N x N window holds one colour: the tabulated result
		uint32_t cols = ring.flat[0];
		bool same = true;
#pragma unroll
		for (int k = 1; k < N; k++) {
			cols &= ring.flat[k];
			same = same && ((ring.flat[k] ^ ring.flat[0]) & 0x00FFFFFFu) == 0;
		}
		cols = same ? (cols >> 24) & risky : 0u;
		if (cols) {
			const uint32_t v = ring.flat[0];
			const uint32_t px_flat = 0xFF000000u | flat_lut[v & 0xFF]
				| (flat_lut[(v >> 8) & 0xFF] << 8) | (flat_lut[(v >> 16) & 0xFF] << 16);
#pragma unroll
			for (int o = 0; o < 4; o++)
				if (cols & (1u << o))
					res[o] = px_flat;
			risky &= ~cols;
		}
	}
	if (risky && wr.lane_out && y < wp.ye) {
		// redo those pixels in the reference's exact operation order
		double tt[N];
#pragma unroll
		for (int k = 0; k < N; k++)
			tt[k] = t[k];
#pragma unroll
		for (int o = 0; o < 4; o++)
			if (risky & (1u << o))
				res[o] = gaussian_exact_pixel_n<N>(in, g.w, g.h, wp.x0 + o, y, tt);
	}
	wr.put(g, wp.ye, res);
}

template <int N, int J>
struct GaussSteps {
	static __device__ __forceinline__ void run(const uint32_t *__restrict__ in,
			const StripGeom &g, const WarpPlace &wp, int s, const FastTaps<N> &tp,
			const unsigned char *__restrict__ flat_lut,
			RowReader<true> &rd, RowWriter<true> &wr, uint4 (&q)[N], GaussRing<N> &ring)
	{
		gauss_step<N, J>(in, g, wp, s + J, tp, flat_lut, rd, wr, q, ring);
		GaussSteps<N, J + 1>::run(in, g, wp, s, tp, flat_lut, rd, wr, q, ring);
	}
};

template <int N>
struct GaussSteps<N, N> {
	static __device__ __forceinline__ void run(const uint32_t *__restrict__,
			const StripGeom &, const WarpPlace &, int, const FastTaps<N> &,
			const unsigned char *__restrict__,
			RowReader<true> &, RowWriter<true> &, uint4 (&)[N], GaussRing<N> &)
	{
	}
};

// Needs rows that can be read as uint4 (width a multiple of 4, aligned pages);
// other shapes take the strict kernel.
template <int N>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
gaussian_fast_kernel(const uint32_t *__restrict__ in, uint32_t *__restrict__ out,
		const StripGeom g, const FastTaps<N> tp)
{
	__shared__ unsigned char flat_lut[256];
	pdl_wait();
	for (int i = threadIdx.x; i < 256; i += blockDim.x)
		flat_lut[i] = tp.flat[i];
	__syncthreads();
	constexpr int C = N / 2;
	const WarpPlace wp = place_warp(g);
	if (!wp.valid)
		return;
	const size_t page = (size_t)blockIdx.y * g.w * g.h;
	in += page;
	out += page;
	RowReader<true> rd;
	RowWriter<true> wr;
	rd.init(in, g, wp.ys - (N - 1 - C), wp.x0);
	wr.init(out, g, wp.ys, wp.x0, wp.out_lane);
	GaussRing<N> ring;
	uint4 q[N];
#pragma unroll
	for (int i = 0; i < N; i++)
		q[i] = rd.next(g);
	// rows_per_strip + N - 1 steps per strip, a multiple of N
	const int steps = g.rows_per_strip + N - 1;
	for (int s = 0; s < steps; s += N)
		GaussSteps<N, 0>::run(in, g, wp, s, tp, flat_lut, rd, wr, q, ring);
}

int strips_for(int w_threads_blocks, int h, int pages, int blocks_per_sm, int min_rows)
{
	// about one full wave of blocks, but strips no shorter than min_rows
	const long want = (long)sm_count() * blocks_per_sm;
	long strips = want / ((long)w_threads_blocks * pages);
	if (strips < 1)
		strips = 1;
	int rows = (int)((h + strips - 1) / strips);
	if (rows < min_rows)
		rows = min_rows;
	return rows;
}

}  // namespace

// ---------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------

template <int N>
static void launch_gaussian_fast(const uint32_t *in, uint32_t *out, int w, int h,
		int pages, const Taps &taps, cudaStream_t st, int y_begin, int y_end)
{
	// the table costs ~13k strict operations on the host: kept per thread for
	// as long as the taps stay the same (the usual case: one sigma per run)
	static thread_local FastTaps<N> tp;
	static thread_local bool tp_valid = false;
	bool same = tp_valid;
	for (int k = 0; k < N && same; k++)
		same = tp.t[k] == taps.t[k];
	if (!same) {
		for (int k = 0; k < N; k++) {
			tp.t[k] = taps.t[k];
			tp.th[k] = taps.t[k] * 0x1p+1000;
			tp.tv[k] = taps.t[k] * 0x1p+34;
		}
		for (int v = 0; v < 256; v++) {
			// the reference's loops on a window of equal values (upstream util.c
			// pf_dbl_matrix_convolution, twice; oracle/pf_oracle.c): every product
			// and every sum rounded, horizontal pass first
			volatile double acc = 0.0;
			for (int ky = 0; ky < N; ky++) {
				volatile double hv = 0.0;
				for (int kx = 0; kx < N; kx++) {
					volatile double prod = (double)v * taps.t[kx];
					hv = hv + prod;
				}
				volatile double prod = hv * taps.t[ky];
				acc = acc + prod;
			}
			const int r = (int)acc;
			tp.flat[v] = (unsigned char)(r > 255 ? 255 : r);
		}
		tp_valid = true;
	}
	const StripGeom g = make_geom(gaussian_fast_kernel<N>, in, out, w, h, pages, N - 1,
			y_begin, y_end, N);
	const int warps = g.warp_cols * g.strips;
	dim3 grid((warps + kWarpsPerBlock - 1) / kWarpsPerBlock, pages);
	launch_pdl(gaussian_fast_kernel<N>, grid, dim3(kWarpsPerBlock * 32), 0, st, in, out, g, tp);
}

template <int N>
static void launch_gaussian_n(const uint32_t *in, uint32_t *out, int w, int h,
		int pages, const Taps &taps, cudaStream_t st, int y_begin, int y_end)
{
	TapsN<N> tp;
	for (int k = 0; k < N; k++)
		tp.t[k] = taps.t[k];
	constexpr int PX = 2;
	const int bx = ((w + PX - 1) / PX + kBlock - 1) / kBlock;
	const int nrows = y_end - y_begin;
	const int rows = strips_for(bx, nrows, pages, 4, 8 * N);
	dim3 grid(bx, (nrows + rows - 1) / rows, pages);
	launch_pdl(gaussian_march_kernel<N, PX>, grid, dim3(kBlock), 0, st, in, out, w, h, y_begin, y_end,
			rows, tp);
}

int launch_gaussian(const uint32_t *in, uint32_t *out, int w, int h, int pages,
		const Taps &taps, cudaStream_t st, int y_begin, int y_end)
{
	if (y_end < 0)
		y_end = h;
	if (y_begin < 0 || y_begin >= y_end || y_end > h)
		return fail(PF_B200_EINVAL, "gaussian: bad row range [%d, %d) of %d", y_begin, y_end, h);
	if (taps.n < 1 || taps.n > kMaxTaps)
		return fail(PF_B200_EINVAL, "gaussian: nb_stddev must be in 1..%d", kMaxTaps);
	// PF_B200_GAUSS_STRICT=1 keeps the all-strict kernel for 3 and 5 taps too
	static const bool strict_env = getenv("PF_B200_GAUSS_STRICT") && atoi(getenv("PF_B200_GAUSS_STRICT"));
	const bool strict = strict_env || (w % 4 != 0) || (((uintptr_t)in | (uintptr_t)out) % 16 != 0);
	switch (strict ? 0 : taps.n) {
	case 3: launch_gaussian_fast<3>(in, out, w, h, pages, taps, st, y_begin, y_end); break;
	case 5: launch_gaussian_fast<5>(in, out, w, h, pages, taps, st, y_begin, y_end); break;
	default: break;
	}
	if (!strict && (taps.n == 3 || taps.n == 5)) {
		PFB_CUDA(cudaGetLastError());
		return PF_B200_OK;
	}
	switch (taps.n) {
	case 3: launch_gaussian_n<3>(in, out, w, h, pages, taps, st, y_begin, y_end); break;
	case 5: launch_gaussian_n<5>(in, out, w, h, pages, taps, st, y_begin, y_end); break;
	case 7: launch_gaussian_n<7>(in, out, w, h, pages, taps, st, y_begin, y_end); break;
	case 9: launch_gaussian_n<9>(in, out, w, h, pages, taps, st, y_begin, y_end); break;
	default: {
		const int nrows = y_end - y_begin;
		dim3 grid((w + kBlock - 1) / kBlock, nrows < 65535 ? nrows : 65535, pages);
		launch_pdl(gaussian_generic_kernel, grid, dim3(kBlock), 0, st, in, out, w, h, y_begin, y_end, taps);
	}
	}
	PFB_CUDA(cudaGetLastError());
	return PF_B200_OK;
}

}  // namespace pfb
