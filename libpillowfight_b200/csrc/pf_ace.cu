// Automatic Color Equalization for sm_100a.
//
// Reference loop (upstream src/pillowfight/_ace.c, ace_thread_adjustment; the
// file is not in the mount): for every pixel p and
// every sample k farther than height/5,
//     score[c] += saturate(slope * (p[c] - sample_k[c]), +-limit) / dist(p, k)
//     denom    += limit / dist(p, k)
// then score[c] /= denom, a page-wide min/max per channel, and a linear
// rescale to 0..255.
//
// B200 formulation.  With w_k = 1/dist and L = limit/slope,
//     score[c] = (slope/limit) * sum_k w_k * clamp(p[c]-s_k[c], -L, L) / sum_k w_k
// so a common factor on w cancels.  The N samples are one table for the whole
// page: a block gathers the sample pixels once into shared memory, and the
// inner loop has no global loads at all.  Per pixel-sample pair the FP64 pipe
// does: d^2 (1 add), three accumulating FMAs and one for the denominator (4);
// the weight is MUFU.RSQ64H's 1/sqrt (optionally one Newton step, 3 more).
// The clamp is one integer instruction (VIADDMNMX.RELU), and its result
// u = c + L is used as it stands as the high word of a subnormal double
// (u * 2^-1042, exact and linear in u), so neither a conversion nor a shuffle
// is issued in the loop and the sums carry no bias: a pixel whose difference
// to every sample is 0, +L or -L scores exactly 0, 1, -1 as in the reference.
// Samples whose skip circle contains the whole tile are dropped per tile,
// those that cannot reach it skip the test.
#include <cfloat>
#include <cmath>
#include <cstdlib>

#include "pf_common.cuh"

namespace pfb {

namespace {

constexpr int kAceBlock = 128;      // threads = columns per block
constexpr int kMaxRowsPerThread = 8; // P: pixels a thread carries (same column), 2 / 4 / 8
#ifndef PF_ACE_UNROLL
#define PF_ACE_UNROLL 4
#endif
constexpr int kAceUnroll = PF_ACE_UNROLL;   // unroll of the sample loop
constexpr int kChunk = 512;         // samples staged in shared memory at a time
// The min/max block of a call: [0..2] min, [3..5] max of the scores as the
// accumulation kernel computed them; [6..8], [9..11] the same over the scores
// ace_refine_kernel recomputed; then 6 doubles per tile of the page-wide grid.
constexpr int kMinmaxHead = 12;
// A score within this fraction of the page's range of its minimum or maximum
// is recomputed in the reference's own arithmetic.  The accumulation kernel's
// weights are good to 2^-19, which bounds its error at 2^-18 of the range
// (measured: 3e-7 on a range of 1.4, tests/test_gpu_parity.py).
constexpr double kNearEps = 0x1p-15;
// Squared distances are carried scaled by 2^-1000, so the weights come out of
// the rsqrt unit scaled by 2^500: the clamp result u = c + L becomes the high
// word of a SUBNORMAL double (value u * 2^-1042, linear in u with no implicit
// leading one), and weight * value must land in the normal range.
// The common factors cancel in the score.
constexpr double kD2Scale = 0x1p-1000;

struct SampleEntry {
	int sx;
	uint32_t bias[3];   // L - s_c (fast path) or s_c (table path) per channel
};

struct AceShared {
	SampleEntry tbl[kChunk];                    // reordered: far samples, then straddling
	double dy2[kChunk * kMaxRowsPerThread];     // [sample][row of the group], stride P
	int sy[kChunk];
	double lut[511];                            // general path only
	int n_far, n_str;
};

__device__ __forceinline__ double rsqrt_approx(double x)
{
	double y;
	asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
	return y;
}

// 2/sqrt(d2) to ~2^-43 (NEWTON) or 1/sqrt(d2) to ~2^-20 relative; a common
// factor on the weights cancels in the score.  MUFU.RSQ64H reads and writes
// only the high word of a double; without the Newton step the low word is
// left as it is (d2's own low word, at most 2^-20 of the value) rather than
// zeroed with an extra instruction -- the same order as the unit's own error,
// and the same weight feeds numerator and denominator.
template <bool NEWTON>
__device__ __forceinline__ double weight_of(double d2)
{
	const double y0 = rsqrt_approx(d2);
	if (!NEWTON)
		return __hiloint2double(__double2hiint(y0), __double2loint(d2));
	const double t = d2 * y0;
	const double q = fma(-t, y0, 3.0);
	return y0 * q;
}

// 1/d to an ulp or two: the unit's 2^-20 estimate and two Newton steps.  The
// quotients that must be exact (the ties) are decided by comparisons, not by
// this; 0 gives NaN, which the caller replaces.
__device__ __forceinline__ double reciprocal(double d)
{
	double y;
	asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
	double e = fma(-d, y, 1.0);
	y = fma(y, e, y);
	e = fma(-d, y, 1.0);
	return fma(y, e, y);
}

__device__ __forceinline__ void atomic_min_double(double *addr, double v)
{
	unsigned long long *a = reinterpret_cast<unsigned long long *>(addr);
	unsigned long long old = *a;
	while (v < __longlong_as_double((long long)old)) {
		const unsigned long long seen = atomicCAS(a, old,
				(unsigned long long)__double_as_longlong(v));
		if (seen == old)
			break;
		old = seen;
	}
}

__device__ __forceinline__ void atomic_max_double(double *addr, double v)
{
	unsigned long long *a = reinterpret_cast<unsigned long long *>(addr);
	unsigned long long old = *a;
	while (v > __longlong_as_double((long long)old)) {
		const unsigned long long seen = atomicCAS(a, old,
				(unsigned long long)__double_as_longlong(v));
		if (seen == old)
			break;
		old = seen;
	}
}

// The exactly tied scores the reference produces (ace_scores_kernel: a pixel whose difference
// to every sample is zero or saturates).
__device__ __forceinline__ bool is_tie_score(double s)
{
	return s == 0.0 || s == 1.0 || s == -1.0;
}

struct AceConst {
	int w, h;
	int row_begin, row_end;
	int nb_samples;
	int too_close2;        // (h/5)^2: pairs with d^2 below this are skipped
	int clamp_l;           // L = limit/slope when it is an integer <= 127 (fast path)
	int tile_rows;         // rows a block walks; tiles sit on a page-wide grid of this pitch
	int score_row0;        // score row stored first in `scores`
	long score_plane;      // doubles between the R, G, B score planes
	const uint32_t *sample_px;   // samples' pixel values, or NULL: read them from the page
	double slope, limit;
};

// Stages one chunk of the sample table for this block's tile.  Samples are
// classified against the tile rectangle: "far" ones can never be skipped,
// "inside" ones always are (dropped here), the rest straddle the skip circle.
// Order inside each class is the sample order, so sums are reproducible.
// 1 = far (never skipped on this tile), 2 = straddles the skip circle, 0 = inside
__device__ __forceinline__ int classify_sample(int sx, int sy, int x_lo, int x_hi,
		int y_lo, int y_hi, int too_close2)
{
	const int dxm = max(max(x_lo - sx, sx - x_hi), 0);
	const int dym = max(max(y_lo - sy, sy - y_hi), 0);
	if (dxm * dxm + dym * dym >= too_close2)
		return 1;
	const int dxM = max(abs(sx - x_lo), abs(sx - x_hi));
	const int dyM = max(abs(sy - y_lo), abs(sy - y_hi));
	return (dxM * dxM + dyM * dyM >= too_close2) ? 2 : 0;
}

template <bool FAST>
__device__ __noinline__ void stage_samples(AceShared &sh, const AceConst &cs,
		const uint32_t *__restrict__ in, const int *__restrict__ samples_xy,
		int k0, int kn, int x_lo, int x_hi, int y_lo, int y_hi)
{
	constexpr int kWarps = kAceBlock / 32;
	constexpr int kPasses = kChunk / kAceBlock;
	__shared__ int counts[kPasses][kWarps][2];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int2 *xy = reinterpret_cast<const int2 *>(samples_xy) + k0;
	const int passes = (kn + kAceBlock - 1) / kAceBlock;      // 1 for the usual 100 samples
	// sweep 1: how many far / straddling samples each warp holds per pass
	for (int p = 0; p < passes; p++) {
		const int k = p * kAceBlock + threadIdx.x;
		int cls = 0;
		if (k < kn) {
			const int2 s = __ldg(xy + k);
			cls = classify_sample(s.x, s.y, x_lo, x_hi, y_lo, y_hi, cs.too_close2);
		}
		const unsigned mf = __ballot_sync(0xFFFFFFFFu, cls == 1);
		const unsigned ms = __ballot_sync(0xFFFFFFFFu, cls == 2);
		if (lane == 0) {
			counts[p][warp][0] = __popc(mf);
			counts[p][warp][1] = __popc(ms);
		}
	}
	__syncthreads();
	int n_far = 0;
	for (int i = 0; i < passes * kWarps; i++)
		n_far += counts[0][i][0];
	// sweep 2: ordered compaction -- far samples first, then straddling, each
	// class in sample order, so sums are reproducible
	int off_far = 0, off_str = n_far;
	for (int p = 0; p < passes; p++) {
		for (int wv = 0; wv < warp; wv++) {
			off_far += counts[p][wv][0];
			off_str += counts[p][wv][1];
		}
		const int k = p * kAceBlock + threadIdx.x;
		int cls = 0;
		int2 s = make_int2(0, 0);
		if (k < kn) {
			s = __ldg(xy + k);
			cls = classify_sample(s.x, s.y, x_lo, x_hi, y_lo, y_hi, cs.too_close2);
		}
		const unsigned mf = __ballot_sync(0xFFFFFFFFu, cls == 1);
		const unsigned ms = __ballot_sync(0xFFFFFFFFu, cls == 2);
		if (cls != 0) {
			const unsigned below = (1u << lane) - 1;
			const int dst = (cls == 1) ? off_far + __popc(mf & below)
			                           : off_str + __popc(ms & below);
			const uint32_t px = cs.sample_px ? __ldg(cs.sample_px + k0 + k)
			                                 : __ldg(in + (size_t)s.y * cs.w + s.x);
			sh.tbl[dst].sx = s.x;
			sh.sy[dst] = s.y;
#pragma unroll
			for (int c = 0; c < 3; c++) {
				const uint32_t v = (px >> (8 * c)) & 0xFF;
				sh.tbl[dst].bias[c] = FAST ? (uint32_t)(cs.clamp_l - (int)v) : v;
			}
		}
		for (int wv = warp; wv < kWarps; wv++) {
			off_far += counts[p][wv][0];
			off_str += counts[p][wv][1];
		}
	}
	if (threadIdx.x == kAceBlock - 1) {     // last thread: offsets are the totals
		sh.n_far = n_far;
		sh.n_str = off_str - n_far;
	}
	__syncthreads();
}

// One sample against the P pixels of a thread.  TEST: the sample straddles the
// skip circle on this tile, so each pair is tested (weight forced to 0).
template <bool FAST, bool NEWTON, bool TEST, int P>
__device__ __forceinline__ void accumulate_sample(const AceShared &sh, int j, int x,
		const uint32_t (&pix)[P], const uint32_t (&pv)[P][3],
		int two_l, double too_close2, double top_value,
		double (&acc)[P][3], double (&top)[P])
{
	const SampleEntry se = sh.tbl[j];
	const int dx = x - se.sx;
	const double dx2 = (double)(dx * dx) * kD2Scale;
	double wgt[P];
#pragma unroll
	for (int r = 0; r < P; r++) {
		const double d2 = dx2 + sh.dy2[j * P + r];
		double wv = weight_of<NEWTON>(d2);
		if (TEST && d2 < too_close2)
			wv = 0.0;
		wgt[r] = wv;
		// the channel that saturates high on every sample (u = 2L): same FMA
		// sequence as acc[], so the ties the reference produces come out
		// exactly: acc == top (saturated high everywhere), acc == top / 2
		// (difference 0 everywhere: scaling every term by two commutes
		// with every rounding) and acc == 0 (saturated low everywhere)
		top[r] = fma(wv, top_value, top[r]);
	}
	if (FAST) {
#pragma unroll
		for (int r = 0; r < P; r++) {
#pragma unroll
			for (int c = 0; c < 3; c++) {
				// clamp(p - s, -L, L) + L = max(min(p + (L - s), 2L), 0) is ONE
				// instruction, VIADDMNMX.RELU, and its result IS the high word
				// of the subnormal double u * 2^-1042: no conversion, no shuffle
				const int u = __viaddmin_s32_relu((int)pv[r][c], (int)se.bias[c], two_l);
				acc[r][c] = fma(wgt[r], __hiloint2double(u, 0), acc[r][c]);
			}
		}
	} else {
#pragma unroll
		for (int r = 0; r < P; r++) {
#pragma unroll
			for (int c = 0; c < 3; c++) {
				const int v = (pix[r] >> (8 * c)) & 0xFF;
				acc[r][c] = fma(wgt[r], sh.lut[v - (int)se.bias[c] + 255], acc[r][c]);
			}
		}
	}
}

// The accumulation pass.  Thread = one column, kRowsPerThread rows at a time.
// FAST: L = limit/slope is an integer <= 127, clamp on the integer pipe.
// !FAST: any slope/limit, saturation values read from a 511-entry table.
#ifndef PF_ACE_MINBLOCKS_P8
#define PF_ACE_MINBLOCKS_P8 3
#endif
template <bool FAST, bool NEWTON, int P>
__global__ void __launch_bounds__(kAceBlock, (P == 2) ? 8 : ((P == 4) ? 6 : ((P == 6) ? 4 : PF_ACE_MINBLOCKS_P8)))
ace_scores_kernel(const uint32_t *__restrict__ in, const int *__restrict__ samples_xy,
		const AceConst cs, double *__restrict__ scores, double *__restrict__ minmax)
{
	extern __shared__ __align__(16) unsigned char smem_raw[];
	AceShared &sh = *reinterpret_cast<AceShared *>(smem_raw);
	pdl_wait();
	const int x = blockIdx.x * kAceBlock + threadIdx.x;
	const int xc = min(x, cs.w - 1);
	// Tiles sit on a page-wide grid of tile_rows rows whatever the band is, so
	// a band split (multi-GPU) classifies and sums exactly like the whole page.
	const int tile_y0 = (cs.row_begin / cs.tile_rows + blockIdx.y) * cs.tile_rows;
	const int tile_y1 = min(cs.h, tile_y0 + cs.tile_rows);      // exclusive
	const int x_lo = blockIdx.x * kAceBlock;
	const int x_hi = min(cs.w - 1, x_lo + kAceBlock - 1);
	const size_t plane = (size_t)cs.score_plane;

	if (!FAST) {
		for (int i = threadIdx.x; i < 511; i += kAceBlock) {
			// the reference's calc_saturation on the integer difference i-255
			double v = (double)(i - 255) * cs.slope;
			if (v > cs.limit)
				v = cs.limit;
			else if (v < -cs.limit)
				v = -cs.limit;
			sh.lut[i] = v;
		}
	}

	const int two_l = 2 * cs.clamp_l;
	const double too_close2 = (double)cs.too_close2 * kD2Scale;
	// what a channel saturated high contributes per unit weight
	const double top_value = FAST ? __hiloint2double(two_l, 0) : cs.limit;

	double mn[3] = { DBL_MAX, DBL_MAX, DBL_MAX };
	double mx[3] = { -DBL_MAX, -DBL_MAX, -DBL_MAX };

	bool staged = false;
	for (int y0 = tile_y0; y0 < tile_y1; y0 += P) {
		if (y0 + P <= cs.row_begin || y0 >= cs.row_end)
			continue;       // row group outside this band (block-uniform)
		// pixel values of this thread's P rows; rows of the group outside the band
		// repeat the nearest one inside (their scores are not stored), so only
		// the band's own rows of `in` are ever read
		uint32_t pix[P];
#pragma unroll
		for (int r = 0; r < P; r++)
			pix[r] = __ldg(in + (size_t)min(max(y0 + r, cs.row_begin), cs.row_end - 1) * cs.w + xc);
		// one channel value per register (the clamp instruction wants them so)
		uint32_t pv[P][3];
#pragma unroll
		for (int r = 0; r < P; r++)
#pragma unroll
			for (int c = 0; c < 3; c++)
				pv[r][c] = (pix[r] >> (8 * c)) & 0xFF;

		double acc[P][3], top[P];
#pragma unroll
		for (int r = 0; r < P; r++) {
			top[r] = 0.0;
#pragma unroll
			for (int c = 0; c < 3; c++)
				acc[r][c] = 0.0;
		}

		for (int k0 = 0; k0 < cs.nb_samples; k0 += kChunk) {
			const int kn = min(kChunk, cs.nb_samples - k0);
			// (re)stage when the chunk or the tile changes; with one chunk the
			// table is staged once per tile
			if (!staged || cs.nb_samples > kChunk) {
				staged = true;
				__syncthreads();
				stage_samples<FAST>(sh, cs, in, samples_xy, k0, kn,
						x_lo, x_hi, tile_y0, tile_y1 - 1);
			}
			const int n_act = sh.n_far + sh.n_str;
			__syncthreads();
			for (int j = threadIdx.x; j < n_act; j += kAceBlock) {
#pragma unroll
				for (int r = 0; r < P; r++) {
					const int dy = y0 + r - sh.sy[j];
					sh.dy2[j * P + r] = (double)(dy * dy) * kD2Scale;
				}
			}
			__syncthreads();

			const int n_far = sh.n_far;
#pragma unroll kAceUnroll
			for (int j = 0; j < n_far; j++)
				accumulate_sample<FAST, NEWTON, false, P>(sh, j, x, pix, pv, two_l,
						too_close2, top_value, acc, top);
			for (int j = n_far; j < n_act; j++)
				accumulate_sample<FAST, NEWTON, true, P>(sh, j, x, pix, pv, two_l,
						too_close2, top_value, acc, top);
		}

		// score = sum(sat/dist) / sum(limit/dist).  `top` is the numerator of a
		// channel saturated high on every sample, i.e. limit * sum(1/dist) up
		// to the common factor, accumulated by the same FMA sequence as acc.
		// The reference gets exactly 1, 0 and -1 for a pixel whose difference
		// to every sample saturates high, is zero, saturates low (flat page
		// areas: large sets of tied scores, also at the extremes the rescale
		// normalises with); the comparisons below reproduce those ties.
#pragma unroll
		for (int r = 0; r < P; r++) {
			const int y = y0 + r;
			if (y >= cs.row_begin && y < cs.row_end && y < tile_y1 && x < cs.w) {
				const size_t o = (size_t)(y - cs.score_row0) * cs.w + x;
				// FAST: acc holds sum(w (c + L)); c / L = (acc - top/2) / (top/2)
				// Every sample skipped (top == 0): the reference divides 0 by 0.
				// Here reciprocal(0) is NaN and so is the product below; only
				// the tie comparisons (0 == 0) have to be kept from firing.
				const bool live = top[r] > 0.0;
				const double den = FAST ? 0.5 * top[r] : top[r];
				const double inv = reciprocal(den);
#pragma unroll
				for (int c = 0; c < 3; c++) {
					const double num = FAST ? acc[r][c] - den : acc[r][c];
					const bool tie_hi = live && num == den, tie_lo = live && num == -den;
					double s = num * inv;           // exactly 0 when num is
					s = tie_hi ? 1.0 : s;
					s = tie_lo ? -1.0 : s;
					scores[c * plane + o] = s;
					if (s < mn[c]) mn[c] = s;     // NaN never enters
					if (s > mx[c]) mx[c] = s;
				}
			}
		}
	}

	// block-wide min / max, then 6 atomics per block
	__shared__ double red[2][3][kAceBlock / 32];
#pragma unroll
	for (int c = 0; c < 3; c++) {
#pragma unroll
		for (int d = 16; d > 0; d >>= 1) {
			mn[c] = fmin(mn[c], __shfl_xor_sync(0xFFFFFFFFu, mn[c], d));
			mx[c] = fmax(mx[c], __shfl_xor_sync(0xFFFFFFFFu, mx[c], d));
		}
		if ((threadIdx.x & 31) == 0) {
			red[0][c][threadIdx.x >> 5] = mn[c];
			red[1][c][threadIdx.x >> 5] = mx[c];
		}
	}
	__syncthreads();
	if (threadIdx.x < 3) {
		const int c = threadIdx.x;
		double a = red[0][c][0], b = red[1][c][0];
		for (int i = 1; i < kAceBlock / 32; i++) {
			a = fmin(a, red[0][c][i]);
			b = fmax(b, red[1][c][i]);
		}
		atomic_min_double(minmax + c, a);
		atomic_max_double(minmax + 3 + c, b);
		// and per tile of the page-wide grid (two bands may share a tile), for
		// ace_refine_kernel to find the tiles that matter
		double *tile = minmax + kMinmaxHead
			+ 6 * ((size_t)(cs.row_begin / cs.tile_rows + blockIdx.y) * gridDim.x + blockIdx.x);
		atomic_min_double(tile + c, a);
		atomic_max_double(tile + 3 + c, b);
	}
}

__global__ void ace_minmax_init_kernel(double *minmax, size_t n)
{
	pdl_wait();
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
			i += (size_t)gridDim.x * blockDim.x)
		minmax[i] = (i % 6 < 3) ? DBL_MAX : -DBL_MAX;
}

constexpr int kRefineThreads = 256;
constexpr int kRefineSlab = 16;      // rows scanned before the candidates are worked off
constexpr int kRefineTerms = 512;    // (candidate, sample) pairs whose terms are in flight
constexpr int kRefineTable = 1024;   // samples kept in shared memory

struct RefineShared {
	double terms[kRefineTerms][4];
	int sx[kRefineTable], sy[kRefineTable];
	uint32_t spx[kRefineTable];
	unsigned short cand[kRefineSlab * kAceBlock];
	int wanted[kRefineThreads];
	int n_wanted, n_cand;
};

// The terms one sample contributes to one pixel in the reference's own
// arithmetic (upstream _ace.c, ace_thread_adjustment; oracle/pf_oracle.c
// ace_adjust_rows): IEEE sqrt and divisions.  A skipped sample contributes
// +0.0, which changes no sum.
// (This code runs once or a few times per page, from a cold instruction
// cache: one copy of it, loops not unrolled.)
__device__ __forceinline__ void ace_exact_terms(const RefineShared &sh, bool table,
		const uint32_t *__restrict__ in, const int *__restrict__ samples_xy,
		const AceConst &cs, int x, int y, int k, double *t)
{
	int sx, sy;
	uint32_t sp;
	if (table) {
		sx = sh.sx[k];
		sy = sh.sy[k];
		sp = sh.spx[k];
	} else {
		const int2 sxy = __ldg(reinterpret_cast<const int2 *>(samples_xy) + k);
		sx = sxy.x;
		sy = sxy.y;
		sp = cs.sample_px ? __ldg(cs.sample_px + k) : __ldg(in + (size_t)sy * cs.w + sx);
	}
	const uint32_t p = __ldg(in + (size_t)y * cs.w + x);
	const double ddx = (double)(x - sx), ddy = (double)(y - sy);
	const double dist = __dsqrt_rn(__dadd_rn(__dmul_rn(ddx, ddx), __dmul_rn(ddy, ddy)));
	const bool skipped = dist < (double)(cs.h / 5);
#pragma unroll 1
	for (int c = 0; c < 4; c++) {
		const int diff = (int)((p >> (8 * c)) & 0xFF) - (int)((sp >> (8 * c)) & 0xFF);
		double v = __dmul_rn((double)diff, cs.slope);
		if (v > cs.limit || c == 3)
			v = cs.limit;
		else if (v < -cs.limit)
			v = -cs.limit;
		t[c] = skipped ? 0.0 : __ddiv_rn(v, dist);
	}
}

// Exact scores of the block's listed candidates sh.cand[0 .. n) (pixels of the
// slab at (x_lo, ys)).  The whole block computes the terms -- one (candidate,
// sample) pair per thread and round, that is where the time goes -- then four
// threads per candidate add them up in sample order: that chain is the
// reference's and cannot be split.
__device__ __forceinline__ void ace_exact_candidates(RefineShared &sh, bool table,
		const uint32_t *__restrict__ in, const int *__restrict__ samples_xy,
		const AceConst &cs, int x_lo, int ys, int n,
		double *__restrict__ scores, double *__restrict__ minmax)
{
	const int chunk = min(cs.nb_samples, kRefineTerms);      // samples per round
	const int batch = kRefineTerms / chunk;                   // candidates per round
	const size_t plane = (size_t)cs.score_plane;
	for (int i0 = 0; i0 < n; i0 += batch) {
		const int nb = min(batch, n - i0);
		double acc = 0.0;
		for (int k0 = 0; k0 < cs.nb_samples; k0 += chunk) {
			const int kn = min(chunk, cs.nb_samples - k0);
			for (int q = threadIdx.x; q < nb * kn; q += kRefineThreads) {
				const int ci = q / kn, j = q - ci * kn;
				const int cd = sh.cand[i0 + ci];
				ace_exact_terms(sh, table, in, samples_xy, cs, x_lo + cd % kAceBlock,
						ys + cd / kAceBlock, k0 + j, sh.terms[q]);
			}
			__syncthreads();
			if (threadIdx.x < 4 * nb) {
				const double (*tm)[4] = sh.terms + (threadIdx.x >> 2) * kn;
				const int a = threadIdx.x & 3;
#pragma unroll 8
				for (int j = 0; j < kn; j++)
					acc = __dadd_rn(acc, tm[j][a]);
			}
			__syncthreads();
		}
		// threads 4i .. 4i+3 hold candidate i's R, G, B sums and the denominator
		const double denom = __shfl_sync(0xFFFFFFFFu, acc, (threadIdx.x & 31) | 3);
		if (threadIdx.x < 4 * nb && (threadIdx.x & 3) < 3) {
			const int cd = sh.cand[i0 + (threadIdx.x >> 2)], c = threadIdx.x & 3;
			const int px = x_lo + cd % kAceBlock, y = ys + cd / kAceBlock;
			const double r = __ddiv_rn(acc, denom);
			scores[c * plane + (size_t)(y - cs.score_row0) * cs.w + px] = r;
			atomic_min_double(minmax + 6 + c, r);      // NaN never enters
			atomic_max_double(minmax + 9 + c, r);
		}
	}
}

// Second pass over the few tiles that reach the page's extremes.  The rescale
// divides by (max - min) and truncates, so a score AT the maximum becomes 255
// and one an ulp below it 254: where many pixels score the same up to rounding
// (a flat mark on a flat page, with no sample on the mark), which of them are
// 254 depends on the last bit of the reference's sums.  Scores within kNearEps
// of either extreme are therefore recomputed exactly as the reference does,
// and the extremes themselves are taken over the recomputed values, which
// makes the rescale of those pixels bit-identical to the reference's.  Tied
// scores (0, +-1) are exact already and are left alone.
//
// A few blocks per SM look through the tile table the accumulation kernel left
// (neighbouring tiles go to different blocks); a block that finds a tile near
// an extreme lists the tile's candidate pixels and works them off with all
// its threads.  tiles_x * tiles_y is the accumulation grid.
__global__ void __launch_bounds__(kRefineThreads)
ace_refine_kernel(const uint32_t *__restrict__ in, const int *__restrict__ samples_xy,
		const AceConst cs, double *__restrict__ scores, double *__restrict__ minmax,
		int tiles_x, int tiles_y)
{
	__shared__ RefineShared sh;
	pdl_wait();
	double lo[3], hi[3];
#pragma unroll
	for (int c = 0; c < 3; c++) {
		const double mn = minmax[c], mx = minmax[3 + c];
		const double thr = (mx - mn) * kNearEps;
		lo[c] = mn + thr;
		hi[c] = mx - thr;
	}
	const int col = threadIdx.x % kAceBlock;
	const int tile_row0 = cs.row_begin / cs.tile_rows;
	const size_t plane = (size_t)cs.score_plane;
	const int n_tiles = tiles_x * tiles_y;
	const bool table = cs.nb_samples <= kRefineTable;
	bool staged = false;

	for (int base = 0; base < n_tiles; base += gridDim.x * kRefineThreads) {
		if (threadIdx.x == 0)
			sh.n_wanted = 0;
		__syncthreads();
		const int t = base + threadIdx.x * gridDim.x + blockIdx.x;
		if (t < n_tiles) {
			const double *tile = minmax + kMinmaxHead
				+ 6 * ((size_t)(tile_row0 + t / tiles_x) * tiles_x + t % tiles_x);
			bool wanted = false;
#pragma unroll
			for (int c = 0; c < 3; c++)
				wanted |= (tile[c] <= lo[c]) || (tile[3 + c] >= hi[c]);
			if (wanted)
				sh.wanted[atomicAdd(&sh.n_wanted, 1)] = t;
		}
		__syncthreads();
		const int n_wanted = sh.n_wanted;
		if (n_wanted && table && !staged) {
			staged = true;
			for (int k = threadIdx.x; k < cs.nb_samples; k += kRefineThreads) {
				const int2 sxy = __ldg(reinterpret_cast<const int2 *>(samples_xy) + k);
				sh.sx[k] = sxy.x;
				sh.sy[k] = sxy.y;
				sh.spx[k] = cs.sample_px ? __ldg(cs.sample_px + k)
				                         : __ldg(in + (size_t)sxy.y * cs.w + sxy.x);
			}
		}
		for (int wi = 0; wi < n_wanted; wi++) {
			const int tw = sh.wanted[wi];
			const int x_lo = (tw % tiles_x) * kAceBlock;
			const int tile_y0 = (tile_row0 + tw / tiles_x) * cs.tile_rows;
			const int y_begin = max(tile_y0, cs.row_begin);
			const int y_end = min(min(cs.h, tile_y0 + cs.tile_rows), cs.row_end);
			for (int ys = y_begin; ys < y_end; ys += kRefineSlab) {
				__syncthreads();
				if (threadIdx.x == 0)
					sh.n_cand = 0;
				__syncthreads();
				// this thread's pixels of the slab: all loads first
				constexpr int kStep = kRefineThreads / kAceBlock;
				double sv[kRefineSlab / kStep][3];
#pragma unroll
				for (int i = 0; i < kRefineSlab / kStep; i++) {
					const int y = ys + threadIdx.x / kAceBlock + i * kStep;
					const bool in_tile = y < y_end && x_lo + col < cs.w;
					const size_t o = in_tile ? (size_t)(y - cs.score_row0) * cs.w + x_lo + col : 0;
#pragma unroll
					for (int c = 0; c < 3; c++)
						sv[i][c] = in_tile ? scores[c * plane + o] : 0.0;   // 0.0: a tie, never listed
				}
#pragma unroll
				for (int i = 0; i < kRefineSlab / kStep; i++) {
					bool need = false;
#pragma unroll
					for (int c = 0; c < 3; c++)
						need |= (sv[i][c] <= lo[c] || sv[i][c] >= hi[c])
							&& !is_tie_score(sv[i][c]);                  // false for NaN
					if (need)
						sh.cand[atomicAdd(&sh.n_cand, 1)] = (unsigned short)
							((threadIdx.x / kAceBlock + i * kStep) * kAceBlock + col);
				}
				__syncthreads();
				const int n = sh.n_cand;
				ace_exact_candidates(sh, table, in, samples_xy, cs, x_lo, ys, n, scores, minmax);
			}
		}
		__syncthreads();
	}
}

// The reference's linear scaling, operation for operation (true FP64 divide).
__device__ __noinline__ uint32_t ace_scale_exact(double s, double mn, double span)
{
	// a flat channel (0/0 = NaN for every pixel) and the pixels at the minimum
	// (areas of a page can be tied there) need no division
	if (!(span > 0.0) || s == mn)
		return 0u;
	const double v = __ddiv_rn(__dmul_rn(255.0, __dadd_rn(s, -mn)), span);
	if (!(v > 0.0))         // also NaN
		return 0u;
	return (uint32_t)__double2int_rz(fmin(v, 255.0));
}

// out[c] = trunc(clamp(255 * (score - min) / (max - min), 0, 255)); alpha 255.
// The quotient is formed with a reciprocal multiply (within 2e-13 of the
// reference's rounded quotient) and truncated by one round-toward-zero add
// against 2^31 that also exposes 21 fraction bits; only a value within 2e-6
// of an integer -- which includes the page minimum and maximum, and NaN --
// goes through the exact divide.  Four pixels per thread, 16-byte stores.
//
// MULTI: the same rows are also stored to further images (`more`), which may
// live on other GPUs (peer memory over NVLink): the band exchange of a
// row-band split rides on the rescale instead of following it.
constexpr int kMaxScaleOuts = 16;
struct ScaleOuts {
	uint32_t *p[kMaxScaleOuts - 1];
	int n;
	int multicast;      // `out` is an NVLS multicast address: one store, the switch delivers it to every GPU
};

// 16 bytes to a multicast address (multimem.st: the NVSwitch replicates the
// store into the buffer of every GPU bound to the multicast object)
__device__ __forceinline__ void multimem_store16(void *addr, uint4 v)
{
	asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
			:: "l"(addr), "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)),
			   "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w)) : "memory");
}

__device__ __forceinline__ void multimem_store4(void *addr, uint32_t v)
{
	asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;"
			:: "l"(addr), "f"(__uint_as_float(v)) : "memory");
}

template <bool MULTI>
__global__ void __launch_bounds__(256)
ace_scale_kernel(const double *__restrict__ scores, const double *__restrict__ minmax,
		size_t npix, size_t plane, uint32_t *__restrict__ out, const ScaleOuts more)
{
	pdl_wait();
	double mn[3], span[3], inv[3];
#pragma unroll
	for (int c = 0; c < 3; c++) {
		// the extremes over the recomputed scores where ace_refine_kernel
		// produced some; a tied extreme (exact as it is) takes part as well
		const double mnf = minmax[c], mxf = minmax[3 + c];
		const double mne = minmax[6 + c], mxe = minmax[9 + c];
		mn[c] = (is_tie_score(mnf) || mne == DBL_MAX) ? fmin(mnf, mne) : mne;
		const double mx = (is_tie_score(mxf) || mxe == -DBL_MAX) ? fmax(mxf, mxe) : mxe;
		span[c] = mx - mn[c];
		inv[c] = 1.0 / span[c];
	}
	bool vec = (npix % 4 == 0) && (plane % 2 == 0)
		&& (((uintptr_t)scores | (uintptr_t)out) % 16 == 0);
	if (MULTI)
		for (int d = 0; d < more.n; d++)
			vec = vec && ((uintptr_t)more.p[d] % 16 == 0);
	const size_t ngroups = vec ? npix / 4 : 0;
	for (size_t gi = (size_t)blockIdx.x * blockDim.x + threadIdx.x; gi < ngroups;
			gi += (size_t)gridDim.x * blockDim.x) {
		uint32_t px[4] = { 0xFF000000u, 0xFF000000u, 0xFF000000u, 0xFF000000u };
#pragma unroll
		for (int c = 0; c < 3; c++) {
			const double2 a = __ldcs(reinterpret_cast<const double2 *>(scores + c * plane) + 2 * gi);
			const double2 b = __ldcs(reinterpret_cast<const double2 *>(scores + c * plane) + 2 * gi + 1);
			const double sv[4] = { a.x, a.y, b.x, b.y };
#pragma unroll
			for (int k = 0; k < 4; k++) {
				const double t = (255.0 * (sv[k] - mn[c])) * inv[c];
				const uint32_t fx = (uint32_t)__double2loint(__dadd_rz(t, 2147483648.0));
				uint32_t q = min(fx >> 21, 255u);
				if (((fx + 4u) & 0x1FFFFFu) < 8u || !(t >= 0.0))
					q = ace_scale_exact(sv[k], mn[c], span[c]);
				px[k] |= q << (8 * c);
			}
		}
		const uint4 v = make_uint4(px[0], px[1], px[2], px[3]);
		if (MULTI && more.multicast)
			multimem_store16(reinterpret_cast<uint4 *>(out) + gi, v);
		else
			reinterpret_cast<uint4 *>(out)[gi] = v;
		if (MULTI)
			for (int d = 0; d < more.n; d++)
				reinterpret_cast<uint4 *>(more.p[d])[gi] = v;
	}
	// shapes that cannot be vectorised: the plain loop
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x + 4 * ngroups; i < npix;
			i += (size_t)gridDim.x * blockDim.x) {
		uint32_t rgba = 0xFF000000u;
#pragma unroll
		for (int c = 0; c < 3; c++)
			rgba |= ace_scale_exact(__ldcs(scores + c * plane + i), mn[c], span[c]) << (8 * c);
		if (MULTI && more.multicast)
			multimem_store4(out + i, rgba);
		else
			out[i] = rgba;
		if (MULTI)
			for (int d = 0; d < more.n; d++)
				more.p[d][i] = rgba;
	}
}

// FP64 FMA rate probe: 8 independent chains per thread.
__global__ void fp64_peak_kernel(double *sink, int iters, double a, double b)
{
	double v0 = threadIdx.x, v1 = v0 + 1, v2 = v0 + 2, v3 = v0 + 3;
	double v4 = v0 + 4, v5 = v0 + 5, v6 = v0 + 6, v7 = v0 + 7;
	for (int i = 0; i < iters; i++) {
		v0 = fma(v0, a, b); v1 = fma(v1, a, b); v2 = fma(v2, a, b); v3 = fma(v3, a, b);
		v4 = fma(v4, a, b); v5 = fma(v5, a, b); v6 = fma(v6, a, b); v7 = fma(v7, a, b);
	}
	const double s = v0 + v1 + v2 + v3 + v4 + v5 + v6 + v7;
	if (s == 12345.678)
		sink[0] = s;
}

}  // namespace

int launch_ace_minmax_init(double *minmax, int w, int h, cudaStream_t st)
{
	const size_t n = ace_minmax_doubles(w, h);
	PFB_CUDA(launch_pdl(ace_minmax_init_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, st,
			minmax, n));
	return PF_B200_OK;
}

namespace {

// Pixels a thread carries and rows a tile holds.  Knobs (read once per
// process; tools/ace_time.py sweeps them); the defaults are the measured best
// on B200 (profiles/README.md):
//   PF_B200_ACE_ROWS = 2 | 4 | 6 | 8   pixels per thread (default 8)
//   PF_B200_ACE_TILE_ROWS = n          rows per tile (default 16)
int ace_rows_per_thread(bool fast)
{
	static const int rows_env = getenv("PF_B200_ACE_ROWS") ? atoi(getenv("PF_B200_ACE_ROWS")) : 8;
	if (!fast)
		return 4;
	return (rows_env == 2 || rows_env == 4 || rows_env == 6) ? rows_env : 8;
}

int ace_tile_rows(bool fast)
{
	static const int tile_env = getenv("PF_B200_ACE_TILE_ROWS") ? atoi(getenv("PF_B200_ACE_TILE_ROWS")) : 0;
	const int rows_p = ace_rows_per_thread(fast);
	// a tile is a whole number of row groups
	return (tile_env >= rows_p && tile_env % rows_p == 0) ? tile_env : (rows_p == 6 ? 18 : 16);
}

// Checks the arguments and lays out the launch both ACE kernels share.
int ace_setup(const AceParams &p, AceConst &cs, dim3 &grid, bool &fast)
{
	if (p.nb_samples < 1)
		return fail(PF_B200_EINVAL, "ace: nb_samples must be >= 1");
	if (p.row_begin < 0 || p.row_end > p.h || p.row_begin >= p.row_end)
		return fail(PF_B200_EINVAL, "ace: bad row band [%d, %d) of %d rows",
				p.row_begin, p.row_end, p.h);
	if (!(p.slope > 0.0) || !(p.limit > 0.0))
		return fail(PF_B200_EINVAL, "ace: slope and limit must be positive");
	if (p.w > 46340 || p.h > 32767)
		return fail(PF_B200_ETOOBIG, "ace: page of %d x %d too large", p.w, p.h);

	cs.w = p.w;
	cs.h = p.h;
	cs.row_begin = p.row_begin;
	cs.row_end = p.row_end;
	cs.nb_samples = p.nb_samples;
	const int tc = p.h / 5;
	cs.too_close2 = tc * tc;
	cs.slope = p.slope;
	cs.limit = p.limit;
	cs.sample_px = p.sample_px;
	cs.score_row0 = p.score_row0 >= 0 ? p.score_row0 : p.row_begin;
	cs.score_plane = p.score_plane >= 0 ? p.score_plane : (long)(p.row_end - p.row_begin) * p.w;
	// fast path: saturate(slope*d) == slope*clamp(d, -L, L) with integer L <= 127
	const double l = p.limit / p.slope;
	fast = l >= 1.0 && l <= 127.0 && l == floor(l) && p.slope * l == p.limit;
	cs.clamp_l = fast ? (int)l : 0;
	cs.tile_rows = ace_tile_rows(fast);
	grid = dim3((p.w + kAceBlock - 1) / kAceBlock,
			(p.row_end - 1) / cs.tile_rows - p.row_begin / cs.tile_rows + 1);
	return PF_B200_OK;
}

}  // namespace

size_t ace_minmax_doubles(int w, int h)
{
	// a band of the page never has more tile rows than the page has, plus one
	const int tile_rows = ace_tile_rows(true) < ace_tile_rows(false) ? ace_tile_rows(true)
	                                                                 : ace_tile_rows(false);
	const size_t tiles = (size_t)((w + kAceBlock - 1) / kAceBlock) * ((h + tile_rows - 1) / tile_rows + 1);
	return kMinmaxHead + 6 * tiles;
}

int launch_ace_scores(const uint32_t *in, const int *samples_xy,
		const AceParams &p, double *scores, double *minmax, cudaStream_t st)
{
	AceConst cs;
	dim3 grid;
	bool fast;
	const int rc = ace_setup(p, cs, grid, fast);
	if (rc)
		return rc;
	const size_t smem = sizeof(AceShared);
	//   PF_B200_ACE_NEWTON = 1         one Newton step on MUFU.RSQ64H's 1/dist:
	//       weights good to ~2^-43 instead of ~2^-22.  Off by default: both
	//       settings sit three orders of magnitude inside the +-1 LSB / 0.1 %
	//       bar (1.1e-4 % vs 3e-5 % of channel values differ on the test page)
	//       and the step costs 3 of 8 FP64 instructions per pair.
	const int rows_env = ace_rows_per_thread(fast);
	static const int newton_env = getenv("PF_B200_ACE_NEWTON") ? atoi(getenv("PF_B200_ACE_NEWTON")) : 0;
	void (*kern)(const uint32_t *, const int *, const AceConst, double *, double *) = nullptr;
	if (!fast)
		kern = ace_scores_kernel<false, true, 4>;
	else if (newton_env)
		kern = rows_env == 2 ? ace_scores_kernel<true, true, 2>
			: rows_env == 8 ? ace_scores_kernel<true, true, 8> : ace_scores_kernel<true, true, 4>;
	else
		kern = rows_env == 2 ? ace_scores_kernel<true, false, 2>
			: rows_env == 8 ? ace_scores_kernel<true, false, 8>
			: rows_env == 6 ? ace_scores_kernel<true, false, 6> : ace_scores_kernel<true, false, 4>;
	PFB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
	PFB_CUDA(launch_pdl(kern, grid, dim3(kAceBlock), smem, st, in, samples_xy, cs, scores, minmax));
	return PF_B200_OK;
}


int launch_ace_refine(const uint32_t *in, const int *samples_xy,
		const AceParams &p, double *scores, double *minmax, cudaStream_t st)
{
	AceConst cs;
	dim3 grid;
	bool fast;
	const int rc = ace_setup(p, cs, grid, fast);
	if (rc)
		return rc;
	PFB_CUDA(launch_pdl(ace_refine_kernel, dim3(sm_count() * 4), dim3(kRefineThreads), 0, st,
			in, samples_xy, cs, scores, minmax, (int)grid.x, (int)grid.y));
	return PF_B200_OK;
}

int launch_ace_scale(const double *scores, const double *minmax, int w,
		int rows, uint32_t *out, cudaStream_t st, long plane)
{
	const size_t npix = (size_t)w * rows;
	const size_t work = (npix + 3) / 4;       // 4 pixels per thread
	const int blocks = (int)((work + 255) / 256 < (size_t)sm_count() * 8
			? (work + 255) / 256 : (size_t)sm_count() * 8);
	ScaleOuts more = {};
	PFB_CUDA(launch_pdl(ace_scale_kernel<false>, dim3(blocks), dim3(256), 0, st,
			scores, minmax, npix, plane >= 0 ? (size_t)plane : npix, out, more));
	return PF_B200_OK;
}

int launch_ace_scale_multi(const double *scores, const double *minmax, int w, int rows,
		uint32_t *const *outs, int nb_outs, cudaStream_t st, bool multicast)
{
	if (nb_outs < 1 || nb_outs > kMaxScaleOuts)
		return fail(PF_B200_EINVAL, "ace: %d destinations (1..%d)", nb_outs, kMaxScaleOuts);
	ScaleOuts more = {};
	for (int d = 1; d < nb_outs; d++)
		more.p[d - 1] = outs[d];
	more.n = nb_outs - 1;
	more.multicast = multicast ? 1 : 0;
	const size_t npix = (size_t)w * rows;
	const size_t work = (npix + 3) / 4;
	const int blocks = (int)((work + 255) / 256 < (size_t)sm_count() * 8
			? (work + 255) / 256 : (size_t)sm_count() * 8);
	PFB_CUDA(launch_pdl(ace_scale_kernel<true>, dim3(blocks), dim3(256), 0, st,
			scores, minmax, npix, npix, outs[0], more));
	return PF_B200_OK;
}

double measure_fp64_peak_tflops()
{
	double *sink = nullptr;
	if (cudaMalloc(&sink, 64) != cudaSuccess)
		return -(double)fail(PF_B200_ECUDA, "fp64 probe: cudaMalloc failed");
	const int blocks = sm_count() * 8, threads = 256, iters = 1 << 15;
	cudaEvent_t e0, e1;
	cudaEventCreate(&e0);
	cudaEventCreate(&e1);
	fp64_peak_kernel<<<blocks, threads>>>(sink, 1024, 1.0000001, 1e-9);   // warm-up
	double best_ms = 1e30;
	for (int rep = 0; rep < 5; rep++) {
		cudaEventRecord(e0);
		fp64_peak_kernel<<<blocks, threads>>>(sink, iters, 1.0000001, 1e-9);
		cudaEventRecord(e1);
		cudaEventSynchronize(e1);
		float ms = 0.f;
		cudaEventElapsedTime(&ms, e0, e1);
		if (ms < best_ms)
			best_ms = ms;
	}
	cudaEventDestroy(e0);
	cudaEventDestroy(e1);
	const cudaError_t err = cudaGetLastError();
	cudaFree(sink);
	if (err != cudaSuccess)
		return -(double)fail(PF_B200_ECUDA, "fp64 probe: %s", cudaGetErrorString(err));
	const double flops = 2.0 * 8.0 * (double)iters * blocks * threads;
	return flops / (best_ms * 1e-3) / 1e12;
}

}  // namespace pfb
