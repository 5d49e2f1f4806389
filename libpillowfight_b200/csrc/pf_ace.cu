// Automatic Color Equalization for sm_100a.
//
// Reference loop (upstream src/pillowfight/_ace.c, ace_thread_adjustment; the
// file is not in the mount): for every pixel p and
// every sample k farther than height/5,
//     score[c] += saturate(slope * (p[c] - sample_k[c]), +-limit) / dist(p, k)
//     denom    += limit / dist(p, k)
// then score[c] /= denom, a page-wide min/max per channel, and a linear
// rescale to 0..255 that truncates.
//
// B200 formulation: a fast pass whose error is BOUNDED, and an exact pass for
// the few pixels where that error could matter.  The result is the reference's
// for every pixel, by construction rather than by measurement.
//
//  1. ace_scores_kernel (all but a few per cent of the time).  With
//     w_k = 1/dist and u_k = clamp((p - s_k) * slope / (2 limit) + 1/2, 0, 1),
//         score = 2 * sum_k w_k u_k / sum_k w_k - 1.
//     Everything is FP32: d^2 is one FADD2 (two rows at a time), the weight one
//     MUFU.RSQ (2^-22.9), the clamp ONE saturating FMA per channel
//     (FFMA.SAT: p * inv + (1/2 - s_k * inv)), the sums FFMA2 over row pairs,
//     flushed every 16 samples into a second set of accumulators (two-level
//     summation).  The N samples are one table for the whole page: a block
//     stages them in shared memory per tile and the loop touches no global
//     memory.  Samples whose skip circle contains the tile are dropped per
//     tile, those that cannot reach it skip the test.  A worst-case bound
//     eps on |score - reference score| follows from the operation counts
//     (ace_score_eps below).
//     Ties -- a pixel whose difference to every sample saturates high, is
//     zero, or saturates low scores exactly 1, 0, -1 in the reference; flat
//     areas of a page are millions of such pixels -- are recognised by exact
//     integer comparisons against the tile's sample extremes, not by the sums.
//  2. ace_refine_kernel: the page extremes exactly.  Every score within
//     2 eps of the approximate minimum or maximum is recomputed in the
//     reference's own arithmetic (IEEE sqrt and divide, sample order) and the
//     extremes are taken over those values.
//  3. ace_scale_kernel: 255 (s - min) / (max - min) in FP32 with the exact
//     extremes.  A value farther from an integer than the bound allows is
//     truncated as it stands; the others (a fraction of a per cent) go on a
//     list.
//  4. ace_fixup_kernel: the listed pixels in the reference's own arithmetic,
//     score and rescale.
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include <cooperative_groups.h>

#include "pf_common.cuh"

namespace pfb {

namespace {

constexpr int kAceBlock = 128;      // threads = columns per block
constexpr int kRows = 8;            // pixels a thread carries (same column)
constexpr int kTileRows = 16;       // rows a block walks; tiles sit on a page-wide grid of this pitch
constexpr int kChunk = 512;         // samples staged in shared memory at a time
constexpr int kFlush = 16;          // samples between flushes of the first-level sums
// The min/max block of a call: [0..2] min, [3..5] max of the scores as the
// accumulation kernel computed them; [6..8], [9..11] the exact extremes
// (ace_refine_kernel), [12..14], [15..17] the extremes of its FP64 candidate
// scores; then 6 doubles per tile of the page-wide grid.
constexpr int kMinmaxHead = 18;

struct SampleEntry {
	float sx;
	float b[3];         // 1/2 - s_c * inv per channel
};

struct AceShared {
	SampleEntry tbl[kChunk + 4];      // far samples, padded to a multiple of 4 with weightless entries, then straddling ones
	float dy2[(kChunk + 4) * kRows];  // [sample][row of the group]
	int sy[kChunk + 4];
	int sxi[kChunk + 4];              // straddling samples are tested in integers
	int n_far4, n_str;
	int smin[3], smax[3];             // channel extremes over the samples staged for this tile (ties)
};

__device__ __forceinline__ float rsqrt_f32(float x)
{
	// every argument is a normal number or +inf, so .ftz changes no result; it
	// only spares the subnormal-input wrapper (FSETP + 2 FMUL) per rsqrt
	float y;
	asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
	return y;
}

__device__ __forceinline__ float fma_sat(float a, float b, float c)
{
	float d;
	asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
	return d;
}

typedef unsigned long long f32x2;     // two floats in a 64-bit register pair

__device__ __forceinline__ f32x2 pack2(float lo, float hi)
{
	f32x2 r;
	asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
	return r;
}

__device__ __forceinline__ void unpack2(f32x2 v, float &lo, float &hi)
{
	asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}

__device__ __forceinline__ f32x2 ffma2(f32x2 a, f32x2 b, f32x2 c)
{
	f32x2 d;
	asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
	return d;
}

__device__ __forceinline__ f32x2 fadd2(f32x2 a, f32x2 b)
{
	f32x2 d;
	asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
	return d;
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

// In the score planes exactly 0, 1 and -1 mean "tie": the reference's score is
// exactly that (ace_scores_kernel moves an approximate score off these values).
__device__ __forceinline__ bool is_tie_score(float s)
{
	return s == 0.0f || s == 1.0f || s == -1.0f;
}

struct AceConst {
	int w, h;
	int row_begin, row_end;
	int nb_samples;
	int too_close2;        // (h/5)^2: pairs with d^2 below this are skipped
	int dsat;              // smallest difference whose saturation reaches the limit (256: none does)
	int score_row0;        // score row stored first in the planes
	long score_plane;      // floats between the R, G, B score planes
	const uint32_t *sample_px;   // samples' pixel values, or NULL: read them from the page
	double slope, limit;
	float inv;             // slope / (2 limit), rounded up a little (ace_setup)
	float eps;             // bound on |approximate score - reference score| (ace_score_eps)
};

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

constexpr int kNoSample = 1 << 30;     // sy of a padding entry: its dy^2 is +inf, its weight 0

// Stages one chunk of the sample table for this block's tile.  Samples are
// classified against the tile rectangle: "far" ones can never be skipped,
// "inside" ones always are (dropped here), the rest straddle the skip circle.
// Also folds the staged samples' channel extremes into sh.smin / sh.smax.
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
	const int n_far4 = (n_far + 3) & ~3;
	// sweep 2: ordered compaction -- far samples first, then straddling
	int off_far = 0, off_str = n_far4;
	int mn[3] = { 255, 255, 255 }, mx[3] = { 0, 0, 0 };
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
			sh.tbl[dst].sx = (float)s.x;
			sh.sxi[dst] = s.x;
			sh.sy[dst] = s.y;
#pragma unroll
			for (int c = 0; c < 3; c++) {
				const int v = (int)((px >> (8 * c)) & 0xFF);
				// (in double: the rounding of b must stay below 2^-26, ace_score_eps)
				sh.tbl[dst].b[c] = (float)(0.5 - (double)v * (double)cs.inv);
				mn[c] = min(mn[c], v);
				mx[c] = max(mx[c], v);
			}
		}
		for (int wv = warp; wv < kWarps; wv++) {
			off_far += counts[p][wv][0];
			off_str += counts[p][wv][1];
		}
	}
	// padding between the two classes: entries without weight
	if ((int)threadIdx.x < n_far4 - n_far) {
		const int dst = n_far + threadIdx.x;
		sh.tbl[dst].sx = 0.f;
		sh.sxi[dst] = 0;
		sh.sy[dst] = kNoSample;
#pragma unroll
		for (int c = 0; c < 3; c++)
			sh.tbl[dst].b[c] = 0.f;
	}
#pragma unroll
	for (int c = 0; c < 3; c++) {
		if (mn[c] <= mx[c]) {
			atomicMin(&sh.smin[c], mn[c]);
			atomicMax(&sh.smax[c], mx[c]);
		}
	}
	if (threadIdx.x == kAceBlock - 1) {     // last thread: offsets are the totals
		sh.n_far4 = n_far4;
		sh.n_str = off_str - n_far4;
	}
	__syncthreads();
}

// One sample against the kRows pixels of a thread (row pairs in f32x2
// registers).  TEST: the sample straddles the skip circle on this tile, so
// each pair is tested (exactly, in integers) and a skipped one gets weight 0.
template <bool TEST>
__device__ __forceinline__ void accumulate_sample(const AceShared &sh, int j, float fx, int x, int y0,
		int too_close2, const float (&pv)[kRows][3], float inv, f32x2 (&acc)[kRows / 2][4])
{
	const SampleEntry se = sh.tbl[j];
	const float dx = fx - se.sx;
	const float dx2 = dx * dx;
	const f32x2 dx22 = pack2(dx2, dx2);
	int dxi2 = 0, dy0 = 0;
	if (TEST) {
		const int dxi = x - sh.sxi[j];
		dxi2 = dxi * dxi;
		dy0 = y0 - sh.sy[j];
	}
#pragma unroll
	for (int r = 0; r < kRows / 2; r++) {
		const f32x2 dy22 = *reinterpret_cast<const f32x2 *>(&sh.dy2[j * kRows + 2 * r]);
		float da, db;
		unpack2(fadd2(dx22, dy22), da, db);
		float wa = rsqrt_f32(da), wb = rsqrt_f32(db);
		if (TEST) {
			const int dya = dy0 + 2 * r, dyb = dya + 1;
			if (dxi2 + dya * dya < too_close2)
				wa = 0.f;
			if (dxi2 + dyb * dyb < too_close2)
				wb = 0.f;
		}
		const f32x2 wv = pack2(wa, wb);
		acc[r][3] = fadd2(acc[r][3], wv);
#pragma unroll
		for (int c = 0; c < 3; c++)
			acc[r][c] = ffma2(wv, pack2(fma_sat(pv[2 * r][c], inv, se.b[c]),
					fma_sat(pv[2 * r + 1][c], inv, se.b[c])), acc[r][c]);
	}
}

__device__ __forceinline__ void zero_sums(f32x2 (&a)[kRows / 2][4])
{
#pragma unroll
	for (int r = 0; r < kRows / 2; r++)
#pragma unroll
		for (int c = 0; c < 4; c++)
			a[r][c] = 0ull;
}

__device__ __forceinline__ void flush_sums(f32x2 (&total)[kRows / 2][4], const f32x2 (&part)[kRows / 2][4])
{
#pragma unroll
	for (int r = 0; r < kRows / 2; r++)
#pragma unroll
		for (int c = 0; c < 4; c++)
			total[r][c] = fadd2(total[r][c], part[r][c]);
}

// The accumulation pass.  Thread = one column, kRows rows at a time; block =
// one tile of 128 columns x kTileRows rows of a page-wide grid (so that a band
// split across GPUs classifies and sums exactly like the whole page).
__global__ void __launch_bounds__(kAceBlock, 4)
ace_scores_kernel(const uint32_t *__restrict__ in, const int *__restrict__ samples_xy,
		const AceConst cs, float *__restrict__ scores, int *__restrict__ fix_counts,
		double *__restrict__ minmax)
{
	extern __shared__ __align__(16) unsigned char smem_raw[];
	AceShared &sh = *reinterpret_cast<AceShared *>(smem_raw);
	pdl_wait();
	const int x = blockIdx.x * kAceBlock + threadIdx.x;
	const int xc = min(x, cs.w - 1);
	const int tile_y0 = (cs.row_begin / kTileRows + blockIdx.y) * kTileRows;
	const int tile_y1 = min(cs.h, tile_y0 + kTileRows);      // exclusive
	const int x_lo = blockIdx.x * kAceBlock;
	const int x_hi = min(cs.w - 1, x_lo + kAceBlock - 1);
	const size_t plane = (size_t)cs.score_plane;
	const float fx = (float)x;
	const float inv = cs.inv;

	// the fix-up lists of this score block start empty (ace_scale_kernel fills them)
	if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x < 3 * kAceFixupSlots)
		fix_counts[threadIdx.x] = 0;      // list lengths, the fix-up pass's tickets, its statistics
	if (threadIdx.x < 3) {
		sh.smin[threadIdx.x] = 255;
		sh.smax[threadIdx.x] = 0;
	}

	float mn[3] = { FLT_MAX, FLT_MAX, FLT_MAX };
	float mx[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };

	bool staged = false;
	for (int y0 = tile_y0; y0 < tile_y1; y0 += kRows) {
		if (y0 + kRows <= cs.row_begin || y0 >= cs.row_end)
			continue;       // row group outside this band (block-uniform)
		// pixel values of this thread's rows; rows of the group outside the band
		// repeat the nearest one inside (their scores are not stored), so only
		// the band's own rows of `in` are ever read
		float pv[kRows][3];
#pragma unroll
		for (int r = 0; r < kRows; r++) {
			const uint32_t pix = __ldg(in + (size_t)min(max(y0 + r, cs.row_begin), cs.row_end - 1) * cs.w + xc);
#pragma unroll
			for (int c = 0; c < 3; c++)
				pv[r][c] = (float)((pix >> (8 * c)) & 0xFF);
		}

		f32x2 total[kRows / 2][4], part[kRows / 2][4];
		zero_sums(total);

		for (int k0 = 0; k0 < cs.nb_samples; k0 += kChunk) {
			const int kn = min(kChunk, cs.nb_samples - k0);
			// (re)stage when the chunk or the tile changes; with one chunk the
			// table is staged once per tile
			if (!staged || cs.nb_samples > kChunk) {
				__syncthreads();
				staged = true;
				stage_samples(sh, cs, in, samples_xy, k0, kn, x_lo, x_hi, tile_y0, tile_y1 - 1);
			}
			const int n_far4 = sh.n_far4;
			const int n_act = n_far4 + sh.n_str;
			__syncthreads();
			for (int j = threadIdx.x; j < n_act; j += kAceBlock) {
				const int sy = sh.sy[j];
#pragma unroll
				for (int r = 0; r < kRows; r++) {
					const float dy = (float)(y0 + r - sy);
					sh.dy2[j * kRows + r] = (sy == kNoSample) ? __int_as_float(0x7f800000) : dy * dy;
				}
			}
			__syncthreads();

			for (int j0 = 0; j0 < n_far4; j0 += kFlush) {
				zero_sums(part);
				const int j1 = min(n_far4, j0 + kFlush);
				for (int j = j0; j < j1; j += 4) {
#pragma unroll
					for (int u = 0; u < 4; u++)
						accumulate_sample<false>(sh, j + u, fx, x, y0, cs.too_close2, pv, inv, part);
				}
				flush_sums(total, part);
			}
			for (int j0 = n_far4; j0 < n_act; j0 += kFlush) {
				zero_sums(part);
				const int j1 = min(n_act, j0 + kFlush);
				for (int j = j0; j < j1; j++)
					accumulate_sample<true>(sh, j, fx, x, y0, cs.too_close2, pv, inv, part);
				flush_sums(total, part);
			}
		}

		// score = 2 * sum(w u) / sum(w) - 1, and the ties: exact comparisons of the
		// pixel with the extremes of the tile's staged samples (a superset of
		// the samples that count for this pixel: a tie that only holds for the
		// subset is not claimed and takes the ordinary route).  All values are
		// small integers, exact as floats.
		float tie_hi[3], tie_lo[3], tie_eq[3];
#pragma unroll
		for (int c = 0; c < 3; c++) {
			const int lo = sh.smin[c], hi = sh.smax[c];
			tie_hi[c] = (float)(hi + cs.dsat);          // p >= this: every difference saturates high
			tie_lo[c] = (float)(lo - cs.dsat);          // p <= this: ... low
			tie_eq[c] = lo == hi ? (float)lo : -1.f;    // p == this: every difference is zero
		}
		if (x < cs.w) {
			float *row = scores + ((size_t)(y0 - cs.score_row0) * cs.w + x);
#pragma unroll
			for (int r = 0; r < kRows; r++, row += cs.w) {
				const int y = y0 + r;
				if (y < cs.row_begin || y >= cs.row_end || y >= tile_y1)
					continue;
				float a[4], lo, hi;
#pragma unroll
				for (int c = 0; c < 4; c++) {
					unpack2(total[r / 2][c], lo, hi);
					a[c] = (r & 1) ? hi : lo;
				}
				// no sample counts for this pixel: the reference divides 0 by 0; a
				// sample AT the pixel (pages of fewer than 5 rows skip nothing)
				// makes its sums infinite or NaN, and its score NaN
				const bool live = a[3] > 0.f && a[3] <= FLT_MAX;
				const float den = 0.5f * a[3];
				float rden;
				asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rden) : "f"(den));     // 2^-23: inside the bound's epilogue share
				float *dst = row;
#pragma unroll
				for (int c = 0; c < 3; c++, dst += plane) {
					const float p = pv[r][c];
					float s = (a[c] - den) * rden;
					// keep approximate scores off the values that mean "tie"
					s = fminf(fmaxf(s, -0.99999994f), 0.99999994f);
					s = s == 0.f ? 1e-30f : s;
					s = p >= tie_hi[c] ? 1.f : s;
					s = p <= tie_lo[c] ? -1.f : s;
					s = p == tie_eq[c] ? 0.f : s;
					s = live ? s : __int_as_float(0x7fc00000);
					*dst = s;
					mn[c] = fminf(mn[c], s);      // NaN never enters: fminf / fmaxf return the other operand
					mx[c] = fmaxf(mx[c], s);
				}
			}
		}
	}

	// block-wide min / max, then 6 atomics per block
	__shared__ float red[2][3][kAceBlock / 32];
#pragma unroll
	for (int c = 0; c < 3; c++) {
#pragma unroll
		for (int d = 16; d > 0; d >>= 1) {
			mn[c] = fminf(mn[c], __shfl_xor_sync(0xFFFFFFFFu, mn[c], d));
			mx[c] = fmaxf(mx[c], __shfl_xor_sync(0xFFFFFFFFu, mx[c], d));
		}
		if ((threadIdx.x & 31) == 0) {
			red[0][c][threadIdx.x >> 5] = mn[c];
			red[1][c][threadIdx.x >> 5] = mx[c];
		}
	}
	__syncthreads();
	if (threadIdx.x < 3) {
		const int c = threadIdx.x;
		float a = red[0][c][0], b = red[1][c][0];
		for (int i = 1; i < kAceBlock / 32; i++) {
			a = fminf(a, red[0][c][i]);
			b = fmaxf(b, red[1][c][i]);
		}
		// FLT_MAX (nothing seen) becomes DBL_MAX, the block's initial value
		const double da = a == FLT_MAX ? DBL_MAX : (double)a;
		const double db = b == -FLT_MAX ? -DBL_MAX : (double)b;
		atomic_min_double(minmax + c, da);
		atomic_max_double(minmax + 3 + c, db);
		// and per tile of the page-wide grid (two bands may share a tile), for
		// ace_refine_kernel to find the tiles that matter
		double *tile = minmax + kMinmaxHead
			+ 6 * ((size_t)(cs.row_begin / kTileRows + blockIdx.y) * gridDim.x + blockIdx.x);
		atomic_min_double(tile + c, da);
		atomic_max_double(tile + 3 + c, db);
	}
}

// The sample table and the samples' pixel values as one message of 3 N ints:
// x0, y0, x1, y1, ..., then the N pixels (what a GPU that holds only a band
// of the page needs besides its rows).
__global__ void ace_gather_samples_kernel(const uint32_t *__restrict__ page, int w,
		const int *__restrict__ samples_xy, int n, int *__restrict__ msg)
{
	pdl_wait();
	for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
		const int2 s = __ldg(reinterpret_cast<const int2 *>(samples_xy) + k);
		msg[2 * k] = s.x;
		msg[2 * k + 1] = s.y;
		msg[2 * n + k] = (int)__ldg(page + (size_t)s.y * w + s.x);
	}
}

__global__ void ace_minmax_init_kernel(double *minmax, size_t n)
{
	pdl_wait();
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
			i += (size_t)gridDim.x * blockDim.x)
		minmax[i] = (i % 6 < 3) ? DBL_MAX : -DBL_MAX;
}

// ---------------------------------------------------------------------------
// The pixels the fast pass cannot decide: first in FP64 (error ~1e-14, a
// thread per pixel), then, for the handful that even that leaves open, in the
// reference's own arithmetic (a warp per pixel).
// ---------------------------------------------------------------------------

// 1/sqrt(x) for an integer-valued x < 2^31: MUFU.RSQ64H looks at the high word
// of x only (20 mantissa bits) and is good to about 2^-20 of such an argument;
// two Newton steps (each squares the error: e = 1 - x y^2, y' = y + (y/2) e;
// y/2 is an exponent decrement of the high word) leave the rounding of the
// last step, a few 2^-53.  pf_b200_selftest_rsqrt_error(1) measures the
// result over every argument; the bound below assumes 2^-48.
__device__ __forceinline__ double rsqrt_f64(double x)
{
	double y;
	asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
#pragma unroll
	for (int i = 0; i < 2; i++) {
		const double e = fma(-(x * y), y, 1.0);
		const double h = __hiloint2double(__double2hiint(y) - 0x00100000, __double2loint(y));
		y = fma(h, e, y);
	}
	return y;
}

// Bound on |FP64 score (ace_accurate_sums) - reference score|.  Both
// approximate the same real number: the reference with one rounding in the
// square root, one in each division and one per addition; this with weights
// within 2^-48 (the score moves by at most twice that: the same weight is in
// numerator and denominator), one rounding per FMA (in any order) and one
// division: 2^-47 + (4 N + 64) * 2^-53, doubled for room.
__device__ __forceinline__ double ace_accurate_eps(int nb_samples)
{
	return 2.0 * (0x1p-47 + (4.0 * nb_samples + 64.0) * 0x1p-53);
}

// What the exact passes read per sample, staged once per block: coordinates
// and pixel value, and the reference's saturation value for every possible
// channel difference (upstream _ace.c calc_saturation: slope * d, clamped).
constexpr int kTableMax = 512;       // samples kept in shared memory; more are read from global memory
struct ExactTable {
	double sat[511];                 // [d + 255]
	int4 smp[kTableMax];             // x, y, pixel, -
	int4 off[kTableMax];             // byte offsets (255 - s_c) * 8 into sat, per channel
};

// Fills the table (all threads of the block).  Samples beyond kTableMax are
// read from global memory by the loops below.
__device__ __forceinline__ void stage_exact_table(ExactTable &tb, const uint32_t *__restrict__ in,
		const int *__restrict__ samples_xy, const AceConst &cs)
{
	for (int i = threadIdx.x; i < 511; i += blockDim.x) {
		double v = __dmul_rn((double)(i - 255), cs.slope);
		tb.sat[i] = v > cs.limit ? cs.limit : (v < -cs.limit ? -cs.limit : v);
	}
	const int n = min(cs.nb_samples, kTableMax);
	for (int k = threadIdx.x; k < n; k += blockDim.x) {
		const int2 sxy = __ldg(reinterpret_cast<const int2 *>(samples_xy) + k);
		const uint32_t px = cs.sample_px ? __ldg(cs.sample_px + k) : __ldg(in + (size_t)sxy.y * cs.w + sxy.x);
		tb.smp[k] = make_int4(sxy.x, sxy.y, (int)px, 0);
		tb.off[k] = make_int4((255 - (int)(px & 0xFF)) * 8, (255 - (int)((px >> 8) & 0xFF)) * 8,
				(255 - (int)((px >> 16) & 0xFF)) * 8, 0);
	}
	__syncthreads();
}

// sample k: coordinates and pixel, from the table or from global memory
__device__ __forceinline__ int4 exact_sample(const ExactTable &tb, const uint32_t *__restrict__ in,
		const int *__restrict__ samples_xy, const AceConst &cs, int k)
{
	if (k < kTableMax)
		return tb.smp[k];
	const int2 sxy = __ldg(reinterpret_cast<const int2 *>(samples_xy) + k);
	const uint32_t px = cs.sample_px ? __ldg(cs.sample_px + k) : __ldg(in + (size_t)sxy.y * cs.w + sxy.x);
	return make_int4(sxy.x, sxy.y, (int)px, 0);
}

// Sums of one pixel (value p at x, y) over samples k_begin, k_begin + k_step,
// ... in FP64: weights 1/dist from the reciprocal square root, saturation
// values exactly the reference's.  acc[0..2]: R, G, B; acc[3]: the denominator.
__device__ __forceinline__ void ace_accurate_sums(const ExactTable &tb,
		const uint32_t *__restrict__ in, const int *__restrict__ samples_xy, const AceConst &cs,
		uint32_t p, int x, int y, int k_begin, int k_step, double (&acc)[4])
{
	const char *sat = reinterpret_cast<const char *>(tb.sat);
	const int po[3] = { (int)(p & 0xFF) * 8, (int)((p >> 8) & 0xFF) * 8, (int)((p >> 16) & 0xFF) * 8 };
	const int too_close2 = cs.too_close2;
	const double limit = cs.limit;
	const int n_tab = min(cs.nb_samples, kTableMax);
#pragma unroll 4
	for (int k = k_begin; k < n_tab; k += k_step) {
		const int4 sm = tb.smp[k], of = tb.off[k];
		const int dx = x - sm.x, dy = y - sm.y;
		const int d2 = dx * dx + dy * dy;
		// (2^52 + d2) - 2^52: the integer as a double without the conversion unit;
		// d2 == 0 (pages under 5 rows) gives inf, and the sums turn NaN as in the reference
		double wgt = rsqrt_f64(__hiloint2double(0x43300000, d2) - 4503599627370496.0);
		wgt = d2 < too_close2 ? 0.0 : wgt;
		acc[0] = fma(wgt, *reinterpret_cast<const double *>(sat + po[0] + of.x), acc[0]);
		acc[1] = fma(wgt, *reinterpret_cast<const double *>(sat + po[1] + of.y), acc[1]);
		acc[2] = fma(wgt, *reinterpret_cast<const double *>(sat + po[2] + of.z), acc[2]);
		acc[3] = fma(wgt, limit, acc[3]);
	}
	for (int k = kTableMax + k_begin; k < cs.nb_samples; k += k_step) {      // (k_step divides kTableMax)
		const int4 sm = exact_sample(tb, in, samples_xy, cs, k);
		const int dx = x - sm.x, dy = y - sm.y;
		const int d2 = dx * dx + dy * dy;
		double wgt = rsqrt_f64(__hiloint2double(0x43300000, d2) - 4503599627370496.0);
		wgt = d2 < too_close2 ? 0.0 : wgt;
#pragma unroll
		for (int c = 0; c < 3; c++)
			acc[c] = fma(wgt, tb.sat[(int)((p >> (8 * c)) & 0xFF) - (int)(((uint32_t)sm.z >> (8 * c)) & 0xFF) + 255], acc[c]);
		acc[3] = fma(wgt, limit, acc[3]);
	}
}

constexpr int kExactChunk = 64;      // samples whose terms a warp holds at a time

// The scores of one pixel in the reference's own arithmetic, by one warp:
// upstream _ace.c ace_thread_adjustment / oracle ace_adjust_rows operation for
// operation -- IEEE sqrt and divide, terms added in sample order, a skipped
// sample adding nothing (+0.0 changes no sum).  The lanes compute the terms of
// 64 samples at a time (that is where the time goes), then lanes 0..3 add them
// up in order: R, G, B and the denominator.  terms: this warp's shared area.
// All lanes return the three scores.
__device__ __forceinline__ void ace_exact_score_warp(double (*terms)[4], const ExactTable &tb,
		const uint32_t *__restrict__ in, const int *__restrict__ samples_xy, const AceConst &cs,
		int x, int y, double (&r)[3])
{
	const int lane = threadIdx.x & 31;
	const uint32_t p = __ldg(in + (size_t)y * cs.w + x);
	const double too_close = (double)(cs.h / 5);
	double acc = 0.0;
	for (int k0 = 0; k0 < cs.nb_samples; k0 += kExactChunk) {
#pragma unroll
		for (int half = 0; half < kExactChunk / 32; half++) {
			const int j = lane + 32 * half, k = k0 + j;
			double t[4] = { 0.0, 0.0, 0.0, 0.0 };
			if (k < cs.nb_samples) {
				const int4 sm = exact_sample(tb, in, samples_xy, cs, k);
				const uint32_t sp = (uint32_t)sm.z;
				const double ddx = (double)(x - sm.x), ddy = (double)(y - sm.y);
				const double dist = __dsqrt_rn(__dadd_rn(__dmul_rn(ddx, ddx), __dmul_rn(ddy, ddy)));
				if (!(dist < too_close)) {
#pragma unroll
					for (int c = 0; c < 3; c++)
						t[c] = __ddiv_rn(tb.sat[(int)((p >> (8 * c)) & 0xFF) - (int)((sp >> (8 * c)) & 0xFF) + 255], dist);
					t[3] = __ddiv_rn(cs.limit, dist);
				}
			}
#pragma unroll
			for (int c = 0; c < 4; c++)
				terms[j][c] = t[c];
		}
		__syncwarp();
		if (lane < 4) {
			const int n = min(kExactChunk, cs.nb_samples - k0);
#pragma unroll 8
			for (int j = 0; j < n; j++)
				acc = __dadd_rn(acc, terms[j][lane]);
		}
		__syncwarp();
	}
	const double denom = __shfl_sync(0xFFFFFFFFu, acc, 3);
	const double q = __ddiv_rn(acc, denom);
#pragma unroll
	for (int c = 0; c < 3; c++)
		r[c] = __shfl_sync(0xFFFFFFFFu, q, c);
}

constexpr int kRefineThreads = 256;
constexpr int kRefineSlab = 16;      // rows scanned before the candidates are worked off

struct RefineShared {
	ExactTable tb;
	double terms[kRefineThreads / 32][kExactChunk][4];
	unsigned short cand[kRefineSlab * kAceBlock];
	unsigned short todo[kRefineSlab * kAceBlock];
	int wanted[kRefineThreads];
	int n_wanted, n_cand, n_todo;
};

// The page's extremes exactly.  The rescale divides by (max - min) and
// truncates, so the extremes must be the reference's to the last bit.
//   The exact minimum is the exact score of SOME pixel whose approximate score
// lies within 2 eps of the approximate minimum: the candidates.  Phase 1
// computes the candidates' scores in FP64 (ace_accurate_sums, error e2) and
// takes the extremes of those; the exact minimum then belongs to a candidate
// whose FP64 score lies within 2 e2 of that FP64 minimum -- a handful of
// pixels -- and phase 2 recomputes those in the reference's own arithmetic.
// Tied scores (0, +-1 in the planes) are exact as they are and take part as
// they stand.  Each phase scans the same tiles (a few: the tile table the
// accumulation kernel left says which can hold a candidate).
template <int PHASE>
__device__ __forceinline__ void refine_scan(RefineShared &sh, const int *__restrict__ samples_xy,
		const uint32_t *__restrict__ in, const AceConst &cs, const float *__restrict__ scores,
		double *__restrict__ minmax, int tiles_x, int tiles_y)
{
	double lo[3], hi[3], alo[3], ahi[3];
	const double thr = 2.0 * (double)cs.eps;
	const double thr2 = 2.0 * ace_accurate_eps(cs.nb_samples);
#pragma unroll
	for (int c = 0; c < 3; c++) {
		lo[c] = minmax[c] + thr;
		hi[c] = minmax[3 + c] - thr;
		alo[c] = (PHASE == 2) ? minmax[12 + c] + thr2 : 0.0;
		ahi[c] = (PHASE == 2) ? minmax[15 + c] - thr2 : 0.0;
	}
	const int col = threadIdx.x % kAceBlock;
	const int tile_row0 = cs.row_begin / kTileRows;
	const size_t plane = (size_t)cs.score_plane;
	const int n_tiles = tiles_x * tiles_y;
	// extremes this thread has seen: phase 1 the FP64 ones (and ties, which are
	// exact, into the exact ones), phase 2 the exact ones
	double mn1[3] = { DBL_MAX, DBL_MAX, DBL_MAX }, mx1[3] = { -DBL_MAX, -DBL_MAX, -DBL_MAX };
	double emn[3] = { DBL_MAX, DBL_MAX, DBL_MAX }, emx[3] = { -DBL_MAX, -DBL_MAX, -DBL_MAX };

	for (int base = 0; base < n_tiles; base += gridDim.x * kRefineThreads) {
		__syncthreads();
		if (threadIdx.x == 0)
			sh.n_wanted = 0;
		__syncthreads();
		const int t = base + threadIdx.x * gridDim.x + blockIdx.x;
		if (t < n_tiles) {
			const double *tile = minmax + kMinmaxHead
				+ 6 * ((size_t)(tile_row0 + t / tiles_x) * tiles_x + t % tiles_x);
			bool wanted = false, uniform = true;
			double tv[3];
#pragma unroll
			for (int c = 0; c < 3; c++) {
				tv[c] = tile[c];
				wanted |= (tile[c] <= lo[c]) || (tile[3 + c] >= hi[c]);
				uniform = uniform && tile[c] == tile[3 + c] && (tv[c] == 0.0 || tv[c] == 1.0 || tv[c] == -1.0);
			}
			if (wanted && uniform) {
				// every score of the tile is the same tie (flat areas of a page):
				// exact as it stands, nothing to look at
				if (PHASE == 1) {
#pragma unroll
					for (int c = 0; c < 3; c++) {
						if (tv[c] <= lo[c] || tv[c] >= hi[c]) {
							emn[c] = fmin(emn[c], tv[c]);
							emx[c] = fmax(emx[c], tv[c]);
						}
					}
				}
				wanted = false;
			}
			if (wanted)
				sh.wanted[atomicAdd(&sh.n_wanted, 1)] = t;
		}
		__syncthreads();
		const int n_wanted = sh.n_wanted;
		for (int wi = 0; wi < n_wanted; wi++) {
			const int tw = sh.wanted[wi];
			const int x_lo = (tw % tiles_x) * kAceBlock;
			const int tile_y0 = (tile_row0 + tw / tiles_x) * kTileRows;
			const int y_begin = max(tile_y0, cs.row_begin);
			const int y_end = min(min(cs.h, tile_y0 + kTileRows), cs.row_end);
			for (int ys = y_begin; ys < y_end; ys += kRefineSlab) {
				__syncthreads();
				if (threadIdx.x == 0) {
					sh.n_cand = 0;
					sh.n_todo = 0;
				}
				__syncthreads();
				constexpr int kStep = kRefineThreads / kAceBlock;
#pragma unroll
				for (int i = 0; i < kRefineSlab / kStep; i++) {
					const int y = ys + threadIdx.x / kAceBlock + i * kStep;
					if (y >= y_end || x_lo + col >= cs.w)
						continue;
					const size_t o = (size_t)(y - cs.score_row0) * cs.w + x_lo + col;
					bool need = false;
#pragma unroll
					for (int c = 0; c < 3; c++) {
						const float sv = scores[c * plane + o];
						if (!((double)sv <= lo[c] || (double)sv >= hi[c]))    // also NaN
							continue;
						if (is_tie_score(sv)) {
							if (PHASE == 1) {
								emn[c] = fmin(emn[c], (double)sv);
								emx[c] = fmax(emx[c], (double)sv);
							}
						} else {
							need = true;
						}
					}
					if (need)
						sh.cand[atomicAdd(&sh.n_cand, 1)] = (unsigned short)
							((threadIdx.x / kAceBlock + i * kStep) * kAceBlock + col);
				}
				__syncthreads();
				const int n = sh.n_cand;
				// The candidates' scores in FP64.  They take part in every channel
				// (true scores of the page can only move the extremes toward the
				// truth).  A few candidates: a warp each, the samples dealt out
				// over the lanes (latency); many: a thread each (throughput).
				const int lane = threadIdx.x & 31;
				const bool by_warp = n <= 2 * (kRefineThreads / 32);
				for (int i = by_warp ? (int)(threadIdx.x >> 5) : (int)threadIdx.x; i < n;
						i += by_warp ? kRefineThreads / 32 : kRefineThreads) {
					const int cd = sh.cand[i];
					const int px = x_lo + cd % kAceBlock, py = ys + cd / kAceBlock;
					double acc[4] = { 0.0, 0.0, 0.0, 0.0 };
					const uint32_t pv = __ldg(in + (size_t)py * cs.w + px);
					if (by_warp) {
						ace_accurate_sums(sh.tb, in, samples_xy, cs, pv, px, py, lane, 32, acc);
#pragma unroll
						for (int c = 0; c < 4; c++)
#pragma unroll
							for (int d = 16; d > 0; d >>= 1)
								acc[c] += __shfl_xor_sync(0xFFFFFFFFu, acc[c], d);
					} else {
						ace_accurate_sums(sh.tb, in, samples_xy, cs, pv, px, py, 0, 1, acc);
					}
					bool exact = false;
#pragma unroll
					for (int c = 0; c < 3; c++) {
						const double sa = acc[c] / acc[3];
						if (PHASE == 1) {
							if (sa < mn1[c]) mn1[c] = sa;       // NaN never enters
							if (sa > mx1[c]) mx1[c] = sa;
						} else {
							exact |= sa <= alo[c] || sa >= ahi[c];
						}
					}
					if (PHASE == 2 && exact && (!by_warp || lane == 0))
						sh.todo[atomicAdd(&sh.n_todo, 1)] = (unsigned short)cd;
				}
				if (PHASE == 2) {
					__syncthreads();
					// a warp per pixel that may hold an extreme: the reference's arithmetic
					const int n_todo = sh.n_todo, warp = threadIdx.x >> 5;
					for (int i = warp; i < n_todo; i += kRefineThreads / 32) {
						const int cd = sh.todo[i];
						double r[3];
						ace_exact_score_warp(sh.terms[warp], sh.tb, in, samples_xy, cs,
								x_lo + cd % kAceBlock, ys + cd / kAceBlock, r);
#pragma unroll
						for (int c = 0; c < 3; c++) {
							if (r[c] < emn[c]) emn[c] = r[c];
							if (r[c] > emx[c]) emx[c] = r[c];
						}
					}
				}
			}
		}
	}
	// fold this thread's extremes into the block of the page
#pragma unroll
	for (int c = 0; c < 3; c++) {
#pragma unroll
		for (int d = 16; d > 0; d >>= 1) {
			mn1[c] = fmin(mn1[c], __shfl_xor_sync(0xFFFFFFFFu, mn1[c], d));
			mx1[c] = fmax(mx1[c], __shfl_xor_sync(0xFFFFFFFFu, mx1[c], d));
			emn[c] = fmin(emn[c], __shfl_xor_sync(0xFFFFFFFFu, emn[c], d));
			emx[c] = fmax(emx[c], __shfl_xor_sync(0xFFFFFFFFu, emx[c], d));
		}
		if ((threadIdx.x & 31) == 0) {
			if (PHASE == 1 && mn1[c] != DBL_MAX)
				atomic_min_double(minmax + 12 + c, mn1[c]);
			if (PHASE == 1 && mx1[c] != -DBL_MAX)
				atomic_max_double(minmax + 15 + c, mx1[c]);
			if (emn[c] != DBL_MAX)
				atomic_min_double(minmax + 6 + c, emn[c]);
			if (emx[c] != -DBL_MAX)
				atomic_max_double(minmax + 9 + c, emx[c]);
		}
	}
}

// The two phases as two launches (the launch boundary is the grid-wide barrier
// between them; with programmatic dependent launch the second is scheduled
// while the first drains).
template <int PHASE>
__global__ void __launch_bounds__(kRefineThreads)
ace_refine_kernel(const uint32_t *__restrict__ in, const int *__restrict__ samples_xy,
		const AceConst cs, const float *__restrict__ scores, double *minmax,
		int tiles_x, int tiles_y)
{
	__shared__ RefineShared sh;
	pdl_wait();
	stage_exact_table(sh.tb, in, samples_xy, cs);
	refine_scan<PHASE>(sh, samples_xy, in, cs, scores, minmax, tiles_x, tiles_y);
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

// Destinations of a rescale: `out`, and optionally further images that may
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
__device__ __forceinline__ void store_pixel(uint32_t *out, const ScaleOuts &more, size_t i, uint32_t rgba)
{
	if (MULTI && more.multicast)
		multimem_store4(out + i, rgba);
	else
		out[i] = rgba;
	if (MULTI)
		for (int d = 0; d < more.n; d++)
			more.p[d][i] = rgba;
}

// The exact extremes of a channel (ace_refine_kernel's; every page has them
// unless all its scores are NaN).
__device__ __forceinline__ void exact_extremes(const double *__restrict__ minmax, int c, double &mn, double &span)
{
	mn = minmax[6 + c];
	const double mx = minmax[9 + c];
	span = (mn == DBL_MAX || mx == -DBL_MAX) ? 0.0 : mx - mn;
}

// out[c] = trunc(clamp(255 * (score - min) / (max - min), 0, 255)); alpha 255.
// T = (s - min) * K + 2^23 + 2^14 with K = 255 / span * 2^14 is an FP32
// subtraction and an FMA rounded toward zero: the low mantissa bits of T are
// the value in units of 2^-14 output steps, shifted up by one step so that a
// value a little below zero stays positive.  The approximate score is within
// eps of the reference's, so T is within G = (eps + 1e-7) * K + 4 units of the
// reference's quotient (1e-7: min rounded to FP32; 4: the roundings of the
// difference, of K and of T); a value farther than that from an integer
// truncates like the reference's.  The others go on the fix-up list, and the
// tied scores (exactly 0, 1, -1) take their three precomputed results.
template <bool MULTI>
__global__ void __launch_bounds__(256)
ace_scale_kernel(const float *__restrict__ scores, const double *__restrict__ minmax,
		size_t npix, size_t plane, size_t pix_base, float eps,
		uint32_t *__restrict__ out, const ScaleOuts more,
		uint32_t *__restrict__ fix_list, int *__restrict__ fix_count)
{
	// per channel: K, min (FP32), guard band, and the results of the three tie values
	__shared__ float s_k[3], s_mn[3];
	__shared__ uint32_t s_guard[3], s_tie[3][3];
	pdl_wait();
	if (threadIdx.x < 9) {
		const int c = threadIdx.x / 3, v = threadIdx.x % 3;
		double mn, span;
		exact_extremes(minmax, c, mn, span);
		s_tie[c][v] = ace_scale_exact((double)(v - 1), mn, span);
		if (v == 0) {
			// a flat channel (span 0: every score equal, or none) comes out 0 through
			// K = 0, with an empty guard band
			const bool flat = !(span > 0.0);
			const double k = flat ? 0.0 : 255.0 / span * 16384.0;
			const double g = ((double)eps + 1e-7) * k + 4.0;
			s_k[c] = (float)k;
			s_mn[c] = flat ? 0.f : (float)mn;
			s_guard[c] = flat ? 0u : (g < 8192.0 ? (uint32_t)g + 1u : 8192u);   // 8192: everything is uncertain
		}
	}
	__syncthreads();
	float kf[3], mnf[3];
	uint32_t guard[3];
#pragma unroll
	for (int c = 0; c < 3; c++) {
		kf[c] = s_k[c];
		mnf[c] = s_mn[c];
		guard[c] = s_guard[c];
	}
	bool vec = (npix % 4 == 0) && (plane % 4 == 0)
		&& (((uintptr_t)scores | (uintptr_t)out) % 16 == 0);
	if (MULTI)
		for (int d = 0; d < more.n; d++)
			vec = vec && ((uintptr_t)more.p[d] % 16 == 0);
	const size_t nwork = vec ? npix / 4 : npix;
	const int per = vec ? 4 : 1;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	// Uncertain pixels are rare (about one in a hundred): each warp collects
	// its own in shared memory and the block appends them to the list with one
	// atomic at the end; a warp whose buffer is full appends directly.
	constexpr int kBuf = 192;
	__shared__ uint32_t s_buf[256 / 32][kBuf];
	__shared__ int s_n[256 / 32], s_base[256 / 32];
	if (lane == 0)
		s_n[warp] = 0;
	__syncwarp();
	for (size_t g0 = (size_t)blockIdx.x * blockDim.x; g0 < nwork; g0 += (size_t)gridDim.x * blockDim.x) {
		const size_t gi = g0 + threadIdx.x;
		if (gi >= nwork)
			continue;
		float sv[3][4];
		if (vec) {
#pragma unroll
			for (int c = 0; c < 3; c++) {
				const float4 a = __ldcs(reinterpret_cast<const float4 *>(scores + c * plane) + gi);
				sv[c][0] = a.x; sv[c][1] = a.y; sv[c][2] = a.z; sv[c][3] = a.w;
			}
		} else {
#pragma unroll
			for (int c = 0; c < 3; c++) {
				sv[c][0] = __ldcs(scores + c * plane + gi);
				sv[c][1] = sv[c][2] = sv[c][3] = 0.5f;
			}
		}
		uint32_t px[4] = { 0xFF000000u, 0xFF000000u, 0xFF000000u, 0xFF000000u };
		unsigned risky = 0;
#pragma unroll
		for (int c = 0; c < 3; c++) {
#pragma unroll
			for (int k = 0; k < 4; k++) {
				const float s = sv[c][k];
				// low 14 bits: the fraction; above them the integer part + 1
				const uint32_t fx = __float_as_uint(__fmaf_rz(s - mnf[c], kf[c], 8404992.0f)) - 0x4B000000u;
				uint32_t q = (uint32_t)__viaddmin_s32_relu((int)(fx >> 14), -1, 255);
				bool unsure = ((fx + guard[c]) & 0x3FFFu) < 2u * guard[c] || fx >= (258u << 14);
				if (!(fabsf(s) < 1.f) || s == 0.f) {
					// a tie (exactly 0, 1, -1) or no score at all (NaN)
					q = s == 0.f ? s_tie[c][1] : (s == 1.f ? s_tie[c][2] : (s == -1.f ? s_tie[c][0] : 0u));
					unsure = false;
				}
				if (unsure)
					risky |= 1u << k;
				px[k] |= q << (8 * c);
			}
		}
		if (vec) {
			const uint4 v = make_uint4(px[0], px[1], px[2], px[3]);
			if (MULTI && more.multicast)
				multimem_store16(reinterpret_cast<uint4 *>(out) + gi, v);
			else
				reinterpret_cast<uint4 *>(out)[gi] = v;
			if (MULTI)
				for (int d = 0; d < more.n; d++)
					reinterpret_cast<uint4 *>(more.p[d])[gi] = v;
		} else {
			store_pixel<MULTI>(out, more, gi, px[0]);
			risky &= 1u;
		}
		while (risky) {
			const int k = __ffs(risky) - 1;
			risky &= risky - 1;
			const uint32_t entry = (uint32_t)(pix_base + (size_t)per * gi + k);
			const int at = atomicAdd(&s_n[warp], 1);
			if (at < kBuf)
				s_buf[warp][at] = entry;
			else
				fix_list[atomicAdd(fix_count, 1)] = entry;
		}
	}
	__syncthreads();
	if (threadIdx.x == 0) {
		int total = 0;
		for (int wv = 0; wv < 256 / 32; wv++) {
			s_n[wv] = min(s_n[wv], kBuf);
			s_base[wv] = total;
			total += s_n[wv];
		}
		const int base = total ? atomicAdd(fix_count, total) : 0;
		for (int wv = 0; wv < 256 / 32; wv++)
			s_base[wv] += base;
	}
	__syncthreads();
	for (int i = lane; i < s_n[warp]; i += 32)
		fix_list[s_base[warp] + i] = s_buf[warp][i];
}

// The listed pixels again, a thread per pixel: scores in FP64
// (ace_accurate_sums, error e2 ~ 1e-14), rescaled with a true division.  That
// decides the truncation unless the value lies within the (now tiny) bound of
// an integer; those last few -- the pixels AT the page's extremes among them
// -- are redone by a whole warp in the reference's own arithmetic.  List
// entries index the score block; `out` points at its pixel 0.
template <bool MULTI>
__global__ void __launch_bounds__(256)
ace_fixup_kernel(const uint32_t *__restrict__ in, const int *__restrict__ samples_xy,
		const AceConst cs, const double *__restrict__ minmax,
		const uint32_t *__restrict__ fix_list, int *fix_count, int *fix_ticket,
		uint32_t *__restrict__ out, const ScaleOuts more)
{
	__shared__ double terms[256 / 32][kExactChunk][4];
	__shared__ ExactTable tb;
	pdl_wait();
	const int count = __ldcg(fix_count);
	if (count == 0)
		return;
	stage_exact_table(tb, in, samples_xy, cs);
	double mn[3], span[3], guard[3];
#pragma unroll
	for (int c = 0; c < 3; c++) {
		exact_extremes(minmax, c, mn[c], span[c]);
		// e2 on the score, and the reference's own three roundings of the rescale
		guard[c] = 255.0 * (ace_accurate_eps(cs.nb_samples) / span[c] + 0x1p-48);
	}
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	for (int i0 = blockIdx.x * blockDim.x; i0 < count; i0 += gridDim.x * blockDim.x) {
		const int i = i0 + threadIdx.x;
		const bool active = i < count;
		const uint32_t idx = active ? fix_list[i] : 0u;
		const int y = cs.score_row0 + (int)(idx / (uint32_t)cs.w), x = (int)(idx % (uint32_t)cs.w);
		bool open = false;
		if (active) {
			double acc[4] = { 0.0, 0.0, 0.0, 0.0 };
			ace_accurate_sums(tb, in, samples_xy, cs, __ldg(in + (size_t)y * cs.w + x), x, y, 0, 1, acc);
			uint32_t rgba = 0xFF000000u;
#pragma unroll
			for (int c = 0; c < 3; c++) {
				uint32_t q = 0u;
				const double sa = acc[c] / acc[3];
				if (span[c] > 0.0 && sa == sa) {
					const double t = 255.0 * (sa - mn[c]) / span[c];
					const double fl = floor(t);
					if (t - fl < guard[c] || fl + 1.0 - t < guard[c])
						open = true;
					q = (uint32_t)fmin(fmax(fl, 0.0), 255.0);
				}
				rgba |= q << (8 * c);
			}
			if (!open)
				store_pixel<MULTI>(out, more, idx, rgba);
		}
		// the undecided ones, one at a time by the whole warp
		unsigned todo = __ballot_sync(0xFFFFFFFFu, open);
		while (todo) {
			const int src = __ffs(todo) - 1;
			todo &= todo - 1;
			const int ex = __shfl_sync(0xFFFFFFFFu, x, src), ey = __shfl_sync(0xFFFFFFFFu, y, src);
			const uint32_t eidx = __shfl_sync(0xFFFFFFFFu, idx, src);
			double r[3];
			ace_exact_score_warp(terms[warp], tb, in, samples_xy, cs, ex, ey, r);
			if (lane == 0) {
				uint32_t rgba = 0xFF000000u;
#pragma unroll
				for (int c = 0; c < 3; c++)
					rgba |= ace_scale_exact(r[c], mn[c], span[c]) << (8 * c);
				store_pixel<MULTI>(out, more, eidx, rgba);
			}
		}
	}
	// the last block to finish empties the list for the next rescale of this
	// slot (every block has read the length before it gets here)
	__syncthreads();
	if (threadIdx.x == 0) {
		if (atomicAdd(fix_ticket, 1) == (int)gridDim.x - 1) {
			*fix_ticket = 0;
			fix_ticket[kAceFixupSlots] = count;       // statistics: the length of the list just worked off
			*reinterpret_cast<volatile int *>(fix_count) = 0;
		}
	}
}

// The two error constants the bounds above rest on, measured on the device at
// hand: the worst relative error of MUFU.RSQ (FP32) over every mantissa of an
// even and an odd binade (the hardware treats all binades alike), and of
// rsqrt_f64 over EVERY argument it is ever given here -- the integers below
// 2^31.  worst: the maximum as the bit pattern of a non-negative double
// (ordered like an integer).
__global__ void rsqrt_selftest_kernel(int which, unsigned long long *worst)
{
	double mx = 0.0;
	const uint64_t n = which == 0 ? (1ull << 24) : (1ull << 31);
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
		double got, x;
		if (which == 0) {
			const float xf = __uint_as_float(0x3F800000u + (uint32_t)i);      // 1 .. 4
			x = (double)xf;
			got = (double)rsqrt_f32(xf);
		} else {
			if (i == 0)
				continue;
			x = (double)i;
			got = rsqrt_f64(__hiloint2double(0x43300000, (int)i) - 4503599627370496.0);
		}
		const double ref = 1.0 / sqrt(x);
		mx = fmax(mx, fabs(got - ref) / ref);
	}
	atomicMax(worst, (unsigned long long)__double_as_longlong(mx));
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

// FP32 FMA rate probe (the accumulation kernel's pipe): 8 independent chains.
__global__ void fp32_peak_kernel(float *sink, int iters, float a, float b)
{
	float v0 = threadIdx.x, v1 = v0 + 1, v2 = v0 + 2, v3 = v0 + 3;
	float v4 = v0 + 4, v5 = v0 + 5, v6 = v0 + 6, v7 = v0 + 7;
	for (int i = 0; i < iters; i++) {
		v0 = fmaf(v0, a, b); v1 = fmaf(v1, a, b); v2 = fmaf(v2, a, b); v3 = fmaf(v3, a, b);
		v4 = fmaf(v4, a, b); v5 = fmaf(v5, a, b); v6 = fmaf(v6, a, b); v7 = fmaf(v7, a, b);
	}
	const float s = v0 + v1 + v2 + v3 + v4 + v5 + v6 + v7;
	if (s == 12345.678f)
		sink[0] = s;
}

}  // namespace

int launch_ace_gather_samples(const uint32_t *page, int w, const int *samples_xy, int n, int *msg,
		cudaStream_t st)
{
	PFB_CUDA(launch_pdl(ace_gather_samples_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, st,
			page, w, samples_xy, n, msg));
	return PF_B200_OK;
}

int launch_ace_minmax_init(double *minmax, int w, int h, cudaStream_t st)
{
	const size_t n = ace_minmax_doubles(w, h);
	PFB_CUDA(launch_pdl(ace_minmax_init_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, st,
			minmax, n));
	return PF_B200_OK;
}

namespace {

// Worst-case bound on |approximate score - reference score| (score range
// [-1, 1]) for nb_samples samples; what the refinement threshold and the
// rescale's guard band are derived from.
//   weights:  MUFU.RSQ is within 2^-22.94 (measured over two binades,
//       tools/ubench/ace_fp32_proto.cu), its argument d^2 within 2^-23 (three
//       FP32 roundings), so a weight is within e_w = 1.25e-7 + 6e-8 of
//       1/dist and the score moves by at most 2 e_w / (1 - e_w) < 3.8e-7
//       (the same weight multiplies numerator and denominator);
//   clamp:    u = sat(p * inv + b) with inv = slope / (2 limit) * (1 + 2^-22)
//       rounded to FP32 and b = 1/2 - s * inv rounded from double: at most
//       2^-23 + 2^-25 + 2^-26 off, i.e. 3.3e-7 on the score;
//   sums:     first-level sums of at most kFlush terms, second level of at
//       most levels2 partial sums, every operation rounding once (2^-24 of
//       the running sum, all terms >= 0): relative error of a sum at most
//       (kFlush - 1 + levels2) * 2^-24 =: g, of the quotient 2 g, of the
//       score 4 g;
//   epilogue: four FP32 operations on values <= 2 and a reciprocal: < 1e-6.
double ace_score_eps(int nb_samples)
{
	const int chunks = (nb_samples + kChunk - 1) / kChunk;
	const int levels2 = (nb_samples + kFlush - 1) / kFlush + 2 * chunks;
	const double g = (double)(kFlush - 1 + levels2) * 0x1p-24 * 1.01;
	return 4.0 * g + 3.8e-7 + 3.3e-7 + 1.0e-6;
}

// Checks the arguments and lays out the launch the ACE kernels share.
int ace_setup(const AceParams &p, AceConst &cs, dim3 &grid)
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
	// inv a little above slope / (2 limit): a difference that saturates in the
	// reference then reaches 0 or 1 before the rounding of the FMA can matter,
	// and equal values give exactly 1/2 (ace_score_eps, "clamp")
	cs.inv = (float)(p.slope / (2.0 * p.limit) * (1.0 + 0x1p-22));
	cs.eps = (float)(ace_score_eps(p.nb_samples) * 1.001);
	// the reference saturates a difference d when slope * d (rounded) exceeds
	// the limit; reaching it exactly gives the same value
	volatile double prod;
	cs.dsat = 256;
	for (int d = 1; d <= 255; d++) {
		prod = (double)d * p.slope;
		if (prod >= p.limit) {
			cs.dsat = d;
			break;
		}
	}
	grid = dim3((p.w + kAceBlock - 1) / kAceBlock,
			(p.row_end - 1) / kTileRows - p.row_begin / kTileRows + 1);
	return PF_B200_OK;
}

ScaleOuts scale_outs(uint32_t *const *outs, int nb_outs, bool multicast)
{
	ScaleOuts more = {};
	for (int d = 1; d < nb_outs; d++)
		more.p[d - 1] = outs[d];
	more.n = nb_outs - 1;
	more.multicast = multicast ? 1 : 0;
	return more;
}

}  // namespace

size_t ace_minmax_doubles(int w, int h)
{
	// a band of the page never has more tile rows than the page has, plus one
	const size_t tiles = (size_t)((w + kAceBlock - 1) / kAceBlock) * ((h + kTileRows - 1) / kTileRows + 1);
	return kMinmaxHead + 6 * tiles;
}

size_t ace_scores_bytes(int w, int rows)
{
	const size_t n = (size_t)w * rows;
	return (kAceScoreHeader + 16 * n + 255) / 256 * 256;
}

AceScores ace_scores_at(void *block, size_t plane)
{
	AceScores s;
	s.counts = static_cast<int *>(block);
	s.planes = reinterpret_cast<float *>(static_cast<char *>(block) + kAceScoreHeader);
	s.list = reinterpret_cast<uint32_t *>(s.planes + 3 * plane);
	s.plane = plane;
	return s;
}

int launch_ace_scores(const uint32_t *in, const int *samples_xy,
		const AceParams &p, const AceScores &sc, double *minmax, cudaStream_t st)
{
	AceConst cs;
	dim3 grid;
	const int rc = ace_setup(p, cs, grid);
	if (rc)
		return rc;
	const size_t smem = sizeof(AceShared);
	PFB_CUDA(cudaFuncSetAttribute(ace_scores_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
	PFB_CUDA(launch_pdl(ace_scores_kernel, grid, dim3(kAceBlock), smem, st, in, samples_xy, cs,
			sc.planes, sc.counts, minmax));
	return PF_B200_OK;
}

int launch_ace_refine(const uint32_t *in, const int *samples_xy,
		const AceParams &p, const AceScores &sc, double *minmax, cudaStream_t st)
{
	AceConst cs;
	dim3 grid;
	const int rc = ace_setup(p, cs, grid);
	if (rc)
		return rc;
	const dim3 blocks(sm_count() * 2);
	PFB_CUDA(launch_pdl(ace_refine_kernel<1>, blocks, dim3(kRefineThreads), 0, st,
			in, samples_xy, cs, (const float *)sc.planes, minmax, (int)grid.x, (int)grid.y));
	PFB_CUDA(launch_pdl(ace_refine_kernel<2>, blocks, dim3(kRefineThreads), 0, st,
			in, samples_xy, cs, (const float *)sc.planes, minmax, (int)grid.x, (int)grid.y));
	return PF_B200_OK;
}

int launch_ace_scale(const AceParams &p, const AceScores &sc, const double *minmax,
		int row0, int rows, int slot, uint32_t *const *outs, int nb_outs, bool multicast,
		cudaStream_t st)
{
	if (nb_outs < 1 || nb_outs > kMaxScaleOuts)
		return fail(PF_B200_EINVAL, "ace: %d destinations (1..%d)", nb_outs, kMaxScaleOuts);
	if (slot < 0 || slot >= kAceFixupSlots)
		return fail(PF_B200_EINVAL, "ace: fix-up slot %d", slot);
	AceConst cs;
	dim3 grid;
	const int rc = ace_setup(p, cs, grid);
	if (rc)
		return rc;
	const size_t npix = (size_t)p.w * rows;
	const size_t pix_base = (size_t)(row0 - cs.score_row0) * p.w;
	const size_t work = (npix + 3) / 4;       // 4 pixels per thread
	const int blocks = (int)((work + 255) / 256 < (size_t)sm_count() * 8
			? (work + 255) / 256 : (size_t)sm_count() * 8);
	const ScaleOuts more = scale_outs(outs, nb_outs, multicast);
	const float *planes = sc.planes + pix_base;
	// the list of a slot starts at the slot's first pixel: slots cover disjoint rows
	uint32_t *list = sc.list + pix_base;
	if (nb_outs == 1 && !multicast)
		PFB_CUDA(launch_pdl(ace_scale_kernel<false>, dim3(blocks), dim3(256), 0, st, planes, minmax,
				npix, sc.plane, pix_base, cs.eps, outs[0], more, list, sc.counts + slot));
	else
		PFB_CUDA(launch_pdl(ace_scale_kernel<true>, dim3(blocks), dim3(256), 0, st, planes, minmax,
				npix, sc.plane, pix_base, cs.eps, outs[0], more, list, sc.counts + slot));
	return PF_B200_OK;
}

int launch_ace_fixup(const uint32_t *in, const int *samples_xy, const AceParams &p,
		const AceScores &sc, const double *minmax, int row0, int slot,
		uint32_t *const *outs, int nb_outs, bool multicast, cudaStream_t st)
{
	if (nb_outs < 1 || nb_outs > kMaxScaleOuts)
		return fail(PF_B200_EINVAL, "ace: %d destinations (1..%d)", nb_outs, kMaxScaleOuts);
	AceConst cs;
	dim3 grid;
	const int rc = ace_setup(p, cs, grid);
	if (rc)
		return rc;
	const size_t pix_base = (size_t)(row0 - cs.score_row0) * p.w;
	// `outs` point at row0; list entries index the score block
	uint32_t *shifted[kMaxScaleOuts];
	for (int d = 0; d < nb_outs; d++)
		shifted[d] = outs[d] - pix_base;
	const ScaleOuts more = scale_outs(shifted, nb_outs, multicast);
	const uint32_t *list = sc.list + pix_base;
	const int blocks = sm_count() * 8;      // of 128 threads: a list of ~1 % of a page spreads evenly over the SMs
	if (nb_outs == 1 && !multicast)
		PFB_CUDA(launch_pdl(ace_fixup_kernel<false>, dim3(blocks), dim3(128), 0, st, in, samples_xy, cs,
				minmax, list, sc.counts + slot, sc.counts + kAceFixupSlots + slot, shifted[0], more));
	else
		PFB_CUDA(launch_pdl(ace_fixup_kernel<true>, dim3(blocks), dim3(128), 0, st, in, samples_xy, cs,
				minmax, list, sc.counts + slot, sc.counts + kAceFixupSlots + slot, shifted[0], more));
	return PF_B200_OK;
}

template <typename T, typename K>
static double measure_fma_peak(K kern, T a, T b, const char *what)
{
	T *sink = nullptr;
	if (cudaMalloc(&sink, 64) != cudaSuccess)
		return -(double)fail(PF_B200_ECUDA, "%s probe: cudaMalloc failed", what);
	const int blocks = sm_count() * 8, threads = 256, iters = 1 << 15;
	cudaEvent_t e0, e1;
	cudaEventCreate(&e0);
	cudaEventCreate(&e1);
	kern<<<blocks, threads>>>(sink, 1024, a, b);   // warm-up
	double best_ms = 1e30;
	for (int rep = 0; rep < 5; rep++) {
		cudaEventRecord(e0);
		kern<<<blocks, threads>>>(sink, iters, a, b);
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
		return -(double)fail(PF_B200_ECUDA, "%s probe: %s", what, cudaGetErrorString(err));
	const double flops = 2.0 * 8.0 * (double)iters * blocks * threads;
	return flops / (best_ms * 1e-3) / 1e12;
}

double measure_rsqrt_error(int which)
{
	unsigned long long *d = nullptr, h = 0;
	if (cudaMalloc(&d, sizeof(h)) != cudaSuccess)
		return -(double)fail(PF_B200_ECUDA, "rsqrt self-test: cudaMalloc failed");
	cudaMemset(d, 0, sizeof(h));
	rsqrt_selftest_kernel<<<sm_count() * 8, 256>>>(which, d);
	const cudaError_t err = cudaMemcpy(&h, d, sizeof(h), cudaMemcpyDeviceToHost);
	cudaFree(d);
	if (err != cudaSuccess)
		return -(double)fail(PF_B200_ECUDA, "rsqrt self-test: %s", cudaGetErrorString(err));
	double v;
	memcpy(&v, &h, sizeof(v));
	return v;
}

double measure_fp64_peak_tflops()
{
	return measure_fma_peak<double>(fp64_peak_kernel, 1.0000001, 1e-9, "fp64");
}

double measure_fp32_peak_tflops()
{
	return measure_fma_peak<float>(fp32_peak_kernel, 1.0000001f, 1e-9f, "fp32");
}

}  // namespace pfb
