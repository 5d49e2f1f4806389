// Gaussian / Sobel / Canny kernels for sm_100a.
//
// All three are "column-march" stencils: a thread owns a few adjacent columns
// of a horizontal strip of the page and walks down the rows, keeping the
// vertical window of the stencil in registers.  Every RGBA8 pixel is read
// from HBM once per kernel (row halos come out of L1/L2) and written once:
// 8 algorithmic bytes per pixel per stage.
//
// Reference semantics (upstream src/pillowfight/util.c pf_dbl_matrix_convolution,
// _gaussian.c, _sobel.c, _canny.c -- not in the mount):
// true convolution with the kernel centre at size/2, taps outside the page
// skipped.  Skipping a tap and adding +0.0 give the same double, so the
// kernels zero-pad.
#include "pf_common.cuh"

namespace pfb {

namespace {

constexpr int kBlock = 128;
constexpr uint32_t kWhite = 0xFFFFFFFFu;   // kept edge
constexpr uint32_t kBlack = 0xFF000000u;   // no edge
constexpr uint32_t kWeak = 0xFF800080u;    // weak candidate, resolved later

__device__ __forceinline__ uint32_t load_px(const uint32_t *row, int x, int w)
{
	return (row != nullptr && x >= 0 && x < w) ? __ldg(row + x) : 0u;
}

// (r + g + b) / 3, the reference's PF_GET_PIXEL_GRAYSCALE
__device__ __forceinline__ int gray_of(uint32_t p)
{
	const int s = (p & 0xFF) + ((p >> 8) & 0xFF) + ((p >> 16) & 0xFF);
	return (s * 43691) >> 17;
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
		int w, int h, int rows_per_strip, const TapsN<N> tp)
{
	constexpr int C = N / 2;             // taps reach x+C ... x-(N-1-C)
	constexpr int L = N - 1 - C;         // left / up reach
	constexpr int NPIX = PX + N - 1;     // input pixels a thread needs per row

	const int x0 = (blockIdx.x * kBlock + threadIdx.x) * PX;
	if (x0 >= w)
		return;
	const size_t page = (size_t)blockIdx.z * w * h;
	in += page;
	out += page;
	const int ys = blockIdx.y * rows_per_strip;
	const int ye = min(h, ys + rows_per_strip);

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

// Any tap count: one thread per pixel, N*N multiply-adds, same rounding order.
__global__ void __launch_bounds__(kBlock)
gaussian_generic_kernel(const uint32_t *__restrict__ in, uint32_t *__restrict__ out,
		int w, int h, const Taps tp)
{
	const int n = tp.n;
	const int x = blockIdx.x * kBlock + threadIdx.x;
	const int y = blockIdx.y;
	if (x >= w)
		return;
	const size_t page = (size_t)blockIdx.z * w * h;
	in += page;
	out += page;
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
			const double t = tp.t[kx];
			for (int c = 0; c < 3; c++) {
				const double m = __dmul_rn(u8_to_double((p >> (8 * c)) & 0xFF), t);
				hv[c] = first ? m : __dadd_rn(hv[c], m);
			}
			first = false;
		}
		const double tv = tp.t[ky];
		for (int c = 0; c < 3; c++) {
			const double m = __dmul_rn(hv[c], tv);
			acc[c] = first_row ? m : __dadd_rn(acc[c], m);
		}
		first_row = false;
	}
	uint32_t rgba = 0xFF000000u;
	for (int c = 0; c < 3; c++)
		rgba |= (uint32_t)min(trunc_nonneg(acc[c]), 255) << (8 * c);
	out[(size_t)y * w + x] = rgba;
}

// ---------------------------------------------------------------------------
// Sobel and Canny share the gradient march: 4 output columns per thread.
// ---------------------------------------------------------------------------

// Grays of columns x0-HALO .. x0+3+HALO of one row (zero outside the page).
template <int HALO>
__device__ __forceinline__ void load_gray_row(const uint32_t *__restrict__ in,
		int w, int h, int r, int x0, bool vec, int (&g)[4 + 2 * HALO])
{
	if (r < 0 || r >= h) {
#pragma unroll
		for (int i = 0; i < 4 + 2 * HALO; i++)
			g[i] = 0;
		return;
	}
	const uint32_t *row = in + (size_t)r * w;
	if (vec && x0 + 3 < w) {
		const uint4 v = __ldg(reinterpret_cast<const uint4 *>(row + x0));
		g[HALO + 0] = gray_of(v.x);
		g[HALO + 1] = gray_of(v.y);
		g[HALO + 2] = gray_of(v.z);
		g[HALO + 3] = gray_of(v.w);
	} else {
#pragma unroll
		for (int i = 0; i < 4; i++)
			g[HALO + i] = gray_of(load_px(row, x0 + i, w));
	}
#pragma unroll
	for (int i = 0; i < HALO; i++) {
		g[i] = gray_of(load_px(row, x0 - HALO + i, w));
		g[HALO + 4 + i] = gray_of(load_px(row, x0 + 4 + i, w));
	}
}

__device__ __forceinline__ int isqrt_cap255(int m)
{
	if (m >= 255 * 255)
		return 255;
	int s = (int)__fsqrt_rn((float)m);
	if (s * s > m)
		s--;
	if ((s + 1) * (s + 1) <= m)
		s++;
	return s;
}

// out = min(255, floor(hypot(g_x, g_y))) on r, g, b; alpha 255.  g_x, g_y are
// small integers, so the FP64 reference is reproduced exactly in int32.
__global__ void __launch_bounds__(kBlock)
sobel_march_kernel(const uint32_t *__restrict__ in, uint32_t *__restrict__ out,
		int w, int h, int rows_per_strip)
{
	const int x0 = (blockIdx.x * kBlock + threadIdx.x) * 4;
	if (x0 >= w)
		return;
	const size_t page = (size_t)blockIdx.z * w * h;
	in += page;
	out += page;
	const bool vec = (w & 3) == 0;
	const int ys = blockIdx.y * rows_per_strip;
	const int ye = min(h, ys + rows_per_strip);

	// hd = g(x-1) - g(x+1), hs = g(x-1) + 2 g(x) + g(x+1) for rows y-1, y, y+1
	int hd0[4], hs0[4], hd1[4], hs1[4];
	int g[6], gn[6];
	load_gray_row<1>(in, w, h, ys - 1, x0, vec, g);
#pragma unroll
	for (int i = 0; i < 4; i++) {
		hd0[i] = g[i] - g[i + 2];
		hs0[i] = g[i] + 2 * g[i + 1] + g[i + 2];
	}
	load_gray_row<1>(in, w, h, ys, x0, vec, g);
#pragma unroll
	for (int i = 0; i < 4; i++) {
		hd1[i] = g[i] - g[i + 2];
		hs1[i] = g[i] + 2 * g[i + 1] + g[i + 2];
	}
	load_gray_row<1>(in, w, h, ys + 1, x0, vec, gn);
	for (int y = ys; y < ye; y++) {
#pragma unroll
		for (int i = 0; i < 6; i++)
			g[i] = gn[i];
		if (y + 1 < ye)
			load_gray_row<1>(in, w, h, y + 2, x0, vec, gn);   // prefetch
		uint32_t res[4];
#pragma unroll
		for (int i = 0; i < 4; i++) {
			const int hd2 = g[i] - g[i + 2];
			const int hs2 = g[i] + 2 * g[i + 1] + g[i + 2];
			const int gx = hd2 + 2 * hd1[i] + hd0[i];
			const int gy = hs0[i] - hs2;
			const uint32_t v = (uint32_t)isqrt_cap255(gx * gx + gy * gy);
			res[i] = 0xFF000000u | v | (v << 8) | (v << 16);
			hd0[i] = hd1[i];
			hs0[i] = hs1[i];
			hd1[i] = hd2;
			hs1[i] = hs2;
		}
		uint32_t *orow = out + (size_t)y * w;
		if (vec && x0 + 3 < w) {
			*reinterpret_cast<uint4 *>(orow + x0) = make_uint4(res[0], res[1], res[2], res[3]);
		} else {
#pragma unroll
			for (int i = 0; i < 4; i++)
				if (x0 + i < w)
					orow[x0 + i] = res[i];
		}
	}
}

// Canny pass A: gradients, non-maximum suppression, double threshold.
// Writes the final pixel for "none" and "strong"; weak candidates get kWeak
// and their index is appended to the work list (hdr[0] = count).
// Thread: 4 output columns; needs g_x^2+g_y^2 on columns x0-1..x0+4 (6), so
// grays on x0-2..x0+5 (8).
struct MagRow {
	int m[6];      // squared gradient magnitude, 0 outside the page
	int gx[4], gy[4];
};

__device__ __forceinline__ void canny_mag_row(const int (&hd0)[6], const int (&hs0)[6],
		const int (&hd1)[6], const int (&hd2)[6], const int (&hs2)[6],
		int x0, int w, bool row_inside, MagRow &o)
{
#pragma unroll
	for (int i = 0; i < 6; i++) {
		const int gx = hd2[i] + 2 * hd1[i] + hd0[i];
		const int gy = hs0[i] - hs2[i];
		const int xc = x0 - 1 + i;
		o.m[i] = (row_inside && xc >= 0 && xc < w) ? gx * gx + gy * gy : 0;
		if (i >= 1 && i <= 4) {
			o.gx[i - 1] = gx;
			o.gy[i - 1] = gy;
		}
	}
}

__global__ void __launch_bounds__(kBlock)
canny_classify_kernel(const uint32_t *__restrict__ in, uint32_t *__restrict__ out,
		int w, int h, int rows_per_strip, CannyThresholds thr,
		int *__restrict__ hdr, int *__restrict__ list)
{
	const int x0 = (blockIdx.x * kBlock + threadIdx.x) * 4;
	const bool active = x0 < w;
	const size_t npix = (size_t)w * h;
	const size_t page = (size_t)blockIdx.z * npix;
	in += page;
	out += page;
	const bool vec = (w & 3) == 0;
	const int ys = blockIdx.y * rows_per_strip;
	const int ye = min(h, ys + rows_per_strip);
	const int lane = threadIdx.x & 31;

	int hd[3][6], hs[3][6];     // horizontal diffs / sums of the last 3 gray rows
	MagRow mr[3];               // magnitudes of the last 3 gradient rows
	int g[8];

	// Gray row r feeds gradient row r-1, which feeds output row r-2.
	// Output rows ys..ye-1 need gray rows ys-2 .. ye+1.
	for (int r = ys - 2; r <= ye + 1; r++) {
		if (active) {
			load_gray_row<2>(in, w, h, r, x0, vec, g);
#pragma unroll
			for (int i = 0; i < 6; i++) {
				hd[0][i] = hd[1][i]; hs[0][i] = hs[1][i];
				hd[1][i] = hd[2][i]; hs[1][i] = hs[2][i];
				hd[2][i] = g[i] - g[i + 2];
				hs[2][i] = g[i] + 2 * g[i + 1] + g[i + 2];
			}
		}
		if (r < ys)
			continue;       // need 3 gray rows (r-2, r-1, r) first
		if (active) {
			mr[0] = mr[1];
			mr[1] = mr[2];
			const int gr = r - 1;   // gradient row
			canny_mag_row(hd[0], hs[0], hd[1], hd[2], hs[2], x0, w,
					gr >= 0 && gr < h, mr[2]);
		}
		const int y = r - 2;
		if (y < ys)
			continue;       // need 3 gradient rows (y-1, y, y+1)
		uint32_t res[4];
		int nweak = 0;
		if (active) {
#pragma unroll
			for (int i = 0; i < 4; i++) {
				const int gx = mr[1].gx[i], gy = mr[1].gy[i];
				const int m = mr[1].m[i + 1];
				const int ax = abs(gx), ay = abs(gy);
				const int s2 = (ax + ay) * (ax + ay);
				int a, b;
				if (s2 < 2 * ax * ax || (ax == 0 && ay == 0)) {   // |angle| < 22.5 deg
					a = mr[1].m[i + 2]; b = mr[1].m[i];
				} else if (s2 < 2 * ay * ay) {                     // near vertical
					a = mr[2].m[i + 1]; b = mr[0].m[i + 1];
				} else if ((gx ^ gy) >= 0) {                        // 45 deg
					a = mr[2].m[i + 2]; b = mr[0].m[i];
				} else {                                            // 135 deg
					a = mr[2].m[i]; b = mr[0].m[i + 2];
				}
				const bool keep = !(m < a || m < b);
				uint32_t v = kBlack;
				if (keep && m >= thr.strong_min) {
					v = kWhite;
				} else if (keep && m >= thr.weak_min) {
					v = kWeak;
					nweak++;
				}
				res[i] = v;
			}
		}
		// warp-aggregated append of the weak pixels of this row
		const unsigned any = __ballot_sync(0xFFFFFFFFu, nweak > 0);
		if (any) {
			int pre = nweak;
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) {
				const int t = __shfl_up_sync(0xFFFFFFFFu, pre, d);
				if (lane >= d)
					pre += t;
			}
			const int total = __shfl_sync(0xFFFFFFFFu, pre, 31);
			int base = 0;
			if (lane == 31)
				base = atomicAdd(&hdr[0], total);
			base = __shfl_sync(0xFFFFFFFFu, base, 31) + pre - nweak;
			if (nweak > 0) {
				const int pix0 = (int)(blockIdx.z * npix) + y * w + x0;
#pragma unroll
				for (int i = 0; i < 4; i++)
					if (res[i] == kWeak)
						list[base++] = pix0 + i;
			}
		}
		if (active) {
			uint32_t *orow = out + (size_t)y * w;
			if (vec && x0 + 3 < w) {
				*reinterpret_cast<uint4 *>(orow + x0) =
					make_uint4(res[0], res[1], res[2], res[3]);
			} else {
#pragma unroll
				for (int i = 0; i < 4; i++)
					if (x0 + i < w)
						orow[x0 + i] = res[i];
			}
		}
	}
}

// ---------------------------------------------------------------------------
// Canny pass B: hysteresis over the weak work list, warp-cooperative.
// ---------------------------------------------------------------------------

constexpr int kRing = 256;      // per-warp queue of freshly kept pixels

struct PageGeom {
	int w, h, npix;
};

__device__ __forceinline__ uint32_t ld_state(const uint32_t *p)
{
	return __ldcg(p);   // L2: other warps promote pixels with atomics
}

// One warp: for each of its list entries, a weak pixel next to a kept one is
// kept (atomicCAS so that exactly one warp follows it), then the warp follows
// the chain: 4 queue entries x 8 neighbours per step, one lane per neighbour.
// A full queue drops the push and raises *dirty: the pixel is kept but its
// neighbours were not visited, so the sweep has to be repeated.
__device__ void hysteresis_warp_sweep(uint32_t *out, PageGeom pg,
		const int *list, int count, int warp, int nwarps,
		int *ring, int *dirty)
{
	const int lane = threadIdx.x & 31;
	const int nb = lane & 7;                 // neighbour handled by this lane
	const int ndx = (nb < 3) ? nb - 1 : (nb == 3 ? -1 : (nb == 4 ? 1 : nb - 6));
	const int ndy = (nb < 3) ? -1 : (nb < 5 ? 0 : 1);

	for (int base = warp * 32; base < count; base += nwarps * 32) {
		int head = 0, tail = 0;
		const int e = base + lane;
		bool got = false;
		int idx = 0;
		if (e < count) {
			idx = list[e];
			if (ld_state(out + idx) == kWeak) {
				const int p = idx % pg.npix;
				const int y = p / pg.w, x = p - y * pg.w;
				bool near_kept = false;
#pragma unroll
				for (int dy = -1; dy <= 1; dy++) {
#pragma unroll
					for (int dx = -1; dx <= 1; dx++) {
						if (dx == 0 && dy == 0)
							continue;
						const int qx = x + dx, qy = y + dy;
						if (qx < 0 || qy < 0 || qx >= pg.w || qy >= pg.h)
							continue;
						near_kept |= ld_state(out + idx + dy * pg.w + dx) == kWhite;
					}
				}
				if (near_kept)
					got = atomicCAS(out + idx, kWeak, kWhite) == kWeak;
			}
		}
		unsigned m = __ballot_sync(0xFFFFFFFFu, got);
		if (got)
			ring[(tail + __popc(m & ((1u << lane) - 1))) % kRing] = idx;
		tail += __popc(m);
		__syncwarp();

		while (head < tail) {
			const int take = min(4, tail - head);
			const int slot = lane >> 3;
			got = false;
			int q = 0;
			if (slot < take) {
				const int cur = ring[(head + slot) % kRing];
				const int p = cur % pg.npix;
				const int y = p / pg.w, x = p - y * pg.w;
				const int qx = x + ndx, qy = y + ndy;
				if (qx >= 0 && qy >= 0 && qx < pg.w && qy < pg.h) {
					q = cur + ndy * pg.w + ndx;
					if (ld_state(out + q) == kWeak)
						got = atomicCAS(out + q, kWeak, kWhite) == kWeak;
				}
			}
			__syncwarp();
			head += take;
			m = __ballot_sync(0xFFFFFFFFu, got);
			const int room = kRing - (tail - head);
			const int rank = __popc(m & ((1u << lane) - 1));
			if (got) {
				if (rank < room)
					ring[(tail + rank) % kRing] = q;
				else
					*dirty = 1;
			}
			tail += min(__popc(m), room);
			__syncwarp();
		}
	}
}

// hdr[0] = list length, hdr[1] = dirty flag of the previous sweep,
// hdr[2] = dirty flag this sweep raises.
__global__ void __launch_bounds__(256)
canny_hysteresis_kernel(uint32_t *out, PageGeom pg, const int *list, int *hdr)
{
	__shared__ int rings[8][kRing];
	const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int nwarps = (gridDim.x * blockDim.x) >> 5;
	hysteresis_warp_sweep(out, pg, list, hdr[0], warp, nwarps,
			rings[threadIdx.x >> 5], &hdr[2]);
}

// Runs only when the sweep above overflowed a queue: one block repeats the
// sweep until a whole pass drops nothing.  Terminates because every pass keeps
// at least one more pixel or raises no flag.
__global__ void __launch_bounds__(1024)
canny_hysteresis_drain_kernel(uint32_t *out, PageGeom pg, const int *list, int *hdr)
{
	__shared__ int rings[32][kRing];
	__shared__ int dirty;
	if (hdr[2] == 0)
		return;
	for (;;) {
		if (threadIdx.x == 0)
			dirty = 0;
		__syncthreads();
		hysteresis_warp_sweep(out, pg, list, hdr[0], threadIdx.x >> 5, 32,
				rings[threadIdx.x >> 5], &dirty);
		__threadfence();
		__syncthreads();
		const int again = dirty;
		__syncthreads();
		if (!again)
			break;
	}
}

// Canny pass C: weak pixels that no kept pixel reached become background.
__global__ void __launch_bounds__(256)
canny_finalize_kernel(uint32_t *out, const int *list, const int *hdr)
{
	const int count = hdr[0];
	for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count;
			i += gridDim.x * blockDim.x) {
		const int idx = list[i];
		if (ld_state(out + idx) == kWeak)
			out[idx] = kBlack;
	}
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
static void launch_gaussian_n(const uint32_t *in, uint32_t *out, int w, int h,
		int pages, const Taps &taps, cudaStream_t st)
{
	TapsN<N> tp;
	for (int k = 0; k < N; k++)
		tp.t[k] = taps.t[k];
	constexpr int PX = 2;
	const int bx = ((w + PX - 1) / PX + kBlock - 1) / kBlock;
	const int rows = strips_for(bx, h, pages, 4, 8 * N);
	dim3 grid(bx, (h + rows - 1) / rows, pages);
	gaussian_march_kernel<N, PX><<<grid, kBlock, 0, st>>>(in, out, w, h, rows, tp);
}

int launch_gaussian(const uint32_t *in, uint32_t *out, int w, int h, int pages,
		const Taps &taps, cudaStream_t st)
{
	if (taps.n < 1 || taps.n > kMaxTaps)
		return fail(PF_B200_EINVAL, "gaussian: nb_stddev must be in 1..%d", kMaxTaps);
	switch (taps.n) {
	case 3: launch_gaussian_n<3>(in, out, w, h, pages, taps, st); break;
	case 5: launch_gaussian_n<5>(in, out, w, h, pages, taps, st); break;
	case 7: launch_gaussian_n<7>(in, out, w, h, pages, taps, st); break;
	case 9: launch_gaussian_n<9>(in, out, w, h, pages, taps, st); break;
	default: {
		dim3 grid((w + kBlock - 1) / kBlock, h, pages);
		gaussian_generic_kernel<<<grid, kBlock, 0, st>>>(in, out, w, h, taps);
	}
	}
	PFB_CUDA(cudaGetLastError());
	return PF_B200_OK;
}

int launch_sobel(const uint32_t *in, uint32_t *out, int w, int h, int pages,
		cudaStream_t st)
{
	const int bx = ((w + 3) / 4 + kBlock - 1) / kBlock;
	const int rows = strips_for(bx, h, pages, 8, 16);
	dim3 grid(bx, (h + rows - 1) / rows, pages);
	sobel_march_kernel<<<grid, kBlock, 0, st>>>(in, out, w, h, rows);
	PFB_CUDA(cudaGetLastError());
	return PF_B200_OK;
}

size_t canny_scratch_bytes(int w, int h)
{
	return 256 + sizeof(int) * (size_t)w * h;
}

int launch_canny(const uint32_t *in, uint32_t *out, int w, int h, int pages,
		CannyThresholds thr, void *scratch, size_t scratch_bytes,
		cudaStream_t st)
{
	const size_t npix = (size_t)w * h;
	if (npix > 0x7FFFFFFFull)
		return fail(PF_B200_ETOOBIG, "canny: page of %d x %d exceeds 2^31 pixels", w, h);
	if (scratch == nullptr || scratch_bytes < canny_scratch_bytes(w, h))
		return fail(PF_B200_EINVAL, "canny: scratch of %zu bytes, need %zu",
				scratch_bytes, canny_scratch_bytes(w, h));
	if (((uintptr_t)scratch & 255) != 0)
		return fail(PF_B200_EINVAL, "canny: scratch must be 256-byte aligned");
	// pages per launch: what the work list can index and hold
	size_t cap_pages = (scratch_bytes - 256) / (sizeof(int) * npix);
	const size_t idx_pages = 0x7FFFFFFFull / npix;
	if (cap_pages > idx_pages)
		cap_pages = idx_pages;
	if (cap_pages > 65535)
		cap_pages = 65535;
	int *hdr = static_cast<int *>(scratch);
	int *list = hdr + 64;
	const PageGeom pg = { w, h, (int)npix };
	const int bx = ((w + 3) / 4 + kBlock - 1) / kBlock;

	for (int p0 = 0; p0 < pages; p0 += (int)cap_pages) {
		const int np = (pages - p0 < (int)cap_pages) ? pages - p0 : (int)cap_pages;
		const uint32_t *pin = in + (size_t)p0 * npix;
		uint32_t *pout = out + (size_t)p0 * npix;
		PFB_CUDA(cudaMemsetAsync(hdr, 0, 256, st));
		const int rows = strips_for(bx, h, np, 6, 32);
		dim3 grid(bx, (h + rows - 1) / rows, np);
		canny_classify_kernel<<<grid, kBlock, 0, st>>>(pin, pout, w, h, rows, thr, hdr, list);
		const int sweep_blocks = sm_count() * 4;
		canny_hysteresis_kernel<<<sweep_blocks, 256, 0, st>>>(pout, pg, list, hdr);
		canny_hysteresis_drain_kernel<<<1, 1024, 0, st>>>(pout, pg, list, hdr);
		canny_finalize_kernel<<<sweep_blocks, 256, 0, st>>>(pout, list, hdr);
		PFB_CUDA(cudaGetLastError());
	}
	return PF_B200_OK;
}

}  // namespace pfb
