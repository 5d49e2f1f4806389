// Sobel and Canny kernels for sm_100a.
//
// Reference semantics (upstream src/pillowfight/_sobel.c, _canny.c, util.c --
// not in the mount): grayscale = (r+g+b)/3, g_x / g_y by
// true convolution with the 3x3 Sobel kernels (taps outside the page skipped
// = zero padding), intensity = hypot(g_x, g_y).  g_x and g_y are small
// integers, so everything the reference does in double is reproduced exactly
// in int32: floor(hypot) = isqrt(g_x^2+g_y^2), intensity comparisons compare
// g_x^2+g_y^2, the direction sector is an exact integer test.
//
// Layout: the warp march of pf_march.cuh (4 columns per lane, halo lanes,
// register-resident 3-row windows).  Per stage a pixel is read once and
// written once.
#include <cooperative_groups.h>

#include <mutex>

#include "pf_march.cuh"

namespace pfb {

namespace {

constexpr uint32_t kWhite = 0xFFFFFFFFu;   // kept edge
constexpr uint32_t kBlack = 0xFF000000u;   // no edge
// A weak candidate is stored as its index in the work list; every index is
// below kBlack, so "value < kBlack" tags a weak pixel until pass C resolves it.

// (r + g + b) / 3: one SAD against zero sums the three bytes, IMAD.HI divides.
__device__ __forceinline__ int gray_of(uint32_t p)
{
	return (int)__umulhi(__vsadu4(p & 0x00FFFFFFu, 0u), 0x55555556u);
}

// Horizontal parts of the two Sobel kernels for the lane's 4 columns:
//   hd = g(x-1) - g(x+1)            (g_x row term, kernel flipped)
//   hs = g(x-1) + 2 g(x) + g(x+1)   (g_y row term)
struct HRow {
	int hd[4], hs[4];
};

__device__ __forceinline__ void hrow_from_pixels(const uint4 px, HRow &o)
{
	int g[6];
	g[1] = gray_of(px.x);
	g[2] = gray_of(px.y);
	g[3] = gray_of(px.z);
	g[4] = gray_of(px.w);
	g[0] = __shfl_up_sync(kFull, g[4], 1);      // left lane's last column
	g[5] = __shfl_down_sync(kFull, g[1], 1);    // right lane's first column
#pragma unroll
	for (int i = 0; i < 4; i++) {
		o.hd[i] = g[i] - g[i + 2];
		o.hs[i] = g[i] + 2 * g[i + 1] + g[i + 2];
	}
}

// 0xFF000000 + v * 0x010101 with v = min(255, floor(sqrt(m))): the gray pixel.
// sqrt(m + 0.5) sits at least 1/(4*255) away from every integer for m < 2^16,
// far more than MUFU.SQRT's error, so truncating it is exact.  The truncation
// is a round-toward-zero add of 2^23 (the integer lands in the low mantissa
// bits), and one IMAD replicates it into R, G, B: the 0x4B exponent byte
// times 0x010101 wraps to itself, the addend turns it into alpha 0xFF.
__device__ __forceinline__ uint32_t gray_pixel_of_mag2(int m)
{
	float r;
	asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"((float)min(m, 65535) + 0.5f));
	return __float_as_uint(__fadd_rz(r, 8388608.0f)) * 0x010101u + 0xB4000000u;
}

// ---------------------------------------------------------------------------
// Sobel
// ---------------------------------------------------------------------------

// Step J of 3 (s = steps since the top of the strip): the gray row at the
// head of the 3-deep row queue is consumed, its slot is refilled with the row
// 3 below, and the output row above leaves.  Horizontal terms of the last
// three rows sit in hr slots (J+1)%3, (J+2)%3, J.  The first two steps of a
// strip only fill the window.
template <bool FAST, int J>
__device__ __forceinline__ void sobel_step(const StripGeom &g, const WarpPlace &wp, int s,
		RowReader<FAST> &rd, RowWriter<FAST> &wr, uint4 (&q)[3], HRow (&hr)[3])
{
	const uint4 cur = q[J];
	q[J] = rd.next(g);
	hrow_from_pixels(cur, hr[J]);
	if (s < 2)
		return;
	const HRow &a = hr[(J + 1) % 3], &b = hr[(J + 2) % 3], &c = hr[J];
	uint32_t res[4];
#pragma unroll
	for (int i = 0; i < 4; i++) {
		const int gx = c.hd[i] + 2 * b.hd[i] + a.hd[i];
		const int gy = a.hs[i] - c.hs[i];
		res[i] = gray_pixel_of_mag2(gx * gx + gy * gy);
	}
	wr.put(g, wp.ye, res);
}

template <bool FAST>
__device__ __forceinline__ void sobel_march(const uint32_t *__restrict__ in,
		uint32_t *__restrict__ out, const StripGeom &g, const WarpPlace &wp)
{
	RowReader<FAST> rd;
	RowWriter<FAST> wr;
	rd.init(in, g, wp.ys - 1, wp.x0);
	wr.init(out, g, wp.ys, wp.x0, wp.out_lane);
	HRow hr[3];
	uint4 q[3];
	q[0] = rd.next(g);
	q[1] = rd.next(g);
	q[2] = rd.next(g);
	// gray rows ys-1 .. ys+rows_per_strip: rows_per_strip + 2 steps, a multiple of 3
	const int steps = g.rows_per_strip + 2;
	for (int s = 0; s < steps; s += 3) {
		sobel_step<FAST, 0>(g, wp, s, rd, wr, q, hr);
		sobel_step<FAST, 1>(g, wp, s + 1, rd, wr, q, hr);
		sobel_step<FAST, 2>(g, wp, s + 2, rd, wr, q, hr);
	}
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32)
sobel_kernel(const uint32_t *__restrict__ in, uint32_t *__restrict__ out,
		const StripGeom g)
{
	pdl_wait();
	const WarpPlace wp = place_warp(g);
	if (!wp.valid)
		return;
	const size_t page = (size_t)blockIdx.y * g.w * g.h;
	if (g.aligned)
		sobel_march<true>(in + page, out + page, g, wp);
	else
		sobel_march<false>(in + page, out + page, g, wp);
}

// ---------------------------------------------------------------------------
// Canny pass A: gradients, non-maximum suppression, double threshold
// ---------------------------------------------------------------------------

// One gradient row for the lane's 4 columns plus the two neighbouring columns.
struct GRow {
	int m[6];          // g_x^2 + g_y^2 at columns x0-1 .. x0+4, 0 outside the page
	int sector[4];     // 0: E-W, 1: SE-NW, 2: N-S, 3: SW-NE neighbours (valid when `any`)
	bool any;          // some pixel of the warp's row segment reaches the low threshold
};

// `row_inside`: the gradient row is inside the page; for FAST launches it
// already includes "this lane's columns are inside the page".  The direction
// sectors are only needed where a pixel can survive the low threshold, and on
// a document most 120-pixel row segments hold no such pixel: the warp votes
// and skips them.
template <bool FAST>
__device__ __forceinline__ void grow_from_hrows(const HRow &a, const HRow &b,
		const HRow &c, bool row_inside, int x0, int w, int weak_min, GRow &o)
{
	int gxv[4], gyv[4], gx2[4], gy2[4];
	bool cand = false;
#pragma unroll
	for (int i = 0; i < 4; i++) {
		gxv[i] = c.hd[i] + 2 * b.hd[i] + a.hd[i];
		gyv[i] = a.hs[i] - c.hs[i];
		gx2[i] = gxv[i] * gxv[i];
		gy2[i] = gyv[i] * gyv[i];
		const int x = x0 + i;
		const bool inside = FAST ? row_inside : (row_inside && x >= 0 && x < w);
		o.m[i + 1] = inside ? gx2[i] + gy2[i] : 0;
		cand |= o.m[i + 1] >= weak_min;
	}
	o.any = __any_sync(kFull, cand);
	if (o.any) {
#pragma unroll
		for (int i = 0; i < 4; i++) {
			// |angle| < 22.5 deg  <=>  (|gx|+|gy|)^2 < 2 gx^2  <=>  2|gx gy| < gx^2 - gy^2
			const int q = gxv[i] * gyv[i];
			const int p2 = 2 * abs(q);
			const int d = gx2[i] - gy2[i];
			// bit 1: the neighbours are not on this row's E-W line / the main
			// diagonal; bit 0 picks within each half
			o.sector[i] = (p2 < d) ? 0 : ((p2 < -d) ? 2 : ((q >= 0) ? 1 : 3));
		}
	}
	o.m[0] = __shfl_up_sync(kFull, o.m[4], 1);
	o.m[5] = __shfl_down_sync(kFull, o.m[1], 1);
}

struct WeakSink {
	int *hdr, *list, *parent;
	int page_base;
};

// Step J of 3 (s = steps since the top of the strip): gray row r = ys-2+s
// leaves the row queue; gradient row r-1 is formed (slot J, from horizontal
// rows r-2, r-1, r); output row y = r-2 is classified from gradient rows r-3,
// r-2, r-1.  Steps 0-1 of a strip only fill the gray window, steps 2-3 the
// gradient window.
template <bool FAST, int J>
__device__ __forceinline__ void canny_step(const StripGeom &g, const WarpPlace &wp, int s,
		CannyThresholds thr, const WeakSink &ws, RowReader<FAST> &rd, RowWriter<FAST> &wr,
		uint4 (&q)[3], HRow (&hr)[3], GRow (&gr)[3])
{
	const uint4 cur = q[J];
	q[J] = rd.next(g);
	hrow_from_pixels(cur, hr[J]);
	if (s < 2)
		return;
	const int r = wp.ys - 2 + s;
	const int grow = r - 1;
	grow_from_hrows<FAST>(hr[(J + 1) % 3], hr[(J + 2) % 3], hr[J],
			grow >= 0 && grow < g.h && (!FAST || rd.lane_in), wp.x0, g.w, thr.weak_min, gr[J]);
	if (s < 4)
		return;
	const int y = r - 2;
	const GRow &up = gr[(J + 1) % 3], &mid = gr[(J + 2) % 3], &dn = gr[J];
	const bool owns = wr.lane_out && y < wp.ye;
	uint32_t res[4] = { kBlack, kBlack, kBlack, kBlack };
	int nweak = 0;
	if (mid.any) {
#pragma unroll
		for (int i = 0; i < 4; i++) {
			const int m = mid.m[i + 1];
			const int sc = mid.sector[i];
			const int ew = max(mid.m[i], mid.m[i + 2]);
			const int ns = max(up.m[i + 1], dn.m[i + 1]);
			const int d1 = max(dn.m[i + 2], up.m[i]);       // (x+1,y+1), (x-1,y-1)
			const int d3 = max(dn.m[i], up.m[i + 2]);       // (x-1,y+1), (x+1,y-1)
			// two-level select on the sector's bits (no jump table)
			const int lo = (sc & 1) ? d1 : ew;
			const int hi = (sc & 1) ? d3 : ns;
			const int nb = (sc & 2) ? hi : lo;
			if (m >= nb && m >= thr.weak_min) {
				res[i] = kWhite;
				if (m < thr.strong_min && owns && (FAST || wp.x0 + i < g.w)) {
					res[i] = 0u;          // weak: index filled in below
					nweak++;
				}
			}
		}
	}
	// warp-aggregated append of this row's weak pixels to the work list
	if (mid.any && __ballot_sync(kFull, nweak > 0)) {
		const int lane = threadIdx.x & 31;
		int pre = nweak;
#pragma unroll
		for (int dlt = 1; dlt < 32; dlt <<= 1) {
			const int t = __shfl_up_sync(kFull, pre, dlt);
			if (lane >= dlt)
				pre += t;
		}
		int base = 0;
		if (lane == 31)
			base = atomicAdd(&ws.hdr[0], pre);
		base = __shfl_sync(kFull, base, 31) + pre - nweak;
		if (nweak > 0) {
			const int pix0 = ws.page_base + y * g.w + wp.x0;
#pragma unroll
			for (int i = 0; i < 4; i++) {
				if (res[i] == 0u) {
					res[i] = (uint32_t)base;
					ws.list[base] = pix0 + i;
					ws.parent[base] = base;
					base++;
				}
			}
		}
	}
	wr.put(g, wp.ye, res);
}

template <bool FAST>
__device__ __forceinline__ void canny_march(const uint32_t *__restrict__ in,
		uint32_t *__restrict__ out, const StripGeom &g, const WarpPlace &wp,
		CannyThresholds thr, const WeakSink &ws)
{
	RowReader<FAST> rd;
	RowWriter<FAST> wr;
	rd.init(in, g, wp.ys - 2, wp.x0);
	wr.init(out, g, wp.ys, wp.x0, wp.out_lane);
	HRow hr[3];
	GRow gr[3];
	uint4 q[3];
	q[0] = rd.next(g);
	q[1] = rd.next(g);
	q[2] = rd.next(g);
	// gray rows ys-2 .. ys+rows_per_strip+1: rows_per_strip + 4 steps, a multiple of 3
	const int steps = g.rows_per_strip + 4;
	for (int s = 0; s < steps; s += 3) {
		canny_step<FAST, 0>(g, wp, s, thr, ws, rd, wr, q, hr, gr);
		canny_step<FAST, 1>(g, wp, s + 1, thr, ws, rd, wr, q, hr, gr);
		canny_step<FAST, 2>(g, wp, s + 2, thr, ws, rd, wr, q, hr, gr);
	}
}

struct PageGeom {
	int w, h, npix;
};

// Scratch header (ints): the work list's length; (unused); a ticket counter of
// the final pass's blocks; the length of the list last resolved (what
// canny_patch_kernel and the host read afterwards).  The header's first three
// words are zero between calls: whoever resolves a list leaves them so.
enum { kHdrCount = 0, kHdrResolveTicket = 2, kHdrLastCount = 3 };

constexpr int kHystThreads = 1024;
constexpr int kSmallList = 2 * kHystThreads;   // weak pixels one block resolves in shared memory

__global__ void __launch_bounds__(kWarpsPerBlock * 32)
canny_classify_kernel(const uint32_t *__restrict__ in, uint32_t *__restrict__ out,
		const StripGeom g, CannyThresholds thr,
		int *__restrict__ hdr, int *__restrict__ list, int *__restrict__ parent)
{
	pdl_wait();
	const WarpPlace wp = place_warp(g);
	if (!wp.valid)
		return;
	const int npix = g.w * g.h;
	const size_t page = (size_t)blockIdx.y * npix;
	const WeakSink ws = { hdr, list, parent, (int)blockIdx.y * npix };
	if (g.aligned)
		canny_march<true>(in + page, out + page, g, wp, thr, ws);
	else
		canny_march<false>(in + page, out + page, g, wp, thr, ws);
}

// Hysteresis for a short work list (the usual case: a page has hundreds of
// weak pixels), by one block of kHystThreads threads, at most two entries per
// thread, everything in shared memory.  Connected components by hooking and
// pointer jumping (Shiloach-Vishkin): all trees are flattened, then every edge
// whose ends have different roots hangs the larger root under the smaller,
// until no edge does.  A chain of n weak pixels takes a few such rounds, where
// concurrent union-find walked paths of length n (weak pixels follow edges, so
// chains are the rule).  The edges between weak pixels are collected once, as
// a list in shared memory (each undirected edge is held by its end with the
// larger index: at most 4 per entry on average), so a round costs a pass over
// the entries and a pass over the edges, not over eight possible neighbours of
// every entry.

// A load that is kept in L1: the list follows edges, so the neighbourhoods of
// its entries overlap.  (The page is written later in the same kernel, so not
// the read-only path.)
__device__ __forceinline__ uint32_t ld_l1(const uint32_t *p)
{
	uint32_t v;
	asm volatile("ld.global.ca.u32 %0, [%1];" : "=r"(v) : "l"(p));
	return v;
}

__device__ __forceinline__ uint4 ld_l1_v4(const uint32_t *p)
{
	uint4 v;
	asm volatile("ld.global.ca.v4.u32 {%0, %1, %2, %3}, [%4];"
			: "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
	return v;
}

#ifdef PF_CANNY_CLOCKS
#define CANNY_STAMP(k) do { if (threadIdx.x == 0) stamps[k] = (int)clock64(); } while (0)
#else
#define CANNY_STAMP(k)
#endif

struct SmallShared {
	int parent[kSmallList];
	unsigned edge[4 * kSmallList];      // (larger index << 16) | smaller index
	unsigned char keep[kSmallList];
	int nedge;
};

// The 8 neighbours of pixel (x, y) (index idx in `out`), kBlack outside the
// page.  Order: the row above (x-1, x, x+1), W, E, the row below.
template <bool ALIGNED>
__device__ __forceinline__ void canny_neighbours(const uint32_t *out, PageGeom pg, int idx, int x, int y,
		uint32_t (&nv)[8])
{
	if (ALIGNED) {
		// rows are whole 16-byte groups of 4 pixels: the group that holds x, and
		// one more pixel when x is the group's first or last
		const int q = x & 3;
		uint4 v[3];
		uint32_t l[3], r[3];
#pragma unroll
		for (int dy = -1; dy <= 1; dy++) {
			const bool row = y + dy >= 0 && y + dy < pg.h;
			const uint32_t *c = out + idx + dy * pg.w;
			v[dy + 1] = make_uint4(kBlack, kBlack, kBlack, kBlack);
			l[dy + 1] = kBlack;
			r[dy + 1] = kBlack;
			if (row)
				v[dy + 1] = ld_l1_v4(c - q);
			if (row && q == 0 && x > 0)
				l[dy + 1] = ld_l1(c - 1);
			if (row && q == 3 && x + 1 < pg.w)
				r[dy + 1] = ld_l1(c + 1);
		}
		uint32_t t[3][3];
#pragma unroll
		for (int k = 0; k < 3; k++) {
			const uint4 a = v[k];
			t[k][0] = q == 0 ? l[k] : (q == 1 ? a.x : (q == 2 ? a.y : a.z));
			t[k][1] = q == 0 ? a.x : (q == 1 ? a.y : (q == 2 ? a.z : a.w));
			t[k][2] = q == 0 ? a.y : (q == 1 ? a.z : (q == 2 ? a.w : r[k]));
		}
		nv[0] = t[0][0]; nv[1] = t[0][1]; nv[2] = t[0][2];
		nv[3] = t[1][0]; nv[4] = t[1][2];
		nv[5] = t[2][0]; nv[6] = t[2][1]; nv[7] = t[2][2];
	} else {
		int n = 0;
#pragma unroll
		for (int dy = -1; dy <= 1; dy++) {
#pragma unroll
			for (int dx = -1; dx <= 1; dx++) {
				if (dx == 0 && dy == 0)
					continue;
				const bool inside = x + dx >= 0 && x + dx < pg.w && y + dy >= 0 && y + dy < pg.h;
				nv[n] = kBlack;
				if (inside)
					nv[n] = ld_l1(out + idx + dy * pg.w + dx);
				n++;
			}
		}
	}
}

template <bool ALIGNED>
__device__ void canny_resolve_small(uint32_t *out, PageGeom pg, const int *list, int count,
		SmallShared &sh, int *stamps)
{
	const int nt = blockDim.x;
	const int lane = threadIdx.x & 31;
	CANNY_STAMP(0);
	for (int i = threadIdx.x; i < count; i += nt) {
		sh.parent[i] = i;
		sh.keep[i] = 0;
	}
	if (threadIdx.x == 0)
		sh.nedge = 0;
	// this thread's entries: their pixel and its eight neighbours (the loads of
	// both entries issued before any is used: one memory round trip)
	int pix[2] = { -1, -1 };
	uint32_t nv[2][8];
#pragma unroll
	for (int e = 0; e < 2; e++) {
		const int i = threadIdx.x + e * nt;
#pragma unroll
		for (int n = 0; n < 8; n++)
			nv[e][n] = kBlack;
		if (i < count)
			pix[e] = __ldcg(list + i);
	}
#pragma unroll
	for (int e = 0; e < 2; e++) {
		if (pix[e] < 0)
			continue;
		const int p = pix[e] % pg.npix;
		const int y = p / pg.w, x = p - y * pg.w;
		canny_neighbours<ALIGNED>(out, pg, pix[e], x, y, nv[e]);
	}
	CANNY_STAMP(1);
	// whether a strong pixel is among the eight, and the edges this entry holds
	bool strong[2] = { false, false };
	unsigned held[2] = { 0u, 0u };
#pragma unroll
	for (int e = 0; e < 2; e++) {
		const int i = threadIdx.x + e * nt;
#pragma unroll
		for (int n = 0; n < 8; n++) {
			strong[e] |= nv[e][n] == kWhite;
			if (nv[e][n] < kBlack && (int)nv[e][n] < i)
				held[e] |= 1u << n;
		}
	}
	const int mine = __popc(held[0]) + __popc(held[1]);
	int pre = mine;
#pragma unroll
	for (int dlt = 1; dlt < 32; dlt <<= 1) {
		const int t = __shfl_up_sync(kFull, pre, dlt);
		if (lane >= dlt)
			pre += t;
	}
	__syncthreads();         // nedge is zero, parent and keep are set
	int at = 0;
	if (lane == 31 && pre > 0)
		at = atomicAdd(&sh.nedge, pre);
	at = __shfl_sync(kFull, at, 31) + pre - mine;
#pragma unroll
	for (int e = 0; e < 2; e++) {
		const int i = threadIdx.x + e * nt;
#pragma unroll
		for (int n = 0; n < 8; n++)
			if (held[e] >> n & 1u)
				sh.edge[at++] = ((unsigned)i << 16) | nv[e][n];
	}
	__syncthreads();
	const int nedge = sh.nedge;
	CANNY_STAMP(2);
	for (;;) {
		// flatten: every entry points at its parent's parent until that is a
		// root.  No barrier between the jumps: the only writes meanwhile are other
		// entries doing the same, which shorten paths, so a chain of n takes each
		// of its entries about log2(n) jumps, all entries at once.
		{
			volatile int *vp = sh.parent;
			const int i0 = threadIdx.x, i1 = threadIdx.x + nt;
			bool busy0 = i0 < count, busy1 = i1 < count;
			while (busy0 || busy1) {
				const int p0 = busy0 ? vp[i0] : 0, p1 = busy1 ? vp[i1] : 0;
				const int g0 = vp[p0], g1 = vp[p1];
				if (busy0) {
					if (g0 == p0)
						busy0 = false;
					else
						vp[i0] = g0;
				}
				if (busy1) {
					if (g1 == p1)
						busy1 = false;
					else
						vp[i1] = g1;
				}
			}
		}
		__syncthreads();
		// hook: an edge between two trees hangs the larger root under the smaller
		bool hooked = false;
		for (int k = threadIdx.x; k < nedge; k += nt) {
			const unsigned ed = sh.edge[k];
			const int ra = sh.parent[ed >> 16], rb = sh.parent[ed & 0xFFFFu];
			if (ra != rb) {
				atomicMin(&sh.parent[max(ra, rb)], min(ra, rb));
				hooked = true;
			}
		}
		if (!__syncthreads_or(hooked))
			break;
	}
	CANNY_STAMP(3);
#pragma unroll
	for (int e = 0; e < 2; e++) {
		const int i = threadIdx.x + e * nt;
		if (i < count && strong[e])
			sh.keep[sh.parent[i]] = 1;
	}
	__syncthreads();
#pragma unroll
	for (int e = 0; e < 2; e++) {
		const int i = threadIdx.x + e * nt;
		if (i < count)
			out[pix[e]] = sh.keep[sh.parent[i]] ? kWhite : kBlack;
	}
	CANNY_STAMP(4);
}

// ---------------------------------------------------------------------------
// Canny pass B: hysteresis as connected components over the weak work list.
// A weak pixel is kept iff its 8-connected component of weak pixels touches a
// strong pixel.  Union-find over list indices (hook the larger root under the
// smaller with atomicCAS, path halving), so the cost does not depend on how
// long a weak chain is and the result does not depend on scheduling.
// ---------------------------------------------------------------------------

// Root of entry i, or -1 when its component is known to touch a strong pixel
// (the mark is a root's parent, so it is found -- and short-cut to -- like any
// other ancestor).
__device__ __forceinline__ int uf_find(int *parent, int i)
{
	// L2 loads/stores: other SMs hook roots with atomics while this runs
	int p = __ldcg(parent + i);
	while (p != i) {
		if (p < 0)
			return -1;
		const int gp = __ldcg(parent + p);
		if (gp != p)
			__stcg(parent + i, gp);      // path halving; races only shorten paths
		i = p;
		p = gp;
	}
	return i;
}

// b = -1 joins a's component with the strong mark.
__device__ __forceinline__ void uf_union(int *parent, int a, int b)
{
	for (;;) {
		a = uf_find(parent, a);
		if (b >= 0)
			b = uf_find(parent, b);
		if (a == b)
			return;
		if (a < b) {
			const int t = a;
			a = b;
			b = t;
		}
		// a > b >= -1: hang root a under b
		const int old = atomicCAS(&parent[a], a, b);
		if (old == a)
			return;
	}
}

// The weak pixels that the classification kernel left on the list, in two
// kernels, the launch boundary being the grid-wide barrier between:
//   1. join every weak pixel with its weak neighbours (each undirected edge
//      once, from the pixel with the smaller address: E, SW, S, SE) and, if
//      one of its eight neighbours is strong, with the strong mark -- the
//      "root" -1, smaller than every index, so it stays on top of whatever is
//      hung under it; a short list is resolved here by block 0 alone, in
//      shared memory;
//   2. every weak pixel takes its component's verdict (its root is the mark or
//      not), and the block that finishes last clears the header for the next
//      call.
// With nothing on the list each is a launch that returns at once.
//
// All neighbour loads of an entry are issued before any is used: these loops
// are bound by memory latency, not by work.  The image is only read until
// kernel 2 writes the verdicts, so the reads may go through L1 (entries that
// follow one another in the list are neighbours in the page and share
// sectors).
#ifndef PF_CANNY_LINK_MINBLOCKS
#define PF_CANNY_LINK_MINBLOCKS 2
#endif
__global__ void __launch_bounds__(kHystThreads, PF_CANNY_LINK_MINBLOCKS)
canny_link_kernel(uint32_t *out, PageGeom pg, int *list, int *parent, int *hdr)
{
	__shared__ SmallShared sh;
	pdl_wait();
	const int count = __ldcg(&hdr[kHdrCount]);
	if (count == 0)
		return;
	if (count <= kSmallList) {
		if (blockIdx.x == 0) {
			if (pg.w % 4 == 0 && ((uintptr_t)out & 15) == 0)
				canny_resolve_small<true>(out, pg, list, count, sh, hdr + 32);
			else
				canny_resolve_small<false>(out, pg, list, count, sh, hdr + 32);
			__syncthreads();
			if (threadIdx.x == 0) {
				hdr[kHdrLastCount] = count;
				hdr[kHdrCount] = 0;
			}
		}
		return;
	}
	const int first = blockIdx.x * blockDim.x + threadIdx.x;
	const int step = gridDim.x * blockDim.x;
	for (int i = first; i < count; i += step) {
		// all eight neighbours first: one memory round trip (one entry per thread
		// and round: the block is full, and sixteen loads in flight per thread
		// would not fit its 32 registers)
		const int idx = list[i];
		const int p = idx % pg.npix;
		const int y = p / pg.w, x = p - y * pg.w;
		const bool has_e = x + 1 < pg.w, has_w = x > 0, has_s = y + 1 < pg.h, has_n = y > 0;
		uint32_t nv[8];
#pragma unroll
		for (int n = 0; n < 8; n++)
			nv[n] = kBlack;
		// 0-3: E, SW, S, SE (the edges this pixel owns); 4-7: W, NW, N, NE
		if (has_e)
			nv[0] = __ldg(out + idx + 1);
		if (has_s && has_w)
			nv[1] = __ldg(out + idx + pg.w - 1);
		if (has_s)
			nv[2] = __ldg(out + idx + pg.w);
		if (has_s && has_e)
			nv[3] = __ldg(out + idx + pg.w + 1);
		if (has_w)
			nv[4] = __ldg(out + idx - 1);
		if (has_n && has_w)
			nv[5] = __ldg(out + idx - pg.w - 1);
		if (has_n)
			nv[6] = __ldg(out + idx - pg.w);
		if (has_n && has_e)
			nv[7] = __ldg(out + idx - pg.w + 1);
		bool strong = false;
#pragma unroll
		for (int n = 0; n < 8; n++)
			strong |= nv[n] == kWhite;
		// (the mark first: joining two components that both carry it is a pair
		// of finds and nothing else -- on a page of noise, where nearly every
		// weak pixel has a strong neighbour, that is most joins)
		if (strong)
			uf_union(parent, i, -1);
#pragma unroll
		for (int n = 0; n < 4; n++)
			if (nv[n] < kBlack)
				uf_union(parent, i, (int)nv[n]);
	}
}

__global__ void __launch_bounds__(kHystThreads)
canny_verdict_kernel(uint32_t *out, int *list, int *parent, int *hdr)
{
	pdl_wait();
	const int count = __ldcg(&hdr[kHdrCount]);
	if (count == 0)
		return;
	const int first = blockIdx.x * blockDim.x + threadIdx.x;
	const int step = gridDim.x * blockDim.x;
	for (int i = first; i < count; i += step)
		out[__ldcg(list + i)] = uf_find(parent, i) < 0 ? kWhite : kBlack;
	// the block that finishes last clears the header (every block has read
	// the length before it gets here)
	__syncthreads();
	if (threadIdx.x == 0 && atomicAdd(&hdr[kHdrResolveTicket], 1) == (int)gridDim.x - 1) {
		hdr[kHdrResolveTicket] = 0;
		hdr[kHdrLastCount] = count;
		hdr[kHdrCount] = 0;
	}
}

// Clears the header of scratch in an unknown state (a caller's): a kernel, not
// a memset, so that the chain of programmatic launches is not broken.
__global__ void canny_begin_kernel(int *hdr)
{
	pdl_wait();
	hdr[threadIdx.x] = 0;
}

// The final values of the weak pixels, copied to a second image (the caller's
// pinned page, which already holds everything else: see pf_capi.cu).
__global__ void canny_patch_kernel(const uint32_t *__restrict__ out, uint32_t *__restrict__ mirror,
		const int *__restrict__ list, const int *__restrict__ hdr)
{
	pdl_wait();
	const int count = hdr[kHdrLastCount];
	for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
		const int idx = list[i];
		mirror[idx] = out[idx];
	}
}

}  // namespace

// Blocks of the two hysteresis kernels (PF_B200_HYST_BLOCKS per SM, default 2).
static int hysteresis_blocks()
{
	static const int want = getenv("PF_B200_HYST_BLOCKS") ? atoi(getenv("PF_B200_HYST_BLOCKS")) : 2;
	return sm_count() * (want < 1 ? 1 : (want > 2 ? 2 : want));
}

static cudaError_t clear_canny_header(int *hdr, cudaStream_t st)
{
	return launch_pdl(canny_begin_kernel, dim3(1), dim3(64), 0, st, hdr);
}

// The weak pixels the classification left on the list: link, verdict.
static cudaError_t launch_hysteresis(uint32_t *out, PageGeom pg, int *list, int *parent,
		int *hdr, int blocks, cudaStream_t st)
{
	cudaError_t e = launch_pdl(canny_link_kernel, dim3(blocks), dim3(kHystThreads), 0, st,
			out, pg, list, parent, hdr);
	if (e == cudaSuccess)
		e = launch_pdl(canny_verdict_kernel, dim3(blocks), dim3(kHystThreads), 0, st,
				out, list, parent, hdr);
	return e;
}

int launch_sobel(const uint32_t *in, uint32_t *out, int w, int h, int pages,
		cudaStream_t st, int y_begin, int y_end)
{
	if (y_end < 0)
		y_end = h;
	if (y_begin < 0 || y_begin >= y_end || y_end > h)
		return fail(PF_B200_EINVAL, "sobel: bad row range [%d, %d) of %d", y_begin, y_end, h);
	const StripGeom g = make_geom(sobel_kernel, in, out, w, h, pages, 2, y_begin, y_end, 3);
	const int warps = g.warp_cols * g.strips;
	dim3 grid((warps + kWarpsPerBlock - 1) / kWarpsPerBlock, pages);
	PFB_CUDA(launch_pdl(sobel_kernel, grid, dim3(kWarpsPerBlock * 32), 0, st, in, out, g));
	return PF_B200_OK;
}

// scratch: 256-byte header, then per pixel of the largest batch the work list
// entry and the union-find parent (2 ints)
size_t canny_scratch_bytes(int w, int h)
{
	return 256 + 2 * sizeof(int) * (size_t)w * h;
}

int launch_canny(const uint32_t *in, uint32_t *out, int w, int h, int pages,
		CannyThresholds thr, void *scratch, size_t scratch_bytes,
		cudaStream_t st, bool header_is_clear)
{
	const size_t npix = (size_t)w * h;
	if (npix >= 0x7F000000ull)
		return fail(PF_B200_ETOOBIG, "canny: page of %d x %d is too large", w, h);
	if (scratch == nullptr || scratch_bytes < canny_scratch_bytes(w, h))
		return fail(PF_B200_EINVAL, "canny: scratch of %zu bytes, need %zu",
				scratch_bytes, canny_scratch_bytes(w, h));
	if (((uintptr_t)scratch & 255) != 0)
		return fail(PF_B200_EINVAL, "canny: scratch must be 256-byte aligned");
	// pages per launch: what the scratch holds and an int (below kBlack) can index
	size_t cap_pages = (scratch_bytes - 256) / (2 * sizeof(int) * npix);
	const size_t idx_pages = 0x7F000000ull / npix;
	if (cap_pages > idx_pages)
		cap_pages = idx_pages;
	if (cap_pages > 65535)
		cap_pages = 65535;
	int *hdr = static_cast<int *>(scratch);
	int *list = hdr + 64;
	PageGeom pg = { w, h, (int)npix };
	const int coop_blocks = hysteresis_blocks();

	for (int p0 = 0; p0 < pages; p0 += (int)cap_pages) {
		const int np = (pages - p0 < (int)cap_pages) ? pages - p0 : (int)cap_pages;
		const uint32_t *pin = in + (size_t)p0 * npix;
		uint32_t *pout = out + (size_t)p0 * npix;
		int *parent = list + (size_t)np * npix;
		// every pass leaves the header clear; only scratch in an unknown state
		// (a caller's, on its first use) has to be cleared here
		if (!header_is_clear && p0 == 0)
			PFB_CUDA(clear_canny_header(hdr, st));
		const StripGeom g = make_geom(canny_classify_kernel, pin, pout, w, h, np, 4, 0, -1, 3);
		const int warps = g.warp_cols * g.strips;
		dim3 grid((warps + kWarpsPerBlock - 1) / kWarpsPerBlock, np);
		PFB_CUDA(launch_pdl(canny_classify_kernel, grid, dim3(kWarpsPerBlock * 32), 0, st,
				pin, pout, g, thr, hdr, list, parent));
		PFB_CUDA(launch_hysteresis(pout, pg, list, parent, hdr, coop_blocks, st));
	}
	return PF_B200_OK;
}

// One page in three steps, for the host-pointer path that classifies row bands
// as they arrive: canny_begin, canny_classify_rows for every band (in any
// order, each needs its rows plus two above and below in `in`), canny_resolve.
static int canny_check(int w, int h, void *scratch, size_t scratch_bytes)
{
	if ((size_t)w * h >= 0x7F000000ull)
		return fail(PF_B200_ETOOBIG, "canny: page of %d x %d is too large", w, h);
	if (scratch == nullptr || scratch_bytes < canny_scratch_bytes(w, h) || ((uintptr_t)scratch & 255) != 0)
		return fail(PF_B200_EINVAL, "canny: scratch of %zu bytes, need %zu, 256-byte aligned",
				scratch_bytes, canny_scratch_bytes(w, h));
	return PF_B200_OK;
}

int launch_canny_begin(int w, int h, void *scratch, size_t scratch_bytes, cudaStream_t st,
		bool header_is_clear)
{
	const int rc = canny_check(w, h, scratch, scratch_bytes);
	if (rc)
		return rc;
	if (!header_is_clear)
		PFB_CUDA(clear_canny_header(static_cast<int *>(scratch), st));
	return PF_B200_OK;
}

int launch_canny_classify_rows(const uint32_t *in, uint32_t *out, int w, int h,
		int y_begin, int y_end, CannyThresholds thr, void *scratch, cudaStream_t st)
{
	if (y_begin < 0 || y_begin >= y_end || y_end > h)
		return fail(PF_B200_EINVAL, "canny: bad row range [%d, %d) of %d", y_begin, y_end, h);
	int *hdr = static_cast<int *>(scratch);
	int *list = hdr + 64;
	int *parent = list + (size_t)w * h;
	const StripGeom g = make_geom(canny_classify_kernel, in, out, w, h, 1, 4, y_begin, y_end, 3);
	const int warps = g.warp_cols * g.strips;
	dim3 grid((warps + kWarpsPerBlock - 1) / kWarpsPerBlock, 1);
	PFB_CUDA(launch_pdl(canny_classify_kernel, grid, dim3(kWarpsPerBlock * 32), 0, st,
			in, out, g, thr, hdr, list, parent));
	return PF_B200_OK;
}

int launch_canny_resolve(uint32_t *out, int w, int h, void *scratch, cudaStream_t st)
{
	int *hdr = static_cast<int *>(scratch);
	int *list = hdr + 64;
	int *parent = list + (size_t)w * h;
	PageGeom pg = { w, h, w * h };
	PFB_CUDA(launch_hysteresis(out, pg, list, parent, hdr, hysteresis_blocks(), st));
	return PF_B200_OK;
}

// After canny_resolve: the weak pixels' final values into `mirror` (a second
// copy of the page that holds the classified rows).
int launch_canny_patch(const uint32_t *out, uint32_t *mirror, void *scratch, cudaStream_t st)
{
	const int *hdr = static_cast<const int *>(scratch);
	PFB_CUDA(launch_pdl(canny_patch_kernel, dim3(sm_count()), dim3(256), 0, st,
			out, mirror, hdr + 64, hdr));
	return PF_B200_OK;
}

}  // namespace pfb

// ---------------------------------------------------------------------------
// Canny without any allocation: the work list lives in a fixed device array.
// ---------------------------------------------------------------------------

namespace pfb {

namespace {

constexpr size_t kStaticMaxPixels = (size_t)4 << 20;       // 4,194,304 pixels per page
constexpr size_t kStaticScratchBytes = 256 + 2 * sizeof(int) * kStaticMaxPixels;   // 32 MiB + 256 B
constexpr int kStaticMaxDevices = 64;

// One copy per GPU that loads this library, held until the process exits.
// In host code this name is only a handle for cudaGetSymbolAddress; it is not
// a device address and must never be handed to a kernel as it stands.
static __device__ __align__(256) unsigned char g_canny_static_scratch[kStaticScratchBytes];

std::mutex g_static_mutex;
void *g_static_addr[kStaticMaxDevices] = {};   // device address of the array, per device

// Looks the array's address up once per device, with that device current.
int static_scratch_for_current_device(void **addr)
{
	int dev = -1;
	PFB_CUDA(cudaGetDevice(&dev));
	if (dev < 0 || dev >= kStaticMaxDevices)
		return fail(PF_B200_EINVAL, "canny (no-alloc): device index %d outside 0..%d",
				dev, kStaticMaxDevices - 1);
	std::lock_guard<std::mutex> lock(g_static_mutex);
	if (g_static_addr[dev] == nullptr) {
		void *p = nullptr;
		PFB_CUDA(cudaGetSymbolAddress(&p, g_canny_static_scratch));
		g_static_addr[dev] = p;
	}
	*addr = g_static_addr[dev];
	return PF_B200_OK;
}

}  // namespace

}  // namespace pfb

// See include/pillowfight_b200.h for the contract of this entry point.
extern "C" int pf_b200_dev_canny_noalloc(const uint32_t *d_in, uint32_t *d_out,
		int width, int height, void *stream)
{
	using namespace pfb;
	clear_error();
	if (!d_in || !d_out)
		return fail(PF_B200_EINVAL, "NULL device pointer");
	if (d_in == d_out)
		return fail(PF_B200_EINVAL, "device entry points do not run in place");
	if (width <= 0 || height <= 0)
		return fail(PF_B200_EINVAL, "size %d x %d", width, height);
	if ((size_t)width * (size_t)height > kStaticMaxPixels)
		return fail(PF_B200_ETOOBIG,
				"canny (no-alloc): %d x %d = %zu pixels, the built-in scratch holds at most %zu; "
				"use pf_b200_dev_canny with caller-provided scratch for larger pages",
				width, height, (size_t)width * (size_t)height, kStaticMaxPixels);
	void *scratch = nullptr;
	const int rc = static_scratch_for_current_device(&scratch);
	if (rc)
		return rc;
	// launch_canny zeroes the 256-byte header (the work-list length) on the
	// stream before the first kernel; list and parent slots are written before
	// they are read, so whatever the previous call left behind is harmless.
	return launch_canny(d_in, d_out, width, height, 1, canny_thresholds(),
			scratch, kStaticScratchBytes, static_cast<cudaStream_t>(stream));
}
