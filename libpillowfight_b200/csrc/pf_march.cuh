// Shared machinery of the warp-march stencil kernels (Gaussian, Sobel, Canny).
//
// A warp marches down a strip of rows.  Each lane owns 4 adjacent columns (one
// 16-byte load / store per row); lanes 0 and 31 are halo lanes whose pixels
// only feed their neighbours through shuffles, so a warp emits 120 columns.
// The vertical window of the stencil lives in registers, rotated by unrolling
// the row loop into whole groups of steps.  Included by pf_edges.cu and
// pf_gaussian.cu; everything here is file-local to the including unit.
#pragma once

#include "pf_common.cuh"

namespace pfb {

namespace {

constexpr int kWarpsPerBlock = 4;
constexpr int kOutLanes = 30;
constexpr int kWarpCols = kOutLanes * 4;        // 120 output columns per warp
constexpr unsigned kFull = 0xFFFFFFFFu;

struct StripGeom {
	int w, h;
	int y_begin, y_end;  // output rows of this launch
	int rows_per_strip;
	int warp_cols;       // ceil(w / 120)
	int strips;          // ceil((y_end - y_begin) / rows_per_strip)
	int aligned;         // rows can be accessed as uint4
};


__device__ __forceinline__ uint4 load_px4(const uint32_t *__restrict__ img,
		const StripGeom &g, int r, int x0)
{
	uint4 v = make_uint4(0u, 0u, 0u, 0u);
	if (r < 0 || r >= g.h)
		return v;
	const uint32_t *row = img + (size_t)r * g.w;
	if (g.aligned && x0 >= 0 && x0 + 3 < g.w)
		return __ldg(reinterpret_cast<const uint4 *>(row + x0));
	if (x0 >= 0 && x0 < g.w) v.x = __ldg(row + x0);
	if (x0 + 1 >= 0 && x0 + 1 < g.w) v.y = __ldg(row + x0 + 1);
	if (x0 + 2 >= 0 && x0 + 2 < g.w) v.z = __ldg(row + x0 + 2);
	if (x0 + 3 >= 0 && x0 + 3 < g.w) v.w = __ldg(row + x0 + 3);
	return v;
}

__device__ __forceinline__ void store_px4(uint32_t *__restrict__ img,
		const StripGeom &g, int r, int x0, const uint32_t (&v)[4])
{
	uint32_t *row = img + (size_t)r * g.w;
	if (g.aligned && x0 + 3 < g.w) {
		*reinterpret_cast<uint4 *>(row + x0) = make_uint4(v[0], v[1], v[2], v[3]);
		return;
	}
#pragma unroll
	for (int i = 0; i < 4; i++)
		if (x0 + i < g.w)
			row[x0 + i] = v[i];
}

// ---------------------------------------------------------------------------
// Row access for the march.  FAST: the width is a multiple of 4 and rows are
// 16-byte aligned, so a lane's 4 columns are inside the page or outside it
// together and a row is one LDG.128 / STG.128 through a pointer that advances
// by the row pitch; the only tests are one per lane (columns) and one per row
// (warp-uniform).  !FAST: odd widths or unaligned pages, every pixel
// bounds-checked.  A kernel launch uses one or the other throughout.
// Rows are read several steps ahead of their use (the kernels keep a small
// queue of them in registers): a warp has one to five 512-byte rows in flight.
// ---------------------------------------------------------------------------

template <bool FAST>
struct RowReader {
	const uint32_t *img;      // page base (generic path)
	const uint4 *p;           // (row, x0) of the next row to load (fast path)
	int pitch4;               // row pitch in uint4
	int r;                    // index of the next row to load
	int x0;
	bool lane_in;             // the lane's 4 columns are inside the page (fast path: all or none)

	__device__ __forceinline__ void init(const uint32_t *page, const StripGeom &g, int first_row, int x0_)
	{
		img = page;
		x0 = x0_;
		r = first_row;
		pitch4 = g.w >> 2;
		lane_in = x0_ >= 0 && x0_ + 3 < g.w;
		p = reinterpret_cast<const uint4 *>(page + (long)first_row * g.w + x0_);
	}
	__device__ __forceinline__ uint4 next(const StripGeom &g)
	{
		uint4 v;
		if (FAST) {
			v = make_uint4(0u, 0u, 0u, 0u);
			if (lane_in && r >= 0 && r < g.h)
				v = __ldg(p);
			p += pitch4;
		} else {
			v = load_px4(img, g, r, x0);
		}
		r++;
		return v;
	}
};

template <bool FAST>
struct RowWriter {
	uint32_t *img;
	uint4 *p;                 // (row, x0) of the next row to store (fast path)
	int pitch4;
	int r;
	int x0;
	bool lane_out;            // this lane stores: an output lane whose columns are inside the page

	__device__ __forceinline__ void init(uint32_t *page, const StripGeom &g, int first_row, int x0_,
			bool out_lane)
	{
		img = page;
		x0 = x0_;
		r = first_row;
		pitch4 = g.w >> 2;
		lane_out = out_lane && x0_ >= 0 && (FAST ? x0_ + 3 < g.w : x0_ < g.w);
		p = reinterpret_cast<uint4 *>(page + (long)first_row * g.w + x0_);
	}
	// stores row r (if it belongs to the strip), then moves down a row
	__device__ __forceinline__ void put(const StripGeom &g, int row_end, const uint32_t (&v)[4])
	{
		if (FAST) {
			if (lane_out && r < row_end)
				*p = make_uint4(v[0], v[1], v[2], v[3]);
			p += pitch4;
		} else if (lane_out && r < row_end) {
			store_px4(img, g, r, x0, v);
		}
		r++;
	}
};

struct WarpPlace {
	int x0;           // first column of this lane (may be -4 for a halo lane)
	int ys, ye;       // output rows of the strip
	bool out_lane;    // lanes 1..30
	bool valid;
};

__device__ __forceinline__ WarpPlace place_warp(const StripGeom &g)
{
	WarpPlace wp;
	const int gw = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
	const int strip = gw / g.warp_cols;
	const int lane = threadIdx.x & 31;
	const int wx = (gw - strip * g.warp_cols) * kWarpCols - 4;
	wp.valid = strip < g.strips;
	wp.x0 = wx + lane * 4;
	wp.out_lane = lane >= 1 && lane <= kOutLanes;
	wp.ys = g.y_begin + strip * g.rows_per_strip;
	wp.ye = min(g.y_end, wp.ys + g.rows_per_strip);
	return wp;
}

// Strip height so that the whole grid is resident at once (one wave): a grid
// a few blocks over the resident capacity would run a second, nearly empty
// wave and double the kernel time.
template <typename Kernel>
StripGeom make_geom(Kernel kern, const uint32_t *in, const uint32_t *out, int w, int h,
		int pages, int halo_rows, int y_begin = 0, int y_end = -1, int group = 6)
{
	StripGeom g;
	g.w = w;
	g.h = h;
	g.y_begin = y_begin;
	g.y_end = y_end < 0 ? h : y_end;
	const int nrows = g.y_end - g.y_begin;
	g.warp_cols = (w + kWarpCols - 1) / kWarpCols;
	int blocks_per_sm = 0;
	if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, kern,
			kWarpsPerBlock * 32, 0) != cudaSuccess || blocks_per_sm < 1) {
		cudaGetLastError();
		blocks_per_sm = 4;
	}
	const long resident_warps = (long)sm_count() * blocks_per_sm * kWarpsPerBlock;
	long strips = resident_warps / ((long)g.warp_cols * pages);
	if (strips < 1)
		strips = 1;
	// rows + halo_rows steps per strip must be a whole number of step groups
	int rows = (int)((nrows + strips - 1) / strips);
	int groups = (rows + halo_rows + group - 1) / group;
	if (groups * group < 18)
		groups = (18 + group - 1) / group;
	rows = groups * group - halo_rows;
	g.rows_per_strip = rows;
	g.strips = (nrows + rows - 1) / rows;
	g.aligned = (w % 4 == 0) && (((uintptr_t)in | (uintptr_t)out) % 16 == 0);
	return g;
}


}  // namespace

}  // namespace pfb
