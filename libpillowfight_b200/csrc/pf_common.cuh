// Shared declarations of the sm_100a build of libpillowfight's hot path.
// Kernels live in pf_gaussian.cu, pf_edges.cu, pf_ace.cu and pf_matrix.cu, the
// C-ABI in pf_capi.cu.
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "../../include/pillowfight_b200.h"

namespace pfb {

// B200: 148 SMs.  Grids of the streaming kernels are sized from the problem;
// the persistent ones from the device's SM count, queried once per device.
int sm_count();

// Records an error for the calling thread (pf_b200_last_error) and returns code.
int fail(int code, const char *fmt, ...);
void clear_error();

#define PFB_CUDA(call)                                                        \
	do {                                                                      \
		cudaError_t e_ = (call);                                              \
		if (e_ != cudaSuccess)                                                \
			return pfb::fail(PF_B200_ECUDA, "%s:%d: %s -> %s", __FILE__,      \
					__LINE__, #call, cudaGetErrorString(e_));                 \
	} while (0)

constexpr int kMaxTaps = 63;            // Gaussian tap count limit (nb_stddev)

struct Taps {
	int n;
	double t[kMaxTaps];
};

// upstream _gaussian.c generate_gaussian_1d_kernel, on the host (pf_capi.cu)
int make_taps(double sigma, int nb_stddev, Taps &t);

// ---- stencil launchers (pf_gaussian.cu, pf_edges.cu); asynchronous on `st` --
// [y_begin, y_end) restricts the OUTPUT rows (y_end < 0: the whole page);
// input rows up to the stencil's reach outside that range are read, so they
// must already be in place -- this is what lets the host-pointer entry points
// overlap the upload of one band with the work on another.
int launch_gaussian(const uint32_t *in, uint32_t *out, int w, int h, int pages,
		const Taps &taps, cudaStream_t st, int y_begin = 0, int y_end = -1);
int launch_sobel(const uint32_t *in, uint32_t *out, int w, int h, int pages,
		cudaStream_t st, int y_begin = 0, int y_end = -1);

struct CannyThresholds {
	int weak_min;    // smallest g_x^2+g_y^2 whose hypot is  > LOW
	int strong_min;  // smallest g_x^2+g_y^2 whose hypot is  > HIGH
};
CannyThresholds canny_thresholds();   // upstream _canny.c LOW / HIGH as integer tests
size_t canny_scratch_bytes(int w, int h);
// header_is_clear: the scratch's 256-byte header is known to be zero (fresh
// from a memset, or left by an earlier call: every call leaves it zero);
// otherwise it is cleared on the stream first.
int launch_canny(const uint32_t *in, uint32_t *out, int w, int h, int pages,
		CannyThresholds thr, void *scratch, size_t scratch_bytes,
		cudaStream_t st, bool header_is_clear = false);
// One page in steps (row bands classified as they arrive): begin, classify
// every band (each needs two rows above and below it in `in`), resolve; patch
// copies the weak pixels' final values into a second image.
int launch_canny_begin(int w, int h, void *scratch, size_t scratch_bytes, cudaStream_t st,
		bool header_is_clear = false);
int launch_canny_classify_rows(const uint32_t *in, uint32_t *out, int w, int h,
		int y_begin, int y_end, CannyThresholds thr, void *scratch, cudaStream_t st);
int launch_canny_resolve(uint32_t *out, int w, int h, void *scratch, cudaStream_t st);
int launch_canny_patch(const uint32_t *out, uint32_t *mirror, void *scratch, cudaStream_t st);

// ---- pf_dbl_matrix launchers (pf_matrix.cu) ------------------------------
int launch_matrix_convolution(const double *img, double *out, int w, int h,
		const double *kernel, int nx, int ny, cudaStream_t st);
int launch_gradient_polar(const double *gx, const double *gy, double *intensity,
		double *direction, size_t n, cudaStream_t st);

// ---- programmatic dependent launch ---------------------------------------
// Every kernel of the bitmap path (pf_gaussian.cu, pf_edges.cu, pf_ace.cu; the
// pf_dbl_matrix kernels of pf_matrix.cu are launched the plain way) begins
// with pdl_wait() and is launched with launch_pdl(): its blocks may be scheduled while the kernel before it on the
// stream is still draining, and wait (griddepcontrol.wait: until that kernel
// has completed and its writes are visible) before they touch memory.  What is
// saved is the launch latency between dependent kernels, a few microseconds
// each -- 10 launches per page.  PF_B200_PDL=0 turns it off.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait()
{
	asm volatile("griddepcontrol.wait;" ::: "memory");
}

bool pdl_enabled();

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem,
		cudaStream_t st, Args... args)
{
	cudaLaunchConfig_t cfg = {};
	cfg.gridDim = grid;
	cfg.blockDim = block;
	cfg.dynamicSmemBytes = smem;
	cfg.stream = st;
	cudaLaunchAttribute attr[1];
	attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
	attr[0].val.programmaticStreamSerializationAllowed = 1;
	cfg.attrs = attr;
	cfg.numAttrs = pdl_enabled() ? 1 : 0;
	return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}
#endif

// ---- ACE launchers (pf_ace.cu) -------------------------------------------
struct AceParams {
	int w, h;
	int row_begin, row_end;
	int nb_samples;
	double slope, limit;
	// Optional (host-pointer streaming path, row bands on other GPUs): the
	// samples' pixel values, so that the passes over a band do not need the
	// rest of the page; NULL = read them from the page.
	const uint32_t *sample_px = nullptr;
	// Which score row is stored first in the score block and how far apart the
	// planes are.  Defaults (< 0): a band-local block, row_begin first, planes
	// of (row_end-row_begin)*w floats.
	int score_row0 = -1;
	long score_plane = -1;
};

// A score block (device memory, ace_scores_bytes(w, rows) bytes, 256-byte
// aligned): a header with the lengths of the fix-up lists, the R, G, B score
// planes (FP32), and room for the fix-up list (one entry per pixel at most).
// A rescale call appends to the list of one slot; slots used on one block must
// cover disjoint rows (a slot's list starts at its first row's pixel index).
constexpr int kAceScoreHeader = 256;     // bytes
constexpr int kAceFixupSlots = 16;
struct AceScores {
	int *counts;         // kAceFixupSlots list lengths, then as many tickets; the accumulation pass zeroes them, the fix-up pass leaves them zero
	float *planes;
	uint32_t *list;
	size_t plane;        // floats between planes = capacity of the list
};
size_t ace_scores_bytes(int w, int rows);
AceScores ace_scores_at(void *block, size_t plane);

// `minmax` is a block of ace_minmax_doubles(w, h) doubles: the page-wide
// extremes first, then one entry per tile (pf_ace.cu, kMinmaxHead).
size_t ace_minmax_doubles(int w, int h);
int launch_ace_minmax_init(double *minmax, int w, int h, cudaStream_t st);
int launch_ace_gather_samples(const uint32_t *page, int w, const int *samples_xy, int n, int *msg,
		cudaStream_t st);
// Pass 1, rows [row_begin, row_end): approximate scores (error bounded) into
// the planes, their extremes folded into minmax[0..5] and the tile table.
int launch_ace_scores(const uint32_t *in, const int *samples_xy,
		const AceParams &p, const AceScores &sc, double *minmax, cudaStream_t st);
// Pass 2, after pass 1 of every band (and, across GPUs, the min/max exchange),
// same parameters: the exact extremes of the band into minmax[6..11].
int launch_ace_refine(const uint32_t *in, const int *samples_xy,
		const AceParams &p, const AceScores &sc, double *minmax, cudaStream_t st);
// Pass 3, after pass 2 of every band (and the second exchange): rows
// [row0, row0 + rows) of the block rescaled to RGBA8 into nb_outs images
// (outs[d] points at row0 of image d); uncertain pixels listed under `slot`.
int launch_ace_scale(const AceParams &p, const AceScores &sc, const double *minmax,
		int row0, int rows, int slot, uint32_t *const *outs, int nb_outs, bool multicast,
		cudaStream_t st);
// Pass 4, same row0 / slot / outs: the listed pixels recomputed in the
// reference's arithmetic and stored again.
int launch_ace_fixup(const uint32_t *in, const int *samples_xy, const AceParams &p,
		const AceScores &sc, const double *minmax, int row0, int slot,
		uint32_t *const *outs, int nb_outs, bool multicast, cudaStream_t st);
double measure_fp64_peak_tflops();
double measure_fp32_peak_tflops();
double measure_rsqrt_error(int which);

}  // namespace pfb
