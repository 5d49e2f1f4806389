// Shared declarations of the sm_100a build of libpillowfight's hot path.
// Kernels live in pf_stencil.cu / pf_ace.cu, the C-ABI in pf_capi.cu.
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

// ---- stencil launchers (pf_stencil.cu); all asynchronous on `st` ----------
int launch_gaussian(const uint32_t *in, uint32_t *out, int w, int h, int pages,
		const Taps &taps, cudaStream_t st);
int launch_sobel(const uint32_t *in, uint32_t *out, int w, int h, int pages,
		cudaStream_t st);

struct CannyThresholds {
	int weak_min;    // smallest g_x^2+g_y^2 whose hypot is  > LOW
	int strong_min;  // smallest g_x^2+g_y^2 whose hypot is  > HIGH
};
CannyThresholds canny_thresholds();   // upstream _canny.c LOW / HIGH as integer tests
size_t canny_scratch_bytes(int w, int h);
int launch_canny(const uint32_t *in, uint32_t *out, int w, int h, int pages,
		CannyThresholds thr, void *scratch, size_t scratch_bytes,
		cudaStream_t st);

// ---- ACE launchers (pf_ace.cu) -------------------------------------------
struct AceParams {
	int w, h;
	int row_begin, row_end;
	int nb_samples;
	double slope, limit;
};
int launch_ace_minmax_init(double *minmax, cudaStream_t st);
int launch_ace_scores(const uint32_t *in, const int *samples_xy,
		const AceParams &p, double *scores, double *minmax, cudaStream_t st);
int launch_ace_scale(const double *scores, const double *minmax, int w,
		int rows, uint32_t *out, cudaStream_t st);
double measure_fp64_peak_tflops();

}  // namespace pfb
