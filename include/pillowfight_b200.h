/*
 * pillowfight_b200.h -- C-ABI of the B200 (sm_100a) build of libpillowfight's
 * data-parallel hot path: ACE and the Gaussian / Sobel / Canny stencil chain.
 *
 * Everything here is plain C: pointers, ints, doubles.  No CUDA or torch types
 * appear in a signature (a CUDA stream is passed as `void *`).
 *
 * REFERENCE CITATIONS.  The snapshot mounted at /root/reference holds only a
 * "moved to GNOME GitLab" README (README.md:1-3) -- no C sources.  Where this
 * header says "replaces <upstream file>", it names the upstream file of the
 * published libpillowfight source tree (src/pillowfight/...) that the entry
 * point stands in for; line numbers cannot be given because the file is not in
 * the mount.  See DESIGN.md section 0 ("parity unpinned").
 */
#ifndef PILLOWFIGHT_B200_H
#define PILLOWFIGHT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------ */
/* Reference data types (upstream src/pillowfight/util.h)                    */
/* ------------------------------------------------------------------------ */

#define PF_NB_COLORS 4      /* RGBA, 32 bits per pixel */
#define PF_NB_RGB_COLORS 3

enum pf_color { COLOR_R = 0, COLOR_G = 1, COLOR_B = 2, COLOR_A = 3 };

union pf_pixel {
	uint32_t whole;
	struct {
		uint8_t r;
		uint8_t g;
		uint8_t b;
		uint8_t a;
	} color;
	uint8_t channels[PF_NB_COLORS];
};

/* Row-major RGBA8 image, no row padding: pixel (x, y) is pixels[y*size.x+x]. */
struct pf_bitmap {
	struct {
		int x;
		int y;
	} size;
	union pf_pixel *pixels;
};

/* Row-major matrix of doubles: value (x, y) is values[y*size.x+x]. */
struct pf_dbl_matrix {
	struct {
		int x;
		int y;
	} size;
	double *values;
};

/* ------------------------------------------------------------------------ */
/* Drop-in entry points: host bitmaps in, host bitmaps out.                  */
/* `out->pixels` is caller-allocated, same size as `in`; in/out may alias.   */
/* These are void like the reference's.  On a CUDA failure (no device, out   */
/* of memory, launch error) they print one line to stderr and record the     */
/* error (pf_b200_last_error); `out` is then undefined (untouched when the   */
/* arguments were refused, possibly partly written when a copy or a kernel   */
/* failed half way), and no work of the call is still in flight on return.   */
/* With the environment variable PF_B200_ABORT_ON_ERROR=1 they abort().      */
/* There is no CPU fallback.                                                 */
/* ------------------------------------------------------------------------ */

/* replaces upstream src/pillowfight/_ace.c: pf_ace().  Draws nb_samples
 * (x, y) sample coordinates from libc rand() (caller seeds with srand()),
 * then runs the ACE adjustment and scaling passes on the GPU.  nb_threads is
 * accepted for signature compatibility and ignored. */
void pf_apply_ace(const struct pf_bitmap *in, struct pf_bitmap *out,
		int nb_samples, double slope, double limit, int nb_threads);

/* replaces upstream src/pillowfight/_gaussian.c: pf_gaussian(). */
void pf_apply_gaussian(const struct pf_bitmap *in, struct pf_bitmap *out,
		double sigma, int nb_stddev);

/* replaces upstream src/pillowfight/_sobel.c: pf_sobel(). */
void pf_apply_sobel(const struct pf_bitmap *in, struct pf_bitmap *out);

/* replaces upstream src/pillowfight/_canny.c: pf_canny(). */
void pf_apply_canny(const struct pf_bitmap *in, struct pf_bitmap *out);

/* Upstream spellings of the same four functions (pillowfight.h). */
void pf_ace(const struct pf_bitmap *in, struct pf_bitmap *out,
		int nb_samples, double slope, double limit, int nb_threads);
void pf_gaussian(const struct pf_bitmap *in, struct pf_bitmap *out,
		double sigma, int nb_stddev);
void pf_sobel(const struct pf_bitmap *in, struct pf_bitmap *out);
void pf_canny(const struct pf_bitmap *in, struct pf_bitmap *out);

/* ------------------------------------------------------------------------ */
/* The pf_dbl_matrix level of the same path (upstream util.c / _gaussian.c / */
/* _sobel.c): host matrices in, newly allocated host matrices out, as in the */
/* reference -- the caller releases them with pf_dbl_matrix_free().  On a    */
/* CUDA failure the returned matrices have values == NULL and size 0 x 0.    */
/* ------------------------------------------------------------------------ */

struct pf_gradient_matrixes {
	struct pf_dbl_matrix g_x;
	struct pf_dbl_matrix g_y;
	struct pf_dbl_matrix intensity;   /* hypot(g_x, g_y) */
	struct pf_dbl_matrix direction;   /* atan2(g_y, g_x) */
};

/* replaces upstream util.c: pf_dbl_matrix_new / pf_dbl_matrix_free (host
 * memory, zero-initialised). */
struct pf_dbl_matrix pf_dbl_matrix_new(int x, int y);
void pf_dbl_matrix_free(struct pf_dbl_matrix *matrix);

/* The reference's 3x3 Sobel kernels (upstream _sobel.c). */
extern const struct pf_dbl_matrix g_pf_kernel_sobel_x;
extern const struct pf_dbl_matrix g_pf_kernel_sobel_y;

/* replaces upstream util.c: pf_dbl_matrix_convolution().  Kernels of up to
 * 64 taps. */
struct pf_dbl_matrix pf_dbl_matrix_convolution(const struct pf_dbl_matrix *img,
		const struct pf_dbl_matrix *kernel);

/* replaces upstream _gaussian.c: pf_gaussian_on_matrix(): horizontal then
 * vertical 1-D convolution with nb_stddev taps. */
struct pf_dbl_matrix pf_gaussian_on_matrix(const struct pf_dbl_matrix *in,
		double sigma, int nb_stddev);

/* replaces upstream _sobel.c: pf_sobel_on_matrix(): g_x, g_y by convolution
 * with the given kernels, optionally blurred (gaussian_stddev > 0), then
 * intensity and direction. */
struct pf_gradient_matrixes pf_sobel_on_matrix(const struct pf_dbl_matrix *in,
		const struct pf_dbl_matrix *kernel_x, const struct pf_dbl_matrix *kernel_y,
		double gaussian_sigma, int gaussian_stddev);

/* ------------------------------------------------------------------------ */
/* Status-returning variants of the same calls (0 = success, else PF_B200_E*) */
/* ------------------------------------------------------------------------ */

#define PF_B200_OK            0
#define PF_B200_EINVAL        1   /* bad argument (NULL, size <= 0, ...)     */
#define PF_B200_ENODEV        2   /* no usable CUDA device                   */
#define PF_B200_ECUDA         3   /* a CUDA runtime call or launch failed    */
#define PF_B200_ETOOBIG       4   /* size beyond a stated maximum            */

int pf_b200_last_error(void);                 /* of the calling thread */
const char *pf_b200_last_error_string(void);  /* of the calling thread */
int pf_b200_device_count(void);               /* 0 when CUDA is unusable */

/* ACE with the sample table given explicitly: sample_xy holds nb_samples
 * pairs (x0, y0, x1, y1, ...).  This is what pf_apply_ace calls after it has
 * drawn the pairs (upstream _ace.c: create_random_pairs + pf_ace body). */
int pf_b200_ace(const struct pf_bitmap *in, struct pf_bitmap *out,
		int nb_samples, const int *sample_xy, double slope, double limit);
int pf_b200_gaussian(const struct pf_bitmap *in, struct pf_bitmap *out,
		double sigma, int nb_stddev);
int pf_b200_sobel(const struct pf_bitmap *in, struct pf_bitmap *out);
int pf_b200_canny(const struct pf_bitmap *in, struct pf_bitmap *out);

/* An extension (no counterpart in the reference's interface; the result is
 * pf_b200_ace's): ACE of one page with its row bands on several GPUs of this
 * process -- BASELINE.json config 4 for a C caller that owns the GPUs itself,
 * without torch or NCCL.  devices: nb_devices distinct device ordinals, or NULL
 * for devices 0 .. nb_devices-1 (nb_devices <= 0: every visible device).  Each
 * GPU is sent its band over its own PCIe link, accumulates it as it arrives
 * and finds the exact extremes of its band; the 12 doubles of every GPU are
 * folded on the host (the one exchange the algorithm has) and each GPU
 * rescales and returns its band.  Needs pinned pages (pf_b200_host_alloc), out
 * of place, and bands of at least 256 rows; anything else is served by
 * pf_b200_ace on the first device.  The calling thread's current device is
 * restored on return. */
int pf_b200_ace_multi_gpu(const struct pf_bitmap *in, struct pf_bitmap *out,
		int nb_samples, const int *sample_xy, double slope, double limit,
		const int *devices, int nb_devices);

/* An extension beside the drop-in calls (no counterpart in the reference's
 * interface; its results are those of the calls it stands for): the whole step
 * on one page with ONE upload.  Any of the four outputs may be NULL:
 *   out_ace                = pf_b200_ace(in, nb_samples, sample_xy, slope, limit)
 *   out_gaussian           = pf_b200_gaussian(in, sigma, nb_stddev)
 *   out_sobel_of_gaussian  = pf_b200_sobel(of that blurred page)
 *   out_canny              = pf_b200_canny(in)
 * The four drop-in calls of such a step move the page to the GPU three times
 * and the blurred page back and forth: 8 page transfers over PCIe; this call
 * moves 1 plus one per result asked for.  The outputs must be distinct pages,
 * none of them the input.  Pinned pages (pf_b200_host_alloc) of 512 rows or
 * more take the overlapped path; anything else is served by the drop-in calls
 * one after the other. */
int pf_b200_page_step(const struct pf_bitmap *in, struct pf_bitmap *out_ace,
		struct pf_bitmap *out_gaussian, struct pf_bitmap *out_sobel_of_gaussian,
		struct pf_bitmap *out_canny, int nb_samples, const int *sample_xy,
		double slope, double limit, double sigma, int nb_stddev);

/* Draw nb_samples pairs the way pf_apply_ace does (x = rand() % width, then
 * y = rand() % height, per sample; upstream _ace.c: create_random_pairs). */
void pf_b200_ace_draw_samples(int nb_samples, int width, int height,
		int *sample_xy);

/* Pinned host memory for callers that want DMA-speed transfers through the
 * entry points above (pageable pointers work too, through a staging copy). */
void *pf_b200_host_alloc(size_t bytes);
void pf_b200_host_free(void *ptr);

/* ------------------------------------------------------------------------ */
/* Device-resident entry points: RGBA8 pages already in HBM.                 */
/* d_in / d_out are device pointers to width*height uint32 RGBA pixels on    */
/* the current device; `stream` is a cudaStream_t (NULL = default stream).   */
/* All work is enqueued on `stream`; nothing synchronises the host.          */
/* They are the bodies the host-pointer entry points run between their       */
/* copies (upstream: the same pf_gaussian/pf_sobel/pf_canny/pf_ace bodies).  */
/* ------------------------------------------------------------------------ */

int pf_b200_dev_gaussian(const uint32_t *d_in, uint32_t *d_out,
		int width, int height, double sigma, int nb_stddev, void *stream);

int pf_b200_dev_sobel(const uint32_t *d_in, uint32_t *d_out,
		int width, int height, void *stream);

/* Canny needs scratch for the hysteresis work list (the weak-edge candidates
 * and their union-find links, 8 bytes per pixel in the worst case):
 * pf_b200_canny_scratch_bytes(width, height) bytes, 256-byte aligned.
 * nb_pages pages of the same size, stored back to back, are processed by one
 * call (page batches, BASELINE.json config 5).  Scratch is per call, not per
 * page: with the minimum size pages go through one at a time; a scratch k
 * times larger lets k pages share each kernel launch. */
size_t pf_b200_canny_scratch_bytes(int width, int height);
int pf_b200_dev_canny(const uint32_t *d_in, uint32_t *d_out,
		int width, int height, int nb_pages,
		void *d_scratch, size_t scratch_bytes, void *stream);

/* Canny for callers that must not allocate.  (No counterpart in the
 * reference: it exists only because of how GPU memory is managed.)
 *
 * Some hosts cannot let a library call cudaMalloc on their behalf: a service
 * that gave the whole GPU to its own memory pool when it started, so a later
 * allocation would fail or eat into that pool; a latency-sensitive loop, where
 * the device-wide synchronisation cudaMalloc/cudaFree imply would stall every
 * stream; or a deployment that freezes its memory footprint after start-up.
 * pf_b200_dev_canny already leaves the scratch to the caller; this variant
 * needs no scratch argument at all.  The work list lives in an array compiled
 * into the library (32 MiB + 256 B of HBM on every GPU that loads the library,
 * held until the process exits), which bounds the page size:
 *
 *     width * height <= 4,194,304 pixels   (e.g. 2048 x 2048, or A4 at 200 dpi)
 *
 * Larger pages are refused with PF_B200_ETOOBIG and a message naming the
 * limit.  One page per call; d_in / d_out / stream as for pf_b200_dev_canny.
 * The call itself performs no cudaMalloc, cudaFree or other allocating call.
 *
 * The first call on each device looks up the array's device address (once
 * per device, with that device current) and keeps it for the life of the
 * process, indexed by device ordinal.  Make that first call during set-up,
 * before any stream capture is started; with CUDA's lazy module loading it is
 * also the moment the array comes into being on that GPU.  The remembered
 * address belongs to the device's primary context: after cudaDeviceReset(),
 * or with a second (driver-API) context current on the same ordinal, it no
 * longer points at the array, so neither is supported with this entry point.
 *
 * Every call on a device shares the one array.  Each call clears the part
 * that has to start at zero and overwrites the rest as it goes, so calls that
 * run one after another (same stream, or ordered by events) are safe; two
 * calls whose kernels can overlap on the GPU (different streams or host
 * threads without ordering) race on the array and give wrong edges. */
int pf_b200_dev_canny_noalloc(const uint32_t *d_in, uint32_t *d_out,
		int width, int height, void *stream);

/* Batched Gaussian / Sobel over nb_pages same-size pages stored back to back. */
int pf_b200_dev_gaussian_batch(const uint32_t *d_in, uint32_t *d_out,
		int width, int height, int nb_pages,
		double sigma, int nb_stddev, void *stream);
int pf_b200_dev_sobel_batch(const uint32_t *d_in, uint32_t *d_out,
		int width, int height, int nb_pages, void *stream);

/* ACE in device passes so that row bands can be split across GPUs
 * (BASELINE.json config 4).  upstream _ace.c: ace_thread_adjustment (passes
 * 1, 3 and the second half of 5), ace_thread_scaling (pass 5).
 *
 * How the result is made the reference's for every pixel: pass 1 computes the
 * scores in FP32 with a worst-case error bound; pass 3 recomputes, in the
 * reference's own arithmetic (IEEE sqrt and divide, sample order), the scores
 * that could be the page's extremes, so the extremes are exact; pass 5
 * rescales with them and recomputes the same way every pixel whose rescaled
 * value lies closer to an integer than the bound allows (a fraction of a per
 * cent of a page).
 *
 * d_minmax: pf_b200_ace_minmax_bytes(width, height) bytes on the device:
 *   doubles [0..2] minimum and [3..5] maximum of R, G, B over the pass-1
 *   scores, [6..8] and [9..11] the exact extremes of pass 3, then one entry
 *   per tile of the page (bookkeeping of the passes).  One per page and GPU.
 * d_scores: a score block of pf_b200_ace_scores_bytes(width, row_end-row_begin)
 *   bytes, 256-byte aligned, per band: three FP32 score planes plus the list
 *   of pixels pass 5 has to recompute (16 bytes per pixel and a header).
 * d_samples: nb_samples * 2 ints (x, y) on the device.
 *
 *   0. pf_b200_dev_ace_minmax_init: once per page and device.
 *   1. pf_b200_dev_ace_scores: rows [row_begin, row_end) of the page.
 *   2. (optional) all-reduce d_minmax[0..5] across GPUs: min over the first
 *      3, max over the last 3.  Pass 3 works from the extremes it finds there:
 *      a band's own are enough, the page's leave it fewer pixels to look at.
 *   3. pf_b200_dev_ace_refine, same arguments as pass 1.  Not optional.
 *   4. (multi-GPU only) all-reduce d_minmax[6..11]: min over [6..8], max over
 *      [9..11] -- the one exchange the algorithm needs.
 *   5. pf_b200_dev_ace_scale, same arguments plus the destination: the
 *      band's RGBA8 rows into d_out_band ((row_end-row_begin)*width pixels).
 */
size_t pf_b200_ace_minmax_bytes(int width, int height);
size_t pf_b200_ace_scores_bytes(int width, int nb_rows);
int pf_b200_dev_ace_minmax_init(double *d_minmax, int width, int height, void *stream);
int pf_b200_dev_ace_scores(const uint32_t *d_in, int width, int height,
		int row_begin, int row_end,
		const int *d_samples, int nb_samples, double slope, double limit,
		void *d_scores, double *d_minmax, void *stream);
int pf_b200_dev_ace_refine(const uint32_t *d_in, int width, int height,
		int row_begin, int row_end,
		const int *d_samples, int nb_samples, double slope, double limit,
		void *d_scores, double *d_minmax, void *stream);
int pf_b200_dev_ace_scale(const uint32_t *d_in, int width, int height,
		int row_begin, int row_end,
		const int *d_samples, int nb_samples, double slope, double limit,
		void *d_scores, double *d_minmax, uint32_t *d_out_band, void *stream);

/* The same three passes for a GPU that holds only its band of the page: d_band
 * points at row row_begin (rows [row_begin, row_end) back to back) and
 * d_sample_px holds the nb_samples sample pixels' values, gathered where the
 * whole page is (page[y_k * width + x_k]).  A row-band split then moves
 * height/N rows to each GPU instead of the whole page. */
int pf_b200_dev_ace_scores_band(const uint32_t *d_band, int width, int height,
		int row_begin, int row_end,
		const int *d_samples, const uint32_t *d_sample_px, int nb_samples,
		double slope, double limit,
		void *d_scores, double *d_minmax, void *stream);
int pf_b200_dev_ace_refine_band(const uint32_t *d_band, int width, int height,
		int row_begin, int row_end,
		const int *d_samples, const uint32_t *d_sample_px, int nb_samples,
		double slope, double limit,
		void *d_scores, double *d_minmax, void *stream);
int pf_b200_dev_ace_scale_band(const uint32_t *d_band, int width, int height,
		int row_begin, int row_end,
		const int *d_samples, const uint32_t *d_sample_px, int nb_samples,
		double slope, double limit,
		void *d_scores, double *d_minmax, uint32_t *d_out_band, void *stream);

/* What goes to such a GPU besides its rows, gathered where the whole page is,
 * as one message of 3 * nb_samples ints: the sample table (x0, y0, x1, ...)
 * followed by the nb_samples sample pixels (d_sample_px for the calls above is
 * d_message + 2 * nb_samples).  The sample coordinates must lie inside the
 * page. */
int pf_b200_dev_ace_gather_samples(const uint32_t *d_page, int width, int height,
		const int *d_samples, int nb_samples, int *d_message, void *stream);

/* Pass 5 with several destinations: the band's rows are stored to nb_outs
 * images at once (d_outs: HOST array of nb_outs device pointers, 1..16, each
 * pointing at the band's first row in that image).  The images may live on
 * other GPUs (peer-mapped memory): in a row-band split every GPU writes its
 * band straight into every GPU's result over NVLink and the all-gather
 * disappears.  d_band / d_sample_px as above; with d_sample_px == NULL d_band
 * must lie inside the whole page (the samples are then read from it). */
int pf_b200_dev_ace_scale_to(const uint32_t *d_band, int width, int height,
		int row_begin, int row_end,
		const int *d_samples, const uint32_t *d_sample_px, int nb_samples,
		double slope, double limit,
		void *d_scores, double *d_minmax,
		uint32_t *const *d_outs, int nb_outs, void *stream);
/* The same with ONE destination that is an NVLS multicast address (a
 * multicast object bound to the result image of every GPU): the rows are
 * stored with multimem.st and the NVSwitch delivers them to all GPUs,
 * the caller's included. */
int pf_b200_dev_ace_scale_multicast(const uint32_t *d_band, int width, int height,
		int row_begin, int row_end,
		const int *d_samples, const uint32_t *d_sample_px, int nb_samples,
		double slope, double limit,
		void *d_scores, double *d_minmax,
		uint32_t *d_multicast_band, void *stream);

/* The passes above for one whole page on one GPU, with one caller-owned
 * work area of pf_b200_ace_workspace_bytes(width, height) bytes (the score
 * block, 16 bytes per pixel, and the min/max block; 256-byte aligned).
 * upstream _ace.c: pf_ace on a page resident in HBM. */
size_t pf_b200_ace_workspace_bytes(int width, int height);
int pf_b200_dev_ace(const uint32_t *d_in, uint32_t *d_out, int width, int height,
		const int *d_samples, int nb_samples, double slope, double limit,
		void *d_workspace, size_t workspace_bytes, void *stream);

/* Measures the FP64 FMA rate of the current device (the ACE roofline
 * denominator): returns TFLOP/s (FMA = 2 flops), or a negative PF_B200_E*. */
double pf_b200_fp64_peak_tflops(void);
/* ACE's exactness rests on two measured constants of the hardware: the worst
 * relative error of the FP32 reciprocal square root the accumulation pass uses
 * (which = 0: every mantissa of two binades; the bound assumes <= 1.25e-7) and
 * of the FP64 one of the exact passes (which = 1: every argument it is ever
 * given, the integers below 2^31; the bound assumes <= 2^-48).  Returns the
 * worst error found on the current device, or a negative PF_B200_E*.
 * tests/test_gpu_parity.py holds every device the tests run on to both. */
double pf_b200_selftest_rsqrt_error(int which);
/* The same for FP32 (the pipe the ACE accumulation pass runs on). */
double pf_b200_fp32_peak_tflops(void);

#ifdef __cplusplus
}
#endif

#endif /* PILLOWFIGHT_B200_H */
