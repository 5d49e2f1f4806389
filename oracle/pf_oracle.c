/*
 * pf_oracle.c -- CPU restatement of libpillowfight's hot path, used ONLY as
 * the checker by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
 * legs.  It is test infrastructure: nothing in libpillowfight_b200/ may link,
 * load or call it.
 *
 * PARITY UNPINNED.  /root/reference holds only a README ("moved to GNOME
 * GitLab", README.md:1-3); the C sources, tests and golden images are not in
 * the mount, nothing can be compiled into oracle/_ref, and no upstream golden
 * vector exists to pin this file against.  Each function below restates the
 * published libpillowfight algorithm for the named upstream file
 * (src/pillowfight/<file>) as far as the task specification (BASELINE.json
 * north_star) and the public description of the algorithms determine it; the
 * choices the specification does not determine are listed as U1..U9 in
 * DESIGN.md section 0 and marked [U#] here.  Line numbers cannot be cited.
 *
 * Plain C99 + libm + pthreads; built by oracle/Makefile with
 * gcc -O2 -ffp-contract=off (no FMA contraction, plain IEEE double).
 */
#include <float.h>
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/pillowfight_b200.h"

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

#define PX(bm, col, row) ((bm)->pixels[(size_t)(row) * (bm)->size.x + (col)])
#define MV(m, col, row) ((m)->values[(size_t)(row) * (m)->size.x + (col)])

/* ---------------------------------------------------------------------- */
/* util.c: matrices, conversions, convolution                              */
/* ---------------------------------------------------------------------- */

struct pf_dbl_matrix pfo_matrix_new(int x, int y)
{
	struct pf_dbl_matrix m;
	m.size.x = x;
	m.size.y = y;
	m.values = calloc((size_t)x * y, sizeof(double));
	return m;
}

void pfo_matrix_free(struct pf_dbl_matrix *m)
{
	free(m->values);
	m->values = NULL;
}

/* util.c: pf_dbl_matrix_convolution.  True convolution (kernel flipped),
 * kernel centre at size/2, taps that fall outside the image are skipped
 * [U1], accumulation order kernel-x outer / kernel-y inner, ascending, from
 * 0.0 -- the order matters for the last bit and the GPU kernels follow it. */
struct pf_dbl_matrix pfo_convolution(const struct pf_dbl_matrix *img,
		const struct pf_dbl_matrix *kernel)
{
	struct pf_dbl_matrix out = pfo_matrix_new(img->size.x, img->size.y);
	const int kcx = kernel->size.x / 2, kcy = kernel->size.y / 2;
	int x, y, kx, ky;

	for (y = 0; y < img->size.y; y++) {
		for (x = 0; x < img->size.x; x++) {
			double val = 0.0;
			for (kx = 0; kx < kernel->size.x; kx++) {
				const int ix = x - kx + kcx;
				if (ix < 0 || ix >= img->size.x)
					continue;
				for (ky = 0; ky < kernel->size.y; ky++) {
					const int iy = y - ky + kcy;
					if (iy < 0 || iy >= img->size.y)
						continue;
					val += MV(img, ix, iy) * MV(kernel, kx, ky);
				}
			}
			MV(&out, x, y) = val;
		}
	}
	return out;
}

/* util.h: PF_GET_PIXEL_GRAYSCALE -- integer mean of R, G, B [U2]. */
static inline int gray_of(union pf_pixel p)
{
	return ((int)p.color.r + (int)p.color.g + (int)p.color.b) / 3;
}

/* util.c: pf_rgb_bitmap_to_grayscale_dbl_matrix */
void pfo_bitmap_to_gray_matrix(const struct pf_bitmap *in,
		struct pf_dbl_matrix *out)
{
	int x, y;
	for (y = 0; y < in->size.y; y++)
		for (x = 0; x < in->size.x; x++)
			MV(out, x, y) = (double)gray_of(PX(in, x, y));
}

/* double -> uint8 store rule shared by every matrix -> bitmap conversion:
 * clamp to [0, 255], then C truncation [U3]. */
static inline uint8_t store_u8(double v)
{
	if (!(v > 0.0))
		return 0; /* also NaN */
	if (v > 255.0)
		v = 255.0;
	return (uint8_t)(int)v;
}

/* util.c: pf_grayscale_dbl_matrix_to_rgb_bitmap */
void pfo_gray_matrix_to_bitmap(const struct pf_dbl_matrix *in,
		struct pf_bitmap *out)
{
	int x, y;
	for (y = 0; y < in->size.y; y++) {
		for (x = 0; x < in->size.x; x++) {
			const uint8_t v = store_u8(MV(in, x, y));
			union pf_pixel p;
			p.color.r = v;
			p.color.g = v;
			p.color.b = v;
			p.color.a = 0xFF;
			PX(out, x, y) = p;
		}
	}
}

/* util.c: pf_bitmap_channel_to_dbl_matrix */
void pfo_channel_to_matrix(const struct pf_bitmap *in,
		struct pf_dbl_matrix *out, int channel)
{
	int x, y;
	for (y = 0; y < in->size.y; y++)
		for (x = 0; x < in->size.x; x++)
			MV(out, x, y) = (double)PX(in, x, y).channels[channel];
}

/* util.c: pf_matrix_to_rgb_bitmap (one channel; alpha forced opaque) */
void pfo_matrix_to_channel(const struct pf_dbl_matrix *in,
		struct pf_bitmap *out, int channel)
{
	int x, y;
	for (y = 0; y < in->size.y; y++) {
		for (x = 0; x < in->size.x; x++) {
			PX(out, x, y).channels[channel] = store_u8(MV(in, x, y));
			PX(out, x, y).color.a = 0xFF;
		}
	}
}

/* ---------------------------------------------------------------------- */
/* _gaussian.c                                                             */
/* ---------------------------------------------------------------------- */

/* _gaussian.c: generate_gaussian_1d_kernel.  nb_stddev is the tap count
 * [U4]; taps are the sampled normal density around index nb_stddev/2,
 * normalised to sum 1 in index order.  sigma == 0 derives sigma from the tap
 * count.  Exported so that the GPU host code is checked against the same
 * taps (it generates its own with the same libm calls). */
void pfo_gaussian_taps(double sigma, int nb_stddev, double *taps)
{
	double sum = 0.0;
	int i;

	if (sigma == 0.0)
		sigma = 0.3 * (((double)(nb_stddev - 1)) * 0.5 - 1.0) + 0.8;
	for (i = 0; i < nb_stddev; i++) {
		const double d = (double)(i - nb_stddev / 2);
		taps[i] = exp(-(d * d) / (2.0 * sigma * sigma))
			/ sqrt(2.0 * M_PI * sigma * sigma);
		sum += taps[i];
	}
	for (i = 0; i < nb_stddev; i++)
		taps[i] = taps[i] / sum;
}

/* _gaussian.c: pf_gaussian_on_matrix -- horizontal 1-D pass, then vertical. */
struct pf_dbl_matrix pfo_gaussian_on_matrix(const struct pf_dbl_matrix *in,
		double sigma, int nb_stddev)
{
	struct pf_dbl_matrix kh, kv, tmp, out;

	kh = pfo_matrix_new(nb_stddev, 1);
	pfo_gaussian_taps(sigma, nb_stddev, kh.values);
	kv = pfo_matrix_new(1, nb_stddev);
	memcpy(kv.values, kh.values, sizeof(double) * nb_stddev);

	tmp = pfo_convolution(in, &kh);
	out = pfo_convolution(&tmp, &kv);

	pfo_matrix_free(&tmp);
	pfo_matrix_free(&kh);
	pfo_matrix_free(&kv);
	return out;
}

/* _gaussian.c: pf_gaussian -- each of R, G, B blurred independently. */
void pfo_gaussian(const struct pf_bitmap *in, struct pf_bitmap *out,
		double sigma, int nb_stddev)
{
	int c;
	struct pf_bitmap dst;
	const size_t n = (size_t)in->size.x * in->size.y;

	dst.size = in->size;
	dst.pixels = malloc(n * sizeof(union pf_pixel));
	for (c = 0; c < PF_NB_RGB_COLORS; c++) {
		struct pf_dbl_matrix m = pfo_matrix_new(in->size.x, in->size.y);
		struct pf_dbl_matrix b;
		pfo_channel_to_matrix(in, &m, c);
		b = pfo_gaussian_on_matrix(&m, sigma, nb_stddev);
		pfo_matrix_to_channel(&b, &dst, c);
		pfo_matrix_free(&m);
		pfo_matrix_free(&b);
	}
	memcpy(out->pixels, dst.pixels, n * sizeof(union pf_pixel));
	free(dst.pixels);
}

/* ---------------------------------------------------------------------- */
/* _sobel.c                                                                */
/* ---------------------------------------------------------------------- */

static const double g_sobel_x[9] = {
	-1.0, 0.0, 1.0,
	-2.0, 0.0, 2.0,
	-1.0, 0.0, 1.0,
};
static const double g_sobel_y[9] = {
	-1.0, -2.0, -1.0,
	 0.0,  0.0,  0.0,
	 1.0,  2.0,  1.0,
};

struct pfo_gradients {
	struct pf_dbl_matrix g_x;
	struct pf_dbl_matrix g_y;
	struct pf_dbl_matrix intensity;
	struct pf_dbl_matrix direction;
};

/* _sobel.c: pf_sobel_on_matrix (without the optional gradient blur):
 * g_x, g_y by convolution with the 3x3 kernels, intensity = hypot(g_x, g_y),
 * direction = atan2(g_y, g_x) [U5]. */
struct pfo_gradients pfo_sobel_on_matrix(const struct pf_dbl_matrix *in)
{
	struct pfo_gradients g;
	struct pf_dbl_matrix kx, ky;
	int x, y;

	kx.size.x = 3; kx.size.y = 3; kx.values = (double *)g_sobel_x;
	ky.size.x = 3; ky.size.y = 3; ky.values = (double *)g_sobel_y;
	g.g_x = pfo_convolution(in, &kx);
	g.g_y = pfo_convolution(in, &ky);
	g.intensity = pfo_matrix_new(in->size.x, in->size.y);
	g.direction = pfo_matrix_new(in->size.x, in->size.y);
	for (y = 0; y < in->size.y; y++) {
		for (x = 0; x < in->size.x; x++) {
			const double vx = MV(&g.g_x, x, y), vy = MV(&g.g_y, x, y);
			MV(&g.intensity, x, y) = hypot(vx, vy);
			MV(&g.direction, x, y) = atan2(vy, vx);
		}
	}
	return g;
}

static void gradients_free(struct pfo_gradients *g)
{
	pfo_matrix_free(&g->g_x);
	pfo_matrix_free(&g->g_y);
	pfo_matrix_free(&g->intensity);
	pfo_matrix_free(&g->direction);
}

/* _sobel.c: pf_sobel -- gradient intensity of the grayscale page, clamped. */
void pfo_sobel(const struct pf_bitmap *in, struct pf_bitmap *out)
{
	struct pf_dbl_matrix gray = pfo_matrix_new(in->size.x, in->size.y);
	struct pfo_gradients g;

	pfo_bitmap_to_gray_matrix(in, &gray);
	g = pfo_sobel_on_matrix(&gray);
	pfo_gray_matrix_to_bitmap(&g.intensity, out);
	gradients_free(&g);
	pfo_matrix_free(&gray);
}

/* ---------------------------------------------------------------------- */
/* _canny.c                                                                */
/* ---------------------------------------------------------------------- */

#define PFO_CANNY_LOW (0.686 * 255.0)   /* [U6] */
#define PFO_CANNY_HIGH (1.372 * 255.0)  /* [U6] */

void pfo_canny_thresholds(double *low, double *high)
{
	*low = PFO_CANNY_LOW;
	*high = PFO_CANNY_HIGH;
}

static inline double mat_or_zero(const struct pf_dbl_matrix *m, int x, int y)
{
	if (x < 0 || y < 0 || x >= m->size.x || y >= m->size.y)
		return 0.0;
	return MV(m, x, y);
}

/* _canny.c: non-maximum suppression.  The direction is quantised to the
 * nearest multiple of pi/4; a pixel survives when its intensity is not below
 * either neighbour along that direction (neighbours outside the page count
 * as 0); out of place [U7]. */
static struct pf_dbl_matrix non_maximum_suppression(
		const struct pf_dbl_matrix *intensity,
		const struct pf_dbl_matrix *direction)
{
	static const int off[4][2] = { {1, 0}, {1, 1}, {0, 1}, {-1, 1} };
	struct pf_dbl_matrix out = pfo_matrix_new(intensity->size.x,
			intensity->size.y);
	int x, y;

	for (y = 0; y < intensity->size.y; y++) {
		for (x = 0; x < intensity->size.x; x++) {
			const double v = MV(intensity, x, y);
			const long q = lround(MV(direction, x, y) * 4.0 / M_PI);
			const int s = (int)((q + 8) % 4);
			const double a = mat_or_zero(intensity,
					x + off[s][0], y + off[s][1]);
			const double b = mat_or_zero(intensity,
					x - off[s][0], y - off[s][1]);
			MV(&out, x, y) = (v < a || v < b) ? 0.0 : v;
		}
	}
	return out;
}

/* _canny.c: double threshold + hysteresis [U8].  strong: > HIGH; weak:
 * > LOW and not strong.  A weak pixel is kept when it is 8-connected to a
 * strong pixel through weak pixels.  The fixed point does not depend on the
 * visiting order.  Result: 255.0 for kept pixels, 0.0 otherwise. */
static void hysteresis(struct pf_dbl_matrix *m)
{
	const int w = m->size.x, h = m->size.y;
	const size_t n = (size_t)w * h;
	uint8_t *state = malloc(n); /* 0 none, 1 weak, 2 kept */
	int32_t *stack = malloc(n * sizeof(int32_t));
	size_t top = 0, i;

	for (i = 0; i < n; i++) {
		const double v = m->values[i];
		if (v > PFO_CANNY_HIGH) {
			state[i] = 2;
			stack[top++] = (int32_t)i;
		} else if (v > PFO_CANNY_LOW) {
			state[i] = 1;
		} else {
			state[i] = 0;
		}
	}
	while (top > 0) {
		const int32_t p = stack[--top];
		const int px = p % w, py = p / w;
		int dx, dy;
		for (dy = -1; dy <= 1; dy++) {
			for (dx = -1; dx <= 1; dx++) {
				const int qx = px + dx, qy = py + dy;
				size_t q;
				if (qx < 0 || qy < 0 || qx >= w || qy >= h)
					continue;
				q = (size_t)qy * w + qx;
				if (state[q] == 1) {
					state[q] = 2;
					stack[top++] = (int32_t)q;
				}
			}
		}
	}
	for (i = 0; i < n; i++)
		m->values[i] = (state[i] == 2) ? 255.0 : 0.0;
	free(stack);
	free(state);
}

/* _canny.c: pf_canny_on_matrix */
struct pf_dbl_matrix pfo_canny_on_matrix(const struct pf_dbl_matrix *in)
{
	struct pfo_gradients g = pfo_sobel_on_matrix(in);
	struct pf_dbl_matrix out = non_maximum_suppression(&g.intensity,
			&g.direction);
	gradients_free(&g);
	hysteresis(&out);
	return out;
}

/* _canny.c: pf_canny */
void pfo_canny(const struct pf_bitmap *in, struct pf_bitmap *out)
{
	struct pf_dbl_matrix gray = pfo_matrix_new(in->size.x, in->size.y);
	struct pf_dbl_matrix edges;

	pfo_bitmap_to_gray_matrix(in, &gray);
	edges = pfo_canny_on_matrix(&gray);
	pfo_gray_matrix_to_bitmap(&edges, out);
	pfo_matrix_free(&edges);
	pfo_matrix_free(&gray);
}

/* ---------------------------------------------------------------------- */
/* _ace.c                                                                  */
/* ---------------------------------------------------------------------- */

/* _ace.c: create_random_pairs -- libc rand(), x then y per sample [U9]. */
void pfo_ace_draw_samples(int nb_samples, int width, int height, int *xy)
{
	int i;
	for (i = 0; i < nb_samples; i++) {
		xy[2 * i] = rand() % width;
		xy[2 * i + 1] = rand() % height;
	}
}

struct ace_job {
	const struct pf_bitmap *in;
	struct pf_bitmap *out;
	const int *xy;
	int nb_samples;
	double slope, limit;
	int row_begin, row_end;
	double *scores; /* 3 planes of width*height */
	double mn[3], mx[3];
};

/* _ace.c: calc_saturation */
static inline double saturation(double diff, double slope, double limit)
{
	double r = diff * slope;
	if (r > limit)
		r = limit;
	else if (r < -limit)
		r = -limit;
	return r;
}

/* _ace.c: ace_thread_adjustment -- the hot loop.  For every pixel, every
 * sample farther than height/5 contributes saturation(pixel - sample)/dist
 * per channel and limit/dist to the shared denominator. */
static void *ace_adjust_rows(void *arg)
{
	struct ace_job *j = arg;
	const int w = j->in->size.x, h = j->in->size.y;
	const size_t plane = (size_t)w * h;
	const int too_close = h / 5;
	int x, y, k, c;

	for (c = 0; c < 3; c++) {
		j->mn[c] = DBL_MAX;
		j->mx[c] = -DBL_MAX;
	}
	for (y = j->row_begin; y < j->row_end; y++) {
		for (x = 0; x < w; x++) {
			const union pf_pixel p = PX(j->in, x, y);
			double acc[3] = { 0.0, 0.0, 0.0 };
			double denom = 0.0;
			for (k = 0; k < j->nb_samples; k++) {
				const int sx = j->xy[2 * k], sy = j->xy[2 * k + 1];
				const double ddx = (double)(x - sx);
				const double ddy = (double)(y - sy);
				const double dist = sqrt(ddx * ddx + ddy * ddy);
				union pf_pixel s;
				if (dist < (double)too_close)
					continue;
				s = PX(j->in, sx, sy);
				for (c = 0; c < 3; c++)
					acc[c] += saturation(
						(double)((int)p.channels[c] - (int)s.channels[c]),
						j->slope, j->limit) / dist;
				denom += j->limit / dist;
			}
			for (c = 0; c < 3; c++) {
				const double r = acc[c] / denom;
				j->scores[c * plane + (size_t)y * w + x] = r;
				if (r > j->mx[c])
					j->mx[c] = r;
				if (r < j->mn[c])
					j->mn[c] = r;
			}
		}
	}
	return NULL;
}

/* _ace.c: ace_thread_scaling -- linear rescale of the scores to 0..255. */
static void *ace_scale_rows(void *arg)
{
	struct ace_job *j = arg;
	const int w = j->in->size.x, h = j->in->size.y;
	const size_t plane = (size_t)w * h;
	int x, y, c;

	for (y = j->row_begin; y < j->row_end; y++) {
		for (x = 0; x < w; x++) {
			union pf_pixel p;
			for (c = 0; c < 3; c++) {
				const double r = j->scores[c * plane + (size_t)y * w + x];
				p.channels[c] = store_u8(255.0 * (r - j->mn[c])
						/ (j->mx[c] - j->mn[c]));
			}
			p.color.a = 0xFF;
			PX(j->out, x, y) = p;
		}
	}
	return NULL;
}

/* _ace.c: pf_ace body with the sample table given.  Rows are split into
 * nb_threads bands; the result does not depend on the split. */
void pfo_ace_with_samples(const struct pf_bitmap *in, struct pf_bitmap *out,
		int nb_samples, const int *xy, double slope, double limit,
		int nb_threads)
{
	const int h = in->size.y;
	const size_t plane = (size_t)in->size.x * h;
	struct ace_job *jobs;
	pthread_t *tids;
	struct pf_bitmap dst;
	double *scores = malloc(3 * plane * sizeof(double));
	double mn[3] = { DBL_MAX, DBL_MAX, DBL_MAX };
	double mx[3] = { -DBL_MAX, -DBL_MAX, -DBL_MAX };
	int t, c;

	if (nb_threads < 1)
		nb_threads = 1;
	if (nb_threads > h)
		nb_threads = h;
	jobs = calloc(nb_threads, sizeof(*jobs));
	tids = calloc(nb_threads, sizeof(*tids));
	dst.size = in->size;
	dst.pixels = malloc(plane * sizeof(union pf_pixel));

	for (t = 0; t < nb_threads; t++) {
		jobs[t].in = in;
		jobs[t].out = &dst;
		jobs[t].xy = xy;
		jobs[t].nb_samples = nb_samples;
		jobs[t].slope = slope;
		jobs[t].limit = limit;
		jobs[t].row_begin = (int)((long)h * t / nb_threads);
		jobs[t].row_end = (int)((long)h * (t + 1) / nb_threads);
		jobs[t].scores = scores;
		pthread_create(&tids[t], NULL, ace_adjust_rows, &jobs[t]);
	}
	for (t = 0; t < nb_threads; t++) {
		pthread_join(tids[t], NULL);
		for (c = 0; c < 3; c++) {
			if (jobs[t].mn[c] < mn[c])
				mn[c] = jobs[t].mn[c];
			if (jobs[t].mx[c] > mx[c])
				mx[c] = jobs[t].mx[c];
		}
	}
	for (t = 0; t < nb_threads; t++) {
		for (c = 0; c < 3; c++) {
			jobs[t].mn[c] = mn[c];
			jobs[t].mx[c] = mx[c];
		}
		pthread_create(&tids[t], NULL, ace_scale_rows, &jobs[t]);
	}
	for (t = 0; t < nb_threads; t++)
		pthread_join(tids[t], NULL);

	memcpy(out->pixels, dst.pixels, plane * sizeof(union pf_pixel));
	free(dst.pixels);
	free(scores);
	free(tids);
	free(jobs);
}

/* _ace.c: pf_ace */
void pfo_ace(const struct pf_bitmap *in, struct pf_bitmap *out,
		int nb_samples, double slope, double limit, int nb_threads)
{
	int *xy = malloc(sizeof(int) * 2 * (size_t)nb_samples);
	pfo_ace_draw_samples(nb_samples, in->size.x, in->size.y, xy);
	pfo_ace_with_samples(in, out, nb_samples, xy, slope, limit, nb_threads);
	free(xy);
}

/* Bounded-sample helper for bench.py's CPU baseline: the adjustment pass on
 * rows [row_begin, row_end) only, nb_threads bands, returning nothing (the
 * work is the measurement).  Same loop as pfo_ace_with_samples. */
void pfo_ace_adjust_band(const struct pf_bitmap *in, int nb_samples,
		const int *xy, double slope, double limit,
		int row_begin, int row_end, int nb_threads, double *scores)
{
	struct ace_job *jobs;
	pthread_t *tids;
	const int rows = row_end - row_begin;
	int t;

	if (nb_threads < 1)
		nb_threads = 1;
	if (nb_threads > rows)
		nb_threads = rows;
	jobs = calloc(nb_threads, sizeof(*jobs));
	tids = calloc(nb_threads, sizeof(*tids));
	for (t = 0; t < nb_threads; t++) {
		jobs[t].in = in;
		jobs[t].xy = xy;
		jobs[t].nb_samples = nb_samples;
		jobs[t].slope = slope;
		jobs[t].limit = limit;
		jobs[t].row_begin = row_begin + (int)((long)rows * t / nb_threads);
		jobs[t].row_end = row_begin + (int)((long)rows * (t + 1) / nb_threads);
		jobs[t].scores = scores;
		pthread_create(&tids[t], NULL, ace_adjust_rows, &jobs[t]);
	}
	for (t = 0; t < nb_threads; t++)
		pthread_join(tids[t], NULL);
	free(tids);
	free(jobs);
}
