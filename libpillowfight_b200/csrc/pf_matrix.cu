// pf_dbl_matrix-level entry points on the GPU: convolution, Gaussian, Sobel.
//
// Reference semantics (upstream src/pillowfight/util.c pf_dbl_matrix_convolution,
// _gaussian.c pf_gaussian_on_matrix, _sobel.c pf_sobel_on_matrix -- not in the
// mount).  The inputs are arbitrary doubles, so unlike the
// bitmap kernels nothing can be reformulated: every output element is the
// reference's sum, term by term in the reference's order (kernel-x outer,
// kernel-y inner, taps outside the matrix skipped), with separate multiply and
// add roundings.  One thread per element; the matrices are streamed once from
// HBM (neighbouring threads share their reads through L1/L2).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "pf_common.cuh"

namespace pfb {

namespace {

constexpr int kMaxKernel = 64;           // taps per axis

struct ConvKernel {
	int nx, ny;
	double v[kMaxKernel];                // nx * ny <= 64, row-major (x fastest)
};

__global__ void __launch_bounds__(256)
matrix_convolution_kernel(const double *__restrict__ img, double *__restrict__ out,
		int w, int h, const ConvKernel k)
{
	const int x = blockIdx.x * blockDim.x + threadIdx.x;
	if (x >= w)
		return;
	const int cx = k.nx / 2, cy = k.ny / 2;
	// grid.y is capped at 65535: taller matrices wrap around
	for (int y = blockIdx.y; y < h; y += gridDim.y) {
		double val = 0.0;
		for (int kx = 0; kx < k.nx; kx++) {
			const int ix = x - kx + cx;
			if (ix < 0 || ix >= w)
				continue;
			for (int ky = 0; ky < k.ny; ky++) {
				const int iy = y - ky + cy;
				if (iy < 0 || iy >= h)
					continue;
				val = __dadd_rn(val, __dmul_rn(__ldg(img + (size_t)iy * w + ix), k.v[ky * k.nx + kx]));
			}
		}
		out[(size_t)y * w + x] = val;
	}
}

// intensity = hypot(g_x, g_y), direction = atan2(g_y, g_x)
__global__ void __launch_bounds__(256)
gradient_polar_kernel(const double *__restrict__ gx, const double *__restrict__ gy,
		double *__restrict__ intensity, double *__restrict__ direction, size_t n)
{
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
			i += (size_t)gridDim.x * blockDim.x) {
		const double a = gx[i], b = gy[i];
		intensity[i] = hypot(a, b);
		direction[i] = atan2(b, a);
	}
}

}  // namespace

int launch_matrix_convolution(const double *img, double *out, int w, int h,
		const double *kernel, int nx, int ny, cudaStream_t st)
{
	if (nx < 1 || ny < 1 || nx * ny > kMaxKernel)
		return fail(PF_B200_EINVAL, "convolution kernel %d x %d: at most %d taps", nx, ny, kMaxKernel);
	ConvKernel k;
	k.nx = nx;
	k.ny = ny;
	for (int i = 0; i < nx * ny; i++)
		k.v[i] = kernel[i];
	dim3 grid((w + 255) / 256, h < 65535 ? h : 65535);
	matrix_convolution_kernel<<<grid, 256, 0, st>>>(img, out, w, h, k);
	PFB_CUDA(cudaGetLastError());
	return PF_B200_OK;
}

int launch_gradient_polar(const double *gx, const double *gy, double *intensity,
		double *direction, size_t n, cudaStream_t st)
{
	const size_t want = (n + 255) / 256;
	const int blocks = (int)(want < (size_t)sm_count() * 8 ? want : (size_t)sm_count() * 8);
	gradient_polar_kernel<<<blocks > 0 ? blocks : 1, 256, 0, st>>>(gx, gy, intensity, direction, n);
	PFB_CUDA(cudaGetLastError());
	return PF_B200_OK;
}

}  // namespace pfb

// ---------------------------------------------------------------------------
// Host side: per-thread stream and grow-only device matrices.
// ---------------------------------------------------------------------------

namespace pfb {

namespace {

struct MatrixContext {
	int dev = -1;
	cudaStream_t st = nullptr;
	double *buf[6] = {};
	size_t cap[6] = {};

	~MatrixContext()
	{
		if (dev < 0)
			return;
		cudaSetDevice(dev);
		for (double *b : buf)
			if (b)
				cudaFree(b);
		if (st)
			cudaStreamDestroy(st);
	}
};

thread_local MatrixContext t_mctx;

int mctx_init(MatrixContext &c)
{
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0)
		return fail(PF_B200_ENODEV, "no CUDA device is usable (%s); this build has no CPU path",
				cudaGetErrorString(cudaGetLastError()));
	int dev = 0;
	PFB_CUDA(cudaGetDevice(&dev));
	if (c.st && c.dev == dev)
		return PF_B200_OK;
	if (c.dev >= 0 && c.dev != dev)
		return fail(PF_B200_EINVAL, "thread first used device %d, now on %d", c.dev, dev);
	c.dev = dev;
	PFB_CUDA(cudaStreamCreateWithFlags(&c.st, cudaStreamNonBlocking));
	return PF_B200_OK;
}

int mbuf(MatrixContext &c, int i, size_t n)
{
	if (c.cap[i] >= n)
		return PF_B200_OK;
	if (c.buf[i])
		cudaFree(c.buf[i]);
	c.buf[i] = nullptr;
	c.cap[i] = 0;
	PFB_CUDA(cudaMalloc(&c.buf[i], n * sizeof(double)));
	c.cap[i] = n;
	return PF_B200_OK;
}

int check_matrix(const struct pf_dbl_matrix *m, const char *what)
{
	if (!m || !m->values || m->size.x <= 0 || m->size.y <= 0)
		return fail(PF_B200_EINVAL, "%s: NULL or empty matrix", what);
	return PF_B200_OK;
}

struct pf_dbl_matrix empty_matrix()
{
	struct pf_dbl_matrix m;
	m.size.x = 0;
	m.size.y = 0;
	m.values = nullptr;
	return m;
}

void report_matrix(const char *who, int rc)
{
	if (rc == PF_B200_OK)
		return;
	fprintf(stderr, "libpillowfight_b200: %s failed (%d): %s\n", who, rc,
			pf_b200_last_error_string());
	const char *e = getenv("PF_B200_ABORT_ON_ERROR");
	if (e && e[0] == '1')
		abort();
}

// device Gaussian: buf[src] -> buf[dst] through buf[tmp]; taps as in make_taps
int dev_gaussian(MatrixContext &c, int src, int tmp, int dst, int w, int h,
		const double *taps, int n)
{
	int rc = launch_matrix_convolution(c.buf[src], c.buf[tmp], w, h, taps, n, 1, c.st);
	if (rc)
		return rc;
	return launch_matrix_convolution(c.buf[tmp], c.buf[dst], w, h, taps, 1, n, c.st);
}

}  // namespace

}  // namespace pfb

// ---------------------------------------------------------------------------
// extern "C" (see include/pillowfight_b200.h)
// ---------------------------------------------------------------------------

using namespace pfb;

static const double k_sobel_x[9] = { -1.0, 0.0, 1.0, -2.0, 0.0, 2.0, -1.0, 0.0, 1.0 };
static const double k_sobel_y[9] = { -1.0, -2.0, -1.0, 0.0, 0.0, 0.0, 1.0, 2.0, 1.0 };

extern "C" {

const struct pf_dbl_matrix g_pf_kernel_sobel_x = { { 3, 3 }, const_cast<double *>(k_sobel_x) };
const struct pf_dbl_matrix g_pf_kernel_sobel_y = { { 3, 3 }, const_cast<double *>(k_sobel_y) };

struct pf_dbl_matrix pf_dbl_matrix_new(int x, int y)
{
	struct pf_dbl_matrix m;
	m.size.x = x;
	m.size.y = y;
	m.values = static_cast<double *>(calloc((size_t)(x > 0 ? x : 0) * (size_t)(y > 0 ? y : 0) + 1,
			sizeof(double)));
	return m;
}

void pf_dbl_matrix_free(struct pf_dbl_matrix *matrix)
{
	if (!matrix)
		return;
	free(matrix->values);
	matrix->values = nullptr;
}

static int convolution_impl(const struct pf_dbl_matrix *img, const struct pf_dbl_matrix *kernel,
		struct pf_dbl_matrix *out)
{
	clear_error();
	int rc = check_matrix(img, "convolution");
	if (rc || (rc = check_matrix(kernel, "convolution kernel")))
		return rc;
	MatrixContext &c = t_mctx;
	if ((rc = mctx_init(c)))
		return rc;
	const int w = img->size.x, h = img->size.y;
	const size_t n = (size_t)w * h;
	if ((rc = mbuf(c, 0, n)) || (rc = mbuf(c, 1, n)))
		return rc;
	PFB_CUDA(cudaMemcpyAsync(c.buf[0], img->values, n * sizeof(double), cudaMemcpyHostToDevice, c.st));
	if ((rc = launch_matrix_convolution(c.buf[0], c.buf[1], w, h, kernel->values,
			kernel->size.x, kernel->size.y, c.st)))
		return rc;
	*out = pf_dbl_matrix_new(w, h);
	if (!out->values)
		return fail(PF_B200_EINVAL, "out of host memory");
	PFB_CUDA(cudaMemcpyAsync(out->values, c.buf[1], n * sizeof(double), cudaMemcpyDeviceToHost, c.st));
	PFB_CUDA(cudaStreamSynchronize(c.st));
	return PF_B200_OK;
}

struct pf_dbl_matrix pf_dbl_matrix_convolution(const struct pf_dbl_matrix *img,
		const struct pf_dbl_matrix *kernel)
{
	struct pf_dbl_matrix out = empty_matrix();
	const int rc = convolution_impl(img, kernel, &out);
	if (rc) {
		pf_dbl_matrix_free(&out);
		out = empty_matrix();
	}
	report_matrix("pf_dbl_matrix_convolution", rc);
	return out;
}

}  // extern "C"

extern "C" {

static int gaussian_on_matrix_impl(const struct pf_dbl_matrix *in, double sigma, int nb_stddev,
		struct pf_dbl_matrix *out)
{
	clear_error();
	int rc = check_matrix(in, "pf_gaussian_on_matrix");
	if (rc)
		return rc;
	Taps taps;
	if ((rc = make_taps(sigma, nb_stddev, taps)))
		return rc;
	MatrixContext &c = t_mctx;
	if ((rc = mctx_init(c)))
		return rc;
	const int w = in->size.x, h = in->size.y;
	const size_t n = (size_t)w * h;
	if ((rc = mbuf(c, 0, n)) || (rc = mbuf(c, 1, n)) || (rc = mbuf(c, 2, n)))
		return rc;
	PFB_CUDA(cudaMemcpyAsync(c.buf[0], in->values, n * sizeof(double), cudaMemcpyHostToDevice, c.st));
	if ((rc = dev_gaussian(c, 0, 1, 2, w, h, taps.t, taps.n)))
		return rc;
	*out = pf_dbl_matrix_new(w, h);
	if (!out->values)
		return fail(PF_B200_EINVAL, "out of host memory");
	PFB_CUDA(cudaMemcpyAsync(out->values, c.buf[2], n * sizeof(double), cudaMemcpyDeviceToHost, c.st));
	PFB_CUDA(cudaStreamSynchronize(c.st));
	return PF_B200_OK;
}

struct pf_dbl_matrix pf_gaussian_on_matrix(const struct pf_dbl_matrix *in,
		double sigma, int nb_stddev)
{
	struct pf_dbl_matrix out = empty_matrix();
	const int rc = gaussian_on_matrix_impl(in, sigma, nb_stddev, &out);
	if (rc) {
		pf_dbl_matrix_free(&out);
		out = empty_matrix();
	}
	report_matrix("pf_gaussian_on_matrix", rc);
	return out;
}

static int sobel_on_matrix_impl(const struct pf_dbl_matrix *in,
		const struct pf_dbl_matrix *kernel_x, const struct pf_dbl_matrix *kernel_y,
		double gaussian_sigma, int gaussian_stddev, struct pf_gradient_matrixes *out)
{
	clear_error();
	int rc = check_matrix(in, "pf_sobel_on_matrix");
	if (rc || (rc = check_matrix(kernel_x, "sobel kernel x"))
			|| (rc = check_matrix(kernel_y, "sobel kernel y")))
		return rc;
	Taps taps;
	if (gaussian_stddev > 0 && (rc = make_taps(gaussian_sigma, gaussian_stddev, taps)))
		return rc;
	MatrixContext &c = t_mctx;
	if ((rc = mctx_init(c)))
		return rc;
	const int w = in->size.x, h = in->size.y;
	const size_t n = (size_t)w * h;
	for (int i = 0; i < 6; i++)
		if ((rc = mbuf(c, i, n)))
			return rc;
	// 0: in, 1: g_x, 2: g_y, 3: intensity / blur scratch, 4: direction / blurred, 5: blurred
	PFB_CUDA(cudaMemcpyAsync(c.buf[0], in->values, n * sizeof(double), cudaMemcpyHostToDevice, c.st));
	if ((rc = launch_matrix_convolution(c.buf[0], c.buf[1], w, h, kernel_x->values,
			kernel_x->size.x, kernel_x->size.y, c.st))
			|| (rc = launch_matrix_convolution(c.buf[0], c.buf[2], w, h, kernel_y->values,
			kernel_y->size.x, kernel_y->size.y, c.st)))
		return rc;
	int gx = 1, gy = 2;
	if (gaussian_stddev > 0) {
		if ((rc = dev_gaussian(c, 1, 3, 4, w, h, taps.t, taps.n))
				|| (rc = dev_gaussian(c, 2, 3, 5, w, h, taps.t, taps.n)))
			return rc;
		gx = 4;
		gy = 5;
	}
	// intensity and direction go to two buffers that are free at this point
	const int bi = (gx == 1) ? 3 : 1, bd = (gx == 1) ? 4 : 2;
	if ((rc = launch_gradient_polar(c.buf[gx], c.buf[gy], c.buf[bi], c.buf[bd], n, c.st)))
		return rc;
	struct pf_dbl_matrix *dst[4] = { &out->g_x, &out->g_y, &out->intensity, &out->direction };
	const int src[4] = { gx, gy, bi, bd };
	for (int i = 0; i < 4; i++) {
		*dst[i] = pf_dbl_matrix_new(w, h);
		if (!dst[i]->values)
			return fail(PF_B200_EINVAL, "out of host memory");
		PFB_CUDA(cudaMemcpyAsync(dst[i]->values, c.buf[src[i]], n * sizeof(double),
				cudaMemcpyDeviceToHost, c.st));
	}
	PFB_CUDA(cudaStreamSynchronize(c.st));
	return PF_B200_OK;
}

struct pf_gradient_matrixes pf_sobel_on_matrix(const struct pf_dbl_matrix *in,
		const struct pf_dbl_matrix *kernel_x, const struct pf_dbl_matrix *kernel_y,
		double gaussian_sigma, int gaussian_stddev)
{
	struct pf_gradient_matrixes out;
	out.g_x = out.g_y = out.intensity = out.direction = empty_matrix();
	const int rc = sobel_on_matrix_impl(in, kernel_x, kernel_y, gaussian_sigma,
			gaussian_stddev, &out);
	if (rc) {
		pf_dbl_matrix_free(&out.g_x);
		pf_dbl_matrix_free(&out.g_y);
		pf_dbl_matrix_free(&out.intensity);
		pf_dbl_matrix_free(&out.direction);
		out.g_x = out.g_y = out.intensity = out.direction = empty_matrix();
	}
	report_matrix("pf_sobel_on_matrix", rc);
	return out;
}

}  // extern "C"
