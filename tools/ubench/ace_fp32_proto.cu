// Prototype of the FP32 ACE accumulation loop (round 2): times variants of the
// inner loop on an A4-shaped problem and reports the worst score error against
// a double-precision evaluation of the same sums on the host.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a ace_fp32_proto.cu -o ace_fp32_proto
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#ifndef RSQ_FTZ
#define RSQ_FTZ 1
#endif

constexpr int kBlock = 128;
constexpr int kChunk = 512;
constexpr float kScale = 0x1p-60f;      // coordinates carry 2^-60, squared distances 2^-120

struct Entry { float fsx; uint32_t bias[3]; };

template <int P>
struct Shared {
	Entry tbl[kChunk];
	float dy2[kChunk * P];
	int sy[kChunk];
};

__device__ __forceinline__ float rsqrt_f32(float x)
{
	float y;
#if RSQ_FTZ
	// every argument here is a normal number (>= 2^-120) or +inf, so ftz changes
	// nothing -- except that the compiler drops its subnormal-input wrapper
	// (FSETP + 2 predicated FMULs per rsqrt)
	asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
#else
	asm("rsqrt.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
#endif
	return y;
}

__device__ __forceinline__ unsigned long long pack2(float lo, float hi)
{
	unsigned long long r;
	asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
	return r;
}

__device__ __forceinline__ void unpack2(unsigned long long v, float &lo, float &hi)
{
	asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}

__device__ __forceinline__ unsigned long long ffma2(unsigned long long a, unsigned long long b, unsigned long long c)
{
	unsigned long long d;
	asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
	return d;
}

__device__ __forceinline__ unsigned long long fadd2(unsigned long long a, unsigned long long b)
{
	unsigned long long d;
	asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
	return d;
}

// MODE 0: scalar FFMA; MODE 1: FFMA2 over row pairs;
// MODE 2: like 1 without the rsqrt (weight = d2): what is left when MUFU is free;
// MODE 3: like 1 without the clamps (u = pixel value)
template <int P, int MODE, int FLUSH, int MINB>
__global__ void __launch_bounds__(kBlock, MINB)
proto_kernel(const uint32_t *__restrict__ in, const int2 *__restrict__ sxy, const uint32_t *__restrict__ spx,
		int w, int h, int n, int clamp_l, float *__restrict__ scores)
{
	extern __shared__ __align__(16) unsigned char raw[];
	Shared<P> &sh = *reinterpret_cast<Shared<P> *>(raw);
	const int x = blockIdx.x * kBlock + threadIdx.x;
	const int xc = min(x, w - 1);
	const int tile_rows = 16;
	const int ty0 = blockIdx.y * tile_rows, ty1 = min(h, ty0 + tile_rows);
	const int npad = (n + FLUSH - 1) / FLUSH * FLUSH;
	for (int k = threadIdx.x; k < npad; k += kBlock) {
		if (k < n) {
			const int2 s = sxy[k];
			const uint32_t px = spx[k];
			sh.tbl[k].fsx = (float)s.x * kScale;
			sh.sy[k] = s.y;
			for (int c = 0; c < 3; c++)
				sh.tbl[k].bias[c] = (uint32_t)(clamp_l - (int)((px >> (8 * c)) & 0xFF));
		} else {
			sh.tbl[k].fsx = 0.f;
			sh.sy[k] = 1 << 30;       // dy2 = +inf below: weight 0
			for (int c = 0; c < 3; c++)
				sh.tbl[k].bias[c] = 0;
		}
	}
	const int two_l = 2 * clamp_l;
	const float top_val = __int_as_float(two_l);
	const float fx = (float)x * kScale;
	const size_t plane = (size_t)w * h;
	for (int y0 = ty0; y0 < ty1; y0 += P) {
		uint32_t pv[P][3];
#pragma unroll
		for (int r = 0; r < P; r++) {
			const uint32_t p = __ldg(in + (size_t)min(y0 + r, h - 1) * w + xc);
#pragma unroll
			for (int c = 0; c < 3; c++)
				pv[r][c] = (p >> (8 * c)) & 0xFF;
		}
		__syncthreads();
		for (int j = threadIdx.x; j < npad; j += kBlock) {
#pragma unroll
			for (int r = 0; r < P; r++) {
				const int sy = sh.sy[j];
				const float dy = (float)(y0 + r - sy) * kScale;
				sh.dy2[j * P + r] = (sy == (1 << 30)) ? __int_as_float(0x7f800000) : dy * dy;
			}
		}
		__syncthreads();
		if (MODE == 0) {
			float S[P][4], c4[P][4];
#pragma unroll
			for (int r = 0; r < P; r++)
#pragma unroll
				for (int c = 0; c < 4; c++)
					S[r][c] = 0.f;
			for (int j0 = 0; j0 < npad; j0 += FLUSH) {
#pragma unroll
				for (int r = 0; r < P; r++)
#pragma unroll
					for (int c = 0; c < 4; c++)
						c4[r][c] = 0.f;
#pragma unroll 4
				for (int j = j0; j < j0 + FLUSH; j++) {
					const Entry se = sh.tbl[j];
					const float dx = fx - se.fsx;
					const float dx2 = dx * dx;
#pragma unroll
					for (int r = 0; r < P; r++) {
						const float wv = rsqrt_f32(dx2 + sh.dy2[j * P + r]);
						c4[r][3] = fmaf(wv, top_val, c4[r][3]);
#pragma unroll
						for (int c = 0; c < 3; c++) {
							const int u = __viaddmin_s32_relu((int)pv[r][c], (int)se.bias[c], two_l);
							c4[r][c] = fmaf(wv, __int_as_float(u), c4[r][c]);
						}
					}
				}
#pragma unroll
				for (int r = 0; r < P; r++)
#pragma unroll
					for (int c = 0; c < 4; c++)
						S[r][c] += c4[r][c];
			}
#pragma unroll
			for (int r = 0; r < P; r++) {
				const int y = y0 + r;
				if (y < ty1 && x < w) {
					const float den = 0.5f * S[r][3];
#pragma unroll
					for (int c = 0; c < 3; c++)
						scores[c * plane + (size_t)y * w + x] = __fdiv_rn(S[r][c] - den, den);
				}
			}
		} else {
			unsigned long long S[P / 2][4], c4[P / 2][4];
#pragma unroll
			for (int r = 0; r < P / 2; r++)
#pragma unroll
				for (int c = 0; c < 4; c++)
					S[r][c] = 0ull;
			const unsigned long long top2 = pack2(top_val, top_val);
			for (int j0 = 0; j0 < npad; j0 += FLUSH) {
#pragma unroll
				for (int r = 0; r < P / 2; r++)
#pragma unroll
					for (int c = 0; c < 4; c++)
						c4[r][c] = 0ull;
#pragma unroll 4
				for (int j = j0; j < j0 + FLUSH; j++) {
					const Entry se = sh.tbl[j];
					const float dx = fx - se.fsx;
					const float dx2 = dx * dx;
					const unsigned long long dx22 = pack2(dx2, dx2);
#pragma unroll
					for (int r = 0; r < P / 2; r++) {
						const unsigned long long dy22 = *reinterpret_cast<const unsigned long long *>(&sh.dy2[j * P + 2 * r]);
						const unsigned long long d2 = fadd2(dx22, dy22);
						float da, db;
						unpack2(d2, da, db);
						const unsigned long long wv = (MODE == 2) ? d2 : pack2(rsqrt_f32(da), rsqrt_f32(db));
						c4[r][3] = ffma2(wv, top2, c4[r][3]);
#pragma unroll
						for (int c = 0; c < 3; c++) {
							int ua, ub;
							if (MODE == 3) {
								ua = (int)pv[2 * r][c];
								ub = (int)pv[2 * r + 1][c];
							} else {
								ua = __viaddmin_s32_relu((int)pv[2 * r][c], (int)se.bias[c], two_l);
								ub = __viaddmin_s32_relu((int)pv[2 * r + 1][c], (int)se.bias[c], two_l);
							}
							c4[r][c] = ffma2(wv, pack2(__int_as_float(ua), __int_as_float(ub)), c4[r][c]);
						}
					}
				}
#pragma unroll
				for (int r = 0; r < P / 2; r++)
#pragma unroll
					for (int c = 0; c < 4; c++)
						S[r][c] = fadd2(S[r][c], c4[r][c]);
			}
#pragma unroll
			for (int r = 0; r < P / 2; r++) {
				float t0, t1;
				unpack2(S[r][3], t0, t1);
#pragma unroll
				for (int c = 0; c < 3; c++) {
					float a0, a1;
					unpack2(S[r][c], a0, a1);
					const int y = y0 + 2 * r;
					if (y < ty1 && x < w)
						scores[c * plane + (size_t)y * w + x] = __fdiv_rn(a0 - 0.5f * t0, 0.5f * t0);
					if (y + 1 < ty1 && x < w)
						scores[c * plane + (size_t)(y + 1) * w + x] = __fdiv_rn(a1 - 0.5f * t1, 0.5f * t1);
				}
			}
		}
	}
}

__device__ __forceinline__ float fma_sat(float a, float b, float c)
{
	float d;
	asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
	return d;
}

struct EntryF { float sx; float b[3]; };

template <int P>
struct SharedF {
	EntryF tbl[kChunk];
	float dy2[kChunk * P];
	int sy[kChunk];
};

// The clamp on the FMA pipe: u' = sat(p * inv + (0.5 - s * inv)) = clamp((p - s + L) / 2L, 0, 1)
// with inv = (1 + 2^-22) / 2L, so that p - s >= L gives exactly 1, p - s <= -L exactly 0 and
// p == s exactly 0.5.  ACC2: accumulate with FFMA2 over row pairs, else scalar FFMA.
// NALU: that many of the three channels use the integer clamp (ALU pipe) instead.
template <int P, bool ACC2, int MINB, int UNROLL, int NALU>
__global__ void __launch_bounds__(kBlock, MINB)
sat_kernel(const uint32_t *__restrict__ in, const int2 *__restrict__ sxy, const uint32_t *__restrict__ spx,
		int w, int h, int n, int clamp_l, float *__restrict__ scores)
{
	extern __shared__ __align__(16) unsigned char raw[];
	SharedF<P> &sh = *reinterpret_cast<SharedF<P> *>(raw);
	const int x = blockIdx.x * kBlock + threadIdx.x;
	const int xc = min(x, w - 1);
	const int tile_rows = 16;
	const int ty0 = blockIdx.y * tile_rows, ty1 = min(h, ty0 + tile_rows);
	const int npad = (n + UNROLL - 1) / UNROLL * UNROLL;
	const float inv = (float)((1.0 + 0x1p-22) / (2.0 * clamp_l));
	for (int k = threadIdx.x; k < npad; k += kBlock) {
		if (k < n) {
			const int2 s = sxy[k];
			const uint32_t px = spx[k];
			sh.tbl[k].sx = (float)s.x;
			sh.sy[k] = s.y;
			for (int c = 0; c < 3; c++)
				sh.tbl[k].b[c] = (float)(0.5 - (double)((px >> (8 * c)) & 0xFF) * (double)inv);
		} else {
			sh.tbl[k].sx = 0.f;
			sh.sy[k] = 1 << 30;
			for (int c = 0; c < 3; c++)
				sh.tbl[k].b[c] = 0.f;
		}
	}
	const float fx = (float)x;
	const size_t plane = (size_t)w * h;
	for (int y0 = ty0; y0 < ty1; y0 += P) {
		float pv[P][3];
#pragma unroll
		for (int r = 0; r < P; r++) {
			const uint32_t p = __ldg(in + (size_t)min(y0 + r, h - 1) * w + xc);
#pragma unroll
			for (int c = 0; c < 3; c++)
				pv[r][c] = (float)((p >> (8 * c)) & 0xFF);
		}
		__syncthreads();
		for (int j = threadIdx.x; j < npad; j += kBlock) {
#pragma unroll
			for (int r = 0; r < P; r++) {
				const int sy = sh.sy[j];
				const float dy = (float)(y0 + r - sy);
				sh.dy2[j * P + r] = (sy == (1 << 30)) ? __int_as_float(0x7f800000) : dy * dy;
			}
		}
		__syncthreads();
		if (!ACC2) {
			float S[P][4], c4[P][4];
#pragma unroll
			for (int r = 0; r < P; r++)
#pragma unroll
				for (int c = 0; c < 4; c++)
					S[r][c] = 0.f;
			for (int j0 = 0; j0 < npad; j0 += 16) {
#pragma unroll
				for (int r = 0; r < P; r++)
#pragma unroll
					for (int c = 0; c < 4; c++)
						c4[r][c] = 0.f;
				const int j1 = min(npad, j0 + 16);
#pragma unroll UNROLL
				for (int j = j0; j < j1; j++) {
					const EntryF se = sh.tbl[j];
					const float dx = fx - se.sx;
					const float dx2 = dx * dx;
#pragma unroll
					for (int r = 0; r < P; r++) {
						const float wv = rsqrt_f32(dx2 + sh.dy2[j * P + r]);
						c4[r][3] += wv;
#pragma unroll
						for (int c = 0; c < 3; c++)
							c4[r][c] = fmaf(wv, fma_sat(pv[r][c], inv, se.b[c]), c4[r][c]);
					}
				}
#pragma unroll
				for (int r = 0; r < P; r++)
#pragma unroll
					for (int c = 0; c < 4; c++)
						S[r][c] += c4[r][c];
			}
#pragma unroll
			for (int r = 0; r < P; r++) {
				const int y = y0 + r;
				if (y < ty1 && x < w) {
					const float den = 0.5f * S[r][3];
#pragma unroll
					for (int c = 0; c < 3; c++)
						scores[c * plane + (size_t)y * w + x] = __fdiv_rn(S[r][c] - den, den);
				}
			}
		} else {
			unsigned long long S[P / 2][4], c4[P / 2][4];
#pragma unroll
			for (int r = 0; r < P / 2; r++)
#pragma unroll
				for (int c = 0; c < 4; c++)
					S[r][c] = 0ull;
			for (int j0 = 0; j0 < npad; j0 += 16) {
#pragma unroll
				for (int r = 0; r < P / 2; r++)
#pragma unroll
					for (int c = 0; c < 4; c++)
						c4[r][c] = 0ull;
				const int j1 = min(npad, j0 + 16);
#pragma unroll UNROLL
				for (int j = j0; j < j1; j++) {
					const EntryF se = sh.tbl[j];
					const float dx = fx - se.sx;
					const float dx2 = dx * dx;
					const unsigned long long dx22 = pack2(dx2, dx2);
#pragma unroll
					for (int r = 0; r < P / 2; r++) {
						const unsigned long long dy22 = *reinterpret_cast<const unsigned long long *>(&sh.dy2[j * P + 2 * r]);
						const unsigned long long d2 = fadd2(dx22, dy22);
						float da, db;
						unpack2(d2, da, db);
						const unsigned long long wv = pack2(rsqrt_f32(da), rsqrt_f32(db));
						c4[r][3] = fadd2(c4[r][3], wv);
#pragma unroll
						for (int c = 0; c < 3; c++) {
							if (c < NALU) {      // timing experiment only: integer clamp on the ALU pipe
								const int ua = __viaddmin_s32_relu(__float_as_int(pv[2 * r][c]), __float_as_int(se.b[c]), clamp_l);
								const int ub = __viaddmin_s32_relu(__float_as_int(pv[2 * r + 1][c]), __float_as_int(se.b[c]), clamp_l);
								c4[r][c] = ffma2(wv, pack2(__int_as_float(ua), __int_as_float(ub)), c4[r][c]);
							} else {
								c4[r][c] = ffma2(wv, pack2(fma_sat(pv[2 * r][c], inv, se.b[c]),
										fma_sat(pv[2 * r + 1][c], inv, se.b[c])), c4[r][c]);
							}
						}
					}
				}
#pragma unroll
				for (int r = 0; r < P / 2; r++)
#pragma unroll
					for (int c = 0; c < 4; c++)
						S[r][c] = fadd2(S[r][c], c4[r][c]);
			}
#pragma unroll
			for (int r = 0; r < P / 2; r++) {
				float t0, t1;
				unpack2(S[r][3], t0, t1);
#pragma unroll
				for (int c = 0; c < 3; c++) {
					float a0, a1;
					unpack2(S[r][c], a0, a1);
					const int y = y0 + 2 * r;
					if (y < ty1 && x < w)
						scores[c * plane + (size_t)y * w + x] = __fdiv_rn(a0 - 0.5f * t0, 0.5f * t0);
					if (y + 1 < ty1 && x < w)
						scores[c * plane + (size_t)(y + 1) * w + x] = __fdiv_rn(a1 - 0.5f * t1, 0.5f * t1);
				}
			}
		}
	}
}

// exhaustive accuracy of rsqrt.approx.f32 over two binades (odd and even exponent)
__global__ void rsqrt_err_kernel(float *worst)
{
	float mx = 0.f;
	for (uint32_t m = blockIdx.x * blockDim.x + threadIdx.x; m < (1u << 24); m += gridDim.x * blockDim.x) {
		const float x = __uint_as_float(0x03000000u + m);      // 2^-121 .. 2^-119
		const double ref = 1.0 / sqrt((double)x);
		const double e = fabs((double)rsqrt_f32(x) - ref) / ref;
		mx = fmaxf(mx, (float)e);
	}
	atomicMax(reinterpret_cast<int *>(worst), __float_as_int(mx));
}

template <typename K>
void run_k(const char *name, K kern, size_t smem, const uint32_t *d_in, const int2 *d_xy, const uint32_t *d_spx, int w, int h, int n,
		float *d_scores, const std::vector<uint32_t> &page, const std::vector<int2> &xy, bool check = true)
{
	cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
	dim3 grid((w + kBlock - 1) / kBlock, (h + 15) / 16);
	cudaFuncAttributes fa;
	cudaFuncGetAttributes(&fa, kern);
	int occ = 0;
	cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kBlock, smem);
	for (int i = 0; i < 2; i++)
		kern<<<grid, kBlock, smem>>>(d_in, d_xy, d_spx, w, h, n, 100, d_scores);
	cudaEvent_t e0, e1;
	cudaEventCreate(&e0);
	cudaEventCreate(&e1);
	cudaEventRecord(e0);
	for (int i = 0; i < 10; i++)
		kern<<<grid, kBlock, smem>>>(d_in, d_xy, d_spx, w, h, n, 100, d_scores);
	cudaEventRecord(e1);
	cudaEventSynchronize(e1);
	float ms;
	cudaEventElapsedTime(&ms, e0, e1);
	ms /= 10;
	cudaError_t err = cudaGetLastError();
	double worst = 0.0;
	if (check) {
		std::vector<float> sc(3 * (size_t)w * h);
		cudaMemcpy(sc.data(), d_scores, sc.size() * sizeof(float), cudaMemcpyDeviceToHost);
		srand(1);
		for (int t = 0; t < 2000; t++) {
			const int x = rand() % w, y = rand() % h;
			const uint32_t p = page[(size_t)y * w + x];
			for (int c = 0; c < 3; c++) {
				double num = 0.0, den = 0.0;
				for (int k = 0; k < n; k++) {
					const double dx = x - xy[k].x, dy = y - xy[k].y;
					const double d = sqrt(dx * dx + dy * dy);
					if (d == 0.0)
						continue;
					const uint32_t s = page[(size_t)xy[k].y * w + xy[k].x];
					int diff = (int)((p >> (8 * c)) & 0xFF) - (int)((s >> (8 * c)) & 0xFF);
					diff = diff > 100 ? 100 : (diff < -100 ? -100 : diff);
					num += diff / 100.0 / d;
					den += 1.0 / d;
				}
				const double e = fabs(num / den - (double)sc[c * (size_t)w * h + (size_t)y * w + x]);
				if (e > worst)
					worst = e;
			}
		}
	}
	printf("%-34s %7.4f ms  regs %3d  blocks/SM %d  max |score err| %.3e  (%s)\n", name, ms, fa.numRegs, occ,
			worst, cudaGetErrorString(err));
}

template <int P, int MODE, int FLUSH, int MINB>
void run(const char *name, const uint32_t *d_in, const int2 *d_xy, const uint32_t *d_spx, int w, int h, int n,
		float *d_scores, const std::vector<uint32_t> &page, const std::vector<int2> &xy, bool check = true)
{
	run_k(name, proto_kernel<P, MODE, FLUSH, MINB>, sizeof(Shared<P>), d_in, d_xy, d_spx, w, h, n, d_scores, page, xy, check);
}

template <int P, bool ACC2, int MINB, int UNROLL, int NALU = 0>
void run_sat(const char *name, const uint32_t *d_in, const int2 *d_xy, const uint32_t *d_spx, int w, int h, int n,
		float *d_scores, const std::vector<uint32_t> &page, const std::vector<int2> &xy, bool check = true)
{
	run_k(name, sat_kernel<P, ACC2, MINB, UNROLL, NALU>, sizeof(SharedF<P>), d_in, d_xy, d_spx, w, h, n, d_scores, page, xy, check);
}

int main(int argc, char **argv)
{
	const int w = 2480, h = 3508;
	const int n = argc > 1 ? atoi(argv[1]) : 100;
	const bool one = argc > 2;      // only the candidate kernel (for ncu)
	std::vector<uint32_t> page((size_t)w * h);
	srand(7);
	for (auto &p : page)
		p = 0xFF000000u | (rand() & 0xFFFFFF);
	std::vector<int2> xy(n);
	std::vector<uint32_t> spx(n);
	for (int k = 0; k < n; k++) {
		xy[k] = make_int2(rand() % w, rand() % h);
		spx[k] = page[(size_t)xy[k].y * w + xy[k].x];
	}
	uint32_t *d_in, *d_spx;
	int2 *d_xy;
	float *d_scores, *d_worst;
	cudaMalloc(&d_in, page.size() * 4);
	cudaMalloc(&d_spx, n * 4);
	cudaMalloc(&d_xy, n * 8);
	cudaMalloc(&d_scores, 3 * page.size() * 4);
	cudaMalloc(&d_worst, 4);
	cudaMemcpy(d_in, page.data(), page.size() * 4, cudaMemcpyHostToDevice);
	cudaMemcpy(d_spx, spx.data(), n * 4, cudaMemcpyHostToDevice);
	cudaMemcpy(d_xy, xy.data(), n * 8, cudaMemcpyHostToDevice);
	if (one) {
		if (atoi(argv[2]) == 1)
			run<8, 1, 16, 4>("P=8 FFMA2 flush16", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy, false);
		else if (atoi(argv[2]) == 2)
			run_sat<8, true, 4, 4>("sat P=8 FFMA2", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy, false);
		else
			run_sat<8, false, 4, 4>("sat P=8 scalar", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy, false);
		return 0;
	}
	cudaMemset(d_worst, 0, 4);
	rsqrt_err_kernel<<<148 * 8, 256>>>(d_worst);
	float worst;
	cudaMemcpy(&worst, d_worst, 4, cudaMemcpyDeviceToHost);
	printf("rsqrt.approx%s.f32, two binades exhaustively: max rel err %.4e = 2^%.2f\n", RSQ_FTZ ? ".ftz" : "", worst, log2(worst));
	printf("samples: %d (no skip circle in the prototype: every pair is accumulated)\n", n);
	run<8, 0, 16, 4>("P=8 scalar FFMA flush16", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy);
	run<8, 1, 16, 4>("P=8 FFMA2 flush16", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy);
	run<8, 1, 16, 3>("P=8 FFMA2 flush16 3 blocks", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy);
	run<4, 0, 16, 6>("P=4 scalar FFMA flush16", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy);
	run<4, 1, 16, 6>("P=4 FFMA2 flush16", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy);
	run<8, 2, 16, 4>("P=8 FFMA2 no rsqrt", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy, false);
	run<8, 3, 16, 4>("P=8 FFMA2 no clamp", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy, false);
	run<4, 2, 16, 6>("P=4 FFMA2 no rsqrt", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy, false);
	run_sat<8, false, 4, 4>("sat clamp P=8 scalar u4", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy);
	run_sat<8, true, 4, 4>("sat clamp P=8 FFMA2 u4", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy);
	run_sat<8, true, 4, 2>("sat clamp P=8 FFMA2 u2", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy);
	run_sat<8, true, 3, 4>("sat clamp P=8 FFMA2 u4 3 blocks", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy);
	run_sat<4, false, 6, 4>("sat clamp P=4 scalar u4", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy);
	run_sat<4, true, 6, 4>("sat clamp P=4 FFMA2 u4", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy);
	run_sat<4, true, 8, 4>("sat clamp P=4 FFMA2 u4 8 blocks", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy);
	run_sat<8, true, 4, 8>("sat clamp P=8 FFMA2 u8", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy);
	run_sat<6, true, 5, 4>("sat clamp P=6 FFMA2 u4 5 blocks", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy, false);
	run_sat<8, true, 4, 4, 1>("sat P=8 FFMA2, 1 channel on ALU", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy, false);
	run_sat<8, true, 4, 4, 2>("sat P=8 FFMA2, 2 channels on ALU", d_in, d_xy, d_spx, w, h, n, d_scores, page, xy, false);
	return 0;
}
