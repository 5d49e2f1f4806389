// Dependent-chain latency of FP64 instructions on one warp, and of the IEEE
// division / square root sequences.  nvcc -O3 -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cuda_runtime.h>

__global__ void k(double *out, long long *cyc, double a, double b, int n)
{
	double x = a + threadIdx.x;
	long long t0 = clock64();
	for (int i = 0; i < n; i++)
		x = __dadd_rn(x, b);
	long long t1 = clock64();
	double y = x;
	for (int i = 0; i < n; i++)
		y = fma(y, b, a);
	long long t2 = clock64();
	double z = y + 3.0;
	for (int i = 0; i < n; i++)
		z = __ddiv_rn(a, z + 1.5);
	long long t3 = clock64();
	double s = z + 7.0;
	for (int i = 0; i < n; i++)
		s = __dsqrt_rn(s + 2.0);
	long long t4 = clock64();
	__shared__ double sm[256];
	sm[threadIdx.x] = s;
	__syncthreads();
	double q = 0.0;
	for (int i = 0; i < n; i++)
		q = __dadd_rn(q, sm[(i + threadIdx.x) & 255]);
	long long t5 = clock64();
	out[threadIdx.x] = x + y + z + s + q;
	if (threadIdx.x == 0) {
		cyc[0] = t1 - t0; cyc[1] = t2 - t1; cyc[2] = t3 - t2; cyc[3] = t4 - t3; cyc[4] = t5 - t4;
	}
}

int main()
{
	double *out; long long *cyc, h[5];
	cudaMalloc(&out, 256 * 8); cudaMalloc(&cyc, 5 * 8);
	const int n = 1024;
	for (int threads : {32, 4}) {
		k<<<1, threads>>>(out, cyc, 1.0, 1e-9, n);
		k<<<1, threads>>>(out, cyc, 1.0, 1e-9, n);
		cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
		printf("threads %d: DADD %.1f  DFMA %.1f  ddiv_rn(+DADD) %.1f  dsqrt_rn(+DADD) %.1f  LDS+DADD %.1f cycles per dependent step\n",
				threads, (double)h[0] / n, (double)h[1] / n, (double)h[2] / n, (double)h[3] / n, (double)h[4] / n);
	}
	return 0;
}
