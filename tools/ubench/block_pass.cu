// What one pass of a block-wide iterative algorithm costs on one SM: a few
// dependent shared-memory accesses per thread, then a barrier, repeated.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o block_pass block_pass.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void pass_kernel(int passes, long long *out, int *sink)
{
	__shared__ int a[4096];
	__shared__ unsigned char k[4096];
	__shared__ int flag;
	for (int i = threadIdx.x; i < 4096; i += blockDim.x) {
		a[i] = (i * 7 + 3) & 4095;
		k[i] = i & 1;
	}
	if (threadIdx.x == 0)
		flag = 0;
	__syncthreads();
	int v = threadIdx.x;
	const long long t0 = clock64();
	for (int p = 0; p < passes; p++) {
		if (MODE == 0) {                       // barrier only
			__syncthreads();
		} else if (MODE == 1) {                // barrier with OR reduction
			v += __syncthreads_or(v == -1);
		} else if (MODE == 2) {                // one load, barrier
			v = a[v & 4095];
			__syncthreads();
		} else if (MODE == 3) {                // three dependent loads, barrier
			v = a[v & 4095];
			v = a[v & 4095];
			v = a[v & 4095];
			__syncthreads();
		} else if (MODE == 4) {                // an edge pass: load, two byte loads, store, OR barrier
			const int e = a[v & 4095];
			const unsigned char x = k[e & 4095], y = k[(e >> 3) & 4095];
			bool ch = false;
			if (x != y) {
				k[e & 4095] = 1;
				ch = (p == -1);
			}
			v += __syncthreads_or(ch) + e;
		} else if (MODE == 5) {                // three dependent loads, atomicMin, OR barrier
			v = a[v & 4095];
			const int r = a[v & 4095], q = a[(v + 1) & 4095];
			if (r != q)
				atomicMin(&a[max(r, q) & 4095], min(r, q));
			v += __syncthreads_or(p == -1);
		}
	}
	const long long t1 = clock64();
	if (threadIdx.x == 0)
		out[0] = t1 - t0;
	if (v == 0x7fffffff)
		sink[0] = v + flag;
}

template <int MODE>
void run(const char *what, long long *d_out, int *d_sink)
{
	const int passes = 200;
	for (int threads = 1024; threads >= 32; threads /= 2) {
		pass_kernel<MODE><<<1, threads>>>(passes, d_out, d_sink);
		pass_kernel<MODE><<<1, threads>>>(passes, d_out, d_sink);
		long long c = 0;
		cudaMemcpy(&c, d_out, sizeof(c), cudaMemcpyDeviceToHost);
		printf("%-58s %4d threads: %6.1f cycles per pass\n", what, threads, (double)c / passes);
	}
}

int main()
{
	long long *d_out;
	int *d_sink;
	cudaMalloc(&d_out, 8);
	cudaMalloc(&d_sink, 4);
	run<0>("barrier", d_out, d_sink);
	run<1>("barrier with OR", d_out, d_sink);
	run<2>("one shared load + barrier", d_out, d_sink);
	run<3>("three dependent shared loads + barrier", d_out, d_sink);
	run<4>("edge pass (load, 2 byte loads, store) + OR barrier", d_out, d_sink);
	run<5>("hook pass (3 loads, atomicMin) + OR barrier", d_out, d_sink);
	return 0;
}
