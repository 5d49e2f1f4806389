// Instruction-throughput probes for the pipes the ACE inner loop uses (B200).
// Each kernel runs 8 independent chains of one instruction; reports warp
// instructions per cycle per SM.  nvcc -arch=sm_100a -O3 pipes.cu -o pipes
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 4096
#define CHAINS 8

#define PROBE(NAME, TYPE, INIT, BODY)                                          \
	__global__ void NAME(TYPE *sink, TYPE a, TYPE b)                           \
	{                                                                          \
		TYPE v[CHAINS];                                                        \
		for (int i = 0; i < CHAINS; i++)                                       \
			v[i] = INIT;                                                       \
		for (int it = 0; it < ITERS; it++) {                                   \
			_Pragma("unroll") for (int i = 0; i < CHAINS; i++) { BODY; }       \
		}                                                                      \
		TYPE s = v[0];                                                         \
		for (int i = 1; i < CHAINS; i++)                                       \
			s = s + v[i];                                                      \
		if (s == (TYPE)123456789)                                              \
			sink[0] = s;                                                       \
	}

PROBE(k_vimnmx16x2, uint32_t, threadIdx.x + i,
	asm volatile("min.u16x2 %0, %0, %1;" : "+r"(v[i]) : "r"(a)); asm volatile("max.u16x2 %0, %0, %1;" : "+r"(v[i]) : "r"(b)))
PROBE(k_vimnmx32, int, threadIdx.x + i,
	asm volatile("min.s32 %0, %0, %1;" : "+r"(v[i]) : "r"(a)); asm volatile("max.s32 %0, %0, %1;" : "+r"(v[i]) : "r"(b)))
PROBE(k_prmt, uint32_t, threadIdx.x + i,
	asm volatile("prmt.b32 %0, %0, %1, 0x7604;" : "+r"(v[i]) : "r"(a)); asm volatile("prmt.b32 %0, %0, %1, 0x3214;" : "+r"(v[i]) : "r"(b)))
PROBE(k_imad, int, threadIdx.x + i,
	asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(v[i]) : "r"(a), "r"(b)); asm volatile("mad.lo.s32 %0, %0, %2, %1;" : "+r"(v[i]) : "r"(a), "r"(b)))
PROBE(k_iadd, int, threadIdx.x + i,
	asm volatile("add.s32 %0, %0, %1;" : "+r"(v[i]) : "r"(a)); asm volatile("sub.s32 %0, %0, %1;" : "+r"(v[i]) : "r"(b)))
PROBE(k_lop3, uint32_t, threadIdx.x + i,
	asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(v[i]) : "r"(a), "r"(b)); asm volatile("lop3.b32 %0, %0, %2, %1, 0xe8;" : "+r"(v[i]) : "r"(a), "r"(b)))
PROBE(k_dfma, double, threadIdx.x + i,
	asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(v[i]) : "d"(a), "d"(b)); asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(v[i]) : "d"(a), "d"(b)))
PROBE(k_dadd, double, threadIdx.x + i,
	asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(v[i]) : "d"(a)); asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(v[i]) : "d"(b)))
PROBE(k_dmul, double, threadIdx.x + i + 1,
	asm volatile("mul.rn.f64 %0, %0, %1;" : "+d"(v[i]) : "d"(a)); asm volatile("mul.rn.f64 %0, %0, %1;" : "+d"(v[i]) : "d"(b)))
PROBE(k_rsq64, double, threadIdx.x + i + 1.0,
	asm volatile("rsqrt.approx.ftz.f64 %0, %0;" : "+d"(v[i])); asm volatile("rsqrt.approx.ftz.f64 %0, %0;" : "+d"(v[i])))
PROBE(k_ffma, float, threadIdx.x + i,
	asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(v[i]) : "f"(a), "f"(b)); asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(v[i]) : "f"(a), "f"(b)))
PROBE(k_fmnmx, float, threadIdx.x + i,
	asm volatile("min.f32 %0, %0, %1;" : "+f"(v[i]) : "f"(a)); asm volatile("max.f32 %0, %0, %1;" : "+f"(v[i]) : "f"(b)))
PROBE(k_hmnmx2, uint32_t, threadIdx.x + i,
	asm volatile("min.f16x2 %0, %0, %1;" : "+r"(v[i]) : "r"(a)); asm volatile("max.f16x2 %0, %0, %1;" : "+r"(v[i]) : "r"(b)))
PROBE(k_hadd2, uint32_t, threadIdx.x + i,
	asm volatile("add.f16x2 %0, %0, %1;" : "+r"(v[i]) : "r"(a)); asm volatile("add.f16x2 %0, %0, %1;" : "+r"(v[i]) : "r"(b)))
PROBE(k_i2f64, double, threadIdx.x + i,
	{ int t = __double2loint(v[i]); double d; asm volatile("cvt.rn.f64.s32 %0, %1;" : "=d"(d) : "r"(t)); v[i] = d; } { int t = __double2hiint(v[i]); double d; asm volatile("cvt.rn.f64.s32 %0, %1;" : "=d"(d) : "r"(t)); v[i] = d; })
PROBE(k_f2f64, double, threadIdx.x + i,
	{ float t = __int_as_float(__double2hiint(v[i])); double d; asm volatile("cvt.f64.f32 %0, %1;" : "=d"(d) : "f"(t)); v[i] = d; } { float t = __int_as_float(__double2hiint(v[i])); double d; asm volatile("cvt.f64.f32 %0, %1;" : "=d"(d) : "f"(t)); v[i] = d; })
PROBE(k_vabsdiff4, uint32_t, threadIdx.x + i,
	asm volatile("vabsdiff4.u32.u32.u32.add %0, %0, %1, %2;" : "+r"(v[i]) : "r"(a), "r"(b)); asm volatile("vabsdiff4.u32.u32.u32.add %0, %0, %2, %1;" : "+r"(v[i]) : "r"(a), "r"(b)))

PROBE(k_rsq32, float, threadIdx.x + i + 1.0f,
	asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(v[i])); asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(v[i])))
PROBE(k_ffma2, unsigned long long, threadIdx.x + i,
	asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(v[i]) : "l"(a), "l"(b)); asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(v[i]) : "l"(a), "l"(b)))
PROBE(k_fadd2, unsigned long long, threadIdx.x + i,
	asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(v[i]) : "l"(a)); asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(v[i]) : "l"(b)))
PROBE(k_fadd, float, threadIdx.x + i,
	asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(v[i]) : "f"(a)); asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(v[i]) : "f"(b)))
PROBE(k_ffma_denorm, float, __int_as_float(threadIdx.x + i),
	asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(v[i]) : "f"(a), "f"(b)); asm volatile("fma.rn.f32 %0, %2, %1, %0;" : "+f"(v[i]) : "f"(a), "f"(b)))
PROBE(k_viaddmnmx, int, threadIdx.x + i,
	asm volatile("{ .reg .s32 t; add.s32 t, %0, %1; min.s32 t, t, %2; max.s32 %0, t, 0; }" : "+r"(v[i]) : "r"(a), "r"(b)); asm volatile("{ .reg .s32 t; add.s32 t, %0, %2; min.s32 t, t, %1; max.s32 %0, t, 0; }" : "+r"(v[i]) : "r"(a), "r"(b)))

// ---- pairs of instruction kinds interleaved: do their pipes overlap? ------
#define MIX(NAME, BODYA, BODYB)                                                \
	__global__ void NAME(double *sink, double a, double b, uint32_t ua, uint32_t ub) \
	{                                                                          \
		double v[4];                                                           \
		uint32_t u[4];                                                         \
		for (int i = 0; i < 4; i++) {                                          \
			v[i] = threadIdx.x + i + 1.0;                                      \
			u[i] = threadIdx.x + i;                                            \
		}                                                                      \
		for (int it = 0; it < ITERS; it++) {                                   \
			_Pragma("unroll") for (int r = 0; r < 2; r++) {                    \
				_Pragma("unroll") for (int i = 0; i < 4; i++) { BODYA; BODYB; } \
			}                                                                  \
		}                                                                      \
		double s = v[0] + v[1] + v[2] + v[3] + (double)(u[0] + u[1] + u[2] + u[3]); \
		if (s == 123456789.0)                                                  \
			sink[0] = s;                                                       \
	}

#define A_DFMA asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(v[i]) : "d"(a), "d"(b))
#define A_RSQ asm volatile("rsqrt.approx.ftz.f64 %0, %0;" : "+d"(v[i]))
#define B_VMNMX asm volatile("min.u16x2 %0, %0, %1;" : "+r"(u[i]) : "r"(ua))
#define B_PRMT asm volatile("prmt.b32 %0, %0, %1, 0x3214;" : "+r"(u[i]) : "r"(ub))
#define B_IMAD asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(u[i]) : "r"(ua), "r"(ub))
#define B_LOP asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u[i]) : "r"(ua), "r"(ub))
#define B_NONE

MIX(m_dfma_only, A_DFMA, B_NONE)
MIX(m_dfma_vmnmx, A_DFMA, B_VMNMX)
MIX(m_dfma_prmt, A_DFMA, B_PRMT)
MIX(m_dfma_imad, A_DFMA, B_IMAD)
MIX(m_rsq_vmnmx, A_RSQ, B_VMNMX)
MIX(m_rsq_dfma, A_RSQ, { asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(a) : "d"(b), "d"(b)); })
__global__ void m_vmnmx_imad(double *sink, double a, double b, uint32_t ua, uint32_t ub)
{
	uint32_t u[4], w[4];
	for (int i = 0; i < 4; i++) { u[i] = threadIdx.x + i; w[i] = threadIdx.x * 3 + i; }
	for (int it = 0; it < ITERS; it++) {
#pragma unroll
		for (int r = 0; r < 2; r++) {
#pragma unroll
			for (int i = 0; i < 4; i++) {
				asm volatile("min.u16x2 %0, %0, %1;" : "+r"(u[i]) : "r"(ua));
				asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(w[i]) : "r"(ua), "r"(ub));
			}
		}
	}
	if (u[0] + u[1] + u[2] + u[3] + w[0] + w[1] + w[2] + w[3] == 123456789u) sink[0] = 1.0;
}

template <typename K>
void run_mix(const char *name, K kern, int sms, double clk_ghz, int instr_per_iter)
{
	double *sink;
	cudaMalloc(&sink, 64);
	const int blocks = sms * 4, threads = 512;
	kern<<<blocks, threads>>>(sink, 1.0000001, 1e-9, 0x04e404e4u, 0x12345678u);
	cudaEvent_t e0, e1;
	cudaEventCreate(&e0);
	cudaEventCreate(&e1);
	float best = 1e30f;
	for (int r = 0; r < 5; r++) {
		cudaEventRecord(e0);
		kern<<<blocks, threads>>>(sink, 1.0000001, 1e-9, 0x04e404e4u, 0x12345678u);
		cudaEventRecord(e1);
		cudaEventSynchronize(e1);
		float ms;
		cudaEventElapsedTime(&ms, e0, e1);
		if (ms < best) best = ms;
	}
	const double cyc = best * 1e-3 * clk_ghz * 1e9;
	const double per_smsp = cyc / ((double)(threads / 32) * 4 / 4 * ITERS);   // 4 blocks/SM, 16 warps each, 4 SMSPs
	printf("%-16s %8.3f ms  %6.2f cycles per SMSP per (warp x iteration of %d instr)\n", name, best,
			per_smsp, instr_per_iter);
	cudaFree(sink);
}


// DFMA with three distinct register-pair operands per instruction (the ACE
// pattern: fresh multiplicand, fresh accumulator), against the probe above
// whose multiplier and addend never change.
__global__ void k_dfma_distinct(double *sink, double a, double b)
{
	double v[CHAINS], x[CHAINS], y[CHAINS];
	for (int i = 0; i < CHAINS; i++) {
		v[i] = threadIdx.x + i;
		x[i] = a + i * 1e-9;
		y[i] = b + i;
	}
	for (int it = 0; it < ITERS; it++) {
#pragma unroll
		for (int i = 0; i < CHAINS; i++) {
			asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(v[i]) : "d"(x[i]), "d"(y[(i + 1) % CHAINS]));
			asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(v[i]) : "d"(x[(i + 3) % CHAINS]), "d"(y[i]));
		}
	}
	double s = v[0];
	for (int i = 1; i < CHAINS; i++)
		s = s + v[i];
	if (s == 123456789.0)
		sink[0] = s;
}

template <typename T, typename K>
void run(const char *name, K kern, T a, T b, int sms, double clk_ghz)
{
	T *sink;
	cudaMalloc(&sink, 64);
	const int blocks = sms * 4, threads = 512;
	kern<<<blocks, threads>>>(sink, a, b);
	cudaEvent_t e0, e1;
	cudaEventCreate(&e0);
	cudaEventCreate(&e1);
	float best = 1e30f;
	for (int r = 0; r < 5; r++) {
		cudaEventRecord(e0);
		kern<<<blocks, threads>>>(sink, a, b);
		cudaEventRecord(e1);
		cudaEventSynchronize(e1);
		float ms;
		cudaEventElapsedTime(&ms, e0, e1);
		if (ms < best) best = ms;
	}
	const double warp_instr = (double)blocks * (threads / 32) * ITERS * CHAINS * 2;
	const double per_sm_per_clk = warp_instr / sms / (best * 1e-3 * clk_ghz * 1e9);
	printf("%-14s %8.3f ms  %6.3f warp-instr/clk/SM  (%5.1f lanes/clk/SM)\n", name, best,
			per_sm_per_clk, per_sm_per_clk * 32);
	cudaFree(sink);
}

int main()
{
	cudaDeviceProp p;
	cudaGetDeviceProperties(&p, 0);
	int clk_khz = 0;
	cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
	const double ghz = clk_khz / 1e6;
	printf("%s, %d SMs, nominal %.3f GHz (rates assume this clock)\n", p.name, p.multiProcessorCount, ghz);
	const int sms = p.multiProcessorCount;
	run<uint32_t>("vimnmx.u16x2", k_vimnmx16x2, 0x04e404e4u, 0x041c041cu, sms, ghz);
	run<int>("vimnmx.s32", k_vimnmx32, 1000, -1000, sms, ghz);
	run<uint32_t>("prmt", k_prmt, 0x40b00000u, 0x12345678u, sms, ghz);
	run<int>("imad", k_imad, 3, 7, sms, ghz);
	run<int>("iadd", k_iadd, 3, 7, sms, ghz);
	run<uint32_t>("lop3", k_lop3, 0x0f0f0f0fu, 0x33333333u, sms, ghz);
	run<double>("dfma", k_dfma, 1.0000001, 1e-9, sms, ghz);
	run<double>("dfma 3 operands", k_dfma_distinct, 1.0000001, 1e-9, sms, ghz);
	run<double>("dadd", k_dadd, 1.5, -1.5, sms, ghz);
	run<double>("dmul", k_dmul, 1.0000001, 0.9999999, sms, ghz);
	run<double>("rsqrt64", k_rsq64, 0.0, 0.0, sms, ghz);
	run<float>("ffma", k_ffma, 1.0000001f, 1e-9f, sms, ghz);
	run<float>("fmnmx", k_fmnmx, 1000.f, -1000.f, sms, ghz);
	run<uint32_t>("hmnmx2", k_hmnmx2, 0x64006400u, 0x60006000u, sms, ghz);
	run<uint32_t>("hadd2", k_hadd2, 0x3c003c00u, 0xbc00bc00u, sms, ghz);
	run<double>("i2f.f64", k_i2f64, 0.0, 0.0, sms, ghz);
	run<double>("f2f.f64.f32", k_f2f64, 0.0, 0.0, sms, ghz);
	run<uint32_t>("vabsdiff4", k_vabsdiff4, 0x01020304u, 0x04030201u, sms, ghz);
	run<float>("rsqrt32.ftz", k_rsq32, 0.f, 0.f, sms, ghz);
	run<unsigned long long>("ffma2", k_ffma2, 0x3f8000013f800001ull, 0x3089705f3089705full, sms, ghz);
	run<unsigned long long>("fadd2", k_fadd2, 0x3fc000003fc00000ull, 0xbfc00000bfc00000ull, sms, ghz);
	run<float>("fadd", k_fadd, 1.5f, -1.5f, sms, ghz);
	run<float>("ffma denormal op", k_ffma_denorm, 1e20f, 1.4e-43f, sms, ghz);
	run<int>("viaddmnmx.relu", k_viaddmnmx, 3, 200, sms, ghz);
	printf("mixed probes: 16 warps per SMSP, each iteration = 8 A + 8 B instructions per warp\n");
	run_mix("dfma only", m_dfma_only, sms, ghz, 8);
	run_mix("dfma+vimnmx", m_dfma_vmnmx, sms, ghz, 16);
	run_mix("dfma+prmt", m_dfma_prmt, sms, ghz, 16);
	run_mix("dfma+imad", m_dfma_imad, sms, ghz, 16);
	run_mix("rsq64+vimnmx", m_rsq_vmnmx, sms, ghz, 16);
	run_mix("rsq64+dfma", m_rsq_dfma, sms, ghz, 16);
	run_mix("vimnmx+imad", m_vmnmx_imad, sms, ghz, 16);
	return 0;
}
