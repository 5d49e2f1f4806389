// C-ABI of libpillowfight_b200.so (see include/pillowfight_b200.h).
//
// Host-pointer entry points (pf_apply_*) keep one context per host thread: a
// stream, grow-only device buffers and a pinned staging buffer.  A call is
//   host bitmap -> (staging) -> HBM -> kernels -> HBM -> (staging) -> host bitmap
// and returns after the result is in the caller's memory.  There is no CPU
// implementation behind any of them: without a CUDA device they fail.
#include <cfloat>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>
#if defined(__x86_64__)
#include <immintrin.h>
#endif

#include "pf_common.cuh"

namespace pfb {

// ---------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------

static thread_local int t_err = PF_B200_OK;
static thread_local char t_msg[512] = "";

int fail(int code, const char *fmt, ...)
{
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(t_msg, sizeof(t_msg), fmt, ap);
	va_end(ap);
	t_err = code;
	return code;
}

void clear_error()
{
	t_err = PF_B200_OK;
	t_msg[0] = '\0';
}

bool pdl_enabled()
{
	static const bool on = !getenv("PF_B200_PDL") || atoi(getenv("PF_B200_PDL"));
	return on;
}

int sm_count()
{
	static thread_local int cached_dev = -1, cached = 148;
	int dev = 0;
	if (cudaGetDevice(&dev) != cudaSuccess)
		return cached;
	if (dev != cached_dev) {
		int n = 0;
		if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
			cached = n;
		cached_dev = dev;
	}
	return cached;
}

// ---------------------------------------------------------------------------
// per-thread context for the host-pointer entry points
// ---------------------------------------------------------------------------

struct Buffer {
	void *p = nullptr;
	size_t cap = 0;
};

constexpr int kMaxBands = 16;      // row bands a streamed call is cut into, at most

// A few host threads that live as long as the calling thread's context: the
// staging copies of pageable pages run on them (starting eight threads per
// call costs more than some of the copies take, and makes the call's duration
// erratic).  start(f) hands f(worker index) to every worker; wait() returns
// when all have finished it.
class CopyPool {
public:
	~CopyPool()
	{
		{
			std::lock_guard<std::mutex> lk(m_);
			stop_ = true;
		}
		cv_job_.notify_all();
		for (std::thread &t : threads_)
			t.join();
	}
	int size() const { return (int)threads_.size(); }
	// Only between jobs (never while one is running): a new worker starts at
	// the generation current now, so it waits for the NEXT start() and never
	// sees the previous call's job.
	void ensure(int n)
	{
		while ((int)threads_.size() < n) {
			const int idx = (int)threads_.size();
			unsigned long gen;
			{
				std::lock_guard<std::mutex> lk(m_);
				gen = generation_;
			}
			threads_.emplace_back([this, idx, gen] { loop(idx, gen); });
		}
	}
	void start(std::function<void(int)> f)
	{
		std::lock_guard<std::mutex> lk(m_);
		job_ = std::move(f);
		pending_ = (int)threads_.size();
		generation_++;
		cv_job_.notify_all();
	}
	void wait()
	{
		std::unique_lock<std::mutex> lk(m_);
		cv_done_.wait(lk, [this] { return pending_ == 0; });
		job_ = nullptr;     // its captures die with the caller's frame
	}

private:
	void loop(int idx, unsigned long seen)
	{
		for (;;) {
			std::function<void(int)> f;
			{
				std::unique_lock<std::mutex> lk(m_);
				cv_job_.wait(lk, [&] { return stop_ || generation_ != seen; });
				if (stop_)
					return;
				seen = generation_;
				f = job_;
			}
			f(idx);
			{
				std::lock_guard<std::mutex> lk(m_);
				if (--pending_ == 0)
					cv_done_.notify_all();
			}
		}
	}
	std::vector<std::thread> threads_;
	std::mutex m_;
	std::condition_variable cv_job_, cv_done_;
	std::function<void(int)> job_;
	unsigned long generation_ = 0;
	int pending_ = 0;
	bool stop_ = false;
};

struct Context {
	int dev = -1;
	CopyPool pool;
	cudaStream_t st = nullptr;           // kernels
	cudaStream_t st_up = nullptr;        // host -> device copies of a streamed call
	cudaStream_t st_down = nullptr;      // device -> host copies of a streamed call
	cudaEvent_t ev_up[kMaxBands] = {};   // band b is in HBM
	cudaEvent_t ev_done[kMaxBands] = {}; // band b's result is ready
	Buffer d_in, d_out, d_scratch, d_scores, d_small, d_blur;
	Buffer h_stage;   // pinned

	~Context()
	{
		// Thread exit: release what this thread holds.  Errors are ignored
		// (the CUDA runtime may already be shutting down).
		if (dev < 0)
			return;
		cudaSetDevice(dev);
		Buffer *bufs[] = { &d_in, &d_out, &d_scratch, &d_scores, &d_small, &d_blur };
		for (Buffer *b : bufs)
			if (b->p)
				cudaFree(b->p);
		if (h_stage.p)
			cudaFreeHost(h_stage.p);
		for (int i = 0; i < kMaxBands; i++) {
			if (ev_up[i])
				cudaEventDestroy(ev_up[i]);
			if (ev_done[i])
				cudaEventDestroy(ev_done[i]);
		}
		if (st_up)
			cudaStreamDestroy(st_up);
		if (st_down)
			cudaStreamDestroy(st_down);
		if (st)
			cudaStreamDestroy(st);
	}
};

static thread_local Context t_ctx;

static int ctx_init(Context &c)
{
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0)
		return fail(PF_B200_ENODEV, "no CUDA device is usable (%s); this build has no CPU path",
				cudaGetErrorString(cudaGetLastError()));
	int dev = 0;
	PFB_CUDA(cudaGetDevice(&dev));
	if (c.dev == dev && c.st)
		return PF_B200_OK;
	if (c.dev >= 0 && c.dev != dev)
		return fail(PF_B200_EINVAL, "thread first used device %d, now on %d: "
				"one device per host thread", c.dev, dev);
	c.dev = dev;
	PFB_CUDA(cudaStreamCreateWithFlags(&c.st, cudaStreamNonBlocking));
	PFB_CUDA(cudaStreamCreateWithFlags(&c.st_up, cudaStreamNonBlocking));
	PFB_CUDA(cudaStreamCreateWithFlags(&c.st_down, cudaStreamNonBlocking));
	for (int i = 0; i < kMaxBands; i++) {
		PFB_CUDA(cudaEventCreateWithFlags(&c.ev_up[i], cudaEventDisableTiming));
		PFB_CUDA(cudaEventCreateWithFlags(&c.ev_done[i], cudaEventDisableTiming));
	}
	return PF_B200_OK;
}

static int grow(Buffer &b, size_t bytes, bool pinned)
{
	if (b.cap >= bytes)
		return PF_B200_OK;
	if (b.p) {
		if (pinned)
			cudaFreeHost(b.p);
		else
			cudaFree(b.p);
		b.p = nullptr;
		b.cap = 0;
	}
	if (pinned)
		PFB_CUDA(cudaMallocHost(&b.p, bytes));
	else
		PFB_CUDA(cudaMalloc(&b.p, bytes));
	b.cap = bytes;
	return PF_B200_OK;
}

// A streamed call that fails half way leaves copies and kernels queued that
// still read the caller's input or write its output; nothing of the call may
// be in flight when it returns (the caller may free the pages, and the next
// call reuses the device buffers and events).
static int drained(Context &c, int rc)
{
	if (rc != PF_B200_OK && c.st) {
		cudaStreamSynchronize(c.st_up);
		cudaStreamSynchronize(c.st);
		cudaStreamSynchronize(c.st_down);
		// (a Canny pass cut short may have left its scratch header in use)
		if (c.d_scratch.p)
			cudaMemset(c.d_scratch.p, 0, 256);
		cudaGetLastError();
	}
	return rc;
}

// The context's Canny scratch: its header is cleared when the buffer is (re)allocated;
// every Canny pass leaves it clear, so no call has to.
static int grow_canny_scratch(Context &c, int w, int h)
{
	const size_t before = c.d_scratch.cap;
	const int rc = grow(c.d_scratch, canny_scratch_bytes(w, h), false);
	if (rc)
		return rc;
	if (c.d_scratch.cap != before)
		PFB_CUDA(cudaMemsetAsync(c.d_scratch.p, 0, 256, c.st));
	return PF_B200_OK;
}

static bool is_pinned(const void *p)
{
	cudaPointerAttributes a;
	if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
		cudaGetLastError();
		return false;
	}
	return a.type == cudaMemoryTypeHost;
}

// The address the kernels can use for a pinned (page-locked, mapped) host
// buffer, or NULL: pageable memory, or pinned memory the device cannot map.
static void *device_alias(const void *p)
{
	cudaPointerAttributes a;
	if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
		cudaGetLastError();
		return nullptr;
	}
	return a.type == cudaMemoryTypeHost ? a.devicePointer : nullptr;
}

// A copy of a large slice with streaming (non-temporal) stores where the CPU
// has AVX2: the destination is written once and not read by this core again,
// so fetching its cache lines first (what an ordinary store does) is wasted
// memory traffic -- a third of the total.  PF_B200_NT_COPY=0 uses memcpy.
#if defined(__x86_64__)
__attribute__((target("avx2")))
static void copy_streaming_avx2(char *dst, const char *src, size_t n)
{
	// head: up to the first 32-byte boundary of the destination
	size_t head = (32 - ((uintptr_t)dst & 31)) & 31;
	if (head > n)
		head = n;
	memcpy(dst, src, head);
	dst += head;
	src += head;
	n -= head;
	size_t i = 0;
	for (; i + 128 <= n; i += 128) {
		const __m256i a = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(src + i));
		const __m256i b = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(src + i + 32));
		const __m256i c = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(src + i + 64));
		const __m256i d = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(src + i + 96));
		_mm256_stream_si256(reinterpret_cast<__m256i *>(dst + i), a);
		_mm256_stream_si256(reinterpret_cast<__m256i *>(dst + i + 32), b);
		_mm256_stream_si256(reinterpret_cast<__m256i *>(dst + i + 64), c);
		_mm256_stream_si256(reinterpret_cast<__m256i *>(dst + i + 96), d);
	}
	_mm_sfence();
	memcpy(dst + i, src + i, n - i);
}
#endif

static void copy_slice(void *dst, const void *src, size_t n)
{
#if defined(__x86_64__)
	static const bool nt = (!getenv("PF_B200_NT_COPY") || atoi(getenv("PF_B200_NT_COPY")))
		&& __builtin_cpu_supports("avx2");
	if (nt && n >= ((size_t)1 << 20)) {
		copy_streaming_avx2(static_cast<char *>(dst), static_cast<const char *>(src), n);
		return;
	}
#endif
	memcpy(dst, src, n);
}

// Large pageable buffers: the staging copy is the slow part (one core moves
// ~10 GB/s, PCIe 5 takes ~55), so it is split over a few host threads (8:
// with streaming stores the copies scale that far on the test box).  Each
// thread copies its slice into the pinned staging buffer and queues that
// slice's DMA itself, so the first DMA starts while the others still copy.
constexpr size_t kParallelMin = (size_t)8 << 20;       // bytes
constexpr size_t kParallelMax = (size_t)1 << 30;
constexpr int kCopyThreadsMax = 16;
static int copy_threads()
{
	static const int n = getenv("PF_B200_COPY_THREADS") ? atoi(getenv("PF_B200_COPY_THREADS")) : 8;
	return n < 1 ? 1 : (n > kCopyThreadsMax ? kCopyThreadsMax : n);
}

static int upload_threads(Context &c, void *dst, const void *src, size_t bytes)
{
	int rc = grow(c.h_stage, bytes, true);
	if (rc)
		return rc;
	const int kCopyThreads = copy_threads();
	const size_t slice = ((bytes + kCopyThreads - 1) / kCopyThreads + 4095) & ~(size_t)4095;
	cudaError_t errs[kCopyThreadsMax] = {};
	auto work = [&](int i) {
		const size_t off = (size_t)i * slice;
		if (off >= bytes)
			return;
		const size_t n = bytes - off < slice ? bytes - off : slice;
		char *stage = static_cast<char *>(c.h_stage.p) + off;
		copy_slice(stage, static_cast<const char *>(src) + off, n);
		cudaSetDevice(c.dev);
		errs[i] = cudaMemcpyAsync(static_cast<char *>(dst) + off, stage, n,
				cudaMemcpyHostToDevice, c.st);
	};
	// slice 0 on the calling thread, the others on the pool
	c.pool.ensure(kCopyThreads - 1);
	c.pool.start([&](int i) {
		if (i + 1 < kCopyThreads)
			work(i + 1);
	});
	work(0);
	c.pool.wait();
	for (cudaError_t e : errs)
		if (e != cudaSuccess)
			return fail(PF_B200_ECUDA, "upload: %s", cudaGetErrorString(e));
	return PF_B200_OK;
}

static int download_threads(Context &c, void *dst, const void *src, size_t bytes)
{
	int rc = grow(c.h_stage, bytes, true);
	if (rc)
		return rc;
	const int kCopyThreads = copy_threads();      // (ev_done[] has kMaxBands = 16 entries)
	const size_t slice = ((bytes + kCopyThreads - 1) / kCopyThreads + 4095) & ~(size_t)4095;
	for (int i = 0; i < kCopyThreads; i++) {
		const size_t off = (size_t)i * slice;
		if (off >= bytes)
			break;
		const size_t n = bytes - off < slice ? bytes - off : slice;
		PFB_CUDA(cudaMemcpyAsync(static_cast<char *>(c.h_stage.p) + off,
				static_cast<const char *>(src) + off, n, cudaMemcpyDeviceToHost, c.st));
		PFB_CUDA(cudaEventRecord(c.ev_done[i], c.st));
	}
	cudaError_t errs[kCopyThreadsMax] = {};
	auto work = [&](int i) {
		const size_t off = (size_t)i * slice;
		if (off >= bytes)
			return;
		const size_t n = bytes - off < slice ? bytes - off : slice;
		cudaSetDevice(c.dev);
		errs[i] = cudaEventSynchronize(c.ev_done[i]);
		copy_slice(static_cast<char *>(dst) + off, static_cast<char *>(c.h_stage.p) + off, n);
	};
	// slice 0 on the calling thread, the others on the pool
	c.pool.ensure(kCopyThreads - 1);
	c.pool.start([&](int i) {
		if (i + 1 < kCopyThreads)
			work(i + 1);
	});
	work(0);
	c.pool.wait();
	for (cudaError_t e : errs)
		if (e != cudaSuccess)
			return fail(PF_B200_ECUDA, "download: %s", cudaGetErrorString(e));
	return PF_B200_OK;
}

// host -> device.  Pinned caller memory goes by DMA directly; pageable memory
// is copied through the pinned staging buffer (small buffers in chunks so that
// the memcpy of chunk i+1 overlaps the DMA of chunk i, large ones on threads).
static int upload(Context &c, void *dst, const void *src, size_t bytes)
{
	if (is_pinned(src)) {
		PFB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c.st));
		return PF_B200_OK;
	}
	if (bytes >= kParallelMin && bytes <= kParallelMax)
		return upload_threads(c, dst, src, bytes);
	const size_t chunk = 4u << 20;
	int rc = grow(c.h_stage, 2 * chunk, true);
	if (rc)
		return rc;
	cudaEvent_t ev[2];
	PFB_CUDA(cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming));
	PFB_CUDA(cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming));
	int slot = 0;
	bool used[2] = { false, false };
	for (size_t off = 0; off < bytes; off += chunk, slot ^= 1) {
		const size_t n = bytes - off < chunk ? bytes - off : chunk;
		char *stage = static_cast<char *>(c.h_stage.p) + slot * chunk;
		if (used[slot])
			cudaEventSynchronize(ev[slot]);
		memcpy(stage, static_cast<const char *>(src) + off, n);
		cudaMemcpyAsync(static_cast<char *>(dst) + off, stage, n,
				cudaMemcpyHostToDevice, c.st);
		cudaEventRecord(ev[slot], c.st);
		used[slot] = true;
	}
	cudaStreamSynchronize(c.st);
	cudaEventDestroy(ev[0]);
	cudaEventDestroy(ev[1]);
	PFB_CUDA(cudaGetLastError());
	return PF_B200_OK;
}

// device -> host, then wait: the result is in the caller's memory on return.
static int download(Context &c, void *dst, const void *src, size_t bytes)
{
	if (is_pinned(dst)) {
		PFB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c.st));
		PFB_CUDA(cudaStreamSynchronize(c.st));
		return PF_B200_OK;
	}
	if (bytes >= kParallelMin && bytes <= kParallelMax)
		return download_threads(c, dst, src, bytes);
	const size_t chunk = 4u << 20;
	int rc = grow(c.h_stage, 2 * chunk, true);
	if (rc)
		return rc;
	cudaEvent_t ev[2];
	PFB_CUDA(cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming));
	PFB_CUDA(cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming));
	// DMA of chunk i+1 runs while chunk i is copied out of the staging buffer
	size_t issued = 0, done = 0;
	int slot = 0;
	size_t len[2] = { 0, 0 }, at[2] = { 0, 0 };
	while (done < bytes) {
		while (issued < bytes && issued - done < 2 * chunk) {
			const size_t n = bytes - issued < chunk ? bytes - issued : chunk;
			const int s = (int)((issued / chunk) & 1);
			cudaMemcpyAsync(static_cast<char *>(c.h_stage.p) + s * chunk,
					static_cast<const char *>(src) + issued, n,
					cudaMemcpyDeviceToHost, c.st);
			cudaEventRecord(ev[s], c.st);
			len[s] = n;
			at[s] = issued;
			issued += n;
		}
		cudaEventSynchronize(ev[slot]);
		memcpy(static_cast<char *>(dst) + at[slot],
				static_cast<char *>(c.h_stage.p) + slot * chunk, len[slot]);
		done += len[slot];
		slot ^= 1;
	}
	cudaEventDestroy(ev[0]);
	cudaEventDestroy(ev[1]);
	PFB_CUDA(cudaGetLastError());
	return PF_B200_OK;
}

static int check_bitmaps(const struct pf_bitmap *in, const struct pf_bitmap *out)
{
	if (!in || !out || !in->pixels || !out->pixels)
		return fail(PF_B200_EINVAL, "NULL bitmap or pixel buffer");
	if (in->size.x <= 0 || in->size.y <= 0)
		return fail(PF_B200_EINVAL, "bitmap size %d x %d", in->size.x, in->size.y);
	if (out->size.x != in->size.x || out->size.y != in->size.y)
		return fail(PF_B200_EINVAL, "output bitmap is %d x %d, input %d x %d",
				out->size.x, out->size.y, in->size.x, in->size.y);
	return PF_B200_OK;
}

// upstream _gaussian.c: generate_gaussian_1d_kernel -- the taps are made on
// the host with glibc's exp/sqrt, in the reference's operation order, so the
// device only ever multiplies and adds.
int make_taps(double sigma, int nb_stddev, Taps &t)
{
	if (nb_stddev < 1 || nb_stddev > kMaxTaps)
		return fail(PF_B200_EINVAL, "gaussian: nb_stddev %d outside 1..%d", nb_stddev, kMaxTaps);
	if (sigma == 0.0)
		sigma = 0.3 * (((double)(nb_stddev - 1)) * 0.5 - 1.0) + 0.8;
	t.n = nb_stddev;
	double sum = 0.0;     // host code is built with -ffp-contract=off
	for (int i = 0; i < nb_stddev; i++) {
		const double d = (double)(i - nb_stddev / 2);
		t.t[i] = exp(-(d * d) / (2.0 * sigma * sigma))
			/ sqrt(2.0 * M_PI * sigma * sigma);
		sum += t.t[i];
	}
	for (int i = 0; i < nb_stddev; i++)
		t.t[i] = t.t[i] / sum;
	return PF_B200_OK;
}

// upstream _canny.c thresholds, turned into exact integer tests on
// g_x^2 + g_y^2 (hypot of two integers is compared against the double).
static int threshold_to_sq(double thr)
{
	long m = (long)floor(thr * thr);
	if (m < 0)
		m = 0;
	while (m > 0 && sqrt((double)(m - 1)) > thr)
		m--;
	while (!(sqrt((double)m) > thr))
		m++;
	return (int)m;
}

CannyThresholds canny_thresholds()
{
	CannyThresholds t;
	t.weak_min = threshold_to_sq(0.686 * 255.0);
	t.strong_min = threshold_to_sq(1.372 * 255.0);
	return t;
}

enum Op { OP_GAUSSIAN, OP_SOBEL, OP_CANNY };

// Host-pointer calls of this process that are between entry and return (all
// threads, all devices).  A call that is alone splits its page into row bands
// so that its own upload, kernels and download overlap; calls that run side by
// side on several host threads already keep both directions of PCIe busy with
// each other's pages, and then whole-page DMAs are the better use of the link
// (measured, 2-4 threads on one B200: 0.75 ms per Gaussian call in all with one
// DMA each way, 0.88 ms with 8 bands; alone 1.30 against 0.96 ms).
static std::atomic<int> g_calls_in_flight{0};
static thread_local int t_calls_beside = 0;      // other calls in flight when this one began
static thread_local int t_call_depth = 0;        // (pf_b200_page_step may go through the single calls)

struct CallScope {
	CallScope()
	{
		if (t_call_depth++ == 0)
			t_calls_beside = g_calls_in_flight.fetch_add(1, std::memory_order_relaxed);
	}
	~CallScope()
	{
		if (--t_call_depth == 0)
			g_calls_in_flight.fetch_sub(1, std::memory_order_relaxed);
	}
};

// PF_B200_BANDS = n: always n bands; otherwise `alone` (8; 3 for Canny, whose
// bands come back by DMA: 1.00 against 1.07 ms per A4 page), or 1 beside other
// calls.
static int bands_wanted(int alone)
{
	static const int env = getenv("PF_B200_BANDS") ? atoi(getenv("PF_B200_BANDS")) : 0;
	if (env >= 1)
		return env;
	return t_calls_beside > 0 ? 1 : alone;
}

// Row bands of a streamed call: at least 256 rows each, at most kMaxBands.
static int band_rows_for(int h, int alone = 8)
{
	const int nbands = bands_wanted(alone);
	int rows = (h + nbands - 1) / nbands;
	if (rows < 256)
		rows = 256;
	if ((h + rows - 1) / rows > kMaxBands)
		rows = (h + kMaxBands - 1) / kMaxBands;
	return rows;
}

// Queues the upload of a pinned page for a call that works in row bands of
// `rows` rows: band b's event (ev_up[b]) says that its rows and the kReach rows
// below them -- further than any stencil here reaches (<= 31 rows) -- are in
// HBM, so band b's kernels wait for nothing but their own event.
constexpr int kReach = 32;

static int upload_bands(Context &c, uint32_t *din, const uint32_t *src, int w, int h, int rows, int nb)
{
	int u0 = 0;
	for (int b = 0; b < nb; b++) {
		int u1 = (b + 1) * rows + kReach;
		if (u1 > h || b == nb - 1)
			u1 = h;
		if (u1 > u0)
			PFB_CUDA(cudaMemcpyAsync(din + (size_t)u0 * w, src + (size_t)u0 * w,
					(size_t)(u1 - u0) * w * sizeof(uint32_t), cudaMemcpyHostToDevice, c.st_up));
		PFB_CUDA(cudaEventRecord(c.ev_up[b], c.st_up));
		u0 = u1;
	}
	return PF_B200_OK;
}

// Gaussian / Sobel on pinned caller memory: the page goes up band by band on
// one stream, each band is filtered on the kernel stream as soon as it and
// the band below it (the stencil's reach) have arrived, and comes back on a
// third stream -- PCIe carries both directions at once instead of one after
// the other.  Returns after the last band is in the caller's memory.
static int run_stencil_streamed(Context &c, Op op, const struct pf_bitmap *in,
		struct pf_bitmap *out, const Taps &taps)
{
	const int w = in->size.x, h = in->size.y;
	const int rows = band_rows_for(h);
	const int nb = (h + rows - 1) / rows;
	const uint32_t *src = reinterpret_cast<const uint32_t *>(in->pixels);
	uint32_t *dst = reinterpret_cast<uint32_t *>(out->pixels);
	const uint32_t *din = static_cast<const uint32_t *>(c.d_in.p);
	uint32_t *dout = static_cast<uint32_t *>(c.d_out.p);
	// The result rows go straight from the kernel into the caller's page
	// (posted writes over PCIe, full 16-byte row segments per lane): the way
	// back needs no copy engine, no staging in HBM and no per-band hand-over.
	static const bool direct_env = !getenv("PF_B200_DIRECT_OUT") || atoi(getenv("PF_B200_DIRECT_OUT"));
	// (one band = a call beside others: the copy engine, as the link is shared)
	uint32_t *direct = (direct_env && nb > 1) ? static_cast<uint32_t *>(device_alias(out->pixels)) : nullptr;
	int rc_up = upload_bands(c, const_cast<uint32_t *>(din), src, w, h, rows, nb);
	if (rc_up)
		return rc_up;
	for (int b = 0; b < nb; b++) {
		const int y0 = b * rows, y1 = (y0 + rows < h) ? y0 + rows : h;
		PFB_CUDA(cudaStreamWaitEvent(c.st, c.ev_up[b], 0));
		uint32_t *to = direct ? direct : dout;
		int rc = (op == OP_GAUSSIAN) ? launch_gaussian(din, to, w, h, 1, taps, c.st, y0, y1)
		                             : launch_sobel(din, to, w, h, 1, c.st, y0, y1);
		if (rc)
			return rc;
		if (direct)
			continue;
		PFB_CUDA(cudaEventRecord(c.ev_done[b], c.st));
		PFB_CUDA(cudaStreamWaitEvent(c.st_down, c.ev_done[b], 0));
		PFB_CUDA(cudaMemcpyAsync(dst + (size_t)y0 * w, dout + (size_t)y0 * w,
				(size_t)(y1 - y0) * w * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.st_down));
	}
	PFB_CUDA(cudaStreamSynchronize(direct ? c.st : c.st_down));
	return PF_B200_OK;
}

// Canny on pinned caller memory.  Classification is row-local, so bands are
// classified as they arrive and their rows go back at once; hysteresis, which
// needs the whole page, then changes only the weak pixels (a few thousand on a
// document page), and those alone are written into the caller's page again by
// a small kernel.  PCIe carries both directions at once and nothing waits for
// a second pass over the page.  A page with very many weak pixels takes a
// second full download instead.
static int run_canny_streamed(Context &c, const struct pf_bitmap *in, struct pf_bitmap *out,
		uint32_t *mirror)
{
	const int w = in->size.x, h = in->size.y;
	const size_t bytes = (size_t)w * h * sizeof(uint32_t);
	const int rows = band_rows_for(h, 3);
	const int nb = (h + rows - 1) / rows;
	int rc;
	if ((rc = grow_canny_scratch(c, w, h)) || (rc = grow(c.h_stage, 64, true)))
		return rc;
	const uint32_t *src = reinterpret_cast<const uint32_t *>(in->pixels);
	uint32_t *dst = reinterpret_cast<uint32_t *>(out->pixels);
	const uint32_t *din = static_cast<const uint32_t *>(c.d_in.p);
	uint32_t *dout = static_cast<uint32_t *>(c.d_out.p);
	int *weak_count = static_cast<int *>(c.h_stage.p);
	if ((rc = launch_canny_begin(w, h, c.d_scratch.p, c.d_scratch.cap, c.st, true)))
		return rc;
	int rc_up = upload_bands(c, const_cast<uint32_t *>(din), src, w, h, rows, nb);
	if (rc_up)
		return rc_up;
	for (int b = 0; b < nb; b++) {
		const int y0 = b * rows, y1 = (y0 + rows < h) ? y0 + rows : h;
		PFB_CUDA(cudaStreamWaitEvent(c.st, c.ev_up[b], 0));
		if ((rc = launch_canny_classify_rows(din, dout, w, h, y0, y1, canny_thresholds(),
				c.d_scratch.p, c.st)))
			return rc;
		PFB_CUDA(cudaEventRecord(c.ev_done[b], c.st));
		PFB_CUDA(cudaStreamWaitEvent(c.st_down, c.ev_done[b], 0));
		PFB_CUDA(cudaMemcpyAsync(dst + (size_t)y0 * w, dout + (size_t)y0 * w,
				(size_t)(y1 - y0) * w * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.st_down));
	}
	if ((rc = launch_canny_resolve(dout, w, h, c.d_scratch.p, c.st)))
		return rc;
	// (header word 3: the length of the list just resolved)
	PFB_CUDA(cudaMemcpyAsync(weak_count, static_cast<int *>(c.d_scratch.p) + 3, sizeof(int),
			cudaMemcpyDeviceToHost, c.st));
	// what follows on the kernel stream comes after the last band's download
	PFB_CUDA(cudaEventRecord(c.ev_up[0], c.st_down));
	PFB_CUDA(cudaStreamWaitEvent(c.st, c.ev_up[0], 0));
	PFB_CUDA(cudaStreamSynchronize(c.st));
	if (*weak_count == 0)
		return PF_B200_OK;
	if ((size_t)*weak_count <= (size_t)w * h / 64) {
		if ((rc = launch_canny_patch(dout, mirror, c.d_scratch.p, c.st)))
			return rc;
	} else {
		PFB_CUDA(cudaMemcpyAsync(dst, dout, bytes, cudaMemcpyDeviceToHost, c.st));
	}
	PFB_CUDA(cudaStreamSynchronize(c.st));
	return PF_B200_OK;
}

// Gaussian / Sobel on ordinary (pageable) caller memory, both sides -- what a
// C program that mallocs its bitmaps gets.  The same band pipeline as on
// pinned pages, fed and drained through two page-sized pinned staging areas:
// host threads copy the input bands in (streaming stores), the calling thread
// queues each band's DMA as soon as it is staged and each band's kernel as
// soon as the band below it is on the device, the kernels write their rows
// straight into the second staging area, and the host threads copy every band
// out to the caller as soon as its kernel has finished.  Host copies, PCIe in
// both directions and the kernels overlap.
static int run_stencil_pageable(Context &c, Op op, const struct pf_bitmap *in,
		struct pf_bitmap *out, const Taps &taps)
{
	const int w = in->size.x, h = in->size.y;
	const size_t bytes = (size_t)w * h * sizeof(uint32_t);
	int rows = (h + kMaxBands - 1) / kMaxBands;
	if (rows < 128)
		rows = 128;
	const int nb = (h + rows - 1) / rows;
	int rc = grow(c.h_stage, 2 * bytes + 256, true);
	if (rc)
		return rc;
	char *stage_in = static_cast<char *>(c.h_stage.p);
	char *stage_out = stage_in + ((bytes + 255) & ~(size_t)255);
	// pinned caller memory on the way back (the Python layer's result arrays):
	// the kernels write into it directly and nothing is staged out
	uint32_t *direct = static_cast<uint32_t *>(device_alias(out->pixels));
	const bool stage_back = direct == nullptr;
	if (stage_back)
		direct = static_cast<uint32_t *>(device_alias(stage_out));
	if (!direct)
		return fail(PF_B200_ECUDA, "the pinned staging area is not mapped for the device");
	const char *src = reinterpret_cast<const char *>(in->pixels);
	char *dst = reinterpret_cast<char *>(out->pixels);
	const uint32_t *din = static_cast<const uint32_t *>(c.d_in.p);

	std::atomic<int> in_ready[kMaxBands], out_ready[kMaxBands];
	for (int b = 0; b < nb; b++) {
		in_ready[b].store(0);
		out_ready[b].store(0);
	}
	std::atomic<bool> failed(false);
	cudaError_t sync_err[kMaxBands] = {};
	auto band_off = [&](int b) { return (size_t)b * rows * w * sizeof(uint32_t); };
	auto band_len = [&](int b) {
		const int y0 = b * rows, y1 = (y0 + rows < h) ? y0 + rows : h;
		return (size_t)(y1 - y0) * w * sizeof(uint32_t);
	};
	const int nthreads = copy_threads();
	auto work = [&](int t) {
		cudaSetDevice(c.dev);
		for (int b = t; b < nb; b += nthreads) {
			copy_slice(stage_in + band_off(b), src + band_off(b), band_len(b));
			in_ready[b].store(1, std::memory_order_release);
		}
		for (int b = t; stage_back && b < nb; b += nthreads) {
			while (!out_ready[b].load(std::memory_order_acquire)) {
				if (failed.load())
					return;
				std::this_thread::yield();
			}
			sync_err[b] = cudaEventSynchronize(c.ev_done[b]);
			if (sync_err[b] == cudaSuccess)
				copy_slice(dst + band_off(b), stage_out + band_off(b), band_len(b));
		}
	};
	c.pool.ensure(nthreads);
	c.pool.start([&](int t) {
		if (t < nthreads)
			work(t);
	});

	// the calling thread drives the device
	auto launch_band = [&](int b) -> int {
		const int y0 = b * rows, y1 = (y0 + rows < h) ? y0 + rows : h;
		cudaError_t e = cudaStreamWaitEvent(c.st, c.ev_up[b + 1 < nb ? b + 1 : b], 0);
		if (e != cudaSuccess)
			return fail(PF_B200_ECUDA, "%s", cudaGetErrorString(e));
		int r = (op == OP_GAUSSIAN) ? launch_gaussian(din, direct, w, h, 1, taps, c.st, y0, y1)
		                            : launch_sobel(din, direct, w, h, 1, c.st, y0, y1);
		if (r)
			return r;
		e = cudaEventRecord(c.ev_done[b], c.st);
		if (e != cudaSuccess)
			return fail(PF_B200_ECUDA, "%s", cudaGetErrorString(e));
		out_ready[b].store(1, std::memory_order_release);
		return PF_B200_OK;
	};
	for (int b = 0; b < nb && !rc; b++) {
		while (!in_ready[b].load(std::memory_order_acquire))
			std::this_thread::yield();
		cudaError_t e = cudaMemcpyAsync(const_cast<uint32_t *>(din) + band_off(b) / sizeof(uint32_t),
				stage_in + band_off(b), band_len(b), cudaMemcpyHostToDevice, c.st_up);
		if (e == cudaSuccess)
			e = cudaEventRecord(c.ev_up[b], c.st_up);
		if (e != cudaSuccess)
			rc = fail(PF_B200_ECUDA, "upload: %s", cudaGetErrorString(e));
		if (!rc && b >= 1)
			rc = launch_band(b - 1);
	}
	if (!rc)
		rc = launch_band(nb - 1);
	if (rc)
		failed.store(true);
	c.pool.wait();
	if (!rc)
		for (int b = 0; b < nb; b++)
			if (sync_err[b] != cudaSuccess)
				rc = fail(PF_B200_ECUDA, "download: %s", cudaGetErrorString(sync_err[b]));
	// nothing of this call may still be running when the staging areas are reused
	cudaStreamSynchronize(c.st);
	cudaStreamSynchronize(c.st_up);
	return rc;
}

static int run_stencil(Op op, const struct pf_bitmap *in, struct pf_bitmap *out,
		double sigma, int nb_stddev)
{
	CallScope scope;
	clear_error();
	int rc = check_bitmaps(in, out);
	if (rc)
		return rc;
	Taps taps;
	if (op == OP_GAUSSIAN && (rc = make_taps(sigma, nb_stddev, taps)))
		return rc;
	Context &c = t_ctx;
	if ((rc = ctx_init(c)))
		return rc;
	const int w = in->size.x, h = in->size.y;
	const size_t bytes = (size_t)w * h * sizeof(uint32_t);
	if ((rc = grow(c.d_in, bytes, false)) || (rc = grow(c.d_out, bytes, false)))
		return rc;
	// (bands are >= 256 rows, further than any stencil here reaches: <= 31 rows)
	if (h >= 512 && is_pinned(in->pixels) && is_pinned(out->pixels)) {
		if (op != OP_CANNY)
			return run_stencil_streamed(c, op, in, out, taps);
		// the weak pixels are written into the caller's page by a kernel
		if (uint32_t *mirror = static_cast<uint32_t *>(device_alias(out->pixels)))
			return run_canny_streamed(c, in, out, mirror);
	}
	static const bool pipe_env = !getenv("PF_B200_PAGEABLE_PIPELINE") || atoi(getenv("PF_B200_PAGEABLE_PIPELINE"));
	if (pipe_env && op != OP_CANNY && h >= 512 && bytes >= kParallelMin && bytes <= kParallelMax
			&& !is_pinned(in->pixels))
		return run_stencil_pageable(c, op, in, out, taps);
	if ((rc = upload(c, c.d_in.p, in->pixels, bytes)))
		return rc;
	const uint32_t *din = static_cast<const uint32_t *>(c.d_in.p);
	uint32_t *dout = static_cast<uint32_t *>(c.d_out.p);
	switch (op) {
	case OP_GAUSSIAN:
		rc = launch_gaussian(din, dout, w, h, 1, taps, c.st);
		break;
	case OP_SOBEL:
		rc = launch_sobel(din, dout, w, h, 1, c.st);
		break;
	case OP_CANNY: {
		if ((rc = grow_canny_scratch(c, w, h)))
			return rc;
		rc = launch_canny(din, dout, w, h, 1, canny_thresholds(),
				c.d_scratch.p, c.d_scratch.cap, c.st, true);
		break;
	}
	}
	if (rc)
		return rc;
	return download(c, out->pixels, dout, bytes);
}

// ACE on ordinary (pageable) memory on both sides: the same pipeline as
// run_stencil_pageable.  Bands are staged in by the pool and accumulated as
// they arrive (the samples' pixel values go ahead); after the page-wide
// extremes are known each band is rescaled and its uncertain pixels redone
// (launch_ace_scale / launch_ace_fixup) into the second staging area, and the
// pool copies each band out as soon as it is there.
static int run_ace_pageable(Context &c, const struct pf_bitmap *in, struct pf_bitmap *out,
		AceParams p, const int *sample_xy, const int *d_xy, uint32_t *d_spx,
		double *minmax, const AceScores &sc)
{
	const int w = in->size.x, h = in->size.y;
	const size_t npix = (size_t)w * h;
	const size_t bytes = npix * sizeof(uint32_t);
	int rows = (h + kMaxBands - 1) / kMaxBands;
	if (rows < 128)
		rows = 128;
	const int nb = (h + rows - 1) / rows;
	int rc = grow(c.h_stage, 2 * bytes + 256, true);
	if (rc)
		return rc;
	char *stage_in = static_cast<char *>(c.h_stage.p);
	char *stage_out = stage_in + ((bytes + 255) & ~(size_t)255);
	// pinned caller memory on the way back (the Python layer's result arrays):
	// the kernels write into it directly and nothing is staged out
	uint32_t *direct = static_cast<uint32_t *>(device_alias(out->pixels));
	const bool stage_back = direct == nullptr;
	if (stage_back)
		direct = static_cast<uint32_t *>(device_alias(stage_out));
	if (!direct)
		return fail(PF_B200_ECUDA, "the pinned staging area is not mapped for the device");
	const uint32_t *src32 = reinterpret_cast<const uint32_t *>(in->pixels);
	const char *src = reinterpret_cast<const char *>(in->pixels);
	char *dst = reinterpret_cast<char *>(out->pixels);
	const uint32_t *din = static_cast<const uint32_t *>(c.d_in.p);

	uint32_t spx[4096];
	for (int i = 0; i < p.nb_samples; i++)
		spx[i] = src32[(size_t)sample_xy[2 * i + 1] * w + sample_xy[2 * i]];
	// (pageable source: the runtime stages it before returning)
	PFB_CUDA(cudaMemcpyAsync(d_spx, spx, sizeof(uint32_t) * (size_t)p.nb_samples,
			cudaMemcpyHostToDevice, c.st));
	p.sample_px = d_spx;

	std::atomic<int> in_ready[kMaxBands], out_ready[kMaxBands];
	for (int b = 0; b < nb; b++) {
		in_ready[b].store(0);
		out_ready[b].store(0);
	}
	std::atomic<bool> failed(false);
	cudaError_t sync_err[kMaxBands] = {};
	auto band_off = [&](int b) { return (size_t)b * rows * w * sizeof(uint32_t); };
	auto band_len = [&](int b) {
		const int y0 = b * rows, y1 = (y0 + rows < h) ? y0 + rows : h;
		return (size_t)(y1 - y0) * w * sizeof(uint32_t);
	};
	const int nthreads = copy_threads();
	auto work = [&](int t) {
		cudaSetDevice(c.dev);
		for (int b = t; b < nb; b += nthreads) {
			copy_slice(stage_in + band_off(b), src + band_off(b), band_len(b));
			in_ready[b].store(1, std::memory_order_release);
		}
		for (int b = t; stage_back && b < nb; b += nthreads) {
			while (!out_ready[b].load(std::memory_order_acquire)) {
				if (failed.load())
					return;
				std::this_thread::yield();
			}
			sync_err[b] = cudaEventSynchronize(c.ev_done[b]);
			if (sync_err[b] == cudaSuccess)
				copy_slice(dst + band_off(b), stage_out + band_off(b), band_len(b));
		}
	};
	c.pool.ensure(nthreads);
	c.pool.start([&](int t) {
		if (t < nthreads)
			work(t);
	});

	auto cuda_rc = [&](cudaError_t e, const char *what) {
		return e == cudaSuccess ? PF_B200_OK : fail(PF_B200_ECUDA, "%s: %s", what, cudaGetErrorString(e));
	};
	for (int b = 0; b < nb && !rc; b++) {
		while (!in_ready[b].load(std::memory_order_acquire))
			std::this_thread::yield();
		rc = cuda_rc(cudaMemcpyAsync(const_cast<uint32_t *>(din) + band_off(b) / sizeof(uint32_t),
				stage_in + band_off(b), band_len(b), cudaMemcpyHostToDevice, c.st_up), "upload");
		if (!rc)
			rc = cuda_rc(cudaEventRecord(c.ev_up[b], c.st_up), "upload");
		if (!rc)
			rc = cuda_rc(cudaStreamWaitEvent(c.st, c.ev_up[b], 0), "upload");
		if (!rc) {
			p.row_begin = b * rows;
			p.row_end = (p.row_begin + rows < h) ? p.row_begin + rows : h;
			rc = launch_ace_scores(din, d_xy, p, sc, minmax, c.st);
		}
	}
	p.row_begin = 0;
	p.row_end = h;
	if (!rc)
		rc = launch_ace_refine(din, d_xy, p, sc, minmax, c.st);
	for (int b = 0; b < nb && !rc; b++) {
		const int y0 = b * rows, y1 = (y0 + rows < h) ? y0 + rows : h;
		uint32_t *to = direct + (size_t)y0 * w;
		rc = launch_ace_scale(p, sc, minmax, y0, y1 - y0, b, &to, 1, false, c.st);
		if (!rc)
			rc = launch_ace_fixup(din, d_xy, p, sc, minmax, y0, b, &to, 1, false, c.st);
		if (!rc)
			rc = cuda_rc(cudaEventRecord(c.ev_done[b], c.st), "rescale");
		if (!rc)
			out_ready[b].store(1, std::memory_order_release);
	}
	if (rc)
		failed.store(true);
	c.pool.wait();
	if (!rc)
		for (int b = 0; b < nb; b++)
			if (sync_err[b] != cudaSuccess)
				rc = fail(PF_B200_ECUDA, "download: %s", cudaGetErrorString(sync_err[b]));
	cudaStreamSynchronize(c.st);
	cudaStreamSynchronize(c.st_up);
	return rc;
}

static int run_ace_impl(const struct pf_bitmap *in, struct pf_bitmap *out,
		int nb_samples, const int *sample_xy, double slope, double limit)
{
	clear_error();
	int rc = check_bitmaps(in, out);
	if (rc)
		return rc;
	if (nb_samples < 1 || !sample_xy)
		return fail(PF_B200_EINVAL, "ace: nb_samples %d", nb_samples);
	const int w = in->size.x, h = in->size.y;
	for (int i = 0; i < nb_samples; i++)
		if (sample_xy[2 * i] < 0 || sample_xy[2 * i] >= w
				|| sample_xy[2 * i + 1] < 0 || sample_xy[2 * i + 1] >= h)
			return fail(PF_B200_EINVAL, "ace: sample %d (%d, %d) outside the page",
					i, sample_xy[2 * i], sample_xy[2 * i + 1]);
	Context &c = t_ctx;
	if ((rc = ctx_init(c)))
		return rc;
	const size_t npix = (size_t)w * h;
	const size_t bytes = npix * sizeof(uint32_t);
	const size_t mm_bytes = (ace_minmax_doubles(w, h) * sizeof(double) + 63) / 64 * 64;
	const size_t small = mm_bytes + sizeof(int) * 3 * (size_t)nb_samples;
	if ((rc = grow(c.d_in, bytes, false)) || (rc = grow(c.d_out, bytes, false))
			|| (rc = grow(c.d_scores, ace_scores_bytes(w, h), false))
			|| (rc = grow(c.d_small, small, false)))
		return rc;
	double *minmax = static_cast<double *>(c.d_small.p);
	int *d_xy = reinterpret_cast<int *>(static_cast<char *>(c.d_small.p) + mm_bytes);
	uint32_t *d_spx = reinterpret_cast<uint32_t *>(d_xy + 2 * (size_t)nb_samples);
	const uint32_t *din = static_cast<const uint32_t *>(c.d_in.p);
	uint32_t *dout = static_cast<uint32_t *>(c.d_out.p);
	// one score block for the page: bands of a call share its planes
	const AceScores sc = ace_scores_at(c.d_scores.p, npix);
	// the table is tiny; pageable copies are staged by the runtime before it
	// returns, so the caller's array is free to reuse
	PFB_CUDA(cudaMemcpyAsync(d_xy, sample_xy, sizeof(int) * 2 * (size_t)nb_samples,
			cudaMemcpyHostToDevice, c.st));
	AceParams p = { w, h, 0, h, nb_samples, slope, limit };
	p.score_row0 = 0;
	p.score_plane = (long)npix;
	if ((rc = launch_ace_minmax_init(minmax, w, h, c.st)))
		return rc;

	if (h >= 512 && nb_samples <= 4096 && is_pinned(in->pixels) && is_pinned(out->pixels)) {
		// Streamed: with the samples' pixel values sent ahead (gathered here
		// from the caller's page), a band of rows can be accumulated as soon
		// as it has arrived.  The rescale needs the page-wide extremes, so it
		// starts after the last band; the finished page goes back in one DMA.
		const uint32_t *src = reinterpret_cast<const uint32_t *>(in->pixels);
		uint32_t *dst = reinterpret_cast<uint32_t *>(out->pixels);
		uint32_t spx[4096];
		for (int i = 0; i < nb_samples; i++)
			spx[i] = src[(size_t)sample_xy[2 * i + 1] * w + sample_xy[2 * i]];
		PFB_CUDA(cudaMemcpyAsync(d_spx, spx, sizeof(uint32_t) * (size_t)nb_samples,
				cudaMemcpyHostToDevice, c.st));
		const int rows = band_rows_for(h);
		const int nb = (h + rows - 1) / rows;
		for (int b = 0; b < nb; b++) {
			const int y0 = b * rows, y1 = (y0 + rows < h) ? y0 + rows : h;
			PFB_CUDA(cudaMemcpyAsync(const_cast<uint32_t *>(din) + (size_t)y0 * w,
					src + (size_t)y0 * w, (size_t)(y1 - y0) * w * sizeof(uint32_t),
					cudaMemcpyHostToDevice, c.st_up));
			PFB_CUDA(cudaEventRecord(c.ev_up[b], c.st_up));
		}
		p.sample_px = d_spx;
		for (int b = 0; b < nb; b++) {
			p.row_begin = b * rows;
			p.row_end = (p.row_begin + rows < h) ? p.row_begin + rows : h;
			PFB_CUDA(cudaStreamWaitEvent(c.st, c.ev_up[b], 0));
			if ((rc = launch_ace_scores(din, d_xy, p, sc, minmax, c.st)))
				return rc;
		}
		p.row_begin = 0;
		p.row_end = h;
		if ((rc = launch_ace_refine(din, d_xy, p, sc, minmax, c.st)))
			return rc;
		// (the rescale could write into the caller's page like the stencils do, but
		// it is so short that its stores would run at PCIe speed with nothing to
		// overlap them: one DMA of the finished page is faster, 1.48 against 1.62 ms)
		if ((rc = launch_ace_scale(p, sc, minmax, 0, h, 0, &dout, 1, false, c.st))
				|| (rc = launch_ace_fixup(din, d_xy, p, sc, minmax, 0, 0, &dout, 1, false, c.st)))
			return rc;
		PFB_CUDA(cudaMemcpyAsync(dst, dout, bytes, cudaMemcpyDeviceToHost, c.st));
		PFB_CUDA(cudaStreamSynchronize(c.st));
		return PF_B200_OK;
	}

	static const bool pipe_env = !getenv("PF_B200_PAGEABLE_PIPELINE") || atoi(getenv("PF_B200_PAGEABLE_PIPELINE"));
	if (pipe_env && h >= 512 && nb_samples <= 4096 && bytes >= kParallelMin && bytes <= kParallelMax
			&& !is_pinned(in->pixels))
		return run_ace_pageable(c, in, out, p, sample_xy, d_xy, d_spx, minmax, sc);
	if ((rc = upload(c, c.d_in.p, in->pixels, bytes)))
		return rc;
	if ((rc = launch_ace_scores(din, d_xy, p, sc, minmax, c.st))
			|| (rc = launch_ace_refine(din, d_xy, p, sc, minmax, c.st))
			|| (rc = launch_ace_scale(p, sc, minmax, 0, h, 0, &dout, 1, false, c.st))
			|| (rc = launch_ace_fixup(din, d_xy, p, sc, minmax, 0, 0, &dout, 1, false, c.st)))
		return rc;
	return download(c, out->pixels, c.d_out.p, bytes);
}

// ---------------------------------------------------------------------------
// The whole step on one page with ONE upload (an extension beside the drop-in
// entry points; include/pillowfight_b200.h, pf_b200_page_step): ACE, Gaussian,
// Sobel of the blurred page and Canny, any subset.  The four drop-in calls of
// such a step move the page up three times, the blurred page down and up
// again: 8 page transfers; this moves 1 + the results asked for.
// ---------------------------------------------------------------------------

struct StepArgs {
	const struct pf_bitmap *in;
	struct pf_bitmap *ace, *gauss, *sobel, *canny;
	int nb_samples;
	const int *sample_xy;
	double slope, limit, sigma;
	int nb_stddev;
};

// Pinned pages of at least 512 rows: bands go up on one stream; each band is
// accumulated (ACE), blurred and classified (Canny) as soon as it and the band
// below it have arrived; blurred and classified bands go back by DMA as they
// finish; Sobel runs on the blurred page in HBM and ACE's rescale after the
// page-wide extremes are known, both writing straight into the caller's pages.
static int run_step_streamed(Context &c, const StepArgs &a, const Taps &taps)
{
	const int w = a.in->size.x, h = a.in->size.y;
	const size_t npix = (size_t)w * h;
	const size_t bytes = npix * sizeof(uint32_t);
	int rc;
	if ((rc = grow(c.d_in, bytes, false)))
		return rc;
	const bool want_blur = a.gauss || a.sobel;
	if (want_blur && (rc = grow(c.d_blur, bytes, false)))
		return rc;
	if (a.canny && ((rc = grow(c.d_out, bytes, false)) || (rc = grow_canny_scratch(c, w, h))
			|| (rc = grow(c.h_stage, 64, true))))
		return rc;
	const size_t mm_bytes = (ace_minmax_doubles(w, h) * sizeof(double) + 63) / 64 * 64;
	if (a.ace && ((rc = grow(c.d_scores, ace_scores_bytes(w, h), false))
			|| (rc = grow(c.d_small, mm_bytes + sizeof(int) * 3 * (size_t)a.nb_samples, false))))
		return rc;
	const uint32_t *src = reinterpret_cast<const uint32_t *>(a.in->pixels);
	const uint32_t *din = static_cast<const uint32_t *>(c.d_in.p);
	uint32_t *dblur = static_cast<uint32_t *>(c.d_blur.p);
	uint32_t *dcanny = static_cast<uint32_t *>(c.d_out.p);
	double *minmax = static_cast<double *>(c.d_small.p);
	int *d_xy = a.ace ? reinterpret_cast<int *>(static_cast<char *>(c.d_small.p) + mm_bytes) : nullptr;
	uint32_t *d_spx = a.ace ? reinterpret_cast<uint32_t *>(d_xy + 2 * (size_t)a.nb_samples) : nullptr;
	AceScores sc = {};
	AceParams p = { w, h, 0, h, a.nb_samples, a.slope, a.limit };
	if (a.ace) {
		sc = ace_scores_at(c.d_scores.p, npix);
		p.score_row0 = 0;
		p.score_plane = (long)npix;
		uint32_t spx[4096];
		for (int i = 0; i < a.nb_samples; i++)
			spx[i] = src[(size_t)a.sample_xy[2 * i + 1] * w + a.sample_xy[2 * i]];
		PFB_CUDA(cudaMemcpyAsync(d_xy, a.sample_xy, sizeof(int) * 2 * (size_t)a.nb_samples,
				cudaMemcpyHostToDevice, c.st));
		PFB_CUDA(cudaMemcpyAsync(d_spx, spx, sizeof(uint32_t) * (size_t)a.nb_samples,
				cudaMemcpyHostToDevice, c.st));
		p.sample_px = d_spx;
		if ((rc = launch_ace_minmax_init(minmax, w, h, c.st)))
			return rc;
	}
	if (a.canny && (rc = launch_canny_begin(w, h, c.d_scratch.p, c.d_scratch.cap, c.st, true)))
		return rc;

	const int rows = band_rows_for(h);
	const int nb = (h + rows - 1) / rows;
	int rc_up = upload_bands(c, const_cast<uint32_t *>(din), src, w, h, rows, nb);
	if (rc_up)
		return rc_up;
	uint32_t *host_gauss = a.gauss ? reinterpret_cast<uint32_t *>(a.gauss->pixels) : nullptr;
	uint32_t *host_canny = a.canny ? reinterpret_cast<uint32_t *>(a.canny->pixels) : nullptr;
	for (int b = 0; b < nb; b++) {
		const int y0 = b * rows, y1 = (y0 + rows < h) ? y0 + rows : h;
		const size_t off = (size_t)y0 * w, len = (size_t)(y1 - y0) * w * sizeof(uint32_t);
		PFB_CUDA(cudaStreamWaitEvent(c.st, c.ev_up[b], 0));
		if (a.ace) {
			p.row_begin = y0;
			p.row_end = y1;
			if ((rc = launch_ace_scores(din, d_xy, p, sc, minmax, c.st)))
				return rc;
		}
		if (want_blur && (rc = launch_gaussian(din, dblur, w, h, 1, taps, c.st, y0, y1)))
			return rc;
		if (a.canny && (rc = launch_canny_classify_rows(din, dcanny, w, h, y0, y1, canny_thresholds(),
				c.d_scratch.p, c.st)))
			return rc;
		if (a.gauss || a.canny) {
			PFB_CUDA(cudaEventRecord(c.ev_done[b], c.st));
			PFB_CUDA(cudaStreamWaitEvent(c.st_down, c.ev_done[b], 0));
			if (a.gauss)
				PFB_CUDA(cudaMemcpyAsync(host_gauss + off, dblur + off, len, cudaMemcpyDeviceToHost, c.st_down));
			if (a.canny)
				PFB_CUDA(cudaMemcpyAsync(host_canny + off, dcanny + off, len, cudaMemcpyDeviceToHost, c.st_down));
		}
	}
	int *weak_count = static_cast<int *>(c.h_stage.p);
	if (a.canny) {
		if ((rc = launch_canny_resolve(dcanny, w, h, c.d_scratch.p, c.st)))
			return rc;
		PFB_CUDA(cudaMemcpyAsync(weak_count, static_cast<int *>(c.d_scratch.p) + 3, sizeof(int),
				cudaMemcpyDeviceToHost, c.st));
		PFB_CUDA(cudaEventRecord(c.ev_done[0], c.st));
	}
	if (a.sobel) {
		uint32_t *to = static_cast<uint32_t *>(device_alias(a.sobel->pixels));
		if (!to)
			return fail(PF_B200_ECUDA, "the sobel page is not mapped for the device");
		if ((rc = launch_sobel(dblur, to, w, h, 1, c.st)))
			return rc;
	}
	if (a.ace) {
		uint32_t *to = static_cast<uint32_t *>(device_alias(a.ace->pixels));
		if (!to)
			return fail(PF_B200_ECUDA, "the ace page is not mapped for the device");
		p.row_begin = 0;
		p.row_end = h;
		if ((rc = launch_ace_refine(din, d_xy, p, sc, minmax, c.st))
				|| (rc = launch_ace_scale(p, sc, minmax, 0, h, 0, &to, 1, false, c.st))
				|| (rc = launch_ace_fixup(din, d_xy, p, sc, minmax, 0, 0, &to, 1, false, c.st)))
			return rc;
	}
	if (a.canny) {
		// the weak pixels' verdicts into the caller's page, after its bands have landed
		PFB_CUDA(cudaEventSynchronize(c.ev_done[0]));
		PFB_CUDA(cudaStreamSynchronize(c.st_down));
		if (*weak_count > 0) {
			uint32_t *mirror = static_cast<uint32_t *>(device_alias(a.canny->pixels));
			if (mirror && (size_t)*weak_count <= npix / 64) {
				if ((rc = launch_canny_patch(dcanny, mirror, c.d_scratch.p, c.st)))
					return rc;
			} else {
				PFB_CUDA(cudaMemcpyAsync(host_canny, dcanny, bytes, cudaMemcpyDeviceToHost, c.st));
			}
		}
	}
	PFB_CUDA(cudaStreamSynchronize(c.st_down));
	PFB_CUDA(cudaStreamSynchronize(c.st));
	return PF_B200_OK;
}

static int run_ace(const struct pf_bitmap *in, struct pf_bitmap *out,
		int nb_samples, const int *sample_xy, double slope, double limit);

static int run_step(const StepArgs &a)
{
	CallScope scope;
	clear_error();
	if (!a.in || !a.in->pixels || a.in->size.x <= 0 || a.in->size.y <= 0)
		return fail(PF_B200_EINVAL, "NULL or empty input bitmap");
	struct pf_bitmap *outs[4] = { a.ace, a.gauss, a.sobel, a.canny };
	bool any = false, pinned = is_pinned(a.in->pixels);
	for (int i = 0; i < 4; i++) {
		if (!outs[i])
			continue;
		any = true;
		int rc = check_bitmaps(a.in, outs[i]);
		if (rc)
			return rc;
		if (outs[i]->pixels == a.in->pixels)
			return fail(PF_B200_EINVAL, "page step: an output page is the input page");
		for (int j = 0; j < i; j++)
			if (outs[j] && outs[j]->pixels == outs[i]->pixels)
				return fail(PF_B200_EINVAL, "page step: two outputs share a page");
		pinned = pinned && is_pinned(outs[i]->pixels);
	}
	if (!any)
		return fail(PF_B200_EINVAL, "page step: no output asked for");
	if (a.ace && (a.nb_samples < 1 || !a.sample_xy))
		return fail(PF_B200_EINVAL, "ace: nb_samples %d", a.nb_samples);
	const int w = a.in->size.x, h = a.in->size.y;
	if (a.ace)
		for (int i = 0; i < a.nb_samples; i++)
			if (a.sample_xy[2 * i] < 0 || a.sample_xy[2 * i] >= w
					|| a.sample_xy[2 * i + 1] < 0 || a.sample_xy[2 * i + 1] >= h)
				return fail(PF_B200_EINVAL, "ace: sample %d (%d, %d) outside the page",
						i, a.sample_xy[2 * i], a.sample_xy[2 * i + 1]);
	Taps taps = {};
	int rc;
	if ((a.gauss || a.sobel) && (rc = make_taps(a.sigma, a.nb_stddev, taps)))
		return rc;
	if (pinned && h >= 512 && (!a.ace || a.nb_samples <= 4096)) {
		Context &c = t_ctx;
		if ((rc = ctx_init(c)))
			return rc;
		return drained(c, run_step_streamed(c, a, taps));
	}
	// small or pageable pages: the drop-in calls one after the other (the blurred
	// page into a scratch page when the caller does not want it)
	if (a.ace && (rc = run_ace(a.in, a.ace, a.nb_samples, a.sample_xy, a.slope, a.limit)))
		return rc;
	if (a.gauss || a.sobel) {
		struct pf_bitmap tmp = *a.in;
		std::vector<uint32_t> scratch_page;
		struct pf_bitmap *blur = a.gauss;
		if (!blur) {
			scratch_page.resize((size_t)w * h);
			tmp.pixels = reinterpret_cast<union pf_pixel *>(scratch_page.data());
			blur = &tmp;
		}
		if ((rc = run_stencil(OP_GAUSSIAN, a.in, blur, a.sigma, a.nb_stddev)))
			return rc;
		if (a.sobel && (rc = run_stencil(OP_SOBEL, blur, a.sobel, 0.0, 0)))
			return rc;
	}
	if (a.canny && (rc = run_stencil(OP_CANNY, a.in, a.canny, 0.0, 0)))
		return rc;
	return PF_B200_OK;
}

static int run_ace(const struct pf_bitmap *in, struct pf_bitmap *out,
		int nb_samples, const int *sample_xy, double slope, double limit)
{
	CallScope scope;
	const int rc = run_ace_impl(in, out, nb_samples, sample_xy, slope, limit);
	return (rc != PF_B200_OK && t_ctx.st) ? drained(t_ctx, rc) : rc;
}

// ---------------------------------------------------------------------------
// ACE of one page with its row bands on several GPUs of this process (an
// extension; include/pillowfight_b200.h, pf_b200_ace_multi_gpu): what
// libpillowfight_b200/distributed.py does across processes, for a C caller
// that owns all the GPUs itself.  Each GPU is sent its band of rows over its
// own PCIe link (and the samples' pixel values), accumulates it as it arrives
// and finds the exact extremes of its band; the host folds the 12 doubles of
// every GPU and sends the result back -- the one exchange the algorithm has --
// and each GPU rescales its band and returns it.
// ---------------------------------------------------------------------------

struct DeviceSlot {
	int dev = -1;
	cudaStream_t st = nullptr, st_up = nullptr;
	cudaEvent_t ev_up[kMaxBands] = {};
	Buffer d_rows, d_out, d_scores, d_small;
	double *h_mm = nullptr;       // pinned, 12 doubles
};

struct MultiContext {
	std::vector<DeviceSlot> slots;
	~MultiContext()
	{
		for (DeviceSlot &s : slots) {
			if (s.dev < 0 || cudaSetDevice(s.dev) != cudaSuccess)
				continue;
			Buffer *bufs[] = { &s.d_rows, &s.d_out, &s.d_scores, &s.d_small };
			for (Buffer *b : bufs)
				if (b->p)
					cudaFree(b->p);
			if (s.h_mm)
				cudaFreeHost(s.h_mm);
			for (int i = 0; i < kMaxBands; i++)
				if (s.ev_up[i])
					cudaEventDestroy(s.ev_up[i]);
			if (s.st_up)
				cudaStreamDestroy(s.st_up);
			if (s.st)
				cudaStreamDestroy(s.st);
		}
	}
};

static thread_local MultiContext t_multi;

static int device_slot(int dev, DeviceSlot **out)
{
	for (DeviceSlot &s : t_multi.slots)
		if (s.dev == dev) {
			*out = &s;
			return PF_B200_OK;
		}
	PFB_CUDA(cudaSetDevice(dev));
	DeviceSlot s;
	s.dev = dev;
	PFB_CUDA(cudaStreamCreateWithFlags(&s.st, cudaStreamNonBlocking));
	PFB_CUDA(cudaStreamCreateWithFlags(&s.st_up, cudaStreamNonBlocking));
	for (int i = 0; i < kMaxBands; i++)
		PFB_CUDA(cudaEventCreateWithFlags(&s.ev_up[i], cudaEventDisableTiming));
	PFB_CUDA(cudaMallocHost(reinterpret_cast<void **>(&s.h_mm), 12 * sizeof(double)));
	t_multi.slots.push_back(s);
	*out = &t_multi.slots.back();
	return PF_B200_OK;
}

static int run_ace_multi_impl(const struct pf_bitmap *in, struct pf_bitmap *out, int nb_samples,
		const int *sample_xy, double slope, double limit, const int *devices, int n)
{
	const int w = in->size.x, h = in->size.y;
	const uint32_t *src = reinterpret_cast<const uint32_t *>(in->pixels);
	uint32_t *dst = reinterpret_cast<uint32_t *>(out->pixels);
	t_multi.slots.reserve(64);       // (pointers into the vector are held below)
	std::vector<DeviceSlot *> slot(n);
	int rc;
	for (int d = 0; d < n; d++)
		if ((rc = device_slot(devices[d], &slot[d])))
			return rc;
	std::vector<uint32_t> spx(nb_samples);
	for (int i = 0; i < nb_samples; i++)
		spx[i] = src[(size_t)sample_xy[2 * i + 1] * w + sample_xy[2 * i]];
	const size_t mm_bytes = (ace_minmax_doubles(w, h) * sizeof(double) + 63) / 64 * 64;
	struct Band {
		int b0, b1;
		AceParams p;
		AceScores sc;
		double *minmax;
		int *d_xy;
		const uint32_t *page;
	};
	std::vector<Band> band(n);

	// every GPU: its rows go up in a few pieces, each accumulated as it arrives
	for (int d = 0; d < n; d++) {
		DeviceSlot &s = *slot[d];
		Band &bd = band[d];
		bd.b0 = (int)((long)h * d / n);
		bd.b1 = (int)((long)h * (d + 1) / n);
		const int rows = bd.b1 - bd.b0;
		PFB_CUDA(cudaSetDevice(s.dev));
		const size_t band_bytes = (size_t)rows * w * sizeof(uint32_t);
		if ((rc = grow(s.d_rows, band_bytes, false)) || (rc = grow(s.d_out, band_bytes, false))
				|| (rc = grow(s.d_scores, ace_scores_bytes(w, rows), false))
				|| (rc = grow(s.d_small, mm_bytes + sizeof(int) * 3 * (size_t)nb_samples, false)))
			return rc;
		bd.minmax = static_cast<double *>(s.d_small.p);
		bd.d_xy = reinterpret_cast<int *>(static_cast<char *>(s.d_small.p) + mm_bytes);
		uint32_t *d_spx = reinterpret_cast<uint32_t *>(bd.d_xy + 2 * (size_t)nb_samples);
		PFB_CUDA(cudaMemcpyAsync(bd.d_xy, sample_xy, sizeof(int) * 2 * (size_t)nb_samples,
				cudaMemcpyHostToDevice, s.st));
		PFB_CUDA(cudaMemcpyAsync(d_spx, spx.data(), sizeof(uint32_t) * (size_t)nb_samples,
				cudaMemcpyHostToDevice, s.st));
		if ((rc = launch_ace_minmax_init(bd.minmax, w, h, s.st)))
			return rc;
		bd.sc = ace_scores_at(s.d_scores.p, (size_t)rows * w);
		// the kernels index the page; with the samples' values given they only
		// ever read this band's rows of it
		bd.page = static_cast<const uint32_t *>(s.d_rows.p) - (size_t)bd.b0 * w;
		bd.p = AceParams{ w, h, bd.b0, bd.b1, nb_samples, slope, limit };
		bd.p.sample_px = d_spx;
		bd.p.score_row0 = bd.b0;
		bd.p.score_plane = (long)rows * w;
		int piece = (rows + 3) / 4;
		if (piece < 256)
			piece = 256;
		int k = 0;
		for (int y0 = bd.b0; y0 < bd.b1; y0 += piece, k++) {
			const int y1 = (y0 + piece < bd.b1) ? y0 + piece : bd.b1;
			PFB_CUDA(cudaMemcpyAsync(static_cast<uint32_t *>(s.d_rows.p) + (size_t)(y0 - bd.b0) * w,
					src + (size_t)y0 * w, (size_t)(y1 - y0) * w * sizeof(uint32_t),
					cudaMemcpyHostToDevice, s.st_up));
			PFB_CUDA(cudaEventRecord(s.ev_up[k], s.st_up));
			PFB_CUDA(cudaStreamWaitEvent(s.st, s.ev_up[k], 0));
			AceParams pp = bd.p;
			pp.row_begin = y0;
			pp.row_end = y1;
			if ((rc = launch_ace_scores(bd.page, bd.d_xy, pp, bd.sc, bd.minmax, s.st)))
				return rc;
		}
		if ((rc = launch_ace_refine(bd.page, bd.d_xy, bd.p, bd.sc, bd.minmax, s.st)))
			return rc;
		PFB_CUDA(cudaMemcpyAsync(s.h_mm, bd.minmax + 6, 6 * sizeof(double), cudaMemcpyDeviceToHost, s.st));
	}
	// the one exchange: the exact extremes of the bands, folded on the host
	double ext[6] = { DBL_MAX, DBL_MAX, DBL_MAX, -DBL_MAX, -DBL_MAX, -DBL_MAX };
	for (int d = 0; d < n; d++) {
		PFB_CUDA(cudaSetDevice(slot[d]->dev));
		PFB_CUDA(cudaStreamSynchronize(slot[d]->st));
		for (int c = 0; c < 3; c++) {
			if (slot[d]->h_mm[c] < ext[c])
				ext[c] = slot[d]->h_mm[c];
			if (slot[d]->h_mm[3 + c] > ext[3 + c])
				ext[3 + c] = slot[d]->h_mm[3 + c];
		}
	}
	for (int d = 0; d < n; d++) {
		DeviceSlot &s = *slot[d];
		Band &bd = band[d];
		PFB_CUDA(cudaSetDevice(s.dev));
		for (int c = 0; c < 6; c++)
			s.h_mm[6 + c] = ext[c];
		PFB_CUDA(cudaMemcpyAsync(bd.minmax + 6, s.h_mm + 6, 6 * sizeof(double), cudaMemcpyHostToDevice, s.st));
		uint32_t *to = static_cast<uint32_t *>(s.d_out.p);
		if ((rc = launch_ace_scale(bd.p, bd.sc, bd.minmax, bd.b0, bd.b1 - bd.b0, 0, &to, 1, false, s.st))
				|| (rc = launch_ace_fixup(bd.page, bd.d_xy, bd.p, bd.sc, bd.minmax, bd.b0, 0, &to, 1, false, s.st)))
			return rc;
		PFB_CUDA(cudaMemcpyAsync(dst + (size_t)bd.b0 * w, to, (size_t)(bd.b1 - bd.b0) * w * sizeof(uint32_t),
				cudaMemcpyDeviceToHost, s.st));
	}
	for (int d = 0; d < n; d++) {
		PFB_CUDA(cudaSetDevice(slot[d]->dev));
		PFB_CUDA(cudaStreamSynchronize(slot[d]->st));
	}
	return PF_B200_OK;
}

static int run_ace_multi(const struct pf_bitmap *in, struct pf_bitmap *out, int nb_samples,
		const int *sample_xy, double slope, double limit, const int *devices, int nb_devices)
{
	CallScope scope;
	clear_error();
	int rc = check_bitmaps(in, out);
	if (rc)
		return rc;
	if (nb_samples < 1 || !sample_xy)
		return fail(PF_B200_EINVAL, "ace: nb_samples %d", nb_samples);
	const int w = in->size.x, h = in->size.y;
	for (int i = 0; i < nb_samples; i++)
		if (sample_xy[2 * i] < 0 || sample_xy[2 * i] >= w
				|| sample_xy[2 * i + 1] < 0 || sample_xy[2 * i + 1] >= h)
			return fail(PF_B200_EINVAL, "ace: sample %d (%d, %d) outside the page",
					i, sample_xy[2 * i], sample_xy[2 * i + 1]);
	int visible = 0;
	if (cudaGetDeviceCount(&visible) != cudaSuccess || visible <= 0)
		return fail(PF_B200_ENODEV, "no CUDA device is usable (%s); this build has no CPU path",
				cudaGetErrorString(cudaGetLastError()));
	std::vector<int> devs;
	if (devices) {
		for (int d = 0; d < nb_devices; d++) {
			if (devices[d] < 0 || devices[d] >= visible)
				return fail(PF_B200_EINVAL, "ace: device %d of %d visible", devices[d], visible);
			for (int e = 0; e < d; e++)
				if (devices[e] == devices[d])
					return fail(PF_B200_EINVAL, "ace: device %d listed twice", devices[d]);
			devs.push_back(devices[d]);
		}
	} else {
		for (int d = 0; d < (nb_devices > 0 && nb_devices < visible ? nb_devices : visible); d++)
			devs.push_back(d);
	}
	if (devs.empty() || devs.size() > 64)
		return fail(PF_B200_EINVAL, "ace: %zu devices", devs.size());
	// every band at least 256 rows; pinned pages (the bands go by DMA); in place is fine
	// only out of place here (a band's result returns while other bands still go up)
	int n = (int)devs.size();
	if (n > h / 256)
		n = h / 256;
	int cur = 0;
	PFB_CUDA(cudaGetDevice(&cur));
	if (n < 2 || !is_pinned(in->pixels) || !is_pinned(out->pixels) || in->pixels == out->pixels) {
		// nothing to split, or memory the DMA engines cannot reach: the one-GPU
		// call, on the calling thread's device
		return run_ace(in, out, nb_samples, sample_xy, slope, limit);
	}
	rc = run_ace_multi_impl(in, out, nb_samples, sample_xy, slope, limit, devs.data(), n);
	if (rc != PF_B200_OK)
		for (DeviceSlot &s : t_multi.slots)       // nothing of a failed call may still be in flight
			if (cudaSetDevice(s.dev) == cudaSuccess) {
				cudaStreamSynchronize(s.st_up);
				cudaStreamSynchronize(s.st);
			}
	cudaSetDevice(cur);
	return rc;
}

static void report(const char *who, int rc)
{
	if (rc == PF_B200_OK)
		return;
	fprintf(stderr, "libpillowfight_b200: %s failed (%d): %s\n", who, rc, t_msg);
	const char *e = getenv("PF_B200_ABORT_ON_ERROR");
	if (e && e[0] == '1')
		abort();
}

}  // namespace pfb

// ---------------------------------------------------------------------------
// extern "C"
// ---------------------------------------------------------------------------

using namespace pfb;

extern "C" {

int pf_b200_last_error(void) { return t_err; }
const char *pf_b200_last_error_string(void) { return t_msg; }

int pf_b200_device_count(void)
{
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess) {
		cudaGetLastError();
		return 0;
	}
	return n;
}

void pf_b200_ace_draw_samples(int nb_samples, int width, int height, int *sample_xy)
{
	for (int i = 0; i < nb_samples; i++) {
		sample_xy[2 * i] = rand() % width;
		sample_xy[2 * i + 1] = rand() % height;
	}
}

int pf_b200_ace(const struct pf_bitmap *in, struct pf_bitmap *out,
		int nb_samples, const int *sample_xy, double slope, double limit)
{
	return run_ace(in, out, nb_samples, sample_xy, slope, limit);
}

int pf_b200_gaussian(const struct pf_bitmap *in, struct pf_bitmap *out,
		double sigma, int nb_stddev)
{
	return run_stencil(OP_GAUSSIAN, in, out, sigma, nb_stddev);
}

int pf_b200_sobel(const struct pf_bitmap *in, struct pf_bitmap *out)
{
	return run_stencil(OP_SOBEL, in, out, 0.0, 0);
}

int pf_b200_canny(const struct pf_bitmap *in, struct pf_bitmap *out)
{
	return run_stencil(OP_CANNY, in, out, 0.0, 0);
}

int pf_b200_ace_multi_gpu(const struct pf_bitmap *in, struct pf_bitmap *out,
		int nb_samples, const int *sample_xy, double slope, double limit,
		const int *devices, int nb_devices)
{
	return run_ace_multi(in, out, nb_samples, sample_xy, slope, limit, devices, nb_devices);
}

int pf_b200_page_step(const struct pf_bitmap *in, struct pf_bitmap *out_ace,
		struct pf_bitmap *out_gaussian, struct pf_bitmap *out_sobel_of_gaussian,
		struct pf_bitmap *out_canny, int nb_samples, const int *sample_xy,
		double slope, double limit, double sigma, int nb_stddev)
{
	const StepArgs a = { in, out_ace, out_gaussian, out_sobel_of_gaussian, out_canny,
		nb_samples, sample_xy, slope, limit, sigma, nb_stddev };
	return run_step(a);
}

void pf_apply_ace(const struct pf_bitmap *in, struct pf_bitmap *out,
		int nb_samples, double slope, double limit, int nb_threads)
{
	(void)nb_threads;
	clear_error();
	if (nb_samples < 1 || !in || in->size.x <= 0 || in->size.y <= 0) {
		report("pf_apply_ace", fail(PF_B200_EINVAL, "ace: bad arguments"));
		return;
	}
	int *xy = static_cast<int *>(malloc(sizeof(int) * 2 * (size_t)nb_samples));
	if (!xy) {
		report("pf_apply_ace", fail(PF_B200_EINVAL, "ace: out of host memory"));
		return;
	}
	pf_b200_ace_draw_samples(nb_samples, in->size.x, in->size.y, xy);
	report("pf_apply_ace", run_ace(in, out, nb_samples, xy, slope, limit));
	free(xy);
}

void pf_apply_gaussian(const struct pf_bitmap *in, struct pf_bitmap *out,
		double sigma, int nb_stddev)
{
	report("pf_apply_gaussian", run_stencil(OP_GAUSSIAN, in, out, sigma, nb_stddev));
}

void pf_apply_sobel(const struct pf_bitmap *in, struct pf_bitmap *out)
{
	report("pf_apply_sobel", run_stencil(OP_SOBEL, in, out, 0.0, 0));
}

void pf_apply_canny(const struct pf_bitmap *in, struct pf_bitmap *out)
{
	report("pf_apply_canny", run_stencil(OP_CANNY, in, out, 0.0, 0));
}

void pf_ace(const struct pf_bitmap *in, struct pf_bitmap *out,
		int nb_samples, double slope, double limit, int nb_threads)
{
	pf_apply_ace(in, out, nb_samples, slope, limit, nb_threads);
}

void pf_gaussian(const struct pf_bitmap *in, struct pf_bitmap *out,
		double sigma, int nb_stddev)
{
	pf_apply_gaussian(in, out, sigma, nb_stddev);
}

void pf_sobel(const struct pf_bitmap *in, struct pf_bitmap *out)
{
	pf_apply_sobel(in, out);
}

void pf_canny(const struct pf_bitmap *in, struct pf_bitmap *out)
{
	pf_apply_canny(in, out);
}

void *pf_b200_host_alloc(size_t bytes)
{
	void *p = nullptr;
	if (cudaMallocHost(&p, bytes) != cudaSuccess) {
		fail(PF_B200_ECUDA, "cudaMallocHost(%zu): %s", bytes,
				cudaGetErrorString(cudaGetLastError()));
		return nullptr;
	}
	return p;
}

void pf_b200_host_free(void *ptr)
{
	if (ptr)
		cudaFreeHost(ptr);
}

// ---- device-resident entry points ----------------------------------------

static int check_dev(const void *a, const void *b, int w, int h, int pages)
{
	if (!a || !b)
		return fail(PF_B200_EINVAL, "NULL device pointer");
	if (w <= 0 || h <= 0 || pages <= 0)
		return fail(PF_B200_EINVAL, "size %d x %d x %d pages", w, h, pages);
	if (a == b)
		return fail(PF_B200_EINVAL, "device entry points do not run in place");
	return PF_B200_OK;
}

int pf_b200_dev_gaussian_batch(const uint32_t *d_in, uint32_t *d_out,
		int width, int height, int nb_pages, double sigma, int nb_stddev, void *stream)
{
	clear_error();
	int rc = check_dev(d_in, d_out, width, height, nb_pages);
	if (rc)
		return rc;
	Taps taps;
	if ((rc = make_taps(sigma, nb_stddev, taps)))
		return rc;
	for (int p0 = 0; p0 < nb_pages; p0 += 65535) {
		const int np = nb_pages - p0 < 65535 ? nb_pages - p0 : 65535;
		const size_t off = (size_t)p0 * width * height;
		if ((rc = launch_gaussian(d_in + off, d_out + off, width, height, np, taps,
				static_cast<cudaStream_t>(stream))))
			return rc;
	}
	return PF_B200_OK;
}

int pf_b200_dev_gaussian(const uint32_t *d_in, uint32_t *d_out,
		int width, int height, double sigma, int nb_stddev, void *stream)
{
	return pf_b200_dev_gaussian_batch(d_in, d_out, width, height, 1, sigma, nb_stddev, stream);
}

int pf_b200_dev_sobel_batch(const uint32_t *d_in, uint32_t *d_out,
		int width, int height, int nb_pages, void *stream)
{
	clear_error();
	int rc = check_dev(d_in, d_out, width, height, nb_pages);
	if (rc)
		return rc;
	for (int p0 = 0; p0 < nb_pages; p0 += 65535) {
		const int np = nb_pages - p0 < 65535 ? nb_pages - p0 : 65535;
		const size_t off = (size_t)p0 * width * height;
		if ((rc = launch_sobel(d_in + off, d_out + off, width, height, np,
				static_cast<cudaStream_t>(stream))))
			return rc;
	}
	return PF_B200_OK;
}

int pf_b200_dev_sobel(const uint32_t *d_in, uint32_t *d_out,
		int width, int height, void *stream)
{
	return pf_b200_dev_sobel_batch(d_in, d_out, width, height, 1, stream);
}

size_t pf_b200_canny_scratch_bytes(int width, int height)
{
	if (width <= 0 || height <= 0)
		return 0;
	return canny_scratch_bytes(width, height);
}

int pf_b200_dev_canny(const uint32_t *d_in, uint32_t *d_out,
		int width, int height, int nb_pages,
		void *d_scratch, size_t scratch_bytes, void *stream)
{
	clear_error();
	int rc = check_dev(d_in, d_out, width, height, nb_pages);
	if (rc)
		return rc;
	return launch_canny(d_in, d_out, width, height, nb_pages, canny_thresholds(),
			d_scratch, scratch_bytes, static_cast<cudaStream_t>(stream));
}

size_t pf_b200_ace_minmax_bytes(int width, int height)
{
	clear_error();
	if (width <= 0 || height <= 0) {
		fail(PF_B200_EINVAL, "size %d x %d", width, height);
		return 0;
	}
	return ace_minmax_doubles(width, height) * sizeof(double);
}

size_t pf_b200_ace_scores_bytes(int width, int nb_rows)
{
	clear_error();
	if (width <= 0 || nb_rows <= 0) {
		fail(PF_B200_EINVAL, "size %d x %d", width, nb_rows);
		return 0;
	}
	return ace_scores_bytes(width, nb_rows);
}

int pf_b200_dev_ace_minmax_init(double *d_minmax, int width, int height, void *stream)
{
	clear_error();
	if (!d_minmax)
		return fail(PF_B200_EINVAL, "NULL device pointer");
	if (width <= 0 || height <= 0)
		return fail(PF_B200_EINVAL, "size %d x %d", width, height);
	return launch_ace_minmax_init(d_minmax, width, height, static_cast<cudaStream_t>(stream));
}

int pf_b200_dev_ace_gather_samples(const uint32_t *d_page, int width, int height,
		const int *d_samples, int nb_samples, int *d_message, void *stream)
{
	clear_error();
	if (!d_page || !d_samples || !d_message)
		return fail(PF_B200_EINVAL, "NULL device pointer");
	if (width <= 0 || height <= 0 || nb_samples < 1)
		return fail(PF_B200_EINVAL, "size %d x %d, %d samples", width, height, nb_samples);
	return launch_ace_gather_samples(d_page, width, d_samples, nb_samples, d_message,
			static_cast<cudaStream_t>(stream));
}

// One pass over a band.  d_band points at row row_begin; with d_sample_px ==
// NULL the memory around it is the whole page (the samples are read from it).
enum AcePass { ACE_SCORES, ACE_REFINE, ACE_SCALE };

static int ace_band_pass(AcePass pass, const uint32_t *d_band, int width, int height,
		int row_begin, int row_end, const int *d_samples, const uint32_t *d_sample_px,
		int nb_samples, double slope, double limit, void *d_scores, double *d_minmax,
		uint32_t *const *d_outs, int nb_outs, bool multicast, void *stream)
{
	clear_error();
	if (!d_band || !d_samples || !d_scores || !d_minmax)
		return fail(PF_B200_EINVAL, "NULL device pointer");
	if (width <= 0 || height <= 0)
		return fail(PF_B200_EINVAL, "size %d x %d", width, height);
	if (row_begin < 0 || row_begin >= row_end || row_end > height)
		return fail(PF_B200_EINVAL, "ace: bad row band [%d, %d) of %d rows", row_begin, row_end, height);
	if ((uintptr_t)d_scores & 255)
		return fail(PF_B200_EINVAL, "ace: the score block must be 256-byte aligned");
	AceParams p = { width, height, row_begin, row_end, nb_samples, slope, limit };
	p.sample_px = d_sample_px;
	// the kernels index the page; with the samples' values given they only
	// ever read rows [row_begin, row_end) of it
	const uint32_t *page = d_band - (size_t)row_begin * width;
	const AceScores sc = ace_scores_at(d_scores, (size_t)(row_end - row_begin) * width);
	cudaStream_t st = static_cast<cudaStream_t>(stream);
	switch (pass) {
	case ACE_SCORES:
		return launch_ace_scores(page, d_samples, p, sc, d_minmax, st);
	case ACE_REFINE:
		return launch_ace_refine(page, d_samples, p, sc, d_minmax, st);
	default:
		break;
	}
	if (!d_outs)
		return fail(PF_B200_EINVAL, "NULL pointer");
	for (int i = 0; i < nb_outs && i < 16; i++)
		if (!d_outs[i])
			return fail(PF_B200_EINVAL, "NULL destination %d", i);
	int rc = launch_ace_scale(p, sc, d_minmax, row_begin, row_end - row_begin, 0, d_outs, nb_outs,
			multicast, st);
	if (rc)
		return rc;
	return launch_ace_fixup(page, d_samples, p, sc, d_minmax, row_begin, 0, d_outs, nb_outs,
			multicast, st);
}

int pf_b200_dev_ace_scores(const uint32_t *d_in, int width, int height,
		int row_begin, int row_end,
		const int *d_samples, int nb_samples, double slope, double limit,
		void *d_scores, double *d_minmax, void *stream)
{
	return ace_band_pass(ACE_SCORES, d_in ? d_in + (size_t)(row_begin > 0 ? row_begin : 0) * width : nullptr,
			width, height, row_begin, row_end, d_samples, nullptr, nb_samples, slope, limit,
			d_scores, d_minmax, nullptr, 0, false, stream);
}

int pf_b200_dev_ace_refine(const uint32_t *d_in, int width, int height,
		int row_begin, int row_end,
		const int *d_samples, int nb_samples, double slope, double limit,
		void *d_scores, double *d_minmax, void *stream)
{
	return ace_band_pass(ACE_REFINE, d_in ? d_in + (size_t)(row_begin > 0 ? row_begin : 0) * width : nullptr,
			width, height, row_begin, row_end, d_samples, nullptr, nb_samples, slope, limit,
			d_scores, d_minmax, nullptr, 0, false, stream);
}

int pf_b200_dev_ace_scale(const uint32_t *d_in, int width, int height,
		int row_begin, int row_end,
		const int *d_samples, int nb_samples, double slope, double limit,
		void *d_scores, double *d_minmax, uint32_t *d_out_band, void *stream)
{
	uint32_t *outs[1] = { d_out_band };
	return ace_band_pass(ACE_SCALE, d_in ? d_in + (size_t)(row_begin > 0 ? row_begin : 0) * width : nullptr,
			width, height, row_begin, row_end, d_samples, nullptr, nb_samples, slope, limit,
			d_scores, d_minmax, outs, 1, false, stream);
}

int pf_b200_dev_ace_scores_band(const uint32_t *d_band, int width, int height,
		int row_begin, int row_end,
		const int *d_samples, const uint32_t *d_sample_px, int nb_samples,
		double slope, double limit, void *d_scores, double *d_minmax, void *stream)
{
	if (!d_sample_px)
		return fail(PF_B200_EINVAL, "NULL device pointer");
	return ace_band_pass(ACE_SCORES, d_band, width, height, row_begin, row_end, d_samples, d_sample_px,
			nb_samples, slope, limit, d_scores, d_minmax, nullptr, 0, false, stream);
}

int pf_b200_dev_ace_refine_band(const uint32_t *d_band, int width, int height,
		int row_begin, int row_end,
		const int *d_samples, const uint32_t *d_sample_px, int nb_samples,
		double slope, double limit, void *d_scores, double *d_minmax, void *stream)
{
	if (!d_sample_px)
		return fail(PF_B200_EINVAL, "NULL device pointer");
	return ace_band_pass(ACE_REFINE, d_band, width, height, row_begin, row_end, d_samples, d_sample_px,
			nb_samples, slope, limit, d_scores, d_minmax, nullptr, 0, false, stream);
}

int pf_b200_dev_ace_scale_band(const uint32_t *d_band, int width, int height,
		int row_begin, int row_end,
		const int *d_samples, const uint32_t *d_sample_px, int nb_samples,
		double slope, double limit, void *d_scores, double *d_minmax,
		uint32_t *d_out_band, void *stream)
{
	if (!d_sample_px)
		return fail(PF_B200_EINVAL, "NULL device pointer");
	uint32_t *outs[1] = { d_out_band };
	return ace_band_pass(ACE_SCALE, d_band, width, height, row_begin, row_end, d_samples, d_sample_px,
			nb_samples, slope, limit, d_scores, d_minmax, outs, 1, false, stream);
}

int pf_b200_dev_ace_scale_to(const uint32_t *d_band, int width, int height,
		int row_begin, int row_end,
		const int *d_samples, const uint32_t *d_sample_px, int nb_samples,
		double slope, double limit, void *d_scores, double *d_minmax,
		uint32_t *const *d_outs, int nb_outs, void *stream)
{
	return ace_band_pass(ACE_SCALE, d_band, width, height, row_begin, row_end, d_samples, d_sample_px,
			nb_samples, slope, limit, d_scores, d_minmax, d_outs, nb_outs, false, stream);
}

int pf_b200_dev_ace_scale_multicast(const uint32_t *d_band, int width, int height,
		int row_begin, int row_end,
		const int *d_samples, const uint32_t *d_sample_px, int nb_samples,
		double slope, double limit, void *d_scores, double *d_minmax,
		uint32_t *d_multicast_band, void *stream)
{
	uint32_t *outs[1] = { d_multicast_band };
	return ace_band_pass(ACE_SCALE, d_band, width, height, row_begin, row_end, d_samples, d_sample_px,
			nb_samples, slope, limit, d_scores, d_minmax, outs, 1, true, stream);
}

size_t pf_b200_ace_workspace_bytes(int width, int height)
{
	clear_error();
	if (width <= 0 || height <= 0) {
		fail(PF_B200_EINVAL, "size %d x %d", width, height);
		return 0;
	}
	return ace_scores_bytes(width, height) + ace_minmax_doubles(width, height) * sizeof(double);
}

int pf_b200_dev_ace(const uint32_t *d_in, uint32_t *d_out, int width, int height,
		const int *d_samples, int nb_samples, double slope, double limit,
		void *d_workspace, size_t workspace_bytes, void *stream)
{
	clear_error();
	if (!d_in || !d_out || !d_samples || !d_workspace)
		return fail(PF_B200_EINVAL, "NULL device pointer");
	if (width <= 0 || height <= 0)
		return fail(PF_B200_EINVAL, "size %d x %d", width, height);
	if (d_in == d_out)
		return fail(PF_B200_EINVAL, "ace: the device entry point is out of place");
	if (workspace_bytes < pf_b200_ace_workspace_bytes(width, height) || ((uintptr_t)d_workspace & 255))
		return fail(PF_B200_EINVAL, "ace: work area of %zu bytes, need %zu, 256-byte aligned",
				workspace_bytes, pf_b200_ace_workspace_bytes(width, height));
	cudaStream_t st = static_cast<cudaStream_t>(stream);
	const AceScores sc = ace_scores_at(d_workspace, (size_t)width * height);
	double *minmax = reinterpret_cast<double *>(static_cast<char *>(d_workspace)
			+ ace_scores_bytes(width, height));
	AceParams p = { width, height, 0, height, nb_samples, slope, limit };
	int rc;
	if ((rc = launch_ace_minmax_init(minmax, width, height, st))
			|| (rc = launch_ace_scores(d_in, d_samples, p, sc, minmax, st))
			|| (rc = launch_ace_refine(d_in, d_samples, p, sc, minmax, st))
			|| (rc = launch_ace_scale(p, sc, minmax, 0, height, 0, &d_out, 1, false, st))
			|| (rc = launch_ace_fixup(d_in, d_samples, p, sc, minmax, 0, 0, &d_out, 1, false, st)))
		return rc;
	return PF_B200_OK;
}

double pf_b200_fp64_peak_tflops(void)
{
	clear_error();
	if (pf_b200_device_count() <= 0)
		return -(double)fail(PF_B200_ENODEV, "no CUDA device");
	return measure_fp64_peak_tflops();
}

double pf_b200_selftest_rsqrt_error(int which)
{
	clear_error();
	if (which != 0 && which != 1)
		return -(double)fail(PF_B200_EINVAL, "rsqrt self-test: which = %d", which);
	if (pf_b200_device_count() <= 0)
		return -(double)fail(PF_B200_ENODEV, "no CUDA device");
	return measure_rsqrt_error(which);
}

double pf_b200_fp32_peak_tflops(void)
{
	clear_error();
	if (pf_b200_device_count() <= 0)
		return -(double)fail(PF_B200_ENODEV, "no CUDA device");
	return measure_fp32_peak_tflops();
}

}  // extern "C"
