// Measurement infrastructure (not part of the library): what the host link of this box takes from 1 / 2 / 4 / 8 GPUs at once.
// Plain cudaMemcpyAsync device-to-host copies into pinned memory, one host thread per GPU, all GPUs started together; one copy
// per block. Varied: block size, one or two streams per GPU, 1-D copies against 2-D copies with an odd row pitch (what
// v2m_cuda_emit_rows issues), default against portable pinned allocations, with and without pinning each GPU's host thread
// (and therefore the first touch of its pinned buffer) to its own share of the cores. Also: what the host's memory takes
// from the CPU side (memset from all cores), to tell a link limit from a host-memory limit.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++20 -Xcompiler -pthread -o build/d2h_ceiling d2h_ceiling.cu
// Run:   build/d2h_ceiling [max_gpus]        (prints one line per configuration)
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include <cuda_runtime.h>
#include <pthread.h>
#include <sched.h>
#include <unistd.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

using u64 = std::uint64_t;
using clk = std::chrono::steady_clock;

struct config
{
	char const	*name;
	u64			block_bytes;	// bytes per cudaMemcpyAsync (1-D) or per 2-D copy
	int			streams;
	bool		two_d;			// rows of `width` bytes at a device pitch of width + 85 rounded to 128, host pitch = device pitch
	bool		portable;
	bool		pin_cores;
};

struct spin_barrier
{
	std::atomic<int> count{0}, generation{0};
	int n;
	explicit spin_barrier(int n_) : n(n_) {}
	void wait()
	{
		int const g(generation.load());
		if (count.fetch_add(1) + 1 == n) { count.store(0); generation.fetch_add(1); }
		else while (generation.load() == g) std::this_thread::yield();
	}
};

static void pin_to_share(int rank, int n)
{
	long const ncpu(sysconf(_SC_NPROCESSORS_ONLN));
	cpu_set_t set;
	CPU_ZERO(&set);
	for (long c(rank * ncpu / n); c < (rank + 1) * ncpu / n; ++c) CPU_SET(int(c), &set);
	pthread_setaffinity_np(pthread_self(), sizeof(set), &set);
}

static void unpin()
{
	long const ncpu(sysconf(_SC_NPROCESSORS_ONLN));
	cpu_set_t set;
	CPU_ZERO(&set);
	for (long c(0); c < ncpu; ++c) CPU_SET(int(c), &set);
	pthread_setaffinity_np(pthread_self(), sizeof(set), &set);
}

int main(int argc, char **argv)
{
	int n_dev(0);
	CK(cudaGetDeviceCount(&n_dev));
	int const max_gpus(argc > 1 ? std::min(n_dev, std::atoi(argv[1])) : n_dev);
	u64 const buf_bytes(u64(2) << 30);			// per GPU: 2 GiB device source, 2 GiB pinned destination
	double const seconds(1.5);
	printf("# d2h_ceiling: %d device(s) visible, using up to %d; %ld host cores; %.1f GiB per GPU per pass\n", n_dev, max_gpus, sysconf(_SC_NPROCESSORS_ONLN), buf_bytes / double(1 << 30));

	config const configs[] = {
		{"1d 256MiB x2 streams, default pinned",            u64(256) << 20, 2, false, false, false},
		{"1d 256MiB x1 stream,  default pinned",            u64(256) << 20, 1, false, false, false},
		{"1d  64MiB x2 streams, default pinned",            u64(64) << 20,  2, false, false, false},
		{"1d   2GiB x1 stream,  default pinned",            u64(2) << 30,   1, false, false, false},
		{"2d 256MiB x2 streams (row 250,293,419 B, odd pitch)", u64(256) << 20, 2, true, false, false},
		{"1d 256MiB x2 streams, portable pinned",           u64(256) << 20, 2, false, true,  false},
		{"1d 256MiB x2 streams, default pinned, threads pinned to core shares", u64(256) << 20, 2, false, false, true},
		{"1d 256MiB x2 streams, portable pinned, threads pinned to core shares", u64(256) << 20, 2, false, true, true},
	};

	for (int n(1); n <= max_gpus; n *= 2)
	{
		for (config const &cfg : configs)
		{
			std::vector<double> gbs(n, 0.0);
			std::vector<double> secs(n, 0.0);
			spin_barrier bar(n);
			auto const worker([&](int const rank) {
				if (cfg.pin_cores) pin_to_share(rank, n); else unpin();
				CK(cudaSetDevice(rank));
				unsigned char *d(nullptr), *h(nullptr);
				CK(cudaMalloc(&d, buf_bytes));
				CK(cudaMemset(d, 0x41, buf_bytes));
				CK(cudaHostAlloc(&h, buf_bytes, cfg.portable ? cudaHostAllocPortable : cudaHostAllocDefault));
				std::memset(h, 0, buf_bytes);			// first touch from this (possibly pinned) thread
				cudaStream_t st[2];
				for (auto &s : st) CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
				u64 const width(250293419), pitch((width + 127) / 128 * 128);
				auto const one_pass([&]() {
					int k(0);
					if (cfg.two_d)
					{
						u64 const rows_total(buf_bytes / pitch), rows_per_copy(std::max<u64>(1, cfg.block_bytes / pitch));
						for (u64 r(0); r < rows_total; r += rows_per_copy, k = (k + 1) % cfg.streams)
							CK(cudaMemcpy2DAsync(h + r * pitch, pitch, d + r * pitch, pitch, width, std::min(rows_per_copy, rows_total - r), cudaMemcpyDeviceToHost, st[k]));
					}
					else
						for (u64 off(0); off < buf_bytes; off += cfg.block_bytes, k = (k + 1) % cfg.streams)
							CK(cudaMemcpyAsync(h + off, d + off, std::min(cfg.block_bytes, buf_bytes - off), cudaMemcpyDeviceToHost, st[k]));
				});
				one_pass();
				CK(cudaDeviceSynchronize());
				bar.wait();
				auto const t0(clk::now());
				u64 passes(0);
				while (std::chrono::duration<double>(clk::now() - t0).count() < seconds)
				{
					one_pass();
					for (int i(0); i < cfg.streams; ++i) CK(cudaStreamSynchronize(st[i]));
					++passes;
				}
				double const t(std::chrono::duration<double>(clk::now() - t0).count());
				u64 const bytes_per_pass(cfg.two_d ? (buf_bytes / pitch) * width : buf_bytes);
				gbs[rank] = passes * bytes_per_pass / t / 1e9;
				secs[rank] = t;
				bar.wait();
				for (auto &s : st) cudaStreamDestroy(s);
				cudaFreeHost(h);
				cudaFree(d);
			});
			std::vector<std::thread> pool;
			for (int r(0); r < n; ++r) pool.emplace_back(worker, r);
			for (auto &t : pool) t.join();
			double total(0), lo(1e30), hi(0);
			for (double g : gbs) { total += g; lo = std::min(lo, g); hi = std::max(hi, g); }
			printf("gpus=%d  %-72s total %7.1f GB/s   per GPU min %6.1f max %6.1f\n", n, cfg.name, total, lo, hi);
			fflush(stdout);
		}
	}

	// Host memory from the CPU side: memset of 1 GiB per thread, all cores.
	{
		long const ncpu(sysconf(_SC_NPROCESSORS_ONLN));
		for (long threads : {1L, ncpu / 2, ncpu})
		{
			if (threads < 1) continue;
			u64 const bytes(u64(1) << 30);
			std::vector<unsigned char *> bufs(threads);
			for (auto &b : bufs) { b = static_cast<unsigned char *>(std::malloc(bytes)); std::memset(b, 1, bytes); }
			spin_barrier bar{int(threads)};
			std::vector<double> t(threads);
			auto const worker([&](long i) {
				bar.wait();
				auto const t0(clk::now());
				for (int it(0); it < 4; ++it) std::memset(bufs[i], it, bytes);
				t[i] = std::chrono::duration<double>(clk::now() - t0).count();
			});
			std::vector<std::thread> pool;
			for (long i(0); i < threads; ++i) pool.emplace_back(worker, i);
			for (auto &th : pool) th.join();
			double worst(0);
			for (double x : t) worst = std::max(worst, x);
			printf("host memset  threads=%ld  %7.1f GB/s\n", threads, threads * 4.0 * bytes / worst / 1e9);
			for (auto b : bufs) std::free(b);
		}
	}
	return 0;
}
