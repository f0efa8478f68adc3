// Measurement infrastructure (not part of the library): which side of K1 (transpose_bits) costs what.
// The transpose reads 128-byte (or 64-byte) pieces of source lines that are only 8-byte aligned and writes 64-byte (or
// 128-byte) runs of destination lines that are only 8-byte aligned. These kernels replay ONE side of that pattern each, with
// no bit work, on the real shapes, so that the read side, the write side and the cost of sector-misaligned runs can be told
// apart:
//   W<RUN_BYTES, VEC>  write-only: every warp store covers 32 * VEC bytes as runs of RUN_BYTES of consecutive destination
//                      lines' tiles, rows vary with j as in K1 (32 stores per thread); `shift` moves each row's run starts
//                      back to the 32-byte sector boundary below (ragged line heads and tails are simply clipped).
//   R<PIECE_BYTES>     read-only: K1's tile load (8-byte loads, 16 or 8 lanes per line piece, row bands fastest in the grid),
//                      reduced by XOR; one word per thread written.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++20 -o build/k1_patterns k1_patterns.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#include <cuda_runtime.h>

using u32 = std::uint32_t;
using u64 = std::uint64_t;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

// ---- write side ----------------------------------------------------------------------------------------------------------
// CTA = 512 threads. A warp store writes RUNS = 32 * VEC / RUN_BYTES runs; run q of warp w, store j goes to row
// band_row0 + 32 * (w * RUNS + q) + j. The CTA covers 32 * 16 * RUNS rows x RUN_BYTES bytes.
template <int RUN_BYTES, int VEC>
__global__ void __launch_bounds__(512, 2)
write_pattern(unsigned char *dst, u64 const line_bytes, u64 const rows, int const shift)
{
	constexpr int RUNS = 32 * VEC / RUN_BYTES;
	constexpr int LANES_PER_RUN = RUN_BYTES / VEC;
	unsigned const lane(threadIdx.x & 31), warp(threadIdx.x >> 5);
	unsigned const q(lane / LANES_PER_RUN), l(lane % LANES_PER_RUN);
	u64 const row0(u64(blockIdx.x) * (32 * 16 * RUNS) + 32 * (warp * RUNS + q));
	u64 const x0(u64(blockIdx.y) * RUN_BYTES + l * VEC);
#pragma unroll
	for (int j = 0; j < 32; ++j)
	{
		u64 const r(row0 + j);
		if (r >= rows) continue;
		u64 const line0(r * line_bytes);
		long long off((long long) x0 - (shift ? (long long) (line0 & 31) : 0));
		if (off < 0 || u64(off) + VEC > line_bytes) continue;
		unsigned char *p(dst + line0 + off);
		if (4 == VEC) *reinterpret_cast<u32 *>(p) = 0x41424344u + j;
		else if (8 == VEC) *reinterpret_cast<uint2 *>(p) = make_uint2(0x41424344u + j, lane);
		else *reinterpret_cast<uint4 *>(p) = make_uint4(0x41424344u + j, lane, warp, 7);	// only valid when 16-byte aligned
	}
}

// ---- read side -----------------------------------------------------------------------------------------------------------
// CTA = 512 threads loads LINES source lines x PIECE bytes: 8-byte loads, PIECE / 8 lanes per line.
template <int PIECE, int LINES>
__global__ void __launch_bounds__(512, 2)
read_pattern(u64 const * __restrict__ src, u64 const R64 /* words per line */, u64 const C /* lines */, u64 *sink)
{
	constexpr int LPL = PIECE / 8;					// lanes per line
	constexpr int LINES_PER_PASS = 512 / LPL;
	unsigned const tid(threadIdx.x);
	unsigned const qq(tid % LPL), lp(tid / LPL);
	u64 const c0(u64(blockIdx.y) * LINES), p0(u64(blockIdx.x) * LPL);
	bool const ok(p0 + qq < R64);
	u64 acc(0);
	u64 vals[LINES / LINES_PER_PASS];
#pragma unroll
	for (int p = 0; p < LINES / LINES_PER_PASS; ++p)
	{
		u64 const line(c0 + LINES_PER_PASS * p + lp);
		vals[p] = (ok && line < C) ? __ldg(src + line * R64 + p0 + qq) : 0;
	}
#pragma unroll
	for (int p = 0; p < LINES / LINES_PER_PASS; ++p) acc ^= vals[p];
	if (acc == 0x123456789abcdefull) sink[tid] = acc;		// never true in practice: keeps the loads alive
}

template <typename F>
static float time_ms(F &&launch, int iters = 10)
{
	cudaEvent_t e0, e1;
	CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
	for (int i(0); i < 3; ++i) launch();
	CK(cudaDeviceSynchronize());
	CK(cudaEventRecord(e0));
	for (int i(0); i < iters; ++i) launch();
	CK(cudaEventRecord(e1));
	CK(cudaEventSynchronize(e1));
	CK(cudaGetLastError());
	float ms(0);
	CK(cudaEventElapsedTime(&ms, e0, e1));
	return ms / iters;
}

template <int RUN_BYTES, int VEC>
static void run_write(unsigned char *buf, u64 rows, u64 cols, int shift)
{
	constexpr int RUNS = 32 * VEC / RUN_BYTES;
	u64 const line_bytes(cols / 8);
	dim3 const grid(unsigned((rows + 32 * 16 * RUNS - 1) / (32 * 16 * RUNS)), unsigned((line_bytes + RUN_BYTES - 1) / RUN_BYTES + (shift ? 1 : 0)));
	float const ms(time_ms([&] { write_pattern<RUN_BYTES, VEC><<<grid, 512>>>(buf, line_bytes, rows, shift); }));
	double const bytes(double(rows) * line_bytes);
	printf("  write  runs of %3d B, %2d B per lane, %s   %8.3f ms  %7.1f GB/s\n", RUN_BYTES, VEC, shift ? "sector-aligned run starts" : "runs as K1 places them  ", ms, bytes / ms / 1e6);
}

template <int PIECE, int LINES>
static void run_read(u64 const *src, u64 rows, u64 cols, u64 *sink)
{
	u64 const R64(rows / 64);
	dim3 const grid(unsigned((R64 + PIECE / 8 - 1) / (PIECE / 8)), unsigned((cols + LINES - 1) / LINES));
	float const ms(time_ms([&] { read_pattern<PIECE, LINES><<<grid, 512>>>(src, R64, cols, sink); }));
	double const bytes(double(rows) / 8 * cols);
	printf("  read   %3d-byte pieces of %4d lines per CTA, bands fastest          %8.3f ms  %7.1f GB/s\n", PIECE, LINES, ms, bytes / ms / 1e6);
}

int main()
{
	struct shape { u64 rows, cols; char const *name; };
	shape const shapes[] = {
		{5056, 6000000, "c5: destination lines 750,000 B (16 mod 32)"},
		{5056, 6300480, "c4: destination lines 787,560 B (8 mod 32)"},
		{5056, 6000128, "aligned: destination lines 750,016 B (0 mod 32)"},
		{5056, 1155008, "c2: destination lines 144,376 B (24 mod 32)"},
	};
	u64 const max_bytes(u64(5056) / 8 * 6300480 + 4096);
	unsigned char *buf(nullptr);
	u64 *sink(nullptr);
	CK(cudaMalloc(&buf, max_bytes));
	CK(cudaMalloc(&sink, 4096 * 8));
	CK(cudaMemset(buf, 0x5a, max_bytes));
	{
		float const ms(time_ms([&] { CK(cudaMemsetAsync(buf, 1, u64(5056) / 8 * 6000000)); }));
		printf("cudaMemsetAsync 3.79 GB: %.3f ms  %.1f GB/s\n", ms, 5056.0 / 8 * 6000000 / ms / 1e6);
	}
	for (shape const &s : shapes)
	{
		printf("%s  (%llu x %llu bits, %.3f GB per side)\n", s.name, (unsigned long long) s.rows, (unsigned long long) s.cols, s.rows / 8.0 * s.cols / 1e9);
		for (int shift(0); shift < 2; ++shift)
		{
			run_write<64, 4>(buf, s.rows, s.cols, shift);
			run_write<128, 4>(buf, s.rows, s.cols, shift);
			run_write<128, 8>(buf, s.rows, s.cols, shift);
			run_write<256, 8>(buf, s.rows, s.cols, shift);
		}
		run_read<128, 512>(reinterpret_cast<u64 const *>(buf), s.rows, s.cols, sink);
		run_read<64, 1024>(reinterpret_cast<u64 const *>(buf), s.rows, s.cols, sink);
		run_read<128, 1024>(reinterpret_cast<u64 const *>(buf), s.rows, s.cols, sink);
	}
	return 0;
}
