// Measurement infrastructure (not part of the library): what a write-only stream reaches on this GPU, by access pattern.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++20 -o build/write_ceiling write_ceiling.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>
#include "../../vcf2multialign_b200/csrc/common.cuh"
using namespace v2m;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

// linear fill, 16 B per thread per iteration, grid-stride
__global__ void __launch_bounds__(256) fill_linear(uint4 *dst, u64 n16, int cs)
{
	uint4 const v(make_uint4(0x41414141u, 0x43434343u, 0x47474747u, 0x54545454u));
	u64 const stride(u64(gridDim.x) * blockDim.x);
	for (u64 i(u64(blockIdx.x) * blockDim.x + threadIdx.x); i < n16; i += stride)
		if (cs) st_cs_v4(dst + i, v); else st_v4(dst + i, v);
}

// emit-like pattern: CTA = tile of TCOL columns x 64 rows; warp writes a row's tile with 128-bit stores
template <int TCOL>
__global__ void __launch_bounds__(256) fill_tiles_stg(u8 *dst, u64 pitch, u32 rows)
{
	unsigned const lane(threadIdx.x & 31), warp(threadIdx.x >> 5);
	u64 const t0(u64(blockIdx.x) * TCOL);
	u32 const rb0(blockIdx.y * 64);
	uint4 const v(make_uint4(0x41414141u, 0x43434343u, 0x47474747u, 0x54545454u));
	for (u32 r(rb0 + warp); r < min(rb0 + 64, rows); r += 8)
	{
		u8 *orow(dst + u64(r) * pitch + t0);
#pragma unroll
		for (int j = 0; j < TCOL / 512; ++j) st_v4(orow + j * 512 + lane * 16, v);
	}
}

// same pattern, rows handed to the TMA engine from shared memory (BUFS buffers per warp, ROWS rows per CTA)
template <int TCOL, int BUFS, int ROWS>
__global__ void __launch_bounds__(256) fill_tiles_bulk(u8 *dst, u64 pitch, u32 rows, int restage)
{
	extern __shared__ __align__(128) u8 sm[];
	unsigned const lane(threadIdx.x & 31), warp(threadIdx.x >> 5);
	u64 const t0(u64(blockIdx.x) * TCOL);
	u32 const rb0(blockIdx.y * ROWS);
	uint4 const v(make_uint4(0x41414141u, 0x43434343u, 0x47474747u, 0x54545454u));
	u8 *mine(sm + warp * BUFS * TCOL);
	for (int b = 0; b < BUFS; ++b)
		for (int j = 0; j < TCOL / 512; ++j) *reinterpret_cast<uint4 *>(mine + b * TCOL + j * 512 + lane * 16) = v;
	fence_proxy_async_smem();
	__syncwarp();
	int k(0);
	for (u32 r(rb0 + warp); r < min(rb0 + ROWS, rows); r += 8, ++k)
	{
		u8 *buf(mine + (k % BUFS) * TCOL);
		if (restage)
		{
			if (k >= BUFS) { if (0 == lane) bulk_wait_read<BUFS - 1>(); __syncwarp(); }
			for (int j = 0; j < TCOL / 512; ++j) *reinterpret_cast<uint4 *>(buf + j * 512 + lane * 16) = v;
			fence_proxy_async_smem();
			__syncwarp();
		}
		if (0 == lane) { bulk_s2g(dst + u64(r) * pitch + t0, buf, TCOL); bulk_commit(); }
	}
	if (0 == lane) bulk_wait_read<0>();
}

// generalised: WARPS warps per CTA; SPLIT > 1: the CTA covers SPLIT adjacent tiles, warp groups take one each, so that
// neighbouring pieces of a row are written at about the same time
template <int TCOL, int BUFS, int ROWS, int WARPS, int SPLIT>
__global__ void __launch_bounds__(WARPS * 32) fill_tiles_gen(u8 *dst, u64 pitch, u32 rows)
{
	extern __shared__ __align__(128) u8 sm[];
	unsigned const lane(threadIdx.x & 31), warp(threadIdx.x >> 5);
	constexpr int GW = WARPS / SPLIT;						// warps per tile
	u64 const t0((u64(blockIdx.x) * SPLIT + warp / GW) * TCOL);
	u32 const rb0(blockIdx.y * ROWS);
	uint4 v[TCOL / 512];
	for (int j = 0; j < TCOL / 512; ++j) v[j] = make_uint4(0x41414141u + j, 0x43434343u, 0x47474747u, 0x54545454u + lane);
	u8 *mine(sm + warp * BUFS * TCOL);
	int k(0);
	for (u32 r(rb0 + warp % GW); r < min(rb0 + ROWS, rows); r += GW, ++k)
	{
		u8 *buf(mine + (k % BUFS) * TCOL);
		if (k >= BUFS) { if (0 == lane) bulk_wait_read<BUFS - 1>(); __syncwarp(); }
#pragma unroll
		for (int j = 0; j < TCOL / 512; ++j) *reinterpret_cast<uint4 *>(buf + j * 512 + lane * 16) = v[j];
		fence_proxy_async_smem();
		__syncwarp();
		if (0 == lane) { bulk_s2g(dst + u64(r) * pitch + t0, buf, TCOL); bulk_commit(); }
	}
	if (0 == lane) bulk_wait_read<0>();
}

// persistent form of the product kernel's store pattern: one CTA per SM, WARPS warps with two TCOL-byte buffers each, units of
// SPAN adjacent tiles x ROWS rows drawn from a counter; a warp writes the SPAN tiles of a row back to back
template <int TCOL, int ROWS, int WARPS, int SPAN>
__global__ void __launch_bounds__(WARPS * 32) fill_tiles_persistent(u8 *dst, u64 pitch, u32 rows, u32 n_units, u32 units_x, unsigned *ticket)
{
	extern __shared__ __align__(128) u8 sm[];
	__shared__ unsigned s_unit;
	unsigned const lane(threadIdx.x & 31), warp(threadIdx.x >> 5);
	uint4 v[TCOL / 512];
	for (int j = 0; j < TCOL / 512; ++j) v[j] = make_uint4(0x41414141u + j, 0x43434343u, 0x47474747u, 0x54545454u + lane);
	u8 *mine(sm + warp * 2 * TCOL);
	u32 k(0);
	for (;;)
	{
		__syncthreads();
		if (0 == threadIdx.x) s_unit = atomicAdd(ticket, 1u);
		__syncthreads();
		u32 const unit(s_unit);
		if (unit >= n_units) break;
		u64 const t0(u64(unit % units_x) * SPAN * TCOL);
		u32 const rb0((unit / units_x) * ROWS);
		for (u32 r(rb0 + warp); r < min(rb0 + ROWS, rows); r += WARPS)
			for (int sp = 0; sp < SPAN; ++sp, ++k)
			{
				u8 *buf(mine + (k & 1) * TCOL);
				if (k >= 2) { if (0 == lane) bulk_wait_read<1>(); __syncwarp(); }
#pragma unroll
				for (int j = 0; j < TCOL / 512; ++j) *reinterpret_cast<uint4 *>(buf + j * 512 + lane * 16) = v[j];
				fence_proxy_async_smem();
				__syncwarp();
				if (0 == lane) { bulk_s2g(dst + u64(r) * pitch + t0 + u64(sp) * TCOL, buf, TCOL); bulk_commit(); }
			}
	}
	if (0 == lane) bulk_wait_read<0>();
}

template <int TCOL, int BUFS, int ROWS, int WARPS, int SPLIT>
static void run_gen(u8 *d, u64 pitch, u32 rows, u64 A);

template <typename F>
static void timeit(char const *name, double bytes, F &&launch)
{
	cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
	for (int i = 0; i < 3; ++i) launch();
	CK(cudaDeviceSynchronize());
	int const iters(10);
	CK(cudaEventRecord(e0));
	for (int i = 0; i < iters; ++i) launch();
	CK(cudaEventRecord(e1));
	CK(cudaEventSynchronize(e1));
	CK(cudaGetLastError());
	float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
	printf("%-44s %8.3f ms  %8.1f GB/s\n", name, ms / iters, bytes / (ms / iters * 1e-3) / 1e9);
}

template <int TCOL, int BUFS, int ROWS, int WARPS, int SPLIT>
static void run_gen(u8 *d, u64 pitch, u32 rows, u64 A)
{
	auto k(fill_tiles_gen<TCOL, BUFS, ROWS, WARPS, SPLIT>);
	int const smem(WARPS * BUFS * TCOL);
	CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
	int occ(0); CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, WARPS * 32, smem));
	u64 const n_ctas_x(A / (u64(TCOL) * SPLIT));
	char nm[96]; snprintf(nm, 96, "gen<%d,bufs %d,rows %d,warps %d,split %d> occ %d", TCOL, BUFS, ROWS, WARPS, SPLIT, occ);
	timeit(nm, double(n_ctas_x) * TCOL * SPLIT * rows, [&]{ k<<<dim3(unsigned(n_ctas_x), (rows + ROWS - 1) / ROWS), WARPS * 32, smem>>>(d, pitch, rows); });
}

template <int TCOL, int ROWS, int WARPS, int SPAN>
static void run_persistent(u8 *d, u64 pitch, u32 rows, u64 A, unsigned *ticket)
{
	auto k(fill_tiles_persistent<TCOL, ROWS, WARPS, SPAN>);
	int const smem(WARPS * 2 * TCOL);
	CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
	u32 const units_x(u32(A / (u64(TCOL) * SPAN))), n_units(units_x * ((rows + ROWS - 1) / ROWS));
	char nm[96]; snprintf(nm, 96, "persistent<%d,rows %d,warps %d,span %d>", TCOL, ROWS, WARPS, SPAN);
	timeit(nm, double(units_x) * TCOL * SPAN * rows, [&]{ CK(cudaMemsetAsync(ticket, 0, 4)); k<<<148, WARPS * 32, smem>>>(d, pitch, rows, n_units, units_x, ticket); });
}

int main()
{
	u64 const A(51237456), pitch((A + 15) / 16 * 16);
	u32 const rows(128);
	u64 const bytes(u64(rows) * pitch);
	u8 *d; CK(cudaMalloc(&d, bytes + (1 << 20)));
	u64 const n_tiles4k(A / 4096), n_tiles2k(A / 2048), n_tiles8k(A / 8192);
	timeit("cudaMemsetAsync", double(bytes), [&]{ CK(cudaMemsetAsync(d, 0x41, bytes)); });
	for (int ctas : {148 * 2, 148 * 4, 148 * 8, 148 * 16, 148 * 64})
	{
		char nm[64]; snprintf(nm, 64, "fill_linear st.v4 grid=%d", ctas);
		timeit(nm, double(bytes), [&]{ fill_linear<<<ctas, 256>>>((uint4 *) d, bytes / 16, 0); });
	}
	timeit("fill_linear st.cs.v4 grid=148*8", double(bytes), [&]{ fill_linear<<<148 * 8, 256>>>((uint4 *) d, bytes / 16, 1); });
	timeit("fill_tiles_stg<4096> 64 rows/CTA", double(n_tiles4k) * 4096 * rows, [&]{ fill_tiles_stg<4096><<<dim3(n_tiles4k, rows / 64), 256>>>(d, pitch, rows); });
	timeit("fill_tiles_stg<2048> 64 rows/CTA", double(n_tiles2k) * 2048 * rows, [&]{ fill_tiles_stg<2048><<<dim3(n_tiles2k, rows / 64), 256>>>(d, pitch, rows); });
	{
		auto k1(fill_tiles_bulk<4096, 2, 64>); CK(cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 2 * 4096));
		for (int restage : {0, 1})
		{
			char nm[64]; snprintf(nm, 64, "fill_tiles_bulk<4096,2,64> restage=%d", restage);
			timeit(nm, double(n_tiles4k) * 4096 * rows, [&]{ k1<<<dim3(n_tiles4k, rows / 64), 256, 8 * 2 * 4096>>>(d, pitch, rows, restage); });
		}
		auto k2(fill_tiles_bulk<2048, 2, 64>); CK(cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 2 * 2048));
		timeit("fill_tiles_bulk<2048,2,64> restage=1", double(n_tiles2k) * 2048 * rows, [&]{ k2<<<dim3(n_tiles2k, rows / 64), 256, 8 * 2 * 2048>>>(d, pitch, rows, 1); });
		auto k3(fill_tiles_bulk<8192, 2, 64>); CK(cudaFuncSetAttribute(k3, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 2 * 8192));
		timeit("fill_tiles_bulk<8192,2,64> restage=1", double(n_tiles8k) * 8192 * rows, [&]{ k3<<<dim3(n_tiles8k, rows / 64), 256, 8 * 2 * 8192>>>(d, pitch, rows, 1); });
		auto k4(fill_tiles_bulk<4096, 2, 128>); CK(cudaFuncSetAttribute(k4, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 2 * 4096));
		timeit("fill_tiles_bulk<4096,2,128> restage=1", double(n_tiles4k) * 4096 * rows, [&]{ k4<<<dim3(n_tiles4k, rows / 128), 256, 8 * 2 * 4096>>>(d, pitch, rows, 1); });
		auto k5(fill_tiles_bulk<4096, 4, 64>); CK(cudaFuncSetAttribute(k5, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 4 * 4096));
		timeit("fill_tiles_bulk<4096,4,64> restage=1", double(n_tiles4k) * 4096 * rows, [&]{ k5<<<dim3(n_tiles4k, rows / 64), 256, 8 * 4 * 4096>>>(d, pitch, rows, 1); });
	}
	unsigned *ticket; CK(cudaMalloc(&ticket, 4));
	run_persistent<4096, 64, 19, 1>(d, pitch, rows, A, ticket);
	run_persistent<4096, 32, 19, 2>(d, pitch, rows, A, ticket);
	run_persistent<4096, 64, 19, 2>(d, pitch, rows, A, ticket);
	run_persistent<4096, 32, 19, 4>(d, pitch, rows, A, ticket);
	run_persistent<4096, 64, 24, 1>(d, pitch, rows, A, ticket);
	run_persistent<4096, 64, 24, 2>(d, pitch, rows, A, ticket);
	run_persistent<8192, 64, 12, 1>(d, pitch, rows, A, ticket);
	run_persistent<8192, 64, 10, 1>(d, pitch, rows, A, ticket);
	run_gen<4096, 2, 64, 8, 1>(d, pitch, rows, A);
	run_gen<4096, 2, 64, 8, 2>(d, pitch, rows, A);
	run_gen<4096, 2, 32, 8, 2>(d, pitch, rows, A);
	run_gen<4096, 2, 128, 8, 2>(d, pitch, rows, A);
	run_gen<4096, 2, 64, 8, 4>(d, pitch, rows, A);
	run_gen<8192, 2, 64, 4, 1>(d, pitch, rows, A);
	run_gen<8192, 2, 32, 4, 1>(d, pitch, rows, A);
	run_gen<8192, 2, 128, 4, 1>(d, pitch, rows, A);
	run_gen<8192, 2, 64, 8, 1>(d, pitch, rows, A);
	run_gen<8192, 2, 128, 8, 1>(d, pitch, rows, A);
	run_gen<8192, 3, 64, 4, 1>(d, pitch, rows, A);
	run_gen<16384, 2, 64, 4, 1>(d, pitch, rows, A);
	run_gen<16384, 2, 64, 2, 1>(d, pitch, rows, A);
	run_gen<16384, 2, 128, 2, 1>(d, pitch, rows, A);
	run_gen<8192, 2, 64, 2, 1>(d, pitch, rows, A);
	// copy for reference (read + write bytes)
	{
		u8 *s; CK(cudaMalloc(&s, bytes / 2));
		timeit("cudaMemcpyAsync D2D (R+W bytes)", double(bytes), [&]{ CK(cudaMemcpyAsync(d, s, bytes / 2, cudaMemcpyDeviceToDevice)); });
		cudaFree(s);
	}
	cudaFree(d);
	return 0;
}
