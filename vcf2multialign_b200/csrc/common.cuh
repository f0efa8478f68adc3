// Shared helpers for the sm_100a kernels of libv2m_cuda.so.
#pragma once

#include <cstdint>
#include <cstdio>
#include <string>

#include <cuda_runtime.h>

#ifndef V2M_T32_MULHI
#define V2M_T32_MULHI 0
#endif

namespace v2m {

using u8 = std::uint8_t;
using u32 = std::uint32_t;
using u64 = std::uint64_t;

constexpr u32 PLOIDY_MAX = 0xffffffffu;

// Thread-local error text behind v2m_cuda_last_error().
std::string &last_error();
int fail(int status, std::string msg);
void count_launch(unsigned n = 1);

#define V2M_CUDA_TRY(expr) do { cudaError_t const err_((expr)); if (cudaSuccess != err_) \
	return ::v2m::fail(2, std::string(#expr) + ": " + cudaGetErrorString(err_)); } while (0)

#define V2M_CHECK_LAUNCH(name) do { cudaError_t const err_(cudaGetLastError()); if (cudaSuccess != err_) \
	return ::v2m::fail(2, std::string("launch of " name ": ") + cudaGetErrorString(err_)); ::v2m::count_launch(); } while (0)

// Per-edge record used by row emission: aligned span [a0, a1) and the label's place in the flattened label store.
struct edge_rec { u32 a0, a1, label_off, label_len; };
static_assert(16 == sizeof(edge_rec));

// 128-bit streaming-friendly global accesses.
__device__ __forceinline__ uint4 ld_nc_v4(void const *p)
{
	uint4 r;
	asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
	return r;
}

__device__ __forceinline__ void st_v4(void *p, uint4 const &v)
{
	asm volatile("st.global.v4.u32 [%0], {%1, %2, %3, %4};" :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

__device__ __forceinline__ void st_cs_v4(void *p, uint4 const &v)
{
	asm volatile("st.global.cs.v4.u32 [%0], {%1, %2, %3, %4};" :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ---- bulk async copies (TMA engine, SASS UBLKCP) and their mbarrier / proxy plumbing ----------------------------------
__device__ __forceinline__ u32 smem_u32(void const *p) { return static_cast<u32>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(u64 *bar, u32 const count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}

__device__ __forceinline__ void mbar_expect_tx(u64 *bar, u32 const bytes)
{
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}

__device__ __forceinline__ void mbar_arrive(u64 *bar)
{
	asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void mbar_wait(u64 *bar, u32 const parity)
{
	asm volatile(
		"{\n"
		".reg .pred p;\n"
		"WAIT_%=:\n"
		"mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
		"@p bra DONE_%=;\n"
		"bra WAIT_%=;\n"
		"DONE_%=:\n"
		"}" :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}

// Makes mbarrier.init visible to the async proxy before the first bulk copy that signals it.
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

// Orders this thread's generic-proxy shared-memory writes before subsequent async-proxy (bulk copy) reads.
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// global → shared, completion counted in bytes on `bar`. 16-byte aligned addresses, size % 16 == 0.
__device__ __forceinline__ void bulk_g2s(void *smem_dst, void const *gsrc, u32 const bytes, u64 *bar)
{
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
		:: "r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// shared → global as part of the calling thread's current bulk group. 16-byte aligned addresses, size % 16 == 0.
__device__ __forceinline__ void bulk_s2g(void *gdst, void const *smem_src, u32 const bytes)
{
	asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
		:: "l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}

// The same with an L2 cache policy (createpolicy) for the written lines.
__device__ __forceinline__ void bulk_s2g_hint(void *gdst, void const *smem_src, u32 const bytes, u64 const policy)
{
	asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;"
		:: "l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes), "l"(policy) : "memory");
}

__device__ __forceinline__ u64 l2_policy_evict_first()
{
	u64 policy;
	asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
	return policy;
}

__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }

// Waits until at most t_pending of the calling thread's bulk groups have not yet finished READING their shared-memory source.
template <int t_pending>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" :: "n"(t_pending) : "memory"); }

// In-register 32×32 bit-matrix transpose, LSB-first on both sides: on return bit i of a[j] is what bit j of a[i] was.
// Five block-swap stages (16, 8, 4, 2, 1); with full unrolling a[] lives in registers. On the device the 16- and 8-bit
// stages move whole bytes, so each pair costs two PRMT instead of five shift / logic operations.
__host__ __device__ __forceinline__ void transpose32(u32 (&a)[32])
{
#ifdef __CUDA_ARCH__
#pragma unroll
	for (int k = 0; k < 16; ++k)
	{
		u32 const x(a[k]), y(a[k + 16]);
		a[k] = __byte_perm(x, y, 0x5410);
		a[k + 16] = __byte_perm(x, y, 0x7632);
	}
#pragma unroll
	for (int k = 0; k < 32; k = (k + 8 + 1) & ~8)
	{
		u32 const x(a[k]), y(a[k + 8]);
		a[k] = __byte_perm(x, y, 0x6240);
		a[k + 8] = __byte_perm(x, y, 0x7351);
	}
	u32 m(0x0f0f0f0fu);
#pragma unroll
	for (int j = 4; j != 0; j >>= 1, m ^= (m << j))
#else
	u32 m(0x0000ffffu);
	for (int j = 16; j != 0; j >>= 1, m ^= (m << j))
#endif
	{
#pragma unroll
		for (int k = 0; k < 32; k = (k + j + 1) & ~j)
		{
#ifdef __CUDA_ARCH__
			// two shifts + two bit selects (LOP3) per pair; m << j == ~m for these masks
			u32 const x(a[k]), y(a[k + j]);
			a[k] = (x & m) | ((y << j) & ~m);
#if V2M_T32_MULHI
			a[k + j] = (__umulhi(x, 1u << (32 - j)) & m) | (y & ~m);	// x >> j on the FMA pipe (IMAD.HI): the ALU pipe is the busy one in K1
#else
			a[k + j] = ((x >> j) & m) | (y & ~m);
#endif
#else
			u32 const t(((a[k] >> j) ^ a[k + j]) & m);
			a[k + j] ^= t;
			a[k] ^= (t << j);
#endif
		}
	}
}

} // namespace v2m
