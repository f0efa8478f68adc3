// K-synth: synthetic path bits for the benchmark shapes — measurement infrastructure, not part of the
// reference's path. bit(h, e) is a pure function of (seed, chromosome copy h, site of e), mirrored in numpy by
// vcf2multialign_b200/synth.py:allele_numpy so that small shapes can be cross-checked on the CPU.
#include "../../include/v2m_cuda.h"
#include "common.cuh"

namespace v2m {

__host__ __device__ __forceinline__ u32 hash32(u32 x)
{
	x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
	return x;
}

__host__ __device__ __forceinline__ u32 mix(u32 seed, u32 a, u32 b) { return hash32(hash32(a ^ seed) + b); }

__host__ __device__ __forceinline__ u32 synth_allele(u32 seed, u32 h, u32 site, u32 n_alts, u32 thr)
{
	u32 const off(mix(seed, h, 23) & 8191u);
	u32 const block((site + off) >> 13);
	u32 const founder(mix(seed, h * 0x9e3779b1u + block, 37) & 63u);
	u32 const fh(mix(seed, site * 64u + founder, 41));
	bool const nonref((fh & 0xffffu) < thr);
	u32 alt(n_alts > 1 ? (fh >> 16) & 1u : 0u);
	bool const mut(0 == (mix(seed, h * 0x85ebca6bu + site, 53) & 1023u));
	if (mut && !nonref) alt = 0;
	return (nonref != mut) ? alt + 1 : 0;
}

__global__ void __launch_bounds__(256)
synth_paths_kernel(u64 * __restrict__ dst, u64 const Wh, u64 const E, u32 const H, u32 const seed,
	u32 const * __restrict__ edge_site, u8 const * __restrict__ edge_alt, u8 const * __restrict__ edge_nalts, u64 const total)
{
	u64 const idx(u64(blockIdx.x) * blockDim.x + threadIdx.x);
	if (idx >= total) return;
	u64 const e(idx / Wh), w(idx % Wh);
	u64 word(0);
	if (e < E)
	{
		u32 const site(edge_site[e]), want(1u + edge_alt[e]), n_alts(edge_nalts[e]);
		u32 const thr(65536u >> (2u + mix(seed, site, 11) % 10u));
		for (u32 b(0); b < 64; ++b)
		{
			u32 const h(u32(64 * w) + b);
			if (h < H && want == synth_allele(seed, h, site, n_alts, thr)) word |= u64(1) << b;
		}
	}
	dst[idx] = word;
}

} // namespace v2m

extern "C" int v2m_cuda_synth_paths(void *d_dst, uint64_t h_pad, uint64_t e_pad, uint32_t H, uint64_t E, uint32_t seed,
	const uint32_t *edge_site, const uint8_t *edge_alt, const uint8_t *edge_nalts)
{
	using namespace v2m;
	if (!d_dst || 0 != h_pad % 64 || 0 != e_pad % 64 || (E && (!edge_site || !edge_alt || !edge_nalts)))
		return fail(V2M_ERR_ARGUMENT, "synth_paths: bad argument");
	u64 const Wh(h_pad / 64), total(Wh * e_pad);
	if (0 == total) return V2M_OK;
	u32 *d_site(nullptr); u8 *d_alt(nullptr), *d_nalts(nullptr);
	int st(V2M_OK);
	do
	{
		if (cudaSuccess != cudaMalloc(&d_site, std::max<u64>(E, 1) * 4) || cudaSuccess != cudaMalloc(&d_alt, std::max<u64>(E, 1)) || cudaSuccess != cudaMalloc(&d_nalts, std::max<u64>(E, 1)))
		{ st = fail(V2M_ERR_CUDA, "synth_paths: cudaMalloc failed"); break; }
		if (E && (cudaSuccess != cudaMemcpy(d_site, edge_site, E * 4, cudaMemcpyHostToDevice) || cudaSuccess != cudaMemcpy(d_alt, edge_alt, E, cudaMemcpyHostToDevice)
			|| cudaSuccess != cudaMemcpy(d_nalts, edge_nalts, E, cudaMemcpyHostToDevice)))
		{ st = fail(V2M_ERR_CUDA, "synth_paths: H2D copy failed"); break; }
		synth_paths_kernel<<<unsigned((total + 255) / 256), 256>>>(static_cast<u64 *>(d_dst), Wh, E, H, seed, d_site, d_alt, d_nalts, total);
		cudaError_t const err(cudaDeviceSynchronize());
		if (cudaSuccess != err) { st = fail(V2M_ERR_CUDA, std::string("synth_paths: ") + cudaGetErrorString(err)); break; }
		count_launch();
	} while (false);
	cudaFree(d_site); cudaFree(d_alt); cudaFree(d_nalts);
	return st;
}
