// K5 — genotype ingestion: the "Paths" loop of build_variant_graph (libvcf2multialign/variant_graph.cc:379-425), the step
// that produces paths_by_edge_and_chrom_copy from the parsed GT fields.
//
// What the reference does per handled record, restated: every included chromosome copy (output row index
// ploidy_csum[sample] + copy) whose GT allele a is neither 0 nor NULL_ALLELE and whose ALT has an edge
// (edges_by_alt[a - 1] != EDGE_MAX; only plain and <DEL> ALTs get one, :328-364) sets bit (row, edge) (:424). Every ALT gets
// a fresh edge (variant_graph::add_edge), so an edge column belongs to exactly one (record, ALT) pair. Independently, per
// chromosome copy and in record order, an overlap is reported when the record starts before the target of the copy's
// previous ALT edge (:405-418); the report does not change the bits.
//
// Kernels:
//   ingest_genotypes_kernel   one warp per (record, 1024 chromosome copies): 32 coalesced 64-byte loads of GT values; for each
//                             ALT of the record a ballot per load gives one 32-bit word of that ALT's edge column, lane k
//                             keeps word k, and the column piece goes out as one 128-byte store. Algorithmic bytes:
//                             2 · records · H read + E · H_pad / 8 written. Roofline: HBM read.
//   genotype_overlaps_kernel  one thread per chromosome copy walks the records (loads coalesced across copies) and appends
//                             (record, row) pairs; state (the copy's current target position) is kept between record blocks.
#include <algorithm>
#include <atomic>
#include <cstring>
#include <limits>
#include <string>
#include <thread>
#include <vector>

#include <cuda_fp16.h>

#include "../../include/v2m_cuda.h"
#include "common.cuh"
#include "device_graph.hh"

namespace v2m {

constexpr unsigned short GT_NULL = 0xffffu;						// V2M_GT_NULL
constexpr u64 EDGE_NONE = 0xffffffffffffffffull;				// variant_graph::EDGE_MAX

__global__ void __launch_bounds__(256)
ingest_genotypes_scalar_kernel(unsigned short const * __restrict__ gt, u64 const n_records, u32 const H,
	u64 const * __restrict__ alt_offsets /* [n_records + 1], relative to edges_by_alt */, u64 const * __restrict__ edges_by_alt,
	u32 const Wh32 /* h_pad / 32 */, u32 const blocks_per_record, u64 const e_pad, u32 * __restrict__ paths32, u32 * __restrict__ bad)
{
	unsigned const lane(threadIdx.x & 31);
	u64 const warp_id((u64(blockIdx.x) * blockDim.x + threadIdx.x) >> 5);
	u64 const r(warp_id / blocks_per_record);
	u32 const wb(u32(warp_id % blocks_per_record));
	if (r >= n_records) return;
	u64 const a0(alt_offsets[r]);
	u32 const n_alts(u32(alt_offsets[r + 1] - a0));
	unsigned short const *row(gt + r * H);
	u32 const word_mine(wb * 32 + lane);

	for (u32 base(0); base < max(n_alts, 1u); base += 4)
	{
		u32 const n(min(4u, n_alts - min(n_alts, base)));
		u32 w[4] = {0, 0, 0, 0};
#pragma unroll 4
		for (u32 k = 0; k < 32; ++k)
		{
			u32 const h((wb * 32 + k) * 32 + lane);
			u32 const g(h < H ? row[h] : 0u);
			if (0 == base && 0 != g && GT_NULL != g && g > n_alts) atomicOr(bad, 1u);		// the reference asserts (:399)
#pragma unroll
			for (u32 j = 0; j < 4; ++j)
				if (j < n)
				{
					u32 const b(__ballot_sync(0xffffffffu, g == base + j + 1));
					if (lane == k) w[j] = b;
				}
		}
		if (word_mine < Wh32)
#pragma unroll
			for (u32 j = 0; j < 4; ++j)
			{
				if (j >= n) break;
				u64 const edge(edges_by_alt[a0 + base + j]);
				if (EDGE_NONE == edge) continue;
				if (edge >= e_pad) { atomicOr(bad, 2u); continue; }
				paths32[edge * Wh32 + word_mine] = w[j];
			}
	}
}


// The same, for H % 8 == 0 (rows of GT values are then 16-byte aligned): each lane loads 8 GT values with one 128-bit load,
// so a warp covers 256 chromosome copies per load and 1,024 with its four loads in flight together. Per ALT, a lane turns
// its 8 values into 8 bits with packed 16-bit compares, four lanes OR their bytes into a word of the edge column (two
// shuffles), one more shuffle hands word k to lane k, and the 32 words leave as one 128-byte store.
// Bit 15 / bit 31 of the result are set iff the low / high 16-bit half of x equals the half-word replicated in `pattern`
// (exact: the addition never carries out of the low 15 bits of a half).
__device__ __forceinline__ u32 eq_halves(u32 const x, u32 const pattern)
{
	u32 const t(x ^ pattern);
	return ~(((t & 0x7fff7fffu) + 0x7fff7fffu) | t) & 0x80008000u;
}

// Both 16-bit halves of x incremented modulo 2^16 (no carry between the halves).
__device__ __forceinline__ u32 inc_halves(u32 const x)
{
	return ((x & 0x7fff7fffu) + 0x00010001u) ^ (x & 0x80008000u);
}

struct eq4 { u32 x, y, z, w; };
__device__ __forceinline__ eq4 eq_halves4(uint4 const &v, u32 const pattern)
{
	return eq4{eq_halves(v.x, pattern), eq_halves(v.y, pattern), eq_halves(v.z, pattern), eq_halves(v.w, pattern)};
}

// The 8 result bits of one lane's 8 GT values (value t of the lane → bit t).
__device__ __forceinline__ u32 gather8(eq4 const &m)
{
	u32 const r((m.x >> 15) | (m.y >> 13) | (m.z >> 11) | (m.w >> 9));		// low halves → bits 0, 2, 4, 6; high halves → 16, 18, 20, 22
	return (r & 0x55u) | ((r >> 15) & 0xaau);
}

// The same 8 bits through the half-precision compare unit: HSET2.BM gives 0xffff per equal half in one instruction instead
// of four. As binary16 numbers two halves compare equal exactly when their bits are equal as long as the pattern is a
// positive finite number, i.e. 1 … 0x7bff (subnormals are not flushed; NaN patterns never match, -0 only matches 0):
// ALT numbers from 0x7c00 on take the integer compares above.
__device__ __forceinline__ u32 heq_mask(u32 const x, u32 const pattern)
{
	__half2 a, b;
	memcpy(&a, &x, 4);
	memcpy(&b, &pattern, 4);
	return __heq2_mask(a, b);
}

template <bool t_half>
__device__ __forceinline__ u32 match8(uint4 const &v, u32 const pattern)
{
	if constexpr (t_half)
	{
		u32 const r((heq_mask(v.x, pattern) & 0x00020001u) | (heq_mask(v.y, pattern) & 0x00080004u)
			| (heq_mask(v.z, pattern) & 0x00200010u) | (heq_mask(v.w, pattern) & 0x00800040u));
		return (r | (r >> 16)) & 0xffu;						// low halves → bits 0, 2, 4, 6; high halves (bits 17 … 23) → 1, 3, 5, 7
	}
	else
		return gather8(eq_halves4(v, pattern));
}

// Word `lane` of the 1,024-copy piece of one ALT's edge column.
template <bool t_half>
__device__ __forceinline__ u32 column_word(uint4 const (&v)[4], u32 const pattern, unsigned const lane)
{
	u32 mine(0);
#pragma unroll
	for (int i = 0; i < 4; ++i)
	{
		u32 w(match8 <t_half>(v[i], pattern) << (8 * (lane & 3)));
		w |= __shfl_xor_sync(0xffffffffu, w, 1);
		w |= __shfl_xor_sync(0xffffffffu, w, 2);						// lanes 4g … 4g+3 hold word 8i + g of the block
		u32 const t(__shfl_sync(0xffffffffu, w, 4 * (lane & 7)));		// lane L reads word 8i + L % 8
		if (int(lane >> 3) == i) mine = t;
	}
	return mine;
}

#ifndef V2M_K5_MIN_CTAS
#define V2M_K5_MIN_CTAS 6		// 40 registers, 48 warps per SM: 0.371 -> 0.360 ms on 200k records x 5,008 copies (8: spills, 0.390 ms)
#endif
__global__ void __launch_bounds__(256, V2M_K5_MIN_CTAS)
ingest_genotypes_kernel(unsigned short const * __restrict__ gt, u64 const n_records, u32 const H,
	u64 const * __restrict__ alt_offsets, u64 const * __restrict__ edges_by_alt,
	u32 const Wh32, u32 const blocks_per_record, u64 const e_pad, u32 * __restrict__ paths32, u32 * __restrict__ bad)
{
	unsigned const lane(threadIdx.x & 31);
	u64 const warp_id((u64(blockIdx.x) * blockDim.x + threadIdx.x) >> 5);
	u64 const r(warp_id / blocks_per_record);
	u32 const wb(u32(warp_id % blocks_per_record));
	if (r >= n_records) return;
	u64 const a0(alt_offsets[r]);
	u32 const n_alts(u32(alt_offsets[r + 1] - a0));
	unsigned short const *row(gt + r * H);

	uint4 v[4];
#pragma unroll
	for (int i = 0; i < 4; ++i)
	{
		u32 const h0(wb * 1024 + i * 256 + lane * 8);
		v[i] = (h0 < H) ? ld_nc_v4(row + h0) : make_uint4(0, 0, 0, 0);
	}
	// :399 — every value must be 0, NULL or one of the record's ALTs, i.e. value + 1 (mod 2^16: NULL becomes 0) must not
	// exceed n_alts + 1. One packed increment per word and a running packed maximum (VIMNMX3.U16x2) cost ~60 instructions
	// per warp; collecting the matched halves per ALT next to tests against 0 and NULL cost ~200 of the kernel's 397.
	u32 vmax(0);
#pragma unroll
	for (int i = 0; i < 4; ++i)
	{
		vmax = __vimax3_u16x2(vmax, inc_halves(v[i].x), inc_halves(v[i].y));
		vmax = __vimax3_u16x2(vmax, inc_halves(v[i].z), inc_halves(v[i].w));
	}
	u32 const word_mine(wb * 32 + lane);
	for (u32 a(0); a < n_alts; ++a)
	{
		u64 const edge(edges_by_alt[a0 + a]);
		u32 const pattern((a + 1) | ((a + 1) << 16));
		u32 const mine(a + 1 < 0x7c00u ? column_word <true>(v, pattern, lane) : column_word <false>(v, pattern, lane));		// warp-uniform
		if (EDGE_NONE == edge) continue;
		if (edge >= e_pad) { atomicOr(bad, 2u); continue; }
		if (word_mine < Wh32) paths32[edge * Wh32 + word_mine] = mine;
	}
	u32 const limit(min(n_alts, 0xfffeu) + 1);
	if ((vmax & 0xffffu) > limit || (vmax >> 16) > limit) atomicOr(bad, 1u);
}


__global__ void __launch_bounds__(128)
genotype_overlaps_kernel(unsigned short const * __restrict__ gt, u64 const n_records, u64 const first_record, u32 const H,
	u64 const * __restrict__ alt_offsets, u64 const * __restrict__ edges_by_alt,
	u64 const * __restrict__ record_ref_pos, u64 const * __restrict__ edge_target_pos, u64 const n_edges,
	u64 * __restrict__ target_by_copy, u64 * __restrict__ out_pairs, u64 const cap, unsigned long long * __restrict__ out_count)
{
	u32 const h(blockIdx.x * blockDim.x + threadIdx.x);
	if (h >= H) return;
	u64 target(target_by_copy[h]);
	constexpr int U = 8;
	for (u64 r0(0); r0 < n_records; r0 += U)
	{
		u32 g[U];
#pragma unroll
		for (int u = 0; u < U; ++u)
			g[u] = (r0 + u < n_records) ? gt[(r0 + u) * H + h] : 0u;
#pragma unroll
		for (int u = 0; u < U; ++u)
		{
			if (0 == g[u] || GT_NULL == g[u]) continue;
			u64 const r(r0 + u);
			u64 const a0(alt_offsets[r]);
			if (g[u] > alt_offsets[r + 1] - a0) continue;					// reported by the ingest kernel
			u64 const edge(edges_by_alt[a0 + g[u] - 1]);
			if (EDGE_NONE == edge || edge >= n_edges) continue;
			if (record_ref_pos[r] < target)
			{
				unsigned long long const slot(atomicAdd(out_count, 1ull));
				if (slot < cap) { out_pairs[2 * slot] = first_record + r; out_pairs[2 * slot + 1] = h; }
			}
			target = edge_target_pos[edge];
		}
	}
	target_by_copy[h] = target;
}

namespace {
	struct dbuf
	{
		void *p{};
		~dbuf() { if (p) cudaFree(p); }
		int alloc(u64 bytes) { V2M_CUDA_TRY(cudaMalloc(&p, std::max<u64>(bytes, 16))); return 0; }
		template <typename T> T *as() const { return static_cast<T *>(p); }
	};
}

namespace {
	void launch_ingest(unsigned short const *d_gt, u64 nr, u32 H, u64 const *d_off, u64 const *d_eba, u32 Wh32, u64 e_pad, u32 *d_paths32, u32 *d_bad, cudaStream_t st)
	{
		bool const vec(0 == H % 8 && 0 == (reinterpret_cast<u64>(d_gt) & 15));
		u32 const blocks_per_record((Wh32 + 31) / 32);			// 1,024 chromosome copies per warp in both kernels
		u64 const warps(nr * blocks_per_record);
		unsigned const ctas(unsigned((warps * 32 + 255) / 256));
		if (vec) ingest_genotypes_kernel<<<ctas, 256, 0, st>>>(d_gt, nr, H, d_off, d_eba, Wh32, blocks_per_record, e_pad, d_paths32, d_bad);
		else ingest_genotypes_scalar_kernel<<<ctas, 256, 0, st>>>(d_gt, nr, H, d_off, d_eba, Wh32, blocks_per_record, e_pad, d_paths32, d_bad);
	}
}

// Core: fills d_paths (device, e_pad × h_pad / 64 words, zeroed here). gt is host memory and is streamed in record blocks.
// d_paths_he (optional): paths_by_chrom_copy_and_edge, the orientation build_variant_graph derives with transpose_matrix at its
// end (variant_graph.cc:453). Here it is born with the bits: after every record block the edge columns that are final — all
// edges below the smallest edge index of any later record, in whole 64-edge words — are transposed into their place in the
// copy-major lines while the block's output is still in the L2 (a block writes ~4 MB of edge columns), so a graph built on
// the device never runs K1 over the whole matrix. Fusing the transposed stores into the ingest kernel itself was considered
// and not done: a warp holds one record's 1 – 2 edge columns, the copy-major lines want 32 consecutive edges per word, and
// K5 is bound by reading 2 bytes per genotype — the transposed orientation is 6 % of its traffic.
int paths_from_genotypes_device(unsigned short const *gt, u64 const n_records, u32 const H, u64 const *alt_offsets, u64 const *edges_by_alt,
	u64 const *record_ref_pos, u64 const *edge_target_pos, u64 const n_edges, u64 const h_pad, u64 const e_pad, void *d_paths,
	std::vector<u64> *overlaps, void *d_paths_he = nullptr)
{
	if (0 != h_pad % 64 || 0 != e_pad % 64 || h_pad < H) return fail(V2M_ERR_ARGUMENT, "paths_from_genotypes: path matrix dimensions must be multiples of 64 covering all chromosome copies");
	if (n_records && (!gt || !alt_offsets)) return fail(V2M_ERR_ARGUMENT, "paths_from_genotypes: NULL argument");
	u64 const n_alts(n_records ? alt_offsets[n_records] : 0);
	if (n_alts && !edges_by_alt) return fail(V2M_ERR_ARGUMENT, "paths_from_genotypes: NULL edges_by_alt");
	bool const want_overlaps(overlaps && record_ref_pos && edge_target_pos);
	if (overlaps) overlaps->clear();
	V2M_CUDA_TRY(cudaMemset(d_paths, 0, e_pad * h_pad / 8));
	if (0 == n_records || 0 == H)
	{
		if (d_paths_he) V2M_CUDA_TRY(cudaMemset(d_paths_he, 0, e_pad * h_pad / 8));
		return V2M_OK;
	}

	u32 const Wh32(u32(h_pad / 32));
	u64 block_records(std::max<u64>(1, std::min<u64>(n_records, (u64(64) << 20) / (u64(H) * 2))));		// ~64 MB of GT values per block
	if (char const *env = std::getenv("V2M_K5_BLOCK_RECORDS"))		// tests force small blocks
		if (long long const v = std::atoll(env); v > 0) block_records = std::min<u64>(n_records, u64(v));
	// Smallest edge index of any ALT of record r or later: the edge columns below it are final once records < r are done.
	std::vector<u64> frontier;
	if (d_paths_he)
	{
		frontier.assign(n_records + 1, e_pad);
		for (u64 r(n_records); r-- > 0; )
		{
			u64 m(frontier[r + 1]);
			for (u64 a(alt_offsets[r]); a < alt_offsets[r + 1]; ++a)
				if (EDGE_NONE != edges_by_alt[a]) m = std::min(m, edges_by_alt[a]);
			frontier[r] = m;
		}
	}
	u64 transposed_upto(0);			// edge columns [0, transposed_upto) are in d_paths_he already (a multiple of 64)

	dbuf d_gt[2], d_off, d_eba, d_ref, d_tgt, d_state, d_pairs, d_count, d_bad;
	if (d_gt[0].alloc(block_records * H * 2) || d_gt[1].alloc(block_records * H * 2) || d_off.alloc((n_records + 1) * 8) || d_eba.alloc(n_alts * 8) || d_bad.alloc(4))
		return V2M_ERR_CUDA;
	V2M_CUDA_TRY(cudaMemcpy(d_off.p, alt_offsets, (n_records + 1) * 8, cudaMemcpyHostToDevice));
	if (n_alts) V2M_CUDA_TRY(cudaMemcpy(d_eba.p, edges_by_alt, n_alts * 8, cudaMemcpyHostToDevice));
	V2M_CUDA_TRY(cudaMemset(d_bad.p, 0, 4));
	u64 pair_cap(0);
	if (want_overlaps)
	{
		pair_cap = std::max<u64>(1024, n_records);
		if (d_ref.alloc(n_records * 8) || d_tgt.alloc(n_edges * 8) || d_state.alloc(u64(H) * 8) || d_pairs.alloc(pair_cap * 16) || d_count.alloc(8)) return V2M_ERR_CUDA;
		V2M_CUDA_TRY(cudaMemcpy(d_ref.p, record_ref_pos, n_records * 8, cudaMemcpyHostToDevice));
		if (n_edges) V2M_CUDA_TRY(cudaMemcpy(d_tgt.p, edge_target_pos, n_edges * 8, cudaMemcpyHostToDevice));
	}

	// GT values arrive in pageable host memory: blocks are copied into two pinned buffers by a few threads and sent with
	// asynchronous copies, so that the host copy of block b + 1 overlaps the transfer and the kernels of block b.
	cudaStream_t streams[2];
	cudaEvent_t done[2];
	// The two pinned staging buffers are kept per calling thread between calls: page-locking 2 × 64 MB costs more than
	// sending them.
	struct pinned_pair
	{
		unsigned short *h[2] = {nullptr, nullptr};
		u64 bytes{};
		~pinned_pair() { for (auto p : h) if (p) cudaFreeHost(p); }
	};
	thread_local pinned_pair staging;
	unsigned short **h_gt(staging.h);
	for (auto &s : streams) V2M_CUDA_TRY(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
	for (auto &e : done) V2M_CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
	struct guard_t
	{
		cudaStream_t *s; cudaEvent_t *e;
		~guard_t() { for (int i(0); i < 2; ++i) { cudaStreamSynchronize(s[i]); cudaStreamDestroy(s[i]); cudaEventDestroy(e[i]); } }
	} const guard{streams, done};
	if (staging.bytes < block_records * H * 2)
	{
		for (auto &h : staging.h) { if (h) cudaFreeHost(h); h = nullptr; }
		staging.bytes = 0;
		for (auto &h : staging.h)
			if (cudaSuccess != cudaMallocHost(reinterpret_cast<void **>(&h), block_records * H * 2)) return fail(V2M_ERR_CUDA, "paths_from_genotypes: cudaMallocHost failed");
		staging.bytes = block_records * H * 2;
	}
	unsigned const copy_threads(std::max(1u, std::min(16u, std::thread::hardware_concurrency())));
	auto const stage([&](unsigned short *dst, unsigned short const *src, u64 const bytes) {
		u64 const piece(((bytes + copy_threads - 1) / copy_threads + 4095) & ~u64(4095));		// rounded up: never 0 for bytes > 0
		std::vector<std::thread> pool;
		for (unsigned t(1); t < copy_threads; ++t)
			if (u64(t) * piece < bytes)
				pool.emplace_back([=]() { std::memcpy(reinterpret_cast<char *>(dst) + t * piece, reinterpret_cast<char const *>(src) + t * piece, std::min(piece, bytes - t * piece)); });
		std::memcpy(dst, src, std::min(piece, bytes));
		for (auto &th : pool) th.join();
	});

	for (int attempt(0); attempt < 2; ++attempt)
	{
		if (want_overlaps)
		{
			V2M_CUDA_TRY(cudaMemset(d_state.p, 0, u64(H) * 8));
			V2M_CUDA_TRY(cudaMemset(d_count.p, 0, 8));
		}
		V2M_CUDA_TRY(cudaDeviceSynchronize());
		int k(0);
		bool first(true);
		for (u64 r0(0); r0 < n_records; r0 += block_records, k ^= 1)
		{
			u64 const nr(std::min(block_records, n_records - r0));
			cudaStream_t const st(streams[k]);
			V2M_CUDA_TRY(cudaStreamSynchronize(st));			// buffers k are free again
			stage(h_gt[k], gt + r0 * H, nr * H * 2);
			V2M_CUDA_TRY(cudaMemcpyAsync(d_gt[k].p, h_gt[k], nr * H * 2, cudaMemcpyHostToDevice, st));
			if (0 == attempt)
			{
				launch_ingest(d_gt[k].as<unsigned short>(), nr, H, d_off.as<u64>() + r0, d_eba.as<u64>(), Wh32, e_pad, static_cast<u32 *>(d_paths), d_bad.as<u32>(), st);
				V2M_CHECK_LAUNCH("ingest_genotypes_kernel");
			}
			if (want_overlaps)
			{
				// The overlap kernel carries per-copy state from block to block: block b follows block b - 1 (on the other stream).
				if (!first) V2M_CUDA_TRY(cudaStreamWaitEvent(st, done[k ^ 1], 0));
				genotype_overlaps_kernel<<<(H + 127) / 128, 128, 0, st>>>(d_gt[k].as<unsigned short>(), nr, r0, H, d_off.as<u64>() + r0, d_eba.as<u64>(),
					d_ref.as<u64>() + r0, d_tgt.as<u64>(), n_edges, d_state.as<u64>(), d_pairs.as<u64>(), pair_cap, d_count.as<unsigned long long>());
				V2M_CHECK_LAUNCH("genotype_overlaps_kernel");
			}
			if (0 == attempt && d_paths_he)
			{
				u64 const upto(std::min(frontier[r0 + nr], e_pad) / 64 * 64);
				if (upto > transposed_upto)
				{
					// the range may reach back into columns the previous block (other stream) wrote
					if (!first && !want_overlaps) V2M_CUDA_TRY(cudaStreamWaitEvent(st, done[k ^ 1], 0));
					if (int const rc = launch_transpose_bits_into(static_cast<u64 const *>(d_paths) + transposed_upto * (h_pad / 64), h_pad, upto - transposed_upto,
							static_cast<u64 *>(d_paths_he) + transposed_upto / 64, e_pad, st)) return rc;
					transposed_upto = upto;
				}
			}
			V2M_CUDA_TRY(cudaEventRecord(done[k], st));
			first = false;
		}
		V2M_CUDA_TRY(cudaStreamSynchronize(streams[0]));
		V2M_CUDA_TRY(cudaStreamSynchronize(streams[1]));
		if (0 == attempt && d_paths_he && transposed_upto < e_pad)
		{
			// what is left: the last partial word's columns and the zero padding up to e_pad
			if (int const rc = launch_transpose_bits_into(static_cast<u64 const *>(d_paths) + transposed_upto * (h_pad / 64), h_pad, e_pad - transposed_upto,
					static_cast<u64 *>(d_paths_he) + transposed_upto / 64, e_pad, nullptr)) return rc;
			transposed_upto = e_pad;
			V2M_CUDA_TRY(cudaDeviceSynchronize());
		}
		u32 bad(0);
		V2M_CUDA_TRY(cudaMemcpy(&bad, d_bad.p, 4, cudaMemcpyDeviceToHost));
		if (bad & 1u) return fail(V2M_ERR_ARGUMENT, "paths_from_genotypes: a genotype refers to an ALT its record does not have (the reference asserts, variant_graph.cc:399)");
		if (bad & 2u) return fail(V2M_ERR_ARGUMENT, "paths_from_genotypes: edge index beyond e_pad");
		if (!want_overlaps) break;
		unsigned long long count(0);
		V2M_CUDA_TRY(cudaMemcpy(&count, d_count.p, 8, cudaMemcpyDeviceToHost));
		if (count <= pair_cap)
		{
			overlaps->resize(2 * count);
			if (count) V2M_CUDA_TRY(cudaMemcpy(overlaps->data(), d_pairs.p, count * 16, cudaMemcpyDeviceToHost));
			// The reference reports in record order, then in output row order within the record.
			std::vector<std::pair<u64, u64>> pairs(count);
			for (u64 i(0); i < count; ++i) pairs[i] = {(*overlaps)[2 * i], (*overlaps)[2 * i + 1]};
			std::sort(pairs.begin(), pairs.end());
			for (u64 i(0); i < count; ++i) { (*overlaps)[2 * i] = pairs[i].first; (*overlaps)[2 * i + 1] = pairs[i].second; }
			break;
		}
		if (attempt) return fail(V2M_ERR_STATE, "paths_from_genotypes: overlap list overflow after resizing");
		pair_cap = count;
		cudaFree(d_pairs.p); d_pairs.p = nullptr;
		if (d_pairs.alloc(pair_cap * 16)) return V2M_ERR_CUDA;
	}
	return V2M_OK;
}

} // namespace v2m


extern "C" int v2m_cuda_paths_from_genotypes_device(const uint16_t *gt, uint64_t n_records, uint32_t n_copies,
	const uint64_t *alt_offsets, const uint64_t *edges_by_alt, const uint64_t *record_ref_pos, const uint64_t *edge_target_pos, uint64_t n_edges,
	uint64_t h_pad, uint64_t e_pad, void *d_paths, uint64_t *overlaps, uint64_t overlap_cap, uint64_t *n_overlaps)
{
	using namespace v2m;
	if (!d_paths) return fail(V2M_ERR_ARGUMENT, "paths_from_genotypes: NULL output");
	std::vector<u64> ov;
	bool const want(nullptr != n_overlaps);
	if (int const st = paths_from_genotypes_device(gt, n_records, n_copies, alt_offsets, edges_by_alt, record_ref_pos, edge_target_pos, n_edges,
			h_pad, e_pad, d_paths, want ? &ov : nullptr)) return st;
	if (want)
	{
		*n_overlaps = ov.size() / 2;
		if (ov.size() / 2 > overlap_cap || (!overlaps && !ov.empty()))
			return fail(V2M_ERR_ARGUMENT, "paths_from_genotypes: overlap capacity too small (needed count is in *n_overlaps)");
		std::copy(ov.begin(), ov.end(), overlaps);
	}
	return V2M_OK;
}

extern "C" int v2m_cuda_paths_from_genotypes_both(const uint16_t *gt, uint64_t n_records, uint32_t n_copies,
	const uint64_t *alt_offsets, const uint64_t *edges_by_alt, const uint64_t *record_ref_pos, const uint64_t *edge_target_pos, uint64_t n_edges,
	uint64_t h_pad, uint64_t e_pad, void *d_paths_by_edge_and_chrom_copy, void *d_paths_by_chrom_copy_and_edge,
	uint64_t *overlaps, uint64_t overlap_cap, uint64_t *n_overlaps)
{
	using namespace v2m;
	if (!d_paths_by_edge_and_chrom_copy || !d_paths_by_chrom_copy_and_edge) return fail(V2M_ERR_ARGUMENT, "paths_from_genotypes_both: NULL output");
	std::vector<u64> ov;
	bool const want(nullptr != n_overlaps);
	if (int const st = paths_from_genotypes_device(gt, n_records, n_copies, alt_offsets, edges_by_alt, record_ref_pos, edge_target_pos, n_edges,
			h_pad, e_pad, d_paths_by_edge_and_chrom_copy, want ? &ov : nullptr, d_paths_by_chrom_copy_and_edge)) return st;
	if (want)
	{
		*n_overlaps = ov.size() / 2;
		if (ov.size() / 2 > overlap_cap || (!overlaps && !ov.empty()))
			return fail(V2M_ERR_ARGUMENT, "paths_from_genotypes: overlap capacity too small (needed count is in *n_overlaps)");
		std::copy(ov.begin(), ov.end(), overlaps);
	}
	return V2M_OK;
}

extern "C" int v2m_cuda_paths_from_genotypes(const uint16_t *gt, uint64_t n_records, uint32_t n_copies,
	const uint64_t *alt_offsets, const uint64_t *edges_by_alt, const uint64_t *record_ref_pos, const uint64_t *edge_target_pos, uint64_t n_edges,
	uint64_t h_pad, uint64_t e_pad, uint64_t *paths_by_edge_and_chrom_copy, uint64_t *overlaps, uint64_t overlap_cap, uint64_t *n_overlaps)
{
	using namespace v2m;
	u64 const bytes(e_pad * h_pad / 8);
	if (bytes && !paths_by_edge_and_chrom_copy) return fail(V2M_ERR_ARGUMENT, "paths_from_genotypes: NULL output");
	void *d_paths(nullptr);
	V2M_CUDA_TRY(cudaMalloc(&d_paths, std::max<u64>(bytes, 16)));
	int st(v2m_cuda_paths_from_genotypes_device(gt, n_records, n_copies, alt_offsets, edges_by_alt, record_ref_pos, edge_target_pos, n_edges,
		h_pad, e_pad, d_paths, overlaps, overlap_cap, n_overlaps));
	if (!st && bytes && cudaSuccess != cudaMemcpy(paths_by_edge_and_chrom_copy, d_paths, bytes, cudaMemcpyDeviceToHost))
		st = fail(V2M_ERR_CUDA, "paths_from_genotypes: D2H copy failed");
	cudaFree(d_paths);
	return st;
}

// Measurement hook: mean device time of `iters` launches of ingest_genotypes_kernel on GT values already resident in HBM.
extern "C" int v2m_cuda_time_ingest_genotypes(const void *d_gt, uint64_t n_records, uint32_t n_copies, const void *d_alt_offsets, const void *d_edges_by_alt,
	uint64_t h_pad, uint64_t e_pad, void *d_paths, int iters, float *ms_mean)
{
	using namespace v2m;
	if (!d_gt || !d_alt_offsets || !d_edges_by_alt || !d_paths || !ms_mean || iters < 1) return fail(V2M_ERR_ARGUMENT, "time_ingest_genotypes: bad argument");
	u32 const Wh32(u32(h_pad / 32));
	u32 *d_bad(nullptr);
	V2M_CUDA_TRY(cudaMalloc(reinterpret_cast<void **>(&d_bad), 4));
	V2M_CUDA_TRY(cudaMemset(d_bad, 0, 4));
	cudaEvent_t e0, e1;
	V2M_CUDA_TRY(cudaEventCreate(&e0)); V2M_CUDA_TRY(cudaEventCreate(&e1));
	cudaEventRecord(e0, nullptr);
	for (int i(0); i < iters; ++i)
	{
		launch_ingest(static_cast<unsigned short const *>(d_gt), n_records, n_copies, static_cast<u64 const *>(d_alt_offsets), static_cast<u64 const *>(d_edges_by_alt),
			Wh32, e_pad, static_cast<u32 *>(d_paths), d_bad, nullptr);
		count_launch();
	}
	cudaEventRecord(e1, nullptr);
	cudaError_t const err(cudaEventSynchronize(e1));
	float ms(0);
	cudaEventElapsedTime(&ms, e0, e1);
	cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d_bad);
	if (cudaSuccess != err) return fail(V2M_ERR_CUDA, std::string("time_ingest_genotypes: ") + cudaGetErrorString(err));
	*ms_mean = ms / float(iters);
	return V2M_OK;
}


// ---- K5 without allocations ---------------------------------------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc (CUDA
// graph capture, tight serving loops, harnesses that forbid it). The scratch the two K5 kernels need — the error flags, the
// overlap counter and the per-chromosome-copy "target of the previous ALT edge" state of genotype_overlaps_kernel — lives in
// one file-scope device array instead of cudaMalloc'ed buffers. Every GPU that loads the library holds the array until the
// process exits; every call on a GPU shares it, so calls that overlap on that GPU race (the header says so).
namespace v2m {

constexpr u64 NO_ALLOC_STATE_OFFSET = 256;			// [0, 4): error flags, [8, 16): overlap count, [256, 256 + 8 * copies): state
static_assert(V2M_NO_ALLOC_SCRATCH_BYTES == NO_ALLOC_STATE_OFFSET + 8ull * V2M_NO_ALLOC_MAX_COPIES, "scratch layout");
static __device__ __align__(256) unsigned char k5_no_alloc_scratch[V2M_NO_ALLOC_SCRATCH_BYTES];

namespace {
	// The array's name is only a stand-in on the host: the device address is asked for once per device, with that device current.
	int no_alloc_scratch(unsigned char **out)
	{
		constexpr int MAX_DEVICES = 64;
		static std::atomic<unsigned char *> by_device[MAX_DEVICES];
		int dev(-1);
		V2M_CUDA_TRY(cudaGetDevice(&dev));
		if (dev < 0 || dev >= MAX_DEVICES) return fail(V2M_ERR_ARGUMENT, "ingest_genotypes_no_alloc: device index beyond 63");
		unsigned char *p(by_device[dev].load(std::memory_order_acquire));
		if (!p)
		{
			void *sym(nullptr);
			cudaError_t const err(cudaGetSymbolAddress(&sym, k5_no_alloc_scratch));
			if (cudaSuccess != err || !sym)
				return fail(V2M_ERR_CUDA, std::string("ingest_genotypes_no_alloc: cudaGetSymbolAddress: ") + cudaGetErrorString(err));
			p = static_cast<unsigned char *>(sym);
			by_device[dev].store(p, std::memory_order_release);		// two threads racing here store the same address
		}
		*out = p;
		return V2M_OK;
	}
}

} // namespace v2m

extern "C" int v2m_cuda_ingest_genotypes_no_alloc(const void *d_gt, uint64_t n_records, uint32_t n_copies,
	const void *d_alt_offsets, const void *d_edges_by_alt, const void *d_record_ref_pos, const void *d_edge_target_pos, uint64_t n_edges,
	uint64_t h_pad, uint64_t e_pad, void *d_paths, void *d_overlaps, uint64_t overlap_cap, uint64_t *n_overlaps, void *stream)
{
	using namespace v2m;
	if (!d_paths) return fail(V2M_ERR_ARGUMENT, "ingest_genotypes_no_alloc: NULL output");
	if (0 != h_pad % 64 || 0 != e_pad % 64 || h_pad < n_copies)
		return fail(V2M_ERR_ARGUMENT, "ingest_genotypes_no_alloc: path matrix dimensions must be multiples of 64 covering all chromosome copies");
	if (n_copies > V2M_NO_ALLOC_MAX_COPIES)
		return fail(V2M_ERR_ARGUMENT, "ingest_genotypes_no_alloc: " + std::to_string(n_copies) + " chromosome copies exceed the static scratch (V2M_NO_ALLOC_MAX_COPIES = "
			+ std::to_string(V2M_NO_ALLOC_MAX_COPIES) + "); use v2m_cuda_paths_from_genotypes_device");
	if (h_pad / 32 > 0xffffffffull) return fail(V2M_ERR_ARGUMENT, "ingest_genotypes_no_alloc: h_pad too large");
	if (n_records && n_copies && (!d_gt || !d_alt_offsets || !d_edges_by_alt)) return fail(V2M_ERR_ARGUMENT, "ingest_genotypes_no_alloc: NULL argument");
	bool const want_overlaps(nullptr != n_overlaps);
	if (want_overlaps && (!d_record_ref_pos || !d_edge_target_pos)) return fail(V2M_ERR_ARGUMENT, "ingest_genotypes_no_alloc: overlaps need record_ref_pos and edge_target_pos");
	if (want_overlaps && overlap_cap && !d_overlaps) return fail(V2M_ERR_ARGUMENT, "ingest_genotypes_no_alloc: NULL overlap buffer");

	cudaStream_t const st(static_cast<cudaStream_t>(stream));
	unsigned char *scratch(nullptr);
	if (int const rc = no_alloc_scratch(&scratch)) return rc;
	u32 *const d_bad(reinterpret_cast<u32 *>(scratch));
	unsigned long long *const d_count(reinterpret_cast<unsigned long long *>(scratch + 8));
	u64 *const d_state(reinterpret_cast<u64 *>(scratch + NO_ALLOC_STATE_OFFSET));

	// The array keeps the last call's contents: flags, counter and the state of the copies of this call start at zero.
	V2M_CUDA_TRY(cudaMemsetAsync(scratch, 0, NO_ALLOC_STATE_OFFSET + 8ull * n_copies, st));
	if (e_pad && h_pad) V2M_CUDA_TRY(cudaMemsetAsync(d_paths, 0, e_pad * h_pad / 8, st));
	if (n_records && n_copies)
	{
		launch_ingest(static_cast<unsigned short const *>(d_gt), n_records, n_copies, static_cast<u64 const *>(d_alt_offsets), static_cast<u64 const *>(d_edges_by_alt),
			u32(h_pad / 32), e_pad, static_cast<u32 *>(d_paths), d_bad, st);
		V2M_CHECK_LAUNCH("ingest_genotypes_kernel (no_alloc)");
		if (want_overlaps)
		{
			genotype_overlaps_kernel<<<(n_copies + 127) / 128, 128, 0, st>>>(static_cast<unsigned short const *>(d_gt), n_records, 0, n_copies,
				static_cast<u64 const *>(d_alt_offsets), static_cast<u64 const *>(d_edges_by_alt), static_cast<u64 const *>(d_record_ref_pos),
				static_cast<u64 const *>(d_edge_target_pos), n_edges, d_state, static_cast<u64 *>(d_overlaps), overlap_cap, d_count);
			V2M_CHECK_LAUNCH("genotype_overlaps_kernel (no_alloc)");
		}
	}
	// Flags and counter come back through a 16-byte copy into this frame; the call returns when the stream has drained.
	u64 head[2] = {0, 0};
	V2M_CUDA_TRY(cudaMemcpyAsync(head, scratch, 16, cudaMemcpyDeviceToHost, st));
	V2M_CUDA_TRY(cudaStreamSynchronize(st));
	u32 const bad{static_cast<u32>(head[0])};
	if (bad & 1u) return fail(V2M_ERR_ARGUMENT, "ingest_genotypes_no_alloc: a genotype refers to an ALT its record does not have (the reference asserts, variant_graph.cc:399)");
	if (bad & 2u) return fail(V2M_ERR_ARGUMENT, "ingest_genotypes_no_alloc: edge index beyond e_pad");
	if (want_overlaps)
	{
		*n_overlaps = head[1];
		if (head[1] > overlap_cap) return fail(V2M_ERR_ARGUMENT, "ingest_genotypes_no_alloc: overlap capacity too small (needed count is in *n_overlaps)");
	}
	return V2M_OK;
}
