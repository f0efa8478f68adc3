// Shared device code of the chunked pBWT sweep (K3 founders.cu, K4 cut_positions.cu): the CTA-wide column step of
// pbwt_context::update_divergence (include/vcf2multialign/pbwt.hh:77-134), column staging by bulk async copies, and the
// first two passes of the three-pass scan (chunk-local order, exact state at every chunk boundary). See founders.cu for the
// description of the scheme.
#pragma once

#include <algorithm>
#include <cstdlib>
#include <vector>

#include "../../include/v2m_cuda.h"
#include "common.cuh"
#include "device_graph.hh"

namespace v2m {

constexpr int K3_THREADS = 1024;
constexpr int K3_WARPS = K3_THREADS / 32;
constexpr int K3_KW_MAX = 16;					// 32-position words per warp → H <= 32 · 16 · 32 = 16,384
constexpr int K3_KW_REGS = 8;					// up to here a thread keeps (position, divergence, zero mask) per word in registers
constexpr int K3_COLB = 16;						// columns per bulk copy
constexpr u32 K3_SENT = 0xffffffffu;			// "differs on an artificial column" marker of pass B
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ u32 lanes_below(unsigned const lane) { return lane ? (FULL >> (32 - lane)) : 0u; }
__device__ __forceinline__ u32 lanes_upto(unsigned const lane) { return FULL >> (31 - lane); }

// Inclusive max-scan over the lanes of a warp, restarted at every lane whose bit is set in `heads`.
__device__ __forceinline__ u32 seg_max_scan(u32 v, u32 const heads, unsigned const lane)
{
#pragma unroll
	for (unsigned d = 1; d < 32; d <<= 1)
	{
		u32 const t(__shfl_up_sync(FULL, v, d));
		// take the partial ending at lane - d iff no segment starts in (lane - d, lane]
		if (lane >= d && 0 == (heads & lanes_upto(lane) & ~lanes_upto(lane - d))) v = max(v, t);
	}
	return v;
}

// Shared-memory state of one sweep.
struct k3_state
{
	unsigned short	*perm;		// Hp positions; values >= H are padding and always sort last
	u32				*div;		// Hp
	u32				*s_zc;		// per-warp exchange: zero count, tail maxima, has-zero / has-one flags
	u32				*s_tz;
	u32				*s_to;
	u32				*s_fl;
	u32				H;
};

// One pBWT column step (pbwt.hh:77-134) on the state in shared memory. bit_of(h) is the column's bit for chromosome copy
// h < H; first_val is the value the reference starts pp / qq with (kk + 1). Returns the number of zeros in the column.
// Every thread of the CTA must call it; it contains two CTA barriers and leaves the state consistent.
template <int KW, typename t_bit>
__device__ __forceinline__ u32 pbwt_step(k3_state const &st, u32 const first_val, t_bit const &bit_of)
{
	unsigned const lane(threadIdx.x & 31), warp(threadIdx.x >> 5);
	u32 const base(warp * KW * 32);
	u32 const below(lanes_below(lane));

	// KW > K3_KW_REGS (more than 8,192 chromosome copies): 3 · KW registers per thread do not fit the 64 a 1,024-thread CTA
	// leaves, so the positions are kept as 16-bit pairs and the zero masks are ballotted again in the second half (bit_of reads
	// the staged column or the class ids, neither changes during the step).
	constexpr bool lean(KW > K3_KW_REGS);
	u32 p[lean ? (KW + 1) / 2 : KW], nd[KW], zm[lean ? 1 : KW];
	u32 open(0);							// bit k: this lane's window reaches back beyond the warp's range
	u32 cz(0), co(0), zc(0);
	bool seen_z(false), seen_o(false);

#pragma unroll
	for (int k = 0; k < KW; ++k)
	{
		u32 const i(base + k * 32 + lane);
		u32 const pk(st.perm[i]);
		u32 const d(st.div[i]);
		bool const bit(pk >= st.H ? true : bit_of(pk));
		u32 const O(__ballot_sync(FULL, bit)), Z(~O);
		u32 v;
		if (0 == O)
		{
			// all zeros: every window has length one, only lane 0 continues the carry
			u32 const wmax(__reduce_max_sync(FULL, d));
			v = (0 == lane) ? max(d, cz) : d;
			if (0 == lane && !seen_z) open |= 1u << k;
			co = max(co, wmax); cz = 0; seen_z = true;
		}
		else if (0 == Z)
		{
			u32 const wmax(__reduce_max_sync(FULL, d));
			v = (0 == lane) ? max(d, co) : d;
			if (0 == lane && !seen_o) open |= 1u << k;
			cz = max(cz, wmax); co = 0; seen_o = true;
		}
		else
		{
			u32 const vz(seg_max_scan(d, Z << 1, lane)), vo(seg_max_scan(d, O << 1, lane));
			u32 const vz31(__shfl_sync(FULL, vz, 31)), vo31(__shfl_sync(FULL, vo, 31));
			if (!bit)
			{
				bool const first(0 == (Z & below));
				v = first ? max(vz, cz) : vz;
				if (first && !seen_z) open |= 1u << k;
			}
			else
			{
				bool const first(0 == (O & below));
				v = first ? max(vo, co) : vo;
				if (first && !seen_o) open |= 1u << k;
			}
			// carries: maximum since the last zero / one of this word (nothing pending if that was lane 31)
			cz = (Z >> 31) ? 0 : vz31;
			co = (O >> 31) ? 0 : vo31;
			seen_z = true; seen_o = true;
		}
		nd[k] = v;
		if constexpr (lean) p[k / 2] = (k & 1) ? (p[k / 2] | (pk << 16)) : pk;
		else { p[k] = pk; zm[k] = Z; }
		zc += __popc(Z);
	}

	if (0 == lane)
	{
		st.s_zc[warp] = zc;
		st.s_tz[warp] = cz;
		st.s_to[warp] = co;
		st.s_fl[warp] = (seen_z ? 1u : 0u) | (seen_o ? 2u : 0u);
	}
	__syncthreads();

	u32 const zc_u(st.s_zc[lane]), tz_u(st.s_tz[lane]), to_u(st.s_to[lane]), fl_u(st.s_fl[lane]);
	u32 const wbelow(lanes_below(warp));
	u32 const zbefore(__reduce_add_sync(FULL, lane < warp ? zc_u : 0u));
	u32 const ztotal(__reduce_add_sync(FULL, zc_u));
	u32 cin_z(first_val), cin_o(first_val);
	{
		u32 const hz(__ballot_sync(FULL, fl_u & 1u) & wbelow), ho(__ballot_sync(FULL, fl_u & 2u) & wbelow);
		if (hz) { unsigned const l(31 - __clz(hz)); cin_z = __reduce_max_sync(FULL, (lane >= l && lane < warp) ? tz_u : 0u); }
		if (ho) { unsigned const l(31 - __clz(ho)); cin_o = __reduce_max_sync(FULL, (lane >= l && lane < warp) ? to_u : 0u); }
	}
	// Every thread read its old entries before the barrier above, so the scatter below may overwrite the state in place.

	u32 zp(zbefore);
#pragma unroll
	for (int k = 0; k < KW; ++k)
	{
		u32 const i(base + k * 32 + lane);
		u32 pk, Z;
		if constexpr (lean)
		{
			pk = (p[k / 2] >> (16 * (k & 1))) & 0xffffu;
			Z = ~__ballot_sync(FULL, pk >= st.H ? true : bit_of(pk));
		}
		else { pk = p[k]; Z = zm[k]; }
		bool const bit(0 == (Z >> lane & 1));
		u32 const zb(zp + __popc(Z & below));
		u32 const np(bit ? ztotal + (i - zb) : zb);
		u32 dv(nd[k]);
		if (open >> k & 1) dv = max(dv, bit ? cin_o : cin_z);
		st.perm[np] = (unsigned short) pk;
		st.div[np] = dv;
		zp += __popc(Z);
	}
	__syncthreads();
	return ztotal;
}



// ---------------------------------------------------------------------------------------------------------
// Parameters, shared-memory layout, column staging
// ---------------------------------------------------------------------------------------------------------
struct k3_params
{
	u64 const		*paths_eh;		// edge-major path words: line e = Wh words
	u32				Wh;
	u32				H;
	u32				Hp;				// positions kept per CTA = 32 · KW · 32 >= H
	u32				n_chunks;
	u32 const		*chunk_cut;		// [n_chunks + 1] cut index at which chunk j starts (chunk_cut[0] = 0)
	u32 const		*cut_edges;		// [n_cuts] edge index of every cut (cut_edges[0] = 0, last = E)
	u32				n_cuts;
	// pass A → B
	unsigned short	*cls;			// [n_chunks][Hp] class id of every chromosome copy on the chunk
	u32				*classdiv;		// [n_chunks][Hp] divergence of the first member of every class against the class before
	u32				*n_classes;		// [n_chunks]
	// pass B → C
	unsigned short	*bperm;			// [n_chunks][Hp] state at the start of every chunk
	u32				*bdiv;			// [n_chunks][Hp]
	// k3_seed_kernel (instead of passes A and B): known states to start from
	unsigned short const	*seed_perm;	// [·][Hp]
	u32 const		*seed_div;
	u32 const		*seed_src;		// [n_chunks] which of them chunk j starts from (chunk 0: none, the identity)
	u32 const		*seed_edge;		// [n_chunks] the edge column that state stands before
	// pass C outputs (indexed by cut - 1)
	u32				*rhs_eq;		// optional [n_cuts - 1][H]
	u32				*distinct;
	u32				*first_class;
	u8				*first_is_ref;
	u32				*joined_count;
	u32				*joined;		// pool of (lhs, rhs, size) triples; chunk j writes from joined_base[j]
	u64 const		*joined_base;	// [n_chunks + 1] triple offsets of the chunks' regions in the pool
};

struct k3_smem
{
	u64				*bar;			// two mbarriers
	u32				*ex;			// 8 × 32 words of per-warp exchange
	unsigned short	*perm;
	u32				*div;
	unsigned short	*rep[2];
	u8				*colbuf[2];
};

__host__ __device__ inline u32 k3_colbuf_bytes(u32 const Wh) { return (K3_COLB * Wh * 8 + 16 + 127) & ~127u; }
__host__ __device__ inline u32 k3_smem_bytes(u32 const Hp, u32 const Wh)
{
	return 128 + 8 * 32 * 4 + Hp * 2 + Hp * 4 + 2 * Hp * 2 + 2 * k3_colbuf_bytes(Wh);
}

__device__ __forceinline__ k3_smem k3_carve(u8 *raw, u32 const Hp, u32 const Wh)
{
	k3_smem s;
	s.bar = reinterpret_cast<u64 *>(raw); raw += 128;
	s.ex = reinterpret_cast<u32 *>(raw); raw += 8 * 32 * 4;
	s.div = reinterpret_cast<u32 *>(raw); raw += Hp * 4;
	s.perm = reinterpret_cast<unsigned short *>(raw); raw += Hp * 2;
	s.rep[0] = reinterpret_cast<unsigned short *>(raw); raw += Hp * 2;
	s.rep[1] = reinterpret_cast<unsigned short *>(raw); raw += Hp * 2;
	s.colbuf[0] = raw; raw += k3_colbuf_bytes(Wh);		// Hp is a multiple of 1024: still 128-byte aligned
	s.colbuf[1] = raw;
	return s;
}

__device__ __forceinline__ k3_state k3_make_state(k3_smem const &s, u32 const H)
{
	return k3_state{s.perm, s.div, s.ex, s.ex + 32, s.ex + 64, s.ex + 96, H};
}

// Columns [eb + batch · K3_COLB, …) of the chunk → colbuf[batch & 1]. Lines are only 8-byte aligned (Wh · 8 bytes each), so
// the copy starts at the 16-byte boundary at or before the first line; k3_col_lead gives the offset of that line in the buffer.
__device__ __forceinline__ u32 k3_col_lead(k3_params const &P, u32 const k0)
{
	return u32((reinterpret_cast<u64>(P.paths_eh) + u64(k0) * P.Wh * 8) & 15);
}

__device__ __forceinline__ void k3_cols_issue(k3_params const &P, k3_smem const &s, u32 const batch, u32 const eb, u32 const ee)
{
	u32 const k0(eb + batch * K3_COLB);
	if (k0 >= ee) return;
	u32 const nb(min(u32(K3_COLB), ee - k0));
	u64 const addr(reinterpret_cast<u64>(P.paths_eh) + u64(k0) * P.Wh * 8);
	u32 const lead(u32(addr & 15));
	u32 const bytes((lead + nb * P.Wh * 8 + 15) & ~15u);
	mbar_expect_tx(&s.bar[batch & 1], bytes);
	bulk_g2s(s.colbuf[batch & 1], reinterpret_cast<void const *>(addr - lead), bytes, &s.bar[batch & 1]);
}

__device__ __forceinline__ void k3_cols_begin(k3_params const &P, k3_smem const &s, u32 const eb, u32 const ee)
{
	if (0 == threadIdx.x)
	{
		mbar_init(&s.bar[0], 1);
		mbar_init(&s.bar[1], 1);
		fence_mbar_init();
		k3_cols_issue(P, s, 0, eb, ee);
		k3_cols_issue(P, s, 1, eb, ee);
	}
	__syncthreads();
}

// The sweep over the chunk's columns [eb, ee). before_column(k) runs before column k is applied (all threads);
// after_column(k, zeros) after it.
template <int KW, typename t_before, typename t_after>
__device__ __forceinline__ void k3_sweep(k3_params const &P, k3_smem const &s, k3_state const &st, u32 const eb, u32 const ee,
	t_before const &before_column, t_after const &after_column)
{
	u32 const words_per_line(P.Wh * 2);
	for (u32 k(eb); k < ee; ++k)
	{
		u32 const rel(k - eb), batch(rel / K3_COLB), in_batch(rel % K3_COLB);
		if (0 == in_batch) mbar_wait(&s.bar[batch & 1], (batch >> 1) & 1);
		before_column(k);
		u32 const *col(reinterpret_cast<u32 const *>(s.colbuf[batch & 1] + k3_col_lead(P, eb + batch * K3_COLB)) + in_batch * words_per_line);
		u32 const zeros(pbwt_step<KW>(st, k + 1, [col](u32 const h) -> bool { return (col[h >> 5] >> (h & 31)) & 1u; }));
		after_column(k, zeros);
		// pbwt_step ended with a CTA barrier: nobody reads this batch's buffer any more once its last column is done
		if ((K3_COLB - 1 == in_batch || k + 1 == ee) && 0 == threadIdx.x) k3_cols_issue(P, s, batch + 2, eb, ee);
	}
}


// ---------------------------------------------------------------------------------------------------------
// Pass A: order and classes of the chromosome copies by the chunk's own columns
// ---------------------------------------------------------------------------------------------------------
template <int KW>
__global__ void __launch_bounds__(K3_THREADS, 1)
k3_local_kernel(k3_params const P)
{
	extern __shared__ __align__(128) u8 k3_raw[];
	k3_smem const s(k3_carve(k3_raw, P.Hp, P.Wh));
	k3_state const st(k3_make_state(s, P.H));
	unsigned const lane(threadIdx.x & 31), warp(threadIdx.x >> 5);
	u32 const j(blockIdx.x);
	u32 const eb(P.cut_edges[P.chunk_cut[j]]), ee(P.cut_edges[P.chunk_cut[j + 1]]);

	for (u32 i(threadIdx.x); i < P.Hp; i += K3_THREADS) { s.perm[i] = (unsigned short) i; s.div[i] = eb; }
	k3_cols_begin(P, s, eb, ee);
	k3_sweep<KW>(P, s, st, eb, ee, [](u32){}, [](u32, u32){});

	// Classes: a position starts a class iff it differs from its predecessor somewhere in the chunk.
	u32 const base(warp * KW * 32);
	u32 hm[KW];
	u32 cnt(0);
#pragma unroll
	for (int k = 0; k < KW; ++k)
	{
		u32 const i(base + k * 32 + lane);
		hm[k] = __ballot_sync(FULL, i < P.H && s.div[i] > eb);
		cnt += __popc(hm[k]);
	}
	if (0 == lane) s.ex[128 + warp] = cnt;
	__syncthreads();
	u32 const cnt_u(s.ex[128 + lane]);
	u32 run(__reduce_add_sync(FULL, lane < warp ? cnt_u : 0u));
	u32 const total(__reduce_add_sync(FULL, cnt_u));
	unsigned short *cls(P.cls + u64(j) * P.Hp);
	u32 *classdiv(P.classdiv + u64(j) * P.Hp);
#pragma unroll
	for (int k = 0; k < KW; ++k)
	{
		u32 const i(base + k * 32 + lane);
		u32 const id(run + __popc(hm[k] & lanes_upto(lane)) - 1);
		if (i < P.H)
		{
			cls[s.perm[i]] = (unsigned short) id;
			if (hm[k] >> lane & 1) classdiv[id] = s.div[i];
		}
		run += __popc(hm[k]);
	}
	if (0 == threadIdx.x) P.n_classes[j] = total;
}



// ---------------------------------------------------------------------------------------------------------
// Pass B: the true state at every chunk boundary
// ---------------------------------------------------------------------------------------------------------
template <int KW>
__global__ void __launch_bounds__(K3_THREADS, 1)
k3_compose_kernel(k3_params const P)
{
	extern __shared__ __align__(128) u8 k3_raw[];
	k3_smem const s(k3_carve(k3_raw, P.Hp, P.Wh));
	k3_state const st(k3_make_state(s, P.H));
	unsigned short *cls(s.rep[0]);

	// pbwt_context's constructor (pbwt.hh:61-74): identity order; its DIVERGENCE_MAX entries never win a maximum in
	// update_divergence (divergence_value::operator<, pbwt.hh:45), i.e. they behave as 0 from the first column on.
	for (u32 i(threadIdx.x); i < P.Hp; i += K3_THREADS)
	{
		s.perm[i] = (unsigned short) i; s.div[i] = 0;
		P.bperm[i] = (unsigned short) i; P.bdiv[i] = 0;
	}
	__syncthreads();

	for (u32 j(0); j + 1 < P.n_chunks; ++j)
	{
		unsigned short const *g_cls(P.cls + u64(j) * P.Hp);
		u32 const *classdiv(P.classdiv + u64(j) * P.Hp);
		for (u32 i(threadIdx.x); i < P.H; i += K3_THREADS) cls[i] = g_cls[i];
		u32 const D(P.n_classes[j]);
		u32 const n_bits(D > 1 ? 32 - __clz(D - 1) : 0);
		__syncthreads();
		for (u32 t(0); t < n_bits; ++t)
			pbwt_step<KW>(st, K3_SENT, [cls, t](u32 const h) -> bool { return (cls[h] >> t) & 1u; });
		// Pairs the chunk separates: the divergence between their (consecutive) classes; position 0: the chunk's end.
		for (u32 i(threadIdx.x); i < P.H; i += K3_THREADS)
			if (K3_SENT == s.div[i] || 0 == i) s.div[i] = classdiv[cls[s.perm[i]]];
		__syncthreads();
		unsigned short *bperm(P.bperm + u64(j + 1) * P.Hp);
		u32 *bdiv(P.bdiv + u64(j + 1) * P.Hp);
		for (u32 i(threadIdx.x); i < P.Hp; i += K3_THREADS) { bperm[i] = s.perm[i]; bdiv[i] = s.div[i]; }
		__syncthreads();
	}
}



// ---------------------------------------------------------------------------------------------------------
// Instead of passes A and B: the state at every chunk boundary from a known state at or before it
// ---------------------------------------------------------------------------------------------------------
template <int KW>
__global__ void __launch_bounds__(K3_THREADS, 1)
k3_seed_kernel(k3_params const P)
{
	extern __shared__ __align__(128) u8 k3_raw[];
	k3_smem const s(k3_carve(k3_raw, P.Hp, P.Wh));
	k3_state const st(k3_make_state(s, P.H));
	u32 const j(blockIdx.x);
	unsigned short *bperm(P.bperm + u64(j) * P.Hp);
	u32 *bdiv(P.bdiv + u64(j) * P.Hp);
	if (0 == j)
	{
		for (u32 i(threadIdx.x); i < P.Hp; i += K3_THREADS) { bperm[i] = (unsigned short) i; bdiv[i] = 0; }		// as pass B starts (pbwt.hh:61-74)
		return;
	}
	u32 const eb(P.seed_edge[j]), ee(P.cut_edges[P.chunk_cut[j]]);
	{
		unsigned short const *sp(P.seed_perm + u64(P.seed_src[j]) * P.Hp);
		u32 const *sd(P.seed_div + u64(P.seed_src[j]) * P.Hp);
		for (u32 i(threadIdx.x); i < P.Hp; i += K3_THREADS) { s.perm[i] = sp[i]; s.div[i] = sd[i]; }
	}
	__syncthreads();
	if (ee > eb)
	{
		k3_cols_begin(P, s, eb, ee);
		k3_sweep<KW>(P, s, st, eb, ee, [](u32){}, [](u32, u32){});
	}
	for (u32 i(threadIdx.x); i < P.Hp; i += K3_THREADS) { bperm[i] = s.perm[i]; bdiv[i] = s.div[i]; }
}



// ---------------------------------------------------------------------------------------------------------
// Host helpers shared by the drivers
// ---------------------------------------------------------------------------------------------------------
struct dev_buf
{
	void *p{};
	~dev_buf() { if (p) cudaFree(p); }
	int alloc(u64 bytes)
	{
		if (p) { cudaFree(p); p = nullptr; }
		V2M_CUDA_TRY(cudaMalloc(&p, std::max<u64>(bytes, 16)));
		return 0;
	}
	template <typename T> T *as() const { return static_cast<T *>(p); }
	void *release() { void *q(p); p = nullptr; return q; }
};


// Pass A (pass == 0) or pass B (pass == 1) for KW 32-position words per warp.
template <int KW>
inline int k3_launch_ab(int const pass, k3_params const &P, u32 const n_ctas, u32 const smem, cudaStream_t stream)
{
	if (0 == pass)
	{
		V2M_CUDA_TRY(cudaFuncSetAttribute(k3_local_kernel<KW>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
		k3_local_kernel<KW><<<n_ctas, K3_THREADS, smem, stream>>>(P);
		V2M_CHECK_LAUNCH("k3_local_kernel");
	}
	else
	{
		V2M_CUDA_TRY(cudaFuncSetAttribute(k3_compose_kernel<KW>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
		k3_compose_kernel<KW><<<1, K3_THREADS, smem, stream>>>(P);
		V2M_CHECK_LAUNCH("k3_compose_kernel");
	}
	return 0;
}

// 32-position words per warp for H chromosome copies: exact up to K3_KW_REGS, the next even count above (fewer instantiations;
// positions beyond H are padding that always sorts last). 0: more than 32 · K3_KW_MAX · 32 copies.
inline u32 k3_words_per_warp(u32 const H)
{
	u32 const words((H + 31) / 32);
	u32 KW(std::max<u32>(1, (words + K3_WARPS - 1) / K3_WARPS));
	if (KW > u32(K3_KW_REGS)) KW = (KW + 1) & ~1u;
	return KW <= u32(K3_KW_MAX) ? KW : 0;
}

// The sweep's shared memory against what a CTA may have on this device.
inline int k3_check_smem(u32 const smem, char const *who)
{
	int dev(0), limit(0);
	V2M_CUDA_TRY(cudaGetDevice(&dev));
	V2M_CUDA_TRY(cudaDeviceGetAttribute(&limit, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
	if (smem > u32(limit))
		return fail(V2M_ERR_ARGUMENT, std::string(who) + ": the pBWT state of this many chromosome copies needs " + std::to_string(smem)
			+ " bytes of shared memory, the device gives a CTA " + std::to_string(limit));
	return 0;
}

// Cut-aligned chunks balanced by columns + cut_weight x cuts: chunk j covers cuts [chunk_cut[j], chunk_cut[j + 1]).
inline std::vector<u32> k3_make_chunks(std::vector<u32> const &cut_edges, u64 const n_seg, u64 P, u64 const cut_weight)
{
	P = std::max<u64>(1, std::min<u64>(P, n_seg));
	std::vector<u32> chunk_cut{0};
	auto const weight([&](u64 c) { return u64(cut_edges[c]) + cut_weight * c; });
	u64 const W(weight(n_seg));
	u64 c(0);
	for (u64 j(1); j < P; ++j)
	{
		u64 const target(W * j / P);
		while (c < n_seg && weight(c) < target) ++c;
		if (c > chunk_cut.back() && c < n_seg) chunk_cut.push_back(u32(c));
	}
	chunk_cut.push_back(u32(n_seg));
	return chunk_cut;
}

} // namespace v2m
