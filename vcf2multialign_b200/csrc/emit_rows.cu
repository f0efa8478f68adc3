// K2 — row emission: the device side of v2m::output_sequence (libvcf2multialign/sequence_writer.cc:22-85) as
// driven by haplotype_output::output_a2m (haplotype_output.cc:38-82) and
// founder_sequence_greedy_output::output_a2m (founder_sequence_greedy_output.cc:515-550).
//
// Data-parallel restatement (SURVEY.md §7 C1–C3, C6):
//   * in aligned mode the step leaving node n starts at column aligned_positions[n]; every row has length
//     A = aligned_positions.back(), so a row is the REF template (what the walk writes when it follows REF edges
//     only, sequence_writer.cc:70-81) overwritten on [aln[src(e)], aln[tgt(e)]) by label(e) + '-' padding for every
//     ALT edge e the walk *takes*;
//   * the walk takes set edge e of a chromosome copy iff src(e) >= target of the previously taken edge
//     (first set edge of a visited node, sequence_writer.cc:51-67; targets are always > sources).
//
// Kernels:
//   build_template_kernel   REF-aligned template, A bytes, once per graph.
//   compose_founder_kernel  path bits of founder rows: per segment the bits of the assigned chromosome copy
//                           (founder_sequence_writing_delegate::handle_node, founder_…cc:106-114).
//   resolve_rows_kernel     K2a: path bits → taken bits, one warp per row, 2048 edges per step.
//   emit_rows_kernel        K2b: rows × columns tile writer; template tile held in registers and streamed to every
//                           row of the block with 128-bit stores, then the row's few taken edges are patched in.
//                           Algorithmic bytes: rows × columns written. Roofline: HBM write.
#include "common.cuh"
#include "device_graph.hh"

namespace v2m {

// ---------------------------------------------------------------------------------------------------------
// REF-aligned template
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
build_template_kernel(u8 const * __restrict__ ref, u32 const * __restrict__ refpos, u32 const * __restrict__ alnpos,
	u64 const node_count, u64 const A, u64 const A_alloc, u8 * __restrict__ tmpl)
{
	u64 const c0((u64(blockIdx.x) * blockDim.x + threadIdx.x) * 16);
	if (c0 >= A_alloc) return;
	__align__(16) u8 bytes[16];
	if (c0 >= A || node_count < 2)
	{
#pragma unroll
		for (int k = 0; k < 16; ++k) bytes[k] = 0;
	}
	else
	{
		// n = last node (excluding the sink) with alnpos[n] <= c0.
		u64 lo(0), hi(node_count - 1);	// invariant: alnpos[lo] <= c0 < alnpos[hi]  (alnpos[0] = 0, alnpos[N-1] = A)
		while (hi - lo > 1)
		{
			u64 const mid((lo + hi) / 2);
			if (alnpos[mid] <= c0) lo = mid; else hi = mid;
		}
		u64 n(lo);
		u32 a_n(alnpos[n]), a_next(alnpos[n + 1]), r_n(refpos[n]), r_len(refpos[n + 1] - r_n);
#pragma unroll
		for (int k = 0; k < 16; ++k)
		{
			u64 const col(c0 + k);
			u8 b(0);
			if (col < A)
			{
				while (col >= a_next)
				{
					++n;
					a_n = a_next; a_next = alnpos[n + 1];
					r_n = refpos[n]; r_len = refpos[n + 1] - r_n;
				}
				u32 const off(u32(col) - a_n);
				b = off < r_len ? ref[r_n + off] : u8('-');
			}
			bytes[k] = b;
		}
	}
	*reinterpret_cast<uint4 *>(tmpl + c0) = *reinterpret_cast<uint4 const *>(bytes);
}


// ---------------------------------------------------------------------------------------------------------
// Founder path bits
// ---------------------------------------------------------------------------------------------------------
// out line f, word w = OR over segments s intersecting edges [64w, 64w + 64) of
// (line assigned(s, f) of the copy-major path matrix, word w) & mask(segment edge range).
__global__ void __launch_bounds__(256)
compose_founder_kernel(u64 const * __restrict__ paths_he, u64 const We,
	u64 const * __restrict__ cut_edges /* [n_segments + 1] */, u64 const n_segments,
	u32 const * __restrict__ assigned /* column-major: [f * n_segments + s] */, u32 const n_founders,
	u64 * __restrict__ out)
{
	u64 const idx(u64(blockIdx.x) * blockDim.x + threadIdx.x);
	if (idx >= u64(n_founders) * We) return;
	u32 const f(idx / We);
	u64 const w(idx % We);
	u64 const e0(64 * w), e1(e0 + 64);
	// First segment whose end is beyond e0.
	u64 lo(0), hi(n_segments);
	while (lo < hi)
	{
		u64 const mid((lo + hi) / 2);
		if (cut_edges[mid + 1] > e0) hi = mid; else lo = mid + 1;
	}
	u64 acc(0);
	for (u64 s(lo); s < n_segments && cut_edges[s] < e1; ++s)
	{
		u64 const b(max(cut_edges[s], e0) - e0), e(min(cut_edges[s + 1], e1) - e0);
		if (b >= e) continue;
		u32 const copy(assigned[u64(f) * n_segments + s]);
		if (PLOIDY_MAX == copy) continue;
		u64 const mask((e - b >= 64 ? ~u64(0) : ((u64(1) << (e - b)) - 1)) << b);
		acc |= paths_he[u64(copy) * We + w] & mask;
	}
	out[idx] = acc;
}


// ---------------------------------------------------------------------------------------------------------
// K2a resolve_rows: path bits → taken bits
// ---------------------------------------------------------------------------------------------------------
struct word_result { u64 taken; u32 cur_out; u32 s_first; };

__device__ __forceinline__ word_result resolve_word(u64 word, u64 const e_base, u32 const * __restrict__ edge_src,
	u32 const * __restrict__ edge_tgt, u32 cur)
{
	word_result r{0, 0, 0xffffffffu};
	bool first(true);
	while (word)
	{
		int const b(__ffsll((long long) word) - 1);
		word &= word - 1;
		u32 const s(edge_src[e_base + b]);
		if (first) { r.s_first = s; first = false; }
		if (s >= cur)
		{
			r.taken |= u64(1) << b;
			cur = edge_tgt[e_base + b];
		}
	}
	r.cur_out = cur;
	return r;
}

// One warp per output row. lines = base of We-word lines; row i reads line line_idx[i] (PLOIDY_MAX → no ALT edges).
__global__ void __launch_bounds__(256)
resolve_rows_kernel(u64 const * __restrict__ lines, u32 const * __restrict__ line_idx, u32 const n_rows, u64 const We,
	u32 const * __restrict__ edge_src, u32 const * __restrict__ edge_tgt, u64 * __restrict__ taken)
{
	unsigned const lane(threadIdx.x & 31);
	u32 const row((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
	if (row >= n_rows) return;
	u32 const li(line_idx ? line_idx[row] : row);
	u64 *out(taken + u64(row) * We);
	if (PLOIDY_MAX == li)
	{
		for (u64 w(lane); w < We; w += 32) out[w] = 0;
		return;
	}
	u64 const *in(lines + u64(li) * We);
	u32 carry(0);	// target node of the last taken edge so far
	for (u64 wbase(0); wbase < We; wbase += 32)
	{
		u64 const w(wbase + lane);
		u64 const word(w < We ? in[w] : 0);
		if (0 == __ballot_sync(0xffffffffu, 0 != word))
		{
			if (w < We) out[w] = 0;
			continue;
		}
		// Optimistic pass: every lane resolves its word as if no earlier taken edge reached into it.
		word_result res(resolve_word(word, 64 * w, edge_src, edge_tgt, 0));
		u32 incl(word ? res.cur_out : 0);
#pragma unroll
		for (int d = 1; d < 32; d <<= 1)
		{
			u32 const o(__shfl_up_sync(0xffffffffu, incl, d));
			if (lane >= unsigned(d)) incl = max(incl, o);
		}
		u32 excl(__shfl_up_sync(0xffffffffu, incl, 1));
		if (0 == lane) excl = 0;
		u32 const incoming(max(carry, excl));
		bool const ok(0 == word || incoming <= res.s_first);
		if (__all_sync(0xffffffffu, ok))
			carry = max(carry, __shfl_sync(0xffffffffu, incl, 31));
		else
		{
			// Some taken edge overlaps later set edges: replay the non-empty words in order with the true state.
			unsigned m(__ballot_sync(0xffffffffu, 0 != word));
			u32 c(carry);
			while (m)
			{
				int const l(__ffs(m) - 1);
				m &= m - 1;
				if (int(lane) == l) res = resolve_word(word, 64 * w, edge_src, edge_tgt, c);
				c = __shfl_sync(0xffffffffu, res.cur_out, l);
			}
			carry = c;
		}
		if (w < We) out[w] = res.taken;
	}
}


// ---------------------------------------------------------------------------------------------------------
// K2b emit_rows
// ---------------------------------------------------------------------------------------------------------
struct emit_params
{
	u8 const		*tmpl;
	edge_rec const	*edges;
	u8 const		*label_chars;
	uint2 const		*tile_edges;	// per absolute tile of K2_TCOL columns: (first edge that can reach into it, first edge starting after it)
	u64 const		*taken;			// plan rows × We
	u64				We;
	u32				row_begin, row_end;
	u64				col_begin, col_end;	// col_begin % 16 == 0
	u64				first_tile;			// col_begin / K2_TCOL
	u8				*out;
	u64				pitch;
};

__device__ __forceinline__ u8 patch_byte(edge_rec const &rec, u8 const * __restrict__ label_chars, u32 const col)
{
	u32 const k(col - rec.a0);
	return k < rec.label_len ? label_chars[rec.label_off + k] : u8('-');
}

template <bool t_streaming>
__global__ void __launch_bounds__(K2_THREADS)
emit_rows_kernel(emit_params const p)
{
	unsigned const lane(threadIdx.x & 31), warp(threadIdx.x >> 5);
	u64 const tile(p.first_tile + blockIdx.x);
	u64 const t0(tile * K2_TCOL);
	u64 const c0(max(t0, p.col_begin)), c1(min(t0 + K2_TCOL, p.col_end));
	u32 const rb0(p.row_begin + blockIdx.y * K2_ROWS_PER_CTA);
	u32 const rb1(min(rb0 + K2_ROWS_PER_CTA, p.row_end));

	// Template tile → registers (the same bytes go to every row of the block).
	uint4 v[K2_TCOL / 512];
#pragma unroll
	for (int j = 0; j < K2_TCOL / 512; ++j)
		v[j] = *reinterpret_cast<uint4 const *>(p.tmpl + t0 + j * 512 + lane * 16);

	uint2 const te(p.tile_edges[tile]);
	bool const has_edges(te.x < te.y);
	u64 const w_lo(te.x / 64), w_hi(has_edges ? (te.y - 1) / 64 : 0);

	for (u32 r(rb0 + warp); r < rb1; r += K2_THREADS / 32)
	{
		u8 *orow(p.out + u64(r - p.row_begin) * p.pitch);	// column c lives at orow[c - col_begin]

#pragma unroll
		for (int j = 0; j < K2_TCOL / 512; ++j)
		{
			u64 const col(t0 + j * 512 + lane * 16);
			if (col >= c0 && col + 16 <= c1)
			{
				if constexpr (t_streaming) st_cs_v4(orow + (col - p.col_begin), v[j]);
				else st_v4(orow + (col - p.col_begin), v[j]);
			}
			else if (col + 16 > c0 && col < c1)
			{
				u32 const wds[4] = {v[j].x, v[j].y, v[j].z, v[j].w};
#pragma unroll
				for (int k = 0; k < 16; ++k)
					if (col + k >= c0 && col + k < c1)
						orow[col + k - p.col_begin] = u8(wds[k >> 2] >> (8 * (k & 3)));
			}
		}

		if (!has_edges) continue;
		__syncwarp();	// template bytes of all lanes are ordered before the patches below

		u64 const *trow(p.taken + u64(r) * p.We);
		for (u64 wbase(w_lo); wbase <= w_hi; wbase += 32)
		{
			u64 const w(wbase + lane);
			u64 word(w <= w_hi ? trow[w] : 0);
			if (w == w_lo) word &= ~u64(0) << (te.x % 64);
			if (w == w_hi && 0 != te.y % 64) word &= (u64(1) << (te.y % 64)) - 1;

			bool pending(false);
			edge_rec prec{};
			u32 plo(0), phi(0);
			auto scan([&]{
				while (word)
				{
					int const b(__ffsll((long long) word) - 1);
					word &= word - 1;
					edge_rec const rec(*reinterpret_cast<edge_rec const *>(&reinterpret_cast<uint4 const *>(p.edges)[64 * w + b]));
					u32 const lo(max(u64(rec.a0), c0)), hi(min(u64(rec.a1), c1));
					if (lo >= hi) continue;
					if (hi - lo <= 32)
					{
						for (u32 col(lo); col < hi; ++col)
							orow[col - p.col_begin] = patch_byte(rec, p.label_chars, col);
					}
					else
					{
						pending = true; prec = rec; plo = lo; phi = hi;
						return;
					}
				}
			});
			scan();
			// Long spans (large deletions / insertions) are written by the whole warp.
			unsigned m;
			while ((m = __ballot_sync(0xffffffffu, pending)))
			{
				int const l(__ffs(m) - 1);
				edge_rec rec;
				rec.a0 = __shfl_sync(0xffffffffu, prec.a0, l);
				rec.a1 = __shfl_sync(0xffffffffu, prec.a1, l);
				rec.label_off = __shfl_sync(0xffffffffu, prec.label_off, l);
				rec.label_len = __shfl_sync(0xffffffffu, prec.label_len, l);
				u32 const lo(__shfl_sync(0xffffffffu, plo, l)), hi(__shfl_sync(0xffffffffu, phi, l));
				for (u32 col(lo + lane); col < hi; col += 32)
					orow[col - p.col_begin] = patch_byte(rec, p.label_chars, col);
				if (int(lane) == l)
				{
					pending = false;
					scan();
				}
			}
		}
	}
}


// ---------------------------------------------------------------------------------------------------------
// Launchers
// ---------------------------------------------------------------------------------------------------------
int launch_build_template(device_graph const &g, cudaStream_t stream)
{
	u64 const threads(g.A_alloc / 16);
	if (0 == threads) return 0;
	build_template_kernel<<<unsigned((threads + 255) / 256), 256, 0, stream>>>(g.ref, g.refpos, g.alnpos, g.N, g.A, g.A_alloc, g.tmpl);
	V2M_CHECK_LAUNCH("build_template_kernel");
	return 0;
}

int launch_compose_founders(device_graph const &g, u64 const *d_cut_edges, u64 n_segments, u32 const *d_assigned, u32 n_founders, u64 *d_out, cudaStream_t stream)
{
	u64 const total(u64(n_founders) * g.We);
	if (0 == total) return 0;
	compose_founder_kernel<<<unsigned((total + 255) / 256), 256, 0, stream>>>(g.paths_he, g.We, d_cut_edges, n_segments, d_assigned, n_founders, d_out);
	V2M_CHECK_LAUNCH("compose_founder_kernel");
	return 0;
}

int launch_resolve_rows(device_graph const &g, u64 const *d_lines, u32 const *d_line_idx, u32 n_rows, u64 *d_taken, cudaStream_t stream)
{
	if (0 == n_rows || 0 == g.We) return 0;
	unsigned const warps_per_cta(8);
	resolve_rows_kernel<<<(n_rows + warps_per_cta - 1) / warps_per_cta, 32 * warps_per_cta, 0, stream>>>(
		d_lines, d_line_idx, n_rows, g.We, g.edge_src, g.edge_tgt, d_taken);
	V2M_CHECK_LAUNCH("resolve_rows_kernel");
	return 0;
}

int launch_emit_rows(device_graph const &g, u64 const *d_taken, u32 row_begin, u32 row_end, u64 col_begin, u64 col_end,
	void *d_out, u64 pitch, bool streaming, cudaStream_t stream)
{
	if (row_begin >= row_end || col_begin >= col_end) return 0;
	emit_params p;
	p.tmpl = g.tmpl; p.edges = g.edges; p.label_chars = g.label_chars; p.tile_edges = g.tile_edges;
	p.taken = d_taken; p.We = g.We;
	p.row_begin = row_begin; p.row_end = row_end; p.col_begin = col_begin; p.col_end = col_end;
	p.first_tile = col_begin / K2_TCOL;
	p.out = static_cast<u8 *>(d_out); p.pitch = pitch;
	u64 const last_tile((col_end - 1) / K2_TCOL);
	u32 const n_rows(row_end - row_begin);
	// grid.y is limited to 65535 row blocks: split if ever needed.
	u32 const row_blocks((n_rows + K2_ROWS_PER_CTA - 1) / K2_ROWS_PER_CTA);
	if (row_blocks > 65535) return fail(1, "too many rows in one emit_rows launch");
	dim3 const grid(unsigned(last_tile - p.first_tile + 1), row_blocks);
	if (streaming) emit_rows_kernel<true><<<grid, K2_THREADS, 0, stream>>>(p);
	else emit_rows_kernel<false><<<grid, K2_THREADS, 0, stream>>>(p);
	V2M_CHECK_LAUNCH("emit_rows_kernel");
	return 0;
}

} // namespace v2m
