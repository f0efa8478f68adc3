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
//   emit_rows_kernel        K2b: rows × columns tile writer; template tile + edge records staged in shared memory by
//                           bulk async copies, rows assembled in shared memory and written by the TMA engine.
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
// One CTA = one tile of K2_TCOL aligned columns × K2_ROWS_PER_CTA rows.
//   stage   one thread issues three bulk async copies (TMA engine) global → shared: the tile of the REF template, the
//           records of the ALT edges that can intersect the tile and those edges' label bytes (contiguous in the
//           flattened label store), so that patching never leaves the SM;
//   rows    each warp owns K2_ROWS_PER_CTA / 8 rows and two K2_TCOL-byte row buffers in shared memory. Per row: the
//           template tile (kept in registers) is written to a buffer with 128-bit STS, the row's taken edges that
//           touch the tile are patched in (label bytes + '-'), and one lane hands the finished buffer to the TMA
//           engine (cp.async.bulk shared → global): the LSU issues no global stores at all, and HBM only ever sees
//           whole 128-byte lines. The taken-edge words of all rows of a warp are fetched up front (independent loads
//           in flight), so no row waits on a dependent global load chain.
// Algorithmic bytes: rows × columns written (+ < 1 % reads). Roofline: HBM write.
constexpr int K2_EDGE_CAP = 256;								// ALT edge records staged per tile (c2 averages ~100)
constexpr int K2_LABEL_CAP = 2048;								// label bytes staged per tile (incl. ≤ 15 bytes of alignment slack)
constexpr int K2_WARPS = K2_THREADS / 32;
constexpr int K2_CHUNKS = K2_TCOL / 512;						// 16-byte chunks of the tile per lane
constexpr u32 K2_SHORT_PATCH = 16;								// longer spans are written by the whole warp

struct emit_smem
{
	alignas(128) u8			tmpl[K2_TCOL];
	alignas(128) edge_rec	edges[K2_EDGE_CAP];
	alignas(128) u8			labels[K2_LABEL_CAP];
	alignas(128) u8			rows[K2_WARPS][2][K2_TCOL];
	alignas(8) u64			bar;
};

struct emit_params
{
	u8 const		*tmpl;
	edge_rec const	*edges;
	u8 const		*label_chars;
	uint2 const		*tile_edges;	// per absolute tile of K2_TCOL columns: (first edge that can reach into it, first edge starting after it)
	uint2 const		*tile_labels;	// per tile: byte range of those edges' labels in label_chars
	u64 const		*taken;			// plan rows × We
	u64				We;
	u32				row_begin, row_end;
	u64				col_begin, col_end;	// col_begin % 16 == 0
	u64				first_tile;			// col_begin / K2_TCOL
	u8				*out;
	u64				pitch;
};

// `labels` is the label store seen from label offset 0: the global array, or the tile's staged window shifted back.
__device__ __forceinline__ u8 patch_byte(edge_rec const &rec, u8 const *labels, u32 const col)
{
	u32 const k(col - rec.a0);
	return k < rec.label_len ? labels[rec.label_off + k] : u8('-');
}

__global__ void __launch_bounds__(K2_THREADS, 3)
emit_rows_kernel(emit_params const p)
{
	extern __shared__ __align__(128) u8 smem_raw[];
	emit_smem &sm(*reinterpret_cast<emit_smem *>(smem_raw));

	unsigned const lane(threadIdx.x & 31), warp(threadIdx.x >> 5);
	u64 const tile(p.first_tile + blockIdx.x);
	u64 const t0(tile * K2_TCOL);
	u64 const c0(max(t0, p.col_begin)), c1(min(t0 + K2_TCOL, p.col_end));
	u32 const rb0(p.row_begin + blockIdx.y * K2_ROWS_PER_CTA);
	u32 const rb1(min(rb0 + K2_ROWS_PER_CTA, p.row_end));

	uint2 const te(p.tile_edges[tile]);
	u32 const n_e(te.y - te.x);
	bool const has_edges(0 != n_e);
	uint2 const tl(p.tile_labels[tile]);
	u32 const lab_base(tl.x & ~15u);											// bulk copies start at 16-byte boundaries
	u32 const lab_bytes(tl.y > tl.x ? ((tl.y - lab_base + 15u) & ~15u) : 0);
	bool const staged(n_e <= K2_EDGE_CAP && lab_bytes <= K2_LABEL_CAP);
	u64 const w_lo(te.x / 64), w_hi(has_edges ? (te.y - 1) / 64 : w_lo);
	u32 const n_words(has_edges ? u32(w_hi - w_lo + 1) : 0);

	// ---- stage the template tile and the tile's edge records ---------------------------------------------------------
	if (0 == threadIdx.x)
	{
		mbar_init(&sm.bar, 1);
		fence_mbar_init();
		u32 const edge_bytes(staged ? n_e * u32(sizeof(edge_rec)) : 0);
		u32 const label_bytes(staged ? lab_bytes : 0);
		mbar_expect_tx(&sm.bar, K2_TCOL + edge_bytes + label_bytes);
		bulk_g2s(sm.tmpl, p.tmpl + t0, K2_TCOL, &sm.bar);
		if (edge_bytes) bulk_g2s(sm.edges, p.edges + te.x, edge_bytes, &sm.bar);
		if (label_bytes) bulk_g2s(sm.labels, p.label_chars + lab_base, label_bytes, &sm.bar);
	}

	// Meanwhile: taken-edge words of this warp's first rows (lane ↔ word; the loop below stays two rows ahead).
	auto const load_word([&](u32 const r, u32 const i) -> u64 {
		if (r >= rb1 || i >= n_words) return 0;
		u64 const w(w_lo + i);
		u64 word(p.taken[u64(r) * p.We + w]);
		if (w == w_lo) word &= ~u64(0) << (te.x % 64);
		if (w == w_hi && 0 != te.y % 64) word &= (u64(1) << (te.y % 64)) - 1;
		return word;
	});
	u64 w_cur(load_word(rb0 + warp, lane));
	u64 w_nxt(load_word(rb0 + warp + K2_WARPS, lane));

	__syncthreads();					// the barrier's initialisation is visible to every waiter
	mbar_wait(&sm.bar, 0);

	uint4 v[K2_CHUNKS];
#pragma unroll
	for (int j = 0; j < K2_CHUNKS; ++j)
		v[j] = *reinterpret_cast<uint4 const *>(sm.tmpl + j * 512 + lane * 16);

	// What one row hands to the TMA engine: the 16-byte aligned part of [c0, c1); a ragged end goes out byte-wise.
	u32 const off0(u32(c0 - t0));
	u32 const n_bytes(u32(c1 - c0));
	u32 const bulk_bytes(n_bytes & ~15u);
	u32 const t0_32 = u32(t0), c0_32 = u32(c0), c1_32 = u32(c1);

	// Edge records and labels as seen from edge 0 / label offset 0 (generic pointers into shared memory when staged).
	edge_rec const *erecs(staged ? sm.edges - te.x : p.edges);
	u8 const *labels(staged ? sm.labels - lab_base : p.label_chars);

	u32 k(0);
#pragma unroll 1
	for (u32 r(rb0 + warp); r < rb1; r += K2_WARPS, ++k)
	{
		u64 const w_nn(load_word(r + 2 * K2_WARPS, lane));
		u8 *buf(sm.rows[warp][k & 1]);
		if (k >= 2)
		{
			if (0 == lane) bulk_wait_read<1>();		// the copy that last read this buffer (two rows ago) is done with it
			__syncwarp();
		}
#pragma unroll
		for (int j = 0; j < K2_CHUNKS; ++j)
			*reinterpret_cast<uint4 *>(buf + j * 512 + lane * 16) = v[j];

		if (has_edges)
		{
			__syncwarp();
#pragma unroll 1
			for (u32 chunk(0); chunk < n_words; chunk += 32)
			{
				u64 const word(0 == chunk ? w_cur : load_word(r, chunk + lane));		// lane ↔ taken word chunk + lane
				unsigned nz(__ballot_sync(0xffffffffu, 0 != word));
#pragma unroll 1
				while (nz)
				{
					// One non-empty word at a time, its 64 edges spread over the lanes: the row's patches in this word are
					// applied side by side (taken edges of one row never overlap, so their order does not matter).
					int const wl(__ffs(nz) - 1);
					nz &= nz - 1;
					u32 const bits_lo(__shfl_sync(0xffffffffu, u32(word), wl)), bits_hi(__shfl_sync(0xffffffffu, u32(word >> 32), wl));
					u64 const e_base(64 * (w_lo + chunk + wl));
#pragma unroll
					for (int half = 0; half < 2; ++half)
					{
						u32 const bits(half ? bits_hi : bits_lo);
						if (0 == bits) continue;
						bool pending(false);
						edge_rec prec{};
						u32 plo(0), phi(0);
						if (bits >> lane & 1u)
						{
							edge_rec const rec(*reinterpret_cast<edge_rec const *>(&reinterpret_cast<uint4 const *>(erecs)[e_base + 32 * half + lane]));
							u32 const lo(max(rec.a0, c0_32)), hi(min(rec.a1, c1_32));
							if (lo < hi)
							{
								if (hi - lo <= K2_SHORT_PATCH)
									for (u32 col(lo); col < hi; ++col) buf[col - t0_32] = patch_byte(rec, labels, col);
								else { pending = true; prec = rec; plo = lo; phi = hi; }
							}
						}
						// Long spans (large deletions / insertions) are written by the whole warp, one at a time.
						unsigned m(__ballot_sync(0xffffffffu, pending));
						while (m)
						{
							int const l(__ffs(m) - 1);
							m &= m - 1;
							edge_rec rec;
							rec.a0 = __shfl_sync(0xffffffffu, prec.a0, l);
							rec.a1 = __shfl_sync(0xffffffffu, prec.a1, l);
							rec.label_off = __shfl_sync(0xffffffffu, prec.label_off, l);
							rec.label_len = __shfl_sync(0xffffffffu, prec.label_len, l);
							u32 const lo(__shfl_sync(0xffffffffu, plo, l)), hi(__shfl_sync(0xffffffffu, phi, l));
							for (u32 col(lo + lane); col < hi; col += 32) buf[col - t0_32] = patch_byte(rec, labels, col);
						}
					}
				}
			}
		}

		fence_proxy_async_smem();
		__syncwarp();
		u8 *orow(p.out + u64(r - p.row_begin) * p.pitch + (c0 - p.col_begin));
		if (0 == lane && bulk_bytes)
		{
			bulk_s2g(orow, buf + off0, bulk_bytes);
			bulk_commit();
		}
		if (lane < n_bytes - bulk_bytes)
			orow[bulk_bytes + lane] = buf[off0 + bulk_bytes + lane];
		w_cur = w_nxt; w_nxt = w_nn;
	}
	if (0 == lane) bulk_wait_read<0>();			// shared memory stays valid until the engine has read it
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
	void *d_out, u64 pitch, cudaStream_t stream)
{
	if (row_begin >= row_end || col_begin >= col_end) return 0;
	emit_params p;
	p.tmpl = g.tmpl; p.edges = g.edges; p.label_chars = g.label_chars; p.tile_edges = g.tile_edges; p.tile_labels = g.tile_labels;
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
	V2M_CUDA_TRY(cudaFuncSetAttribute(emit_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(emit_smem))));	// per device
	emit_rows_kernel<<<grid, K2_THREADS, sizeof(emit_smem), stream>>>(p);
	V2M_CHECK_LAUNCH("emit_rows_kernel");
	return 0;
}

} // namespace v2m
