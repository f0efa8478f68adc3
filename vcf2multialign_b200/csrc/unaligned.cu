// K2u — unaligned rows (--unaligned: v2m::output_sequence with should_output_unaligned, sequence_writer.cc:80:
// no gap fill). A row is then the reference with, for every taken ALT edge e (src → tgt), the stretch
// ref[refpos[src], refpos[tgt]) replaced by label(e). Output offset of reference byte p in row h is p + D_h(p), where
// D_h(p) = Σ over taken edges that end at or before p of (|label| − REF span).
//
//   unaligned_prefix_kernel  one warp per row: per 64-edge word the exclusive prefix of those deltas (We + 1 entries, the
//                            last being the row total), and the row length refpos.back() + total.
//   emit_unaligned_kernel    reference-coordinate tiles of 4096 bytes × row blocks; a warp walks the row's taken edges
//                            that touch the tile and copies reference runs / labels to their shifted offsets.
// Secondary mode (not the headline metric): reference runs go out as 128-bit stores realigned with funnel shifts.
#include <cstdint>

#include "common.cuh"
#include "device_graph.hh"

namespace v2m {

__global__ void __launch_bounds__(256)
unaligned_prefix_kernel(u64 const * __restrict__ taken, u32 const n_rows, u64 const We, edge_rec const * __restrict__ edges,
	uint2 const * __restrict__ edge_ref, u32 const ref_end, int * __restrict__ prefix, u64 * __restrict__ lengths)
{
	unsigned const lane(threadIdx.x & 31);
	u32 const row((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
	if (row >= n_rows) return;
	u64 const *in(taken + u64(row) * We);
	int *out(prefix + u64(row) * (We + 1));			// We + 1 entries per row; the last one is the row total
	int carry(0);
	for (u64 wbase(0); wbase < We; wbase += 32)
	{
		u64 const w(wbase + lane);
		u64 word(w < We ? in[w] : 0);
		int delta(0);
		while (word)
		{
			int const b(__ffsll((long long) word) - 1);
			word &= word - 1;
			u64 const e(64 * w + b);
			uint2 const rr(edge_ref[e]);
			delta += int(edges[e].label_len) - int(rr.y - rr.x);
		}
		int incl(delta);
#pragma unroll
		for (int d = 1; d < 32; d <<= 1)
		{
			int const o(__shfl_up_sync(0xffffffffu, incl, d));
			if (lane >= unsigned(d)) incl += o;
		}
		if (w < We) out[w] = carry + incl - delta;
		carry += __shfl_sync(0xffffffffu, incl, 31);
	}
	if (0 == lane)
	{
		out[We] = carry;
		lengths[row] = u64((long long) ref_end + carry);
	}
}


struct unaligned_params
{
	u8 const		*ref;
	edge_rec const	*edges;
	uint2 const		*edge_ref;
	u8 const		*label_chars;
	uint2 const		*tile_edges;	// per reference tile: (first edge that can reach into it, first edge starting after it)
	u64 const		*taken;
	int const		*prefix;
	u64				We;
	u32				row_begin, row_end;
	u32				ref_end;		// reference_positions.back(): where the walk stops
	u64 const		*offsets;		// [row_end - row_begin] byte offsets into out
	u8				*out;
};

// Warp-cooperative copy of n reference bytes src[0, n) to dst, both at arbitrary byte addresses: the 16-byte aligned part
// of the destination goes out as 128-bit stores whose bytes are cut from aligned 32-bit source words with funnel shifts,
// the (at most 15-byte) ends byte-wise. 512 bytes per warp step instead of 32.
__device__ __forceinline__ void copy_shifted(u8 *dst, u8 const *src, u32 const n, unsigned const lane)
{
	u32 const head(min(n, u32((16 - (reinterpret_cast<std::uintptr_t>(dst) & 15)) & 15)));
	if (lane < head) dst[lane] = src[lane];
	u32 const chunks((n - head) / 16);
	u8 *d(dst + head);
	u8 const *s(src + head);
	if (chunks)
	{
		u32 const shift(u32(reinterpret_cast<std::uintptr_t>(s) & 3) * 8);
		u32 const *sw(reinterpret_cast<u32 const *>(reinterpret_cast<std::uintptr_t>(s) & ~std::uintptr_t(3)));
		for (u32 c(lane); c < chunks; c += 32)
		{
			u32 const *w(sw + 4 * c);
			u32 const w0(__ldg(w)), w1(__ldg(w + 1)), w2(__ldg(w + 2)), w3(__ldg(w + 3));
			u32 const w4(shift ? __ldg(w + 4) : 0);				// not read when the source is word-aligned: it may lie past the array
			uint4 v;
			v.x = __funnelshift_r(w0, w1, shift);
			v.y = __funnelshift_r(w1, w2, shift);
			v.z = __funnelshift_r(w2, w3, shift);
			v.w = __funnelshift_r(w3, w4, shift);
			st_v4(d + 16 * c, v);
		}
	}
	u32 const done(head + 16 * chunks);
	if (lane < n - done) dst[done + lane] = src[done + lane];
}

__global__ void __launch_bounds__(K2_THREADS)
emit_unaligned_kernel(unaligned_params const p)
{
	unsigned const lane(threadIdx.x & 31), warp(threadIdx.x >> 5);
	u64 const tile(blockIdx.x);
	u32 const t0(u32(tile) * K2_TCOL), t1(min(u64(t0) + K2_TCOL, u64(p.ref_end)));
	u32 const rb0(p.row_begin + blockIdx.y * K2_ROWS_PER_CTA);
	u32 const rb1(min(rb0 + K2_ROWS_PER_CTA, p.row_end));
	uint2 const te(p.tile_edges[tile]);
	u64 const w_lo(te.x / 64);

	for (u32 r(rb0 + warp); r < rb1; r += K2_THREADS / 32)
	{
		u8 *orow(p.out + p.offsets[r - p.row_begin]);
		u64 const *trow(p.taken + u64(r) * p.We);
		long long D(p.prefix[u64(r) * (p.We + 1) + w_lo]);	// w_lo <= We
		u32 cur(t0);
		if (te.x < te.y)
		{
			u64 const w_hi((te.y - 1) / 64);
			for (u64 wbase(w_lo); wbase <= w_hi; wbase += 32)
			{
				u64 const w(wbase + lane);
				u64 word(w <= w_hi ? trow[w] : 0);
				if (w == w_hi && 0 != te.y % 64) word &= (u64(1) << (te.y % 64)) - 1;
				unsigned m(__ballot_sync(0xffffffffu, 0 != word));
				while (m)
				{
					int const l(__ffs(m) - 1);
					m &= m - 1;
					u64 wl(__shfl_sync(0xffffffffu, word, l));
					while (wl)
					{
						int const b(__ffsll((long long) wl) - 1);
						wl &= wl - 1;
						u64 const e(64 * (wbase + l) + b);
						edge_rec const rec(p.edges[e]);
						uint2 const rr(p.edge_ref[e]);
						// reference run before the edge
						u32 const run_end(min(rr.x, t1));
						if (cur < run_end) copy_shifted(orow + ((long long) cur + D), p.ref + cur, run_end - cur, lane);
						// the label belongs to the tile that holds the edge's source position
						if (rr.x >= t0 && rr.x < t1)
							for (u32 k(lane); k < rec.label_len; k += 32) orow[(long long) rr.x + D + k] = p.label_chars[rec.label_off + k];
						D += (long long) rec.label_len - (long long) (rr.y - rr.x);
						cur = max(cur, rr.y);
					}
				}
			}
		}
		else if (w_lo < p.We)
		{
			// No edge touches the tile, but edges of the same word that end before it still shift it.
			u64 word(trow[w_lo] & ((0 == te.x % 64) ? 0 : ((u64(1) << (te.x % 64)) - 1)));
			while (word)
			{
				int const b(__ffsll((long long) word) - 1);
				word &= word - 1;
				u64 const e(64 * w_lo + b);
				uint2 const rr(p.edge_ref[e]);
				D += (long long) p.edges[e].label_len - (long long) (rr.y - rr.x);
			}
		}
		if (cur < t1) copy_shifted(orow + ((long long) cur + D), p.ref + cur, t1 - cur, lane);
	}
}


int launch_unaligned_prefix(device_graph const &g, u64 const *d_taken, u32 n_rows, int *d_prefix, u64 *d_lengths, cudaStream_t stream)
{
	if (0 == n_rows) return 0;
	unsigned const warps_per_cta(8);
	unaligned_prefix_kernel<<<(n_rows + warps_per_cta - 1) / warps_per_cta, 32 * warps_per_cta, 0, stream>>>(
		d_taken, n_rows, g.We, g.edges, g.edge_ref, g.ref_end, d_prefix, d_lengths);
	V2M_CHECK_LAUNCH("unaligned_prefix_kernel");
	return 0;
}

int launch_emit_unaligned(device_graph const &g, u64 const *d_taken, int const *d_prefix, u32 row_begin, u32 row_end,
	u64 const *d_offsets, void *d_out, cudaStream_t stream)
{
	if (row_begin >= row_end || 0 == g.ref_end) return 0;
	unaligned_params p;
	p.ref = g.ref; p.edges = g.edges; p.edge_ref = g.edge_ref; p.label_chars = g.label_chars; p.tile_edges = g.ref_tile_edges;
	p.taken = d_taken; p.prefix = d_prefix; p.We = g.We; p.row_begin = row_begin; p.row_end = row_end; p.ref_end = g.ref_end;
	p.offsets = d_offsets; p.out = static_cast<u8 *>(d_out);
	u32 const row_blocks((row_end - row_begin + K2_ROWS_PER_CTA - 1) / K2_ROWS_PER_CTA);
	if (row_blocks > 65535) return fail(1, "too many rows in one emit_unaligned launch");
	dim3 const grid(unsigned(g.n_ref_tiles), row_blocks);
	emit_unaligned_kernel<<<grid, K2_THREADS, 0, stream>>>(p);
	V2M_CHECK_LAUNCH("emit_unaligned_kernel");
	return 0;
}

} // namespace v2m
