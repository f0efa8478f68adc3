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
//   resolve_rows_kernel     K2a: path bits → taken bits, one warp per row, 16,384 edges per step; per-edge loads only for set
//                           bits of edges that an earlier edge can reach over.
//   emit_rows_kernel        K2b: rows × columns tile writer, persistent CTAs; template tile + edge records staged in shared
//                           memory by bulk async copies one tile ahead, rows assembled in shared memory and written by
//                           the TMA engine.
//                           Algorithmic bytes: rows × columns written. Roofline: HBM write.
#include <algorithm>
#include <cstdlib>
#include <map>
#include <mutex>
#include <thread>
#include <tuple>
#include <type_traits>

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
// assigned_samples entries must be chromosome copies of the graph or PLOIDY_MAX (founder_sequence_greedy_output.cc:170-172).
__global__ void __launch_bounds__(256)
check_assigned_kernel(u32 const * __restrict__ assigned, u64 const n, u32 const H, u32 * __restrict__ bad)
{
	u64 const stride(u64(gridDim.x) * blockDim.x);
	bool any(false);
	for (u64 i(u64(blockIdx.x) * blockDim.x + threadIdx.x); i < n; i += stride)
	{
		u32 const v(assigned[i]);
		any |= PLOIDY_MAX != v && v >= H;
	}
	if (any) *bad = 1;
}

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
// The walk takes set edge e iff src(e) >= target of the last taken edge (sequence_writer.cc:51-67; targets of taken edges
// grow along the walk, so "the last taken edge" is the taken edge with the largest index). Most edges can be decided without
// looking at the walk at all. Call a node *clean* when no edge of an earlier node reaches beyond it. graph_upload classifies
// every edge (two bits per edge, E_pad / 4 bytes, L2-resident):
//   simple   first ALT edge of a clean node: taken whenever its bit is set;
//   sibling  2nd … 4th ALT edge of a clean node, in the same 64-edge word as the first: taken iff set and no earlier edge of
//            the node is set (the first set edge of a visited node wins, :51-66) — three shift-and-mask terms per word;
//   complex  everything else (edges of nodes under an earlier edge's span — overlapping records —, deeper siblings, sibling
//            runs that cross a word): decided by the walk's rule, in order, with src(e) and the target of the last taken edge
//            loaded only then.
// The common case is therefore `taken = path word & ~blocked siblings` with no per-edge load. The first version loaded edge_src
// / edge_tgt for every set bit (~1.9 G scattered 4-byte loads on c4: L2 sector throughput, 520 GB/s of algorithmic traffic);
// with only first ALTs exempted the ~5 k second-ALT bits per row still cost a dependent-load chain each (750 GB/s).
// One warp per output row, K2A_UNROLL × 32 words (16,384 edges at 8) per step with all loads of a step issued together.
// Algorithmic bytes: rows × E_pad / 8 read + the same written. Roofline: HBM.
#ifndef V2M_K2A_UNROLL
#define V2M_K2A_UNROLL 8		// c4: 4 → 2.21 ms, 8 → 1.75 ms, 16 → 1.98 ms (only 34 warps per SM exist: each must keep more bytes in flight)
#endif
constexpr int K2A_UNROLL = V2M_K2A_UNROLL;

__global__ void __launch_bounds__(256)
resolve_rows_kernel(u64 const * __restrict__ lines, u32 const * __restrict__ line_idx, u32 const n_rows, u64 const We,
	u64 const * __restrict__ edge_simple, u64 const * __restrict__ edge_sibling,
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
	int hist(-1);		// this lane's part of "index of the last taken edge so far" (the warp-wide maximum is the edge itself)
	for (u64 wbase(0); wbase < We; wbase += 32 * K2A_UNROLL)
	{
		u64 word[K2A_UNROLL], simple[K2A_UNROLL], sib[K2A_UNROLL];
#pragma unroll
		for (int k = 0; k < K2A_UNROLL; ++k)
		{
			u64 const w(wbase + 32 * k + lane);
			word[k] = w < We ? __ldcs(in + w) : 0;					// read once
			simple[k] = w < We ? __ldg(edge_simple + w) : ~u64(0);	// shared by all rows
			sib[k] = w < We ? __ldg(edge_sibling + w) : 0;
		}
#pragma unroll
		for (int k = 0; k < K2A_UNROLL; ++k)
		{
			u64 const w(wbase + 32 * k + lane);
			u64 const set(word[k]);
			// a sibling is blocked by a set edge of its node one, two or three places below it
			u64 const m1(sib[k]), m2(m1 & (m1 << 1)), m3(m2 & (m1 << 2));
			u64 const fast(set & ~(((set << 1) & m1) | ((set << 2) & m2) | ((set << 3) & m3)));
			u64 tk(fast);
			unsigned m(__ballot_sync(0xffffffffu, 0 != (set & ~simple[k])));
			int last(tk ? int(64 * u32(w)) + 63 - __clzll((long long) tk) : -1);	// last taken edge of this word unless a complex edge says otherwise
			while (m)
			{
				// Lanes whose word has a set complex edge, in edge order: each needs the last taken edge before its word.
				int const l(__ffs(m) - 1);
				m &= m - 1;
				int const prev(__reduce_max_sync(0xffffffffu, max(hist, int(lane) < l ? last : -1)));
				if (int(lane) == l)
				{
					int cur_e(prev);
					bool known(false);
					u32 cur_t(0);
					u64 bits(set), t(0);
					last = -1;
					while (bits)
					{
						int const b(__ffsll((long long) bits) - 1);
						bits &= bits - 1;
						u32 const e(64 * u32(w) + u32(b));
						bool take;
						if (simple[k] >> b & 1) take = 0 != (fast >> b & 1);
						else
						{
							if (!known) { cur_t = cur_e >= 0 ? edge_tgt[cur_e] : 0; known = true; }
							take = edge_src[e] >= cur_t;
						}
						if (take) { t |= u64(1) << b; cur_e = int(e); known = false; last = int(e); }
					}
					tk = t;
				}
			}
			hist = max(hist, last);
			if (w < We) __stcs(out + w, tk);
		}
	}
}


// ---------------------------------------------------------------------------------------------------------
// K2b emit_rows
// ---------------------------------------------------------------------------------------------------------
// Persistent CTAs, one per SM: K2P_CWARPS row-writing warps and K2P_SWARPS staging warps.
//   units   the launch is cut into units of one tile of K2_TCOL aligned columns × rows_per_unit (<= K2P_UNIT_ROWS) rows,
//           numbered tile-fastest. CTAs draw unit numbers from a ticket counter, so an SM that is served more slowly by
//           the memory system simply takes fewer units (with a fixed assignment single SMs ran 20 % longer than the
//           rest); the counter is zeroed on the stream before each launch.
//   stage   per unit the staging warp (a) has one lane issue three bulk async copies (TMA engine) global → shared: the
//           tile of the REF template, the records of the ALT edges that can intersect the tile and those edges' label
//           bytes (contiguous in the flattened label store), so that patching never leaves the SM; (b) gathers the
//           rows' taken-edge words that cover the tile's edges into shared memory. K2P_STAGES stage buffers with full /
//           empty mbarriers: the next unit is staged while the rows of the current one are written, so the row warps
//           never issue a global load and never wait for a tile (with one tile per CTA and no prefetch all CTAs of a
//           wave waited together: 15 % of the kernel; reads queued behind the write stream take microseconds).
//   rows    the (unit, row) items of a CTA are dealt to its row warps round-robin across unit boundaries, so every warp
//           writes the same number of rows. Each warp owns two K2_TCOL-byte row buffers in shared memory. Per row: the
//           template tile (kept in registers while the warp stays in a unit) is written to a buffer with 128-bit STS,
//           the row's taken edges that touch the tile are patched in (label bytes + '-'), and one lane hands the
//           finished buffer to the TMA engine (cp.async.bulk shared → global): the LSU issues no global stores at all,
//           and HBM only ever sees whole 128-byte lines.
// Algorithmic bytes: rows × columns written (+ < 1 % reads). Roofline: HBM write.
constexpr int K2_EDGE_CAP = 256;								// ALT edge records staged per tile (c2 averages ~100)
constexpr int K2_LABEL_CAP = 2048;								// label bytes staged per tile (incl. ≤ 15 bytes of alignment slack)
// Two shapes of the CTA, both 20 warps (5 per scheduler, which leaves 96 registers per thread; 21 warps leave 80 and spill):
//   19 row warps + 1 staging warp   launches whose units are long (>= K2P_LONG_UNIT rows): the row warps hide the staging, and
//                                   the 19th row warp is worth ~1 % (c4: 6,526 against 6,472 GB/s);
//   18 row warps + 2 staging warps  launches of few rows, where a unit is one or two rows per warp and the CTA runs at the
//                                   staging warps' pace: one memory round trip per unit and staging warp (8-row c4 launch:
//                                   2.74 -> 4.34 TB/s, 1 row: 0.43 -> 0.80 TB/s).
constexpr int K2P_MAX_CWARPS = 19;
constexpr u32 K2P_LONG_UNIT = 48;		// 32-row units: 6.16 TB/s with 18 + 2 against 5.74 with 19 + 1; 64-row units: 6.30 against 6.39
constexpr u32 K2P_DONE = 0xffffffffu;								// header word 3: no unit left
constexpr int K2_CHUNKS = K2_TCOL / 512;						// 16-byte chunks of the tile per lane
constexpr u32 K2_SHORT_PATCH = 16;								// longer spans are written by the whole warp

#ifndef V2M_K2P_UNIT_ROWS
#define V2M_K2P_UNIT_ROWS 64
#endif
constexpr int K2P_UNIT_ROWS = V2M_K2P_UNIT_ROWS;				// rows per unit at most
constexpr int K2P_WORDS = 4;									// taken-edge words staged per row (a tile of c2 spans 2 – 3)

struct emit_stage
{
	alignas(128) u8			tmpl[K2_TCOL];
	alignas(128) edge_rec	edges[K2_EDGE_CAP];
	alignas(128) u8			labels[K2_LABEL_CAP];
	alignas(128) u64		words[K2P_UNIT_ROWS][K2P_WORDS];	// the rows' taken-edge words that cover the tile's edges
};

// K2P_CWARPS row warps, K2P_SWARPS staging warps (unit k of a CTA is staged by staging warp k % K2P_SWARPS into stage buffer
// k % K2P_STAGES: each staging warp runs one of its own units ahead).
template <int K2P_CWARPS, int K2P_SWARPS>
struct emit_smem
{
	static constexpr int K2P_STAGES = 2 * K2P_SWARPS;
	emit_stage				stage[K2P_STAGES];
	alignas(128) u8			rows[K2P_CWARPS][2][K2_TCOL];
	alignas(16) uint4		hdr[K2P_STAGES];	// per stage: tile_edges.x, tile_edges.y, label window base, unit << 2 | words staged? << 1 | edges staged? (K2P_DONE: none)
	alignas(8) u64			full[K2P_STAGES];	// the stage's copies have landed (transaction bytes)
	alignas(8) u64			empty[K2P_STAGES];	// every row warp is done with the stage
};
static_assert(sizeof(emit_smem<19, 1>) <= 227 * 1024 && sizeof(emit_smem<18, 2>) <= 227 * 1024);

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
	u32				n_tiles;			// tiles the column window touches
	u32				rows_per_unit, n_row_chunks;
	u32				*tickets;			// [0] next unit; zeroed on the launch stream before every launch
	u8				*out;
	u64				pitch;
};

template <int K2P_CWARPS, int K2P_SWARPS>
__global__ void __launch_bounds__((K2P_CWARPS + K2P_SWARPS) * 32, 1)
emit_rows_kernel(emit_params const p)
{
	constexpr int K2P_STAGES = emit_smem<K2P_CWARPS, K2P_SWARPS>::K2P_STAGES;
	extern __shared__ __align__(128) u8 smem_raw[];
	emit_smem<K2P_CWARPS, K2P_SWARPS> &sm(*reinterpret_cast<emit_smem<K2P_CWARPS, K2P_SWARPS> *>(smem_raw));

	unsigned const lane(threadIdx.x & 31), warp(threadIdx.x >> 5);
	u32 const n_units(p.n_tiles * p.n_row_chunks);
	u32 const RC(p.rows_per_unit);					// <= K2P_UNIT_ROWS

	if (0 == threadIdx.x)
	{
		for (int i = 0; i < K2P_STAGES; ++i) { mbar_init(&sm.full[i], 2); mbar_init(&sm.empty[i], K2P_CWARPS); }
		fence_mbar_init();
	}
	__syncthreads();

	// ---- staging warp: template tile, edge records and labels of the units, one unit ahead of the rows ------------------
	if (warp >= K2P_CWARPS)
	{
		// The units of a CTA form a sequence k = 0, 1, …; staging warp w fills the stages of k = w, w + K2P_SWARPS, … with units
		// it draws itself. With one staging warp a launch of a few rows (units too short for the row warps to hide the
		// staging behind) ran at one staging round trip per unit; two warps halve that.
		u32 const sw(warp - K2P_CWARPS);
		// Two units ahead of itself: the ticket of unit k + 2 is drawn (one lane's atomic, its result not looked at) while unit
		// k is staged, and the edge / label ranges of unit k + 1 are loaded while the words of unit k are on their way. The warp
		// then waits for one memory round trip per unit — the gather — instead of three (ticket, label range, gather), which is
		// what bounds launches of a few rows: their units are too short for the row warps to hide the staging behind.
		auto const draw_ticket([&]() -> u32 { return 0 == lane ? atomicAdd(p.tickets, 1u) : 0u; });
		auto const edges_of([&](u32 const u) -> uint2 { return u < n_units ? p.tile_edges[p.first_tile + u / p.n_row_chunks] : make_uint2(0, 0); });
		auto const labels_of([&](u32 const u) -> uint2 { return u < n_units ? p.tile_labels[p.first_tile + u / p.n_row_chunks] : make_uint2(0, 0); });
		u32 unit(__shfl_sync(0xffffffffu, draw_ticket(), 0));
		u32 ticket_after(draw_ticket());				// lane 0: the unit after `unit`; read at the top of the next iteration
		uint2 te(edges_of(unit)), tl(labels_of(unit));
#pragma unroll 1
		for (u32 k(sw); ; k += K2P_SWARPS)
		{
			u32 const s(k % K2P_STAGES);
			if (k >= K2P_STAGES) mbar_wait(&sm.empty[s], (k / K2P_STAGES - 1) & 1);		// the rows of unit k - K2P_STAGES are done with this stage
			if (unit >= n_units)
			{
				if (0 == lane)
				{
					sm.hdr[s] = make_uint4(0, 0, 0, K2P_DONE);
					mbar_arrive(&sm.full[s]);
					mbar_arrive(&sm.full[s]);
				}
				break;
			}
			u64 const tile(p.first_tile + unit / p.n_row_chunks);
			u64 const w_lo(te.x / 64);
			u32 const n_words(te.y != te.x ? u32((te.y - 1) / 64 - w_lo + 1) : 0);
			bool const words_staged(n_words <= K2P_WORDS);
			if (0 == lane)
			{
				u32 const n_e(te.y - te.x);
				u32 const lab_base(tl.x & ~15u);										// bulk copies start at 16-byte boundaries
				u32 const lab_bytes(tl.y > tl.x ? ((tl.y - lab_base + 15u) & ~15u) : 0);
				bool const staged(n_e <= K2_EDGE_CAP && lab_bytes <= K2_LABEL_CAP);
				u32 const edge_bytes(staged ? n_e * u32(sizeof(edge_rec)) : 0);
				u32 const label_bytes(staged ? lab_bytes : 0);
				sm.hdr[s] = make_uint4(te.x, te.y, lab_base, unit << 2 | (words_staged ? 2u : 0u) | (staged ? 1u : 0u));
				mbar_expect_tx(&sm.full[s], K2_TCOL + edge_bytes + label_bytes);
				bulk_g2s(sm.stage[s].tmpl, p.tmpl + tile * K2_TCOL, K2_TCOL, &sm.full[s]);
				if (edge_bytes) bulk_g2s(sm.stage[s].edges, p.edges + te.x, edge_bytes, &sm.full[s]);
				if (label_bytes) bulk_g2s(sm.stage[s].labels, p.label_chars + lab_base, label_bytes, &sm.full[s]);
			}
			// The rows' taken-edge words for this tile, gathered by the whole warp: all loads first, then the stores.
			constexpr int PER_LANE = K2P_UNIT_ROWS * K2P_WORDS / 32;
			u64 got[PER_LANE];
			bool const gather(words_staged && 0 != n_words);
			if (gather)
			{
				u32 const rb0(p.row_begin + (unit % p.n_row_chunks) * RC);
				u32 const rb1(min(rb0 + RC, p.row_end));
#pragma unroll
				for (int j = 0; j < PER_LANE; ++j)
				{
					u32 const idx(j * 32 + lane), row(idx / K2P_WORDS), i(idx % K2P_WORDS);
					got[j] = (rb0 + row < rb1 && i < n_words) ? p.taken[u64(rb0 + row) * p.We + w_lo + i] : 0;
				}
			}
			u32 const unit_after(__shfl_sync(0xffffffffu, ticket_after, 0));	// drawn one iteration ago
			ticket_after = draw_ticket();
			uint2 const te_after(edges_of(unit_after)), tl_after(labels_of(unit_after));
			if (gather)
			{
#pragma unroll
				for (int j = 0; j < PER_LANE; ++j)
				{
					u32 const idx(j * 32 + lane);
					sm.stage[s].words[idx / K2P_WORDS][idx % K2P_WORDS] = got[j];
				}
			}
			__syncwarp();
			if (0 == lane) mbar_arrive(&sm.full[s]);		// release: the header and the words are visible to the waiters
			unit = unit_after;
			te = te_after; tl = tl_after;
		}
		return;
	}

	// ---- row warps ----------------------------------------------------------------------------------------------------------
	u64 const store_policy(l2_policy_evict_first());		// rows are written once and never read back on the device (+1 %)
	u32 item(warp);					// this warp's next item in the CTA's flattened (unit, row in unit) sequence
	u32 n_issued(0);				// rows handed to the TMA engine so far: buffer n_issued & 1 is the one to fill
	u32 finished(0);				// bit w: staging warp w has run out of units (it posts K2P_DONE once and stops)
#pragma unroll 1
	for (u32 k(0); ; ++k)
	{
		u32 const s(k % K2P_STAGES);
		u32 const item_end((k + 1) * RC);
		u32 const sw(k % K2P_SWARPS);
		bool skip(0 != (finished >> sw & 1u));
		uint4 hdr(make_uint4(0, 0, 0, K2P_DONE));
		if (!skip)
		{
			mbar_wait(&sm.full[s], (k / K2P_STAGES) & 1);
			hdr = sm.hdr[s];
			if (K2P_DONE == hdr.w) { finished |= 1u << sw; skip = true; }
		}
		if (skip)
		{
			if (finished == (1u << K2P_SWARPS) - 1) break;
			while (item < item_end) item += K2P_CWARPS;			// this k carries no unit: its item numbers are not used
			continue;
		}
		if (item < item_end)
		{
			u32 const unit(hdr.w >> 2);
			bool const words_staged(0 != (hdr.w & 2u));
			u64 const tile(p.first_tile + unit / p.n_row_chunks);
			u64 const t0(tile * K2_TCOL);
			u64 const c0(max(t0, p.col_begin)), c1(min(t0 + K2_TCOL, p.col_end));
			u32 const rb0(p.row_begin + (unit % p.n_row_chunks) * RC);
			u32 const rb1(min(rb0 + RC, p.row_end));
			emit_stage const &st(sm.stage[s]);
			u32 const te_x(hdr.x), te_y(hdr.y), lab_base(hdr.z);
			bool const staged(0 != (hdr.w & 1u)), has_edges(te_y != te_x);
			u64 const w_lo(te_x / 64), w_hi(has_edges ? (te_y - 1) / 64 : w_lo);
			u32 const n_words(has_edges ? u32(w_hi - w_lo + 1) : 0);

			u32 r(rb0 + (item - k * RC));

			uint4 v[K2_CHUNKS];
#pragma unroll
			for (int j = 0; j < K2_CHUNKS; ++j)
				v[j] = *reinterpret_cast<uint4 const *>(st.tmpl + j * 512 + lane * 16);

			// What one row hands to the TMA engine: the 16-byte aligned part of [c0, c1); a ragged end goes out byte-wise.
			u32 const off0(u32(c0 - t0));
			u32 const n_bytes(u32(c1 - c0));
			u32 const bulk_bytes(n_bytes & ~15u);
			u32 const t0_32 = u32(t0), c0_32 = u32(c0), c1_32 = u32(c1);

			// Which bits of taken word i (lane ↔ word) belong to edges of this tile.
			auto const tile_bits([&](u32 const i) -> u64 {
				if (i >= n_words) return 0;
				u64 keep(~u64(0));
				u64 const w(w_lo + i);
				if (w == w_lo) keep &= ~u64(0) << (te_x % 64);
				if (w == w_hi && 0 != te_y % 64) keep &= (u64(1) << (te_y % 64)) - 1;
				return keep;
			});
			u64 const keep0(tile_bits(lane));

			// The rows of this warp in this unit. FAST: everything the patches read was staged (the common case), so it is
			// addressed as shared memory; otherwise edge records, labels and / or taken words come from the global arrays.
			auto const write_rows([&](auto const fast_tag) {
				constexpr bool FAST = decltype(fast_tag)::value;
				auto const edge_record([&](u32 const e) -> edge_rec {
					uint4 const q(FAST ? reinterpret_cast<uint4 const *>(st.edges)[e - te_x]
						: (staged ? reinterpret_cast<uint4 const *>(st.edges)[e - te_x] : reinterpret_cast<uint4 const *>(p.edges)[e]));
					return edge_rec{q.x, q.y, q.z, q.w};
				});
				auto const patch([&](edge_rec const &rec, u32 const col) -> u8 {
					u32 const k(col - rec.a0);
					if (k >= rec.label_len) return u8('-');
					if (FAST || staged) return st.labels[rec.label_off - lab_base + k];
					return p.label_chars[rec.label_off + k];
				});

#pragma unroll 1
				for (; item < item_end; item += K2P_CWARPS, r += K2P_CWARPS)
				{
					if (r >= rb1) continue;			// the last row chunk may be short
					u8 *buf(sm.rows[warp][n_issued & 1]);
					if (n_issued >= 2)
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
							// lane ↔ taken word chunk + lane
							u64 word;
							if (FAST) word = st.words[r - rb0][lane % K2P_WORDS] & keep0;
							else
							{
								u32 const i(chunk + lane);
								word = i < n_words ? (words_staged ? st.words[r - rb0][i % K2P_WORDS] : p.taken[u64(r) * p.We + w_lo + i]) & tile_bits(i) : 0;
							}
							unsigned nz(__ballot_sync(0xffffffffu, 0 != word));
#pragma unroll 1
							while (nz)
							{
								// One non-empty word at a time, its 64 edges spread over the lanes: the row's patches in this word are
								// applied side by side (taken edges of one row never overlap, so their order does not matter).
								int const wl(__ffs(nz) - 1);
								nz &= nz - 1;
								u32 const bits_lo(__shfl_sync(0xffffffffu, u32(word), wl)), bits_hi(__shfl_sync(0xffffffffu, u32(word >> 32), wl));
								u32 const e_base(64 * (u32(w_lo) + chunk + u32(wl)));			// edge indices fit 32 bits (tile_edges)
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
										edge_rec const rec(edge_record(e_base + 32 * half + lane));
										u32 const lo(max(rec.a0, c0_32)), hi(min(rec.a1, c1_32));
										if (lo < hi)
										{
											if (hi - lo <= K2_SHORT_PATCH)
												for (u32 col(lo); col < hi; ++col) buf[col - t0_32] = patch(rec, col);
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
										for (u32 col(lo + lane); col < hi; col += 32) buf[col - t0_32] = patch(rec, col);
									}
								}
							}
						}
					}

					fence_proxy_async_smem();
					__syncwarp();
					u8 *orow(p.out + u64(r - p.row_begin) * p.pitch + (c0 - p.col_begin));
					if (0 == lane)
					{
						// One bulk group per row, empty when the row's window is under 16 bytes: n_issued picks the buffer and
						// wait_group.read<1> counts groups, so the two must advance together (otherwise two such rows in a row
						// would let the second one refill a buffer the engine may still be reading).
						if (bulk_bytes) bulk_s2g_hint(orow, buf + off0, bulk_bytes, store_policy);
						bulk_commit();
					}
					if (lane < n_bytes - bulk_bytes)
						orow[bulk_bytes + lane] = buf[off0 + bulk_bytes + lane];
					++n_issued;
				}
			});
			if (staged && words_staged) write_rows(std::true_type{});
			else write_rows(std::false_type{});
		}
		__syncwarp();
		if (0 == lane) mbar_arrive(&sm.empty[s]);		// release: this warp's reads of the stage are done
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

int launch_check_assigned(u32 const *d_assigned, u64 n, u32 H, u32 *d_bad, cudaStream_t stream)
{
	if (0 == n) return 0;
	check_assigned_kernel<<<unsigned(std::min<u64>((n + 255) / 256, 148 * 8)), 256, 0, stream>>>(d_assigned, n, H, d_bad);
	V2M_CHECK_LAUNCH("check_assigned_kernel");
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
		d_lines, d_line_idx, n_rows, g.We, g.edge_simple, g.edge_sibling, g.edge_src, g.edge_tgt, d_taken);
	V2M_CHECK_LAUNCH("resolve_rows_kernel");
	return 0;
}

// The unit ticket of emit_rows_kernel: one counter per (device, stream), zeroed by an 8-byte memset on the launch stream
// before every launch (so an aborted launch cannot poison the next one on the same or a recycled stream handle);
// cudaStreamPerThread names a different stream in every host thread, hence the thread id in the key.
static int unit_tickets(int const dev, cudaStream_t const stream, u32 *&tickets)
{
	static std::mutex mtx;
	static std::map<std::tuple<int, cudaStream_t, std::thread::id>, u32 *> table;
	std::lock_guard<std::mutex> const lock(mtx);
	u32 *&slot(table[{dev, stream, cudaStreamPerThread == stream ? std::this_thread::get_id() : std::thread::id()}]);
	if (!slot)
		V2M_CUDA_TRY(cudaMalloc(&slot, 2 * sizeof(u32)));
	tickets = slot;
	V2M_CUDA_TRY(cudaMemsetAsync(slot, 0, 2 * sizeof(u32), stream));
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
	p.n_tiles = u32(last_tile - p.first_tile + 1);
	int dev(0), sms(0);
	V2M_CUDA_TRY(cudaGetDevice(&dev));
	V2M_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
	// Units: whole tiles (all rows) when there are enough of them to keep every SM equally busy; otherwise the rows are cut
	// as well, down to one row per row warp.
	u32 chunks((n_rows + K2P_UNIT_ROWS - 1) / K2P_UNIT_ROWS);
	if (u64(p.n_tiles) * chunks < 4 * u64(sms))
		chunks = std::max<u32>(chunks, std::min<u32>((n_rows + K2P_MAX_CWARPS - 1) / K2P_MAX_CWARPS, u32((4 * u64(sms) + p.n_tiles - 1) / p.n_tiles)));
	if (char const *env = std::getenv("V2M_K2_UNIT_ROWS"))			// kernel experiments
		if (int const v = std::atoi(env); v > 0 && v <= K2P_UNIT_ROWS) chunks = (n_rows + u32(v) - 1) / u32(v);
	p.rows_per_unit = (n_rows + chunks - 1) / chunks;
	p.n_row_chunks = (n_rows + p.rows_per_unit - 1) / p.rows_per_unit;
	u64 const n_units(u64(p.n_tiles) * p.n_row_chunks);
	if (n_units * p.rows_per_unit >= (u64(1) << 30)) return fail(1, "too many rows x tiles in one emit_rows launch");
	if (int const st = unit_tickets(dev, stream, p.tickets)) return st;
	unsigned const grid(unsigned(std::min<u64>(n_units, u64(sms))));
	bool long_units(p.rows_per_unit >= K2P_LONG_UNIT);
	if (char const *env = std::getenv("V2M_K2_STAGING_WARPS"))		// kernel experiments: 1 / 2
		if (int const v = std::atoi(env); 1 == v || 2 == v) long_units = 1 == v;
	if (long_units)
	{
		V2M_CUDA_TRY(cudaFuncSetAttribute(emit_rows_kernel<19, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(emit_smem<19, 1>))));	// per device
		emit_rows_kernel<19, 1><<<grid, 20 * 32, sizeof(emit_smem<19, 1>), stream>>>(p);
	}
	else
	{
		V2M_CUDA_TRY(cudaFuncSetAttribute(emit_rows_kernel<18, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(emit_smem<18, 2>))));
		emit_rows_kernel<18, 2><<<grid, 20 * 32, sizeof(emit_smem<18, 2>), stream>>>(p);
	}
	V2M_CHECK_LAUNCH("emit_rows_kernel");
	return 0;
}

} // namespace v2m
