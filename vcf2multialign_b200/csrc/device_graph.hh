// Device-resident form of v2m::variant_graph (include/vcf2multialign/variant_graph.hh:57-66) + the reference.
// HBM layout (all arrays 256-byte aligned by cudaMalloc):
//   ref          L bytes                               reference sequence, verbatim
//   refpos       N × u32                               reference_positions
//   alnpos       N × u32                               aligned_positions (A = alnpos[N-1] < 2^32)
//   edge_src     E_pad × u32                           source node of each ALT edge (from alt_edge_count_csum)
//   edge_tgt     E_pad × u32                           alt_edge_targets
//   edge_simple  E_pad bits                            1 = K2a decides the edge without the walk's state: first ALT edge of a node no
//                                                      earlier node's edge reaches beyond (taken whenever set), or a sibling:
//   edge_sibling E_pad bits                            1 = 2nd … 4th ALT edge of such a node, same 64-edge word as the first (taken
//                                                      iff set and no earlier edge of the node is set)
//   edges        E_pad × 16 B  (edge_rec)              aligned span [aln[src], aln[tgt]) + label offset/length
//   label_chars  Σ|label| bytes                        alt_edge_labels, flattened
//   paths_eh     E_pad × (H_pad/64) × u64              paths_by_edge_and_chrom_copy.values(): one line per edge
//   paths_he     H_pad × (E_pad/64) × u64              paths_by_chrom_copy_and_edge.values(): one line per chromosome copy
//   tmpl         A_alloc bytes                         REF-aligned template (A rounded up to a whole K2 tile)
//   tile_edges   ⌈A / K2_TCOL⌉ × uint2                 per column tile: (first edge whose span can reach into the tile,
//                                                      first edge that starts at or after the tile's end)
//   tile_labels  ⌈A / K2_TCOL⌉ × uint2                 per column tile: byte range of those edges' labels in label_chars
//   edge_ref     E_pad × uint2                         reference span [refpos[src], refpos[tgt]) of each ALT edge (--unaligned)
//   ref_tile_edges ⌈ref_end / K2_TCOL⌉ × uint2         tile_edges in reference coordinates (--unaligned)
#pragma once

#include <mutex>
#include <vector>

#include <cuda_runtime.h>

#include "common.cuh"

namespace v2m {

constexpr int K2_TCOL = 4096;			// columns per emit_rows tile
constexpr int K2_THREADS = 256;
#ifndef V2M_K2_ROWS_PER_CTA
#define V2M_K2_ROWS_PER_CTA 64
#endif
constexpr int K2_ROWS_PER_CTA = V2M_K2_ROWS_PER_CTA;

struct device_graph
{
	int			device{};
	u64			N{}, E{}, L{}, A{}, A_alloc{};
	u64			h_pad{}, e_pad{};
	u64			Wh{}, We{};					// words per edge line / per chromosome-copy line
	u32			H{};
	u64			n_tiles{};
	u64			label_bytes{};

	u8			*ref{};
	u32			*refpos{}, *alnpos{};
	u32			*edge_src{}, *edge_tgt{};
	u64			*edge_simple{};				// E_pad bits: decided without the walk's state (first ALT of a clean node, or a sibling below) (K2a)
	u64			*edge_sibling{};			// E_pad bits: 2nd … 4th ALT of a clean node whose earlier ALTs lie in the same 64-edge word (K2a)
	edge_rec	*edges{};
	u8			*label_chars{};
	u64			*paths_eh{}, *paths_he{};
	u8			*tmpl{};
	uint2		*tile_edges{};
	uint2		*tile_labels{};				// per column tile: [begin, end) of its edges' bytes in label_chars
	uint2		*edge_ref{};
	uint2		*ref_tile_edges{};
	u64			n_ref_tiles{};
	u32			ref_end{};					// reference_positions.back(): where every walk stops
	u64			*edge_of_node{};			// N + 1: alt_edge_count_csum (host copy kept for cut → edge index)
};

// A set of output rows resolved on the device: taken-edge bit lines, one per row.
struct device_plan
{
	device_graph const	*graph{};
	u32					n_rows{};
	u64					*taken{};			// n_rows × We
	u64 const			*lines{};			// K2a input: We-word path lines (the graph's paths_he, or the composed founder lines below)
	u32					*line_idx{};		// haplotype plans: row i reads line line_idx[i] (PLOIDY_MAX: the REF row); founders: NULL
	u64					*own_lines{};		// founder plans: n_rows composed lines (compose_founder_kernel)
	int					*prefix{};			// --unaligned: n_rows × (We + 1) exclusive prefix sums of (|label| − REF span) over taken edges
	u64					*lengths{};			// --unaligned: n_rows row lengths
};

int launch_transpose_bits(void const *d_src, u64 rows, u64 cols, void *d_dst, cudaStream_t stream);
int launch_transpose_bits_into(void const *d_src, u64 rows, u64 cols, void *d_dst, u64 dst_line_bits, cudaStream_t stream);
int launch_build_template(device_graph const &g, cudaStream_t stream);
int launch_check_assigned(u32 const *d_assigned, u64 n, u32 H, u32 *d_bad, cudaStream_t stream);
int launch_compose_founders(device_graph const &g, u64 const *d_cut_edges, u64 n_segments, u32 const *d_assigned, u32 n_founders, u64 *d_out, cudaStream_t stream);
int launch_resolve_rows(device_graph const &g, u64 const *d_lines, u32 const *d_line_idx, u32 n_rows, u64 *d_taken, cudaStream_t stream);
int launch_emit_rows(device_graph const &g, u64 const *d_taken, u32 row_begin, u32 row_end, u64 col_begin, u64 col_end,
	void *d_out, u64 pitch, cudaStream_t stream);
int launch_unaligned_prefix(device_graph const &g, u64 const *d_taken, u32 n_rows, int *d_prefix, u64 *d_lengths, cudaStream_t stream);
int launch_emit_unaligned(device_graph const &g, u64 const *d_taken, int const *d_prefix, u32 row_begin, u32 row_end,
	u64 const *d_offsets, void *d_out, cudaStream_t stream);


// What a v2m_graph_t points to: the device arrays plus host copies of the small node / edge arrays the host-side halves
// need (alt_edge_count_csum: node id → first edge index; aligned_positions and alt_edge_targets: candidate cut positions).
// The pBWT state (permutation, divergence) before a set of edge columns, left behind by the last find_cut_positions on the
// graph (its chunk boundaries after pass B): find_matchings on the same graph starts its chunks from these states instead of
// running the chunk-local sweep and the boundary composition again (the reference, too, sweeps the same columns twice:
// find_cut_positions.cc:170-176 and founder_sequence_greedy_output.cc:461-469).
struct sweep_states
{
	std::vector<u32>	edges;				// ascending edge indices, edges[0] = 0
	unsigned short		*perm{};			// [edges.size()][Hp]
	u32					*div{};				// [edges.size()][Hp]
	u32					Hp{};

	void clear() { cudaFree(perm); cudaFree(div); perm = nullptr; div = nullptr; edges.clear(); Hp = 0; }
};

struct graph_holder
{
	device_graph		g;
	std::vector<u64>	csum;
	std::vector<u64>	alnpos;
	std::vector<u64>	targets;
	mutable std::mutex	sweep_mutex;
	mutable sweep_states sweep;

	~graph_holder()
	{
		sweep.clear();
		cudaFree(g.edge_simple); cudaFree(g.edge_sibling); cudaFree(g.ref); cudaFree(g.refpos); cudaFree(g.alnpos); cudaFree(g.edge_src); cudaFree(g.edge_tgt);
		cudaFree(g.edges); cudaFree(g.label_chars); cudaFree(g.paths_eh); cudaFree(g.paths_he);
		cudaFree(g.tmpl); cudaFree(g.tile_edges); cudaFree(g.tile_labels); cudaFree(g.edge_ref); cudaFree(g.ref_tile_edges);
	}
};

} // namespace v2m

struct v2m_graph_s : v2m::graph_holder {};
