/*
 * v2m_cuda.h — C ABI of libv2m_cuda.so, the B200 (sm_100a) implementation of vcf2multialign's
 * aligned-sequence materialisation path.
 *
 * The reference is a C++ static library (libvcf2multialign.a) without any FFI; "drop-in" means link-time
 * substitution of the symbols listed below. Each entry point cites the reference interface it stands behind
 * (paths relative to the reference checkout). The C++ adapters that re-create those interfaces on top of this
 * ABI live in vcf2multialign_b200/host/ (see INTEGRATION.md).
 *
 * Conventions: every call returns 0 on success and a non-zero status otherwise; v2m_cuda_last_error() returns
 * the message of the calling thread's last failure. Plain pointers and sizes only. There is no CPU fallback:
 * without a CUDA device every compute call fails with V2M_ERR_CUDA.
 *
 * Bit-matrix layout (libbio int_matrix, lib/libbio/include/libbio/int_matrix/int_matrix.hh:144-175,
 * int_vector/int_vector.hh:253-261): column-major, element (y, x) is bit ((x*rows + y) % 64) of
 * 64-bit word ((x*rows + y) / 64); rows and cols are multiples of 64 for every matrix passed here.
 */
#ifndef V2M_CUDA_H
#define V2M_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define V2M_OK            0
#define V2M_ERR_ARGUMENT  1
#define V2M_ERR_CUDA      2
#define V2M_ERR_STATE     3

/* variant_graph::PLOIDY_MAX (include/vcf2multialign/variant_graph.hh:55): "follow REF edges only". */
#define V2M_PLOIDY_MAX 0xffffffffu

typedef struct v2m_graph_s *v2m_graph_t;   /* device-resident variant graph + reference */
typedef struct v2m_plan_s *v2m_plan_t;     /* a set of output rows (haplotypes or founders) resolved on device */

/* Flat view of v2m::variant_graph (include/vcf2multialign/variant_graph.hh:57-66) + the reference sequence
 * (sequence_type, variant_graph.hh:33). std::vector<std::string> alt_edge_labels is flattened into offsets + chars. */
typedef struct v2m_graph_view
{
	uint64_t        node_count;                    /* reference_positions.size() */
	uint64_t        edge_count;                    /* alt_edge_targets.size() */
	const uint64_t *reference_positions;           /* [node_count] */
	const uint64_t *aligned_positions;             /* [node_count] */
	const uint64_t *alt_edge_targets;              /* [edge_count] */
	const uint64_t *alt_edge_count_csum;           /* [node_count + 1] */
	const uint64_t *label_offsets;                 /* [edge_count + 1] into label_chars */
	const char     *label_chars;
	/* paths_by_edge_and_chrom_copy: rows = h_pad chromosome copies, cols = e_pad edges (edge-major lines).
	 * paths_by_chrom_copy_and_edge: rows = e_pad, cols = h_pad (copy-major lines); may be NULL, in which case it is
	 * produced on the device with the K1 transpose, as variant_graph.cc:453 does on the host. */
	const uint64_t *paths_by_edge_and_chrom_copy;
	const uint64_t *paths_by_chrom_copy_and_edge;
	uint64_t        h_pad;
	uint64_t        e_pad;
	uint32_t        total_chromosome_copies;       /* ploidy_csum.back() */
	const char     *ref_seq;
	uint64_t        ref_len;
	/* Non-zero: the two path pointers above are DEVICE pointers (e.g. filled by v2m_cuda_synth_paths or by an upstream
	 * kernel); they are copied device-to-device. */
	int             paths_on_device;
} v2m_graph_view;

/* --- library / device ------------------------------------------------------------------------------------ */
const char *v2m_cuda_last_error(void);
int v2m_cuda_init(int device);                      /* cudaSetDevice + context warm-up */
int v2m_cuda_device_count(int *count);
/* Counts of kernels this library has launched since load (for bench.py's gpu_launches). */
uint64_t v2m_cuda_launch_count(void);

/* Pinned host / device buffers and plain copies, so that callers (tests, bench, host adapters) need no CUDA headers. */
void *v2m_cuda_host_alloc(uint64_t bytes);
void v2m_cuda_host_free(void *p);
void *v2m_cuda_device_alloc(uint64_t bytes);
void v2m_cuda_device_free(void *p);
int v2m_cuda_memcpy_d2h(void *dst, const void *d_src, uint64_t bytes);
int v2m_cuda_memcpy_h2d(void *d_dst, const void *src, uint64_t bytes);
int v2m_cuda_synchronize(void);
/* CUDA events on the library's launching stream (NULL = default stream). */
void *v2m_cuda_event_create(void);
void v2m_cuda_event_destroy(void *ev);
int v2m_cuda_event_record(void *ev, void *stream);
int v2m_cuda_event_elapsed_ms(void *e0, void *e1, float *ms);   /* synchronises on e1 */

/* --- K1: v2m::transpose_matrix (include/vcf2multialign/transpose_matrix.hh:14,
 *         libvcf2multialign/transpose_matrix.cc:41-109) ---------------------------------------------------- */
/* Host buffers; dst has the same word count as src. rows, cols multiples of 64 (else V2M_ERR_ARGUMENT, the
 * reference asserts); cols == 0 is a no-op (the reference returns a default matrix). */
int v2m_cuda_transpose_bits(const uint64_t *src, uint64_t rows, uint64_t cols, uint64_t *dst);
/* Device buffers, asynchronous on `stream` (a cudaStream_t cast to void*; NULL = default stream). */
int v2m_cuda_transpose_bits_device(const void *d_src, uint64_t rows, uint64_t cols, void *d_dst, void *stream);

/* --- graph residency ------------------------------------------------------------------------------------- */
/* view->ref_seq may be NULL: the graph then serves the sweeps (v2m_cuda_find_cut_positions, v2m_cuda_find_matchings,
 * v2m_cuda_segment_classes — the reference's find_cut_positions(graph, …) / find_matchings(graph, …) take no reference
 * sequence either, include/vcf2multialign/output.hh:123,126) and plans; calls that produce rows return V2M_ERR_STATE until
 * v2m_cuda_graph_attach_reference has supplied it. */
int v2m_cuda_graph_upload(const v2m_graph_view *view, v2m_graph_t *out);
/* Uploads the reference sequence of a graph that was uploaded without one and builds the REF-aligned template; a no-op when
 * the graph has it already. */
int v2m_cuda_graph_attach_reference(v2m_graph_t g, const char *ref_seq, uint64_t ref_len);
int v2m_cuda_graph_free(v2m_graph_t g);
int v2m_cuda_graph_aligned_length(v2m_graph_t g, uint64_t *out);   /* aligned_positions.back() */
/* Copy-major path words as resident on the device (rows = e_pad, cols = h_pad): lets callers fetch the K1 result
 * when the view did not supply paths_by_chrom_copy_and_edge. */
int v2m_cuda_graph_download_paths_by_chrom_copy_and_edge(v2m_graph_t g, uint64_t *dst);
/* Lines [first_copy, first_copy + n_copies) of that matrix only (e_pad / 64 words per chromosome copy). */
int v2m_cuda_graph_download_path_lines(v2m_graph_t g, uint32_t first_copy, uint32_t n_copies, uint64_t *dst);
/* The REF-aligned template (= the bytes of the REF row), A bytes. */
int v2m_cuda_graph_download_template(v2m_graph_t g, uint8_t *dst);

/* --- K2a: which ALT edges each output row follows
 *          (the edge test + jump in v2m::output_sequence, libvcf2multialign/sequence_writer.cc:44-67) ------ */
/* Haplotype rows: row i follows chromosome copy copy_idx[i] (sequence_writing_delegate::chromosome_copy_index,
 * include/vcf2multialign/sequence_writer.hh:16-36; haplotype_output.cc:22-32). V2M_PLOIDY_MAX = the REF row. */
int v2m_cuda_plan_haplotypes(v2m_graph_t g, const uint32_t *copy_idx, uint32_t n_rows, v2m_plan_t *out);
/* Founder rows: row f switches to chromosome copy assigned[f*n_segments + s] at node cut_positions[s]
 * (founder_sequence_writing_delegate, libvcf2multialign/founder_sequence_greedy_output.cc:78-115; `assigned` is the
 * column-major lb::matrix of founder_sequence_greedy_output::assigned_samples(), segments × founders).
 * cut_positions has n_segments + 1 node ids, first 0, last the sink. */
int v2m_cuda_plan_founders(v2m_graph_t g, const uint64_t *cut_positions, uint64_t n_segments,
                           const uint32_t *assigned, uint32_t n_founders, v2m_plan_t *out);
/* Runs K2a again for every row of the plan, asynchronously on `stream` (NULL = default stream): the per-row edge test + jump
 * of sequence_writer.cc:44-67 is part of every pass over the rows, so callers that time a whole pass (bench.py) put it inside
 * the timed step; the plan's taken bits are rewritten with the same values. */
int v2m_cuda_plan_resolve(v2m_plan_t p, void *stream);
int v2m_cuda_plan_free(v2m_plan_t p);
int v2m_cuda_plan_row_count(v2m_plan_t p, uint32_t *out);
/* Taken-edge bits, one line of e_pad bits per row (bit e set iff the row's walk follows ALT edge e). */
int v2m_cuda_plan_download_taken(v2m_plan_t p, uint64_t *dst);
/* Per-row output length: aligned → A for every row; unaligned → Σ piece lengths (sequence_writer.cc:80). */
int v2m_cuda_plan_row_lengths(v2m_plan_t p, int unaligned, uint64_t *out /* [n_rows] */);

/* --- K2b: the bytes of v2m::output_sequence (libvcf2multialign/sequence_writer.cc:70-83) ------------------- */
/* Aligned rows [row_begin, row_end) × columns [col_begin, col_end) of the plan, sequence bytes only (no FASTA
 * header, no newline): out[(r - row_begin) * pitch + (c - col_begin)].
 * Device variant: d_out must be 16-byte aligned, pitch % 16 == 0, col_begin % 16 == 0; asynchronous on `stream`.
 * Fastest with rows starting on 128-byte lines (pitch % 128 == 0, but not a multiple of 4096) and 64 or more rows per
 * call. */
int v2m_cuda_emit_rows_device(v2m_plan_t p, uint32_t row_begin, uint32_t row_end, uint64_t col_begin, uint64_t col_end,
                              void *d_out, uint64_t pitch, void *stream);
/* Host variant: kernels fill device tiles that are copied into `out` (any host memory; pinned is faster). Bytes of `out`
 * between the end of a row's columns and the start of the next row (pitch > col_end - col_begin) are unspecified after the
 * call: when pitch equals v2m_cuda_row_pitch(col_end - col_begin) and col_begin % 16 == 0, whole row blocks cross the host
 * link as single 1-D copies (4 % faster than strided ones on PCIe Gen5), padding included. */
int v2m_cuda_emit_rows(v2m_plan_t p, uint32_t row_begin, uint32_t row_end, uint64_t col_begin, uint64_t col_end,
                       uint8_t *out, uint64_t pitch);
/* The row pitch the library's own device buffers use for rows of row_bytes bytes: a multiple of 128 that is not a multiple
 * of 4096 (rows start on a 128-byte line; the same columns of all rows do not fall on the same DRAM channels). */
uint64_t v2m_cuda_row_pitch(uint64_t row_bytes);
/* The host variants keep two device staging tiles and two streams per calling host thread between calls; this frees them
 * (they are also freed when the thread ends). */
void v2m_cuda_release_staging(void);
/* Unaligned rows (--unaligned: no gap fill): row r is written at out + offsets[r - row_begin], length from
 * v2m_cuda_plan_row_lengths(…, 1, …). */
int v2m_cuda_emit_rows_unaligned(v2m_plan_t p, uint32_t row_begin, uint32_t row_end, const uint64_t *offsets, uint8_t *out);

/* --- whole-stream drivers: haplotype_output::output_a2m / founder_sequence_greedy_output::output_a2m
 *     (libvcf2multialign/haplotype_output.cc:38-82, founder_sequence_greedy_output.cc:515-550) -------------- */
/* Writes, for every row i of the plan, '>' ids[i] '\n' row-bytes '\n' to file descriptor fd, in row order, streaming
 * through pinned double buffers. ids: n_rows NUL-terminated strings, concatenated. row_done (may be NULL) is called
 * from the calling thread after each row has been handed to write(2) — the hook behind output_delegate::handled_sequences. */
typedef void (*v2m_row_callback)(uint32_t row, void *ctx);
int v2m_cuda_write_a2m(v2m_plan_t p, const char *ids, int unaligned, int fd, v2m_row_callback row_done, void *ctx);
/* The same stream handed to a caller-supplied sink (e.g. an std::ostream behind the C++ adapter, or the stdin of a --pipe
 * subprocess, output.cc:49-67). The sink returns 0 on success; anything else aborts the stream with V2M_ERR_STATE. */
typedef int (*v2m_sink)(const void *bytes, uint64_t n, void *ctx);
int v2m_cuda_stream_a2m(v2m_plan_t p, const char *ids, int unaligned, v2m_sink sink, void *sink_ctx, v2m_row_callback row_done, void *ctx);

/* The rows alone — what haplotype_output::output_separate / output::output_sequence_file (libvcf2multialign/output.cc:18-44)
 * write one file or --pipe subprocess per row from: sink(row, bytes, n, ctx) is called once per row of the plan, in row
 * order, from the calling thread, with the row's bytes in a pinned buffer that is reused after the call returns. Same block
 * pipeline as v2m_cuda_stream_a2m (the next block is produced and copied while the sink consumes this one). */
typedef int (*v2m_row_sink)(uint32_t row, const void *bytes, uint64_t n, void *ctx);
int v2m_cuda_stream_rows(v2m_plan_t p, int unaligned, v2m_row_sink sink, void *sink_ctx);

/* --- K3: per-cut haplotype classes for founder_sequence_greedy_output::find_matchings
 *         (libvcf2multialign/founder_sequence_greedy_output.cc:213-253, 460-469; pbwt.hh:77-134) ----------- */
/* cut_positions: n_cuts node ids as in founder_sequence_greedy_output::cut_positions() (first 0, last the sink). Every
 * segment must contain at least one ALT edge (true for cuts from find_initial_cut_positions_lambda_min,
 * find_cut_positions.cc:129), otherwise V2M_ERR_ARGUMENT. For each cut c in 1 … n_cuts-1, with segment edges [s, e) between
 * cut c-1 and cut c:
 *   rhs_eq[(c-1)*H + h]   representative of h's class on [s, e) (the member first in the pBWT order at s); may be NULL
 *   distinct[c-1]         number of classes                    (rhs_distinct_eq_classes)
 *   first_class[c-1]      permutation.front() at the cut       (rhs_first_path_eq_class)
 *   first_is_ref[c-1]     rhs_first_path_is_ref accumulated over [s, e)
 *   joined triples (lhs_rep, rhs_rep, size) in permutation order: joined_offsets[c-1] … joined_offsets[c] index into joined
 *   (for c = 1 the lhs is V2M_PLOIDY_MAX: the single-segment case of :474-509). joined holds joined_cap triples; if that is
 *   too small the call returns V2M_ERR_ARGUMENT with the needed count in joined_offsets[n_cuts-1].
 * LIMIT: at most 16,384 chromosome copies — the sweep keeps the pBWT state (permutation as 16-bit positions, divergence) of
 * all copies in one CTA's shared memory (10 bytes per copy + two column buffers: 225 of the 227 KiB a CTA may have at
 * 16,384 copies); the reference's pbwt_context (include/vcf2multialign/pbwt.hh:61-74) has no such limit. 1000 Genomes phase 3
 * (5,008 copies) and 1000G + HGDP (~8,400) fit; above 8,192 copies the kernels spill a few registers and run slower per
 * column. Larger cohorts get V2M_ERR_ARGUMENT from the three founder-mode calls (K3, find_matchings, find_cut_positions).
 * Row emission has no such limit. */
int v2m_cuda_segment_classes(v2m_graph_t g, const uint64_t *cut_positions, uint64_t n_cuts,
                             uint32_t *rhs_eq, uint32_t *distinct, uint32_t *first_class, uint8_t *first_is_ref,
                             uint64_t *joined_offsets /* [n_cuts] */, uint32_t *joined, uint64_t joined_cap);

/* founder_sequence_greedy_output::find_matchings(graph, founder_count) (founder_sequence_greedy_output.cc:154-512; declared
 * include/vcf2multialign/output.hh:126): K3 on the device, then the greedy assignment on the host (same std::sort /
 * std::multimap as the reference). assigned = m_assigned_samples: column-major lb::matrix, (n_cuts - 1) segments × founder_count,
 * unassigned entries V2M_PLOIDY_MAX. *found = what the reference returns (0: fewer than two cut positions or no chromosome
 * copies). Inputs on which the reference asserts (:355) or does not terminate (:309-319, :429-438) give V2M_ERR_STATE. */
int v2m_cuda_find_matchings(v2m_graph_t g, const uint64_t *cut_positions, uint64_t n_cuts, uint32_t founder_count,
                            int keep_ref_edges, uint32_t *assigned, int *found);

/* --- K4: founder_sequence_greedy_output::find_cut_positions(graph, minimum_distance) (include/vcf2multialign/output.hh:123,
 *         libvcf2multialign/founder_sequence_greedy_output.cc:139-152), i.e. find_initial_cut_positions_lambda_min
 *         (include/vcf2multialign/find_cut_positions.hh:22-27, libvcf2multialign/find_cut_positions.cc:93-211) ------ */
/* The pBWT sweep and, at every candidate cut position, the (candidate index, equivalence class count) pairs the reference
 * derives from its divergence value counts run on the device; the recurrence over candidates and the trace-back run on the
 * host. cut_positions receives *n_cuts node ids (first 0, last the sink, as cut_positions() returns them); *score =
 * max_segmentation_height() (the CLI prints 1 + score, main.cc:520); *found = what the reference's member returns (0 iff
 * the score is CUT_POSITION_SCORE_MAX). If cap is too small the call returns V2M_ERR_ARGUMENT with the needed count in
 * *n_cuts. At most 16,384 chromosome copies (see v2m_cuda_segment_classes). */
int v2m_cuda_find_cut_positions(v2m_graph_t g, uint64_t min_distance, uint64_t *cut_positions, uint64_t cap,
                                uint64_t *n_cuts, uint32_t *score, int *found);
/* Timing of the calling thread's last v2m_cuda_find_cut_positions: ms4 = {local sweep, boundary composition, list extraction
 * (device, CUDA events), whole call (host clock)}; counts6 = {chunks, candidates, ALT edges, list pairs read by the recurrence,
 * list capacity per candidate, list extraction attempts}. */
int v2m_cuda_last_cut_positions_timing(float *ms4, uint64_t *counts6);

/* --- K5: the "Paths" loop of build_variant_graph (libvcf2multialign/variant_graph.cc:379-425): path bits from parsed GT
 *         fields (the step before the path) --------------------------------------------------------- */
#define V2M_GT_NULL 0xffffu   /* vcf::sample_genotype::NULL_ALLELE as passed here ('.' in a GT field) */
/* gt[r * n_copies + row]: the GT allele (sample_gt.alt) of handled record r for output row `row` = ploidy_csum[sample] + copy
 * (0 = REF, V2M_GT_NULL = null, a > 0 = ALT a). ALT a of record r has edge edges_by_alt[alt_offsets[r] + a - 1], or EDGE_MAX
 * (all ones) if it got none (variant_graph.cc:328-364). The result is paths_by_edge_and_chrom_copy (rows = h_pad, cols =
 * e_pad), zero outside the set bits. A genotype naming an ALT its record does not have gives V2M_ERR_ARGUMENT (the reference
 * asserts, :399).
 * Overlap reports (:405-418), optional (n_overlaps != NULL, record_ref_pos and edge_target_pos given): for every chromosome
 * copy, in record order, (record, row) is reported when record_ref_pos[r] < the target reference position of the copy's
 * previous ALT edge (edge_target_pos[edge], n_edges entries); pairs come back sorted by record, then row, i.e. in the
 * reference's reporting order. If overlap_cap is too small: V2M_ERR_ARGUMENT with the needed count in *n_overlaps.
 * gt may be pageable memory: it is staged in ~64 MB record blocks through two pinned buffers, which the calling thread keeps
 * until it exits (so repeated calls, one per block of records, do not page-lock memory again). */
int v2m_cuda_paths_from_genotypes(const uint16_t *gt, uint64_t n_records, uint32_t n_copies,
                                  const uint64_t *alt_offsets /* [n_records + 1] */, const uint64_t *edges_by_alt,
                                  const uint64_t *record_ref_pos, const uint64_t *edge_target_pos, uint64_t n_edges,
                                  uint64_t h_pad, uint64_t e_pad, uint64_t *paths_by_edge_and_chrom_copy,
                                  uint64_t *overlaps /* [overlap_cap][2] */, uint64_t overlap_cap, uint64_t *n_overlaps);
/* Same, the matrix stays on the device (d_paths: e_pad * h_pad / 8 bytes), ready for v2m_graph_view.paths_on_device. */
int v2m_cuda_paths_from_genotypes_device(const uint16_t *gt, uint64_t n_records, uint32_t n_copies,
                                         const uint64_t *alt_offsets, const uint64_t *edges_by_alt,
                                         const uint64_t *record_ref_pos, const uint64_t *edge_target_pos, uint64_t n_edges,
                                         uint64_t h_pad, uint64_t e_pad, void *d_paths,
                                         uint64_t *overlaps, uint64_t overlap_cap, uint64_t *n_overlaps);

/* Both orientations born on the device: d_paths_by_edge_and_chrom_copy as above and d_paths_by_chrom_copy_and_edge (the matrix
 * build_variant_graph derives at its end with transpose_matrix, variant_graph.cc:447-453), filled a range of final edge columns
 * at a time after every record block, while the block's columns are still in the L2 — a graph built this way (both pointers in
 * v2m_graph_view with paths_on_device = 1) never runs K1 over the whole matrix. */
int v2m_cuda_paths_from_genotypes_both(const uint16_t *gt, uint64_t n_records, uint32_t n_copies,
                                       const uint64_t *alt_offsets, const uint64_t *edges_by_alt,
                                       const uint64_t *record_ref_pos, const uint64_t *edge_target_pos, uint64_t n_edges,
                                       uint64_t h_pad, uint64_t e_pad, void *d_paths_by_edge_and_chrom_copy, void *d_paths_by_chrom_copy_and_edge,
                                       uint64_t *overlaps, uint64_t overlap_cap, uint64_t *n_overlaps);

/* K5 for callers that cannot allocate. No counterpart in the reference: added for code
 * under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc. Same result as
 * v2m_cuda_paths_from_genotypes_device, but EVERY array is device memory the caller owns (d_gt: n_records * n_copies uint16;
 * d_alt_offsets: n_records + 1 uint64; d_edges_by_alt; d_record_ref_pos: n_records uint64 and d_edge_target_pos: n_edges
 * uint64, both only needed with n_overlaps != NULL; d_paths: e_pad * h_pad / 8 bytes, zeroed here; d_overlaps: overlap_cap
 * (record, row) uint64 pairs, written in NO particular order — sort by (record, row) for the reference's reporting order),
 * and the call makes no cudaMalloc / cudaFree / stream or event creation: the kernels' scratch (error flags, overlap counter,
 * 8 bytes of state per chromosome copy) is a static device array of V2M_NO_ALLOC_SCRATCH_BYTES that every GPU which loads
 * the library holds until the process exits.
 * LIMIT: n_copies <= V2M_NO_ALLOC_MAX_COPIES, else V2M_ERR_ARGUMENT (use v2m_cuda_paths_from_genotypes_device).
 * The array is shared by every call on a GPU and keeps the last call's contents (the part a call uses is cleared on `stream`
 * first): calls that overlap on one GPU — two host threads, or two streams — race; serialise them.
 * The first call on a device looks the array's device address up (cudaGetSymbolAddress) and, under CUDA's lazy module loading,
 * loads the kernels: make one warm-up call per device before entering a region where that is not allowed.
 * Work is enqueued on `stream` (NULL = the default stream); the call returns after the stream has drained, with
 * V2M_ERR_ARGUMENT for a genotype naming an ALT its record does not have, or when more than overlap_cap pairs were found
 * (needed count in *n_overlaps; the first overlap_cap pairs are valid). */
#define V2M_NO_ALLOC_MAX_COPIES 65536u
#define V2M_NO_ALLOC_SCRATCH_BYTES (256u + 8u * V2M_NO_ALLOC_MAX_COPIES)   /* 524,544 bytes per GPU */
int v2m_cuda_ingest_genotypes_no_alloc(const void *d_gt, uint64_t n_records, uint32_t n_copies,
                                       const void *d_alt_offsets, const void *d_edges_by_alt,
                                       const void *d_record_ref_pos, const void *d_edge_target_pos, uint64_t n_edges,
                                       uint64_t h_pad, uint64_t e_pad, void *d_paths,
                                       void *d_overlaps /* device, [overlap_cap][2] */, uint64_t overlap_cap, uint64_t *n_overlaps,
                                       void *stream);

/* --- synthetic workloads (measurement infrastructure) ------------------------------------------------ */
/* Fills d_dst (device, e_pad × h_pad/64 words = paths_by_edge_and_chrom_copy) with LD-structured path bits that are a
 * pure function of (seed, chromosome copy, site); per edge: its site, its 0-based ALT number, the site's ALT count. */
int v2m_cuda_synth_paths(void *d_dst, uint64_t h_pad, uint64_t e_pad, uint32_t H, uint64_t E, uint32_t seed,
                         const uint32_t *edge_site, const uint8_t *edge_alt, const uint8_t *edge_nalts);

/* --- measurement hooks (bench.py) ------------------------------------------------------------------------- */
/* Mean device time in milliseconds of `iters` back-to-back launches of the named stage on resident data,
 * measured with CUDA events on the launching stream. */
int v2m_cuda_time_emit_rows(v2m_plan_t p, uint32_t row_begin, uint32_t row_end, void *d_out, uint64_t pitch,
                            int iters, float *ms_mean);
/* Device times (CUDA events) of the three K3 passes of the calling thread's last v2m_cuda_segment_classes /
 * v2m_cuda_find_matchings call: ms3 = {local sweep, boundary composition, per-cut emission};
 * counts4 = {chunks, cut positions, ALT edges, joined triples}. */
int v2m_cuda_last_segment_classes_timing(float *ms3, uint64_t *counts4);
int v2m_cuda_time_transpose_bits(const void *d_src, uint64_t rows, uint64_t cols, void *d_dst, int iters, float *ms_mean);
/* The host link's ceiling for the end-to-end roofline: `iters` passes of plain cudaMemcpyAsync device-to-host copies of `bytes`
 * in pieces of block_bytes, alternating between the calling thread's two streams (no kernel runs); *seconds = host time from
 * the first copy to the last stream synchronisation. h_dst should be pinned (v2m_cuda_host_alloc). */
int v2m_cuda_time_d2h_copies(void *h_dst, const void *d_src, uint64_t bytes, uint64_t block_bytes, int iters, double *seconds);
/* K5 alone on GT values, offsets and edge indices already resident in HBM. */
int v2m_cuda_time_ingest_genotypes(const void *d_gt, uint64_t n_records, uint32_t n_copies, const void *d_alt_offsets, const void *d_edges_by_alt,
                                   uint64_t h_pad, uint64_t e_pad, void *d_paths, int iters, float *ms_mean);

#ifdef __cplusplus
}
#endif
#endif /* V2M_CUDA_H */
