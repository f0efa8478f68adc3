// ORACLE — TEST INFRASTRUCTURE ONLY.
//
// CPU restatement (plain C++20, no Boost/Ragel/libbio) of the aligned-sequence
// materialisation path of tsnorri/vcf2multialign. Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may use anything under oracle/.
// The product path (vcf2multialign_b200/) never includes, links or calls this.
//
// The real reference cannot be compiled in this image (Boost, Ragel, gengetopt and libbsd
// are absent; SURVEY.md header), so parity is pinned by the reference's own golden
// vectors instead: tests/transpose_matrix.cc:188-279, tests/variant_graph.cc:249-338,
// tests/founder_sequences.cc:118-188 (see tests/test_oracle_golden.py).
//
// Every function cites the reference lines it follows (paths relative to /root/reference).
#pragma once

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <functional>
#include <limits>
#include <map>
#include <numeric>
#include <ostream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <string_view>
#include <utility>
#include <vector>

namespace v2m_oracle {

using u64 = std::uint64_t;
using u32 = std::uint32_t;

constexpr u32 PLOIDY_MAX = std::numeric_limits<u32>::max();   // variant_graph.hh:55
constexpr u64 EDGE_MAX = std::numeric_limits<u64>::max();     // variant_graph.hh:53
constexpr u32 CUT_POSITION_SCORE_MAX = PLOIDY_MAX;            // find_cut_positions.hh:17

struct oracle_error : std::runtime_error { using std::runtime_error::runtime_error; };

#define ORACLE_ASSERT(cond) do { if (!(cond)) throw ::v2m_oracle::oracle_error( \
	std::string("assertion failed: ") + #cond + " (" + __FILE__ + ":" + std::to_string(__LINE__) + ")"); } while (0)


// Packed bit matrix with libbio's layout: column-major, element i of the underlying
// int_vector<1> is bit (i % 64) of word (i / 64), idx(y, x) = x * rows + y
// (lib/libbio/include/libbio/int_vector/int_vector.hh:253-261, int_matrix/int_matrix.hh:144-175,
// matrix/indexing.hh:15-18).
struct bit_matrix
{
	std::vector<u64>	words;
	u64					bits{};		// element count of the underlying vector
	u64					stride{1};	// = number of rows

	bit_matrix() = default;
	bit_matrix(u64 rows, u64 cols): words((rows * cols + 63) / 64, 0), bits(rows * cols), stride(rows) {}

	u64 number_of_rows() const { return stride; }
	u64 number_of_columns() const { return bits / stride; }
	u64 size() const { return bits; }
	bool get(u64 y, u64 x) const { u64 const i(x * stride + y); return (words[i / 64] >> (i % 64)) & 1; }
	void set(u64 y, u64 x) { u64 const i(x * stride + y); words[i / 64] |= u64(1) << (i % 64); }
	// int_vector::resize(new_size, 0): keeps the stride, new elements are zero.
	void resize_bits(u64 new_bits) { words.resize((new_bits + 63) / 64, 0); bits = new_bits; }
	u64 const *column_words(u64 x) const { return words.data() + (x * stride) / 64; } // stride % 64 == 0
	bool operator==(bit_matrix const &o) const { return words == o.words && stride == o.stride; } // int_matrix.hh:251-254
};


// variant_graph.hh:36-79
struct variant_graph
{
	std::vector<u64>			reference_positions;
	std::vector<u64>			aligned_positions;
	std::vector<u64>			alt_edge_targets;
	std::vector<u64>			alt_edge_count_csum;
	std::vector<std::string>	alt_edge_labels;
	bit_matrix					paths_by_chrom_copy_and_edge;	// rows = edges, cols = chromosome copies
	bit_matrix					paths_by_edge_and_chrom_copy;	// rows = chromosome copies, cols = edges
	std::vector<std::string>	sample_names;
	std::vector<u32>			ploidy_csum;

	u64 node_count() const { return reference_positions.size(); }
	u64 edge_count() const { return alt_edge_targets.size(); }
	u32 sample_ploidy(u32 s) const { return ploidy_csum[1 + s] - ploidy_csum[s]; }
	u32 total_chromosome_copies() const { return ploidy_csum.empty() ? 0 : ploidy_csum.back(); }
	u64 aligned_length(u64 lhs, u64 rhs) const { return aligned_positions[rhs] - aligned_positions[lhs]; }

	// variant_graph.cc:77-105
	u64 add_node(u64 ref_pos, u64 aln_pos)
	{
		reference_positions.push_back(ref_pos);
		aligned_positions.push_back(aln_pos);
		alt_edge_count_csum.push_back(alt_edge_count_csum.back());
		return reference_positions.size() - 1;
	}

	u64 add_or_update_node(u64 ref_pos, u64 aln_pos)
	{
		u64 const last(reference_positions.back());
		ORACLE_ASSERT(last <= ref_pos);
		if (last < ref_pos)
			return add_node(ref_pos, aln_pos);
		aligned_positions.back() = std::max(aligned_positions.back(), aln_pos);
		return reference_positions.size() - 1;
	}

	u64 add_edge(std::string_view label = {})
	{
		++alt_edge_count_csum.back();
		alt_edge_targets.emplace_back();
		alt_edge_labels.emplace_back(label);
		return alt_edge_targets.size() - 1;
	}
};


// ---------------------------------------------------------------------------------------
// transpose_matrix — libvcf2multialign/transpose_matrix.cc:18-109
// ---------------------------------------------------------------------------------------

// 8×8 bit-block transpose inside one word (transpose_matrix.cc:18-38). Restated as the
// three-stage masked swap network; identical function to the reference's 15-term
// mask-and-shift sum (bit (8r + c) ↔ bit (8c + r)).
inline u64 transpose8x8(u64 x)
{
	u64 t;
	t = (x ^ (x >> 7)) & 0x00AA00AA00AA00AAULL;  x ^= t ^ (t << 7);
	t = (x ^ (x >> 14)) & 0x0000CCCC0000CCCCULL; x ^= t ^ (t << 14);
	t = (x ^ (x >> 28)) & 0x00000000F0F0F0F0ULL; x ^= t ^ (t << 28);
	return x;
}

// Literal 15-term form of transpose_matrix.cc:21-37, kept to cross-check the network above.
inline u64 transpose8x8_terms(u64 const w)
{
	static constexpr u64 diag(0x8040201008040201ULL);
	u64 r(w & diag);
	for (unsigned k(1); k < 8; ++k)
	{
		r |= (w & (diag << (8 * k))) >> (7 * k);	// below the diagonal
		r |= (w & (diag >> (8 * k))) << (7 * k);	// above the diagonal
	}
	return r;
}

// transpose_matrix.cc:41-109. Same traversal (64-row word band × byte-in-word × 64-column
// group; eight 8×8 blocks per step), same preconditions (dims % 64 == 0, empty → default).
inline bit_matrix transpose_matrix(bit_matrix const &src)
{
	u64 const nrow(src.number_of_rows()), ncol(src.number_of_columns());
	if (0 == ncol)
		return bit_matrix{};

	bit_matrix dst(ncol, nrow);
	ORACLE_ASSERT(0 == nrow % 64);
	ORACLE_ASSERT(0 == ncol % 64);
	u64 const col_groups(ncol / 64), src_col_words(nrow / 64), dst_col_words(ncol / 64);
	ORACLE_ASSERT(src.words.size() == dst.words.size());

	for (u64 rw(0); rw < src_col_words; ++rw)
		for (unsigned byte(0); byte < 8; ++byte)
			for (u64 cg(0); cg < col_groups; ++cg)
			{
				u64 blocks[8];
				for (unsigned b(0); b < 8; ++b)
				{
					u64 acc(0);
					for (unsigned k(0); k < 8; ++k)
					{
						u64 const col(64 * cg + 8 * b + k);
						acc |= ((src.words[col * src_col_words + rw] >> (8 * byte)) & 0xff) << (8 * k);
					}
					blocks[b] = transpose8x8(acc);
				}
				for (unsigned k(0); k < 8; ++k)
				{
					u64 &dw(dst.words[(64 * rw + 8 * byte + k) * dst_col_words + cg]);
					for (unsigned b(0); b < 8; ++b)
						dw |= ((blocks[b] >> (8 * k)) & 0xff) << (8 * b);
				}
			}
	return dst;
}


// ---------------------------------------------------------------------------------------
// FASTA: first (or named) record, sequence bytes verbatim, newlines dropped
// (lib/libbio/src/fasta_reader.rl:30-73, 202-223, 284-295).
// ---------------------------------------------------------------------------------------
inline bool read_single_fasta_sequence(char const *path, std::vector<char> &seq, char const *seq_name = nullptr)
{
	std::ifstream is(path, std::ios::binary);
	if (!is) throw oracle_error(std::string("unable to open ") + path);
	seq.clear();
	std::string line;
	bool copying(false), seen_header(false);
	while (std::getline(is, line))
	{
		if (!line.empty() && '\r' == line.back()) line.pop_back();
		if (line.empty()) continue;
		if (';' == line[0]) continue;
		if ('>' == line[0])
		{
			if (copying) return true;	// handle_sequence_end() cancels the parse
			seen_header = true;
			auto const end(line.find_first_of(" \t", 1));
			std::string_view const id(std::string_view(line).substr(1, (std::string::npos == end ? line.size() : end) - 1));
			if (!seq_name || id == seq_name) copying = true;
			continue;
		}
		if (!seen_header) copying = true;	// "sequence_only" input
		if (copying) seq.insert(seq.end(), line.begin(), line.end());
	}
	return true;
}


// ---------------------------------------------------------------------------------------
// Minimal VCF reader: CHROM POS ID REF ALT … FORMAT(GT first or anywhere) samples.
// ALT kinds per lib/libbio/src/vcf_reader_parser.rl:395-445; GT per
// lib/libbio/src/vcf_genotype_field_gt_parser.rl:52-107.
// ---------------------------------------------------------------------------------------
enum class sv_type { NONE, DEL, OTHER_SV, UNKNOWN };

struct vcf_alt { std::string alt; sv_type type{sv_type::NONE}; };

constexpr std::uint16_t NULL_ALLELE = 0xffff;

struct vcf_record
{
	u64								lineno{};
	std::string						chrom;
	u64								pos{};		// 1-based
	std::string						id;
	std::string						ref;
	std::vector<vcf_alt>			alts;
	std::vector<std::vector<std::uint16_t>> gt;	// per sample, per chromosome copy
	bool							has_gt{};
};

inline std::vector<std::string_view> split(std::string_view sv, char sep)
{
	std::vector<std::string_view> out;
	std::size_t b(0);
	while (true)
	{
		auto const e(sv.find(sep, b));
		if (std::string_view::npos == e) { out.push_back(sv.substr(b)); break; }
		out.push_back(sv.substr(b, e - b));
		b = e + 1;
	}
	return out;
}

inline vcf_alt classify_alt(std::string_view a)
{
	vcf_alt r; r.alt = std::string(a);
	if ("." == a) r.type = sv_type::UNKNOWN;			// alt_unknown
	else if (!a.empty() && '<' == a.front())
	{
		// <DEL> and <DEL:subtype…> are DEL; everything else is some other SV kind.
		// (<DEL:ME…> is DEL_ME in the reference grammar; treated as OTHER_SV here and avoided in data.)
		auto const body(a.substr(1, a.size() - 2));
		auto const head(body.substr(0, body.find(':')));
		if ("DEL" == head && 0 != body.rfind("DEL:ME", 0)) r.type = sv_type::DEL;
		else r.type = sv_type::OTHER_SV;
	}
	else r.type = sv_type::NONE;							// bases or '*'
	return r;
}

struct vcf_file
{
	std::vector<std::string>	sample_names;
	std::vector<vcf_record>		records;
	u64							last_header_lineno{};
};

inline vcf_file read_vcf(char const *path)
{
	std::ifstream is(path, std::ios::binary);
	if (!is) throw oracle_error(std::string("unable to open ") + path);
	vcf_file vf;
	std::string line;
	u64 lineno(0);
	while (std::getline(is, line))
	{
		++lineno;
		if (!line.empty() && '\r' == line.back()) line.pop_back();
		if (line.empty()) continue;
		if ('#' == line[0])
		{
			vf.last_header_lineno = lineno;
			if (0 == line.rfind("#CHROM", 0))
			{
				auto const f(split(line, '\t'));
				for (std::size_t i(9); i < f.size(); ++i) vf.sample_names.emplace_back(f[i]);
			}
			continue;
		}
		auto const f(split(line, '\t'));
		if (f.size() < 8) throw oracle_error("malformed VCF line " + std::to_string(lineno));
		vcf_record rec;
		rec.lineno = lineno - vf.last_header_lineno;
		rec.chrom = std::string(f[0]);
		rec.pos = std::stoull(std::string(f[1]));
		rec.id = std::string(f[2]);
		rec.ref = std::string(f[3]);
		for (auto const a : split(f[4], ',')) rec.alts.push_back(classify_alt(a));
		if (f.size() > 9)
		{
			auto const fmt(split(f[8], ':'));
			std::size_t gt_idx(fmt.size());
			for (std::size_t i(0); i < fmt.size(); ++i) if ("GT" == fmt[i]) gt_idx = i;
			rec.has_gt = gt_idx < fmt.size();
			if (rec.has_gt)
				for (std::size_t s(9); s < f.size(); ++s)
				{
					auto const sub(split(f[s], ':'));
					std::vector<std::uint16_t> gts;
					if (gt_idx < sub.size())
					{
						std::string_view const g(sub[gt_idx]);
						std::size_t i(0);
						while (i <= g.size())
						{
							std::size_t j(i);
							while (j < g.size() && '|' != g[j] && '/' != g[j]) ++j;
							auto const part(g.substr(i, j - i));
							std::uint16_t idx(0);
							if ("." == part) idx = NULL_ALLELE;
							else { for (char const c : part) idx = idx * 10 + (c - '0'); }
							if (NULL_ALLELE != idx && idx > rec.alts.size()) idx = 0;	// gt parser end_part: substitute with zero
							gts.push_back(idx);
							i = j + 1;
						}
					}
					rec.gt.push_back(std::move(gts));
				}
		}
		vf.records.push_back(std::move(rec));
	}
	return vf;
}


// ---------------------------------------------------------------------------------------
// build_variant_graph — libvcf2multialign/variant_graph.cc:108-454
// ---------------------------------------------------------------------------------------
struct overlap_report { u64 lineno; u64 ref_pos; std::string var_id; std::string sample; u32 chrom_copy; u32 gt; };

// What the "Paths" loop of build_variant_graph (variant_graph.cc:379-425) consumes and produces, per handled record, in
// flat form: the input and expected output of the device's genotype ingestion. Filled only when
// build_delegate::trace is set.
struct build_trace
{
	u32							copies{};			// ploidy_csum.back()
	std::vector<std::uint16_t>	gt;					// [records][copies]: sample_gt.alt by output row index (0 = REF, NULL_ALLELE = '.')
	std::vector<u64>			alt_offsets{0};		// [records + 1] into edges_by_alt
	std::vector<u64>			edges_by_alt;		// edge index of every ALT, EDGE_MAX where the ALT has no edge
	std::vector<u64>			ref_pos;			// [records]
	std::vector<u64>			edge_target_pos;	// [edges] reference position of the edge's target node
	std::vector<u64>			overlaps;			// (record, row) pairs in the order reported (:405-418)
};

struct build_delegate
{
	std::function<bool(std::string_view, u32)>	should_include = [](std::string_view, u32){ return true; };
	std::vector<overlap_report>				overlaps;
	std::vector<u64>							ref_mismatches;	// 1-based variant indices; building stops at the first
	build_trace									*trace{};
};

inline void build_variant_graph(
	std::vector<char> const &ref_seq, vcf_file const &vf, char const *chr_id,
	variant_graph &graph, build_delegate &delegate, u64 *handled_variants = nullptr, u64 *chr_id_mismatches = nullptr)
{
	struct edge_destination { u64 edge_index; u64 position; };
	constexpr u64 row_col_divisor(64);		// variant_graph.cc:127
	constexpr u64 column_allocation(512);	// variant_graph.cc:128
	std::string_view const ref_sv(ref_seq.data(), ref_seq.size());

	graph = variant_graph{};
	graph.sample_names = vf.sample_names;
	graph.alt_edge_count_csum.push_back(0);
	graph.add_node(0, 0);

	u64 var_idx(0), aln_pos(0), prev_ref_pos(0);
	bool is_first(true);
	std::vector<u64> edges_by_alt, target_ref_positions_by_chrom_copy, current_edge_targets;
	std::multimap<u64, edge_destination> next_aligned_positions;
	struct sample_chromosome_index { u32 sample_vcf, sample_out, copy_vcf, copy_out; };
	std::vector<sample_chromosome_index> included;

	// variant_graph.cc:157-179
	auto add_target_nodes([&](u64 const ref_pos){
		auto it(next_aligned_positions.begin());
		auto const begin(it);
		for (; it != next_aligned_positions.end(); ++it)
		{
			if (ref_pos < it->first) break;
			u64 const dist(it->first - prev_ref_pos);
			aln_pos = std::max(aln_pos + dist, it->second.position);
			u64 const node_idx(graph.add_or_update_node(it->first, aln_pos));
			ORACLE_ASSERT(it->second.edge_index < graph.alt_edge_targets.size());
			graph.alt_edge_targets[it->second.edge_index] = node_idx;
			prev_ref_pos = it->first;
		}
		next_aligned_positions.erase(begin, it);
	});

	for (auto const &var : vf.records)
	{
		++var_idx;
		if (var.chrom != chr_id) { if (chr_id_mismatches) ++*chr_id_mismatches; continue; }
		if (!var.has_gt) throw oracle_error("Variant " + std::to_string(var_idx) + " does not have a genotype.");

		if (is_first)
		{
			// variant_graph.cc:215-289 — ploidy from the first record; sample filtering.
			is_first = false;
			graph.ploidy_csum.assign(1 + graph.sample_names.size(), 0);
			std::vector<u32> removed;
			u32 sample_out(0);
			for (u32 s(0); s < var.gt.size(); ++s)
			{
				u32 inc(0);
				for (u32 c(0); c < var.gt[s].size(); ++c)
					if (delegate.should_include(vf.sample_names[s], c))
					{
						included.push_back({s, sample_out, c, inc});
						++inc;
					}
				if (inc)
				{
					graph.ploidy_csum[1 + sample_out] = graph.ploidy_csum[sample_out] + inc;
					++sample_out;
				}
				else removed.push_back(s);
			}
			if (!removed.empty())
			{
				graph.ploidy_csum.resize(graph.ploidy_csum.size() - removed.size());
				removed.push_back(UINT32_MAX);
				auto it(removed.begin());
				std::vector<std::string> kept;
				for (u32 i(0); i < graph.sample_names.size(); ++i)
				{
					if (i < *it) kept.push_back(std::move(graph.sample_names[i]));
					else ++it;
				}
				graph.sample_names.swap(kept);
			}
			u64 const rows(row_col_divisor * ((graph.ploidy_csum.back() + row_col_divisor - 1) / row_col_divisor));
			graph.paths_by_edge_and_chrom_copy = rows ? bit_matrix(rows, column_allocation) : bit_matrix(1, 0);
			target_ref_positions_by_chrom_copy.assign(graph.ploidy_csum.back(), 0);
		}

		if (handled_variants) ++*handled_variants;
		u64 const ref_pos(var.pos - 1);
		if (!(prev_ref_pos <= ref_pos))
			throw oracle_error("Variant " + std::to_string(var_idx) + " has non-increasing position");

		add_target_nodes(ref_pos);													// :303
		aln_pos += ref_pos - prev_ref_pos;											// :306-307
		graph.add_or_update_node(ref_pos, aln_pos);								// :308

		if (var.ref != ref_sv.substr(ref_pos, var.ref.size()))						// :313-317
		{
			delegate.ref_mismatches.push_back(var_idx);
			break;	// the test delegates return false → parsing stops
		}

		// Edges (:319-366): only plain (NONE) and <DEL> ALTs get one.
		edges_by_alt.assign(var.alts.size(), EDGE_MAX);
		u64 min_edge(0), max_edge(0);
		current_edge_targets.clear();
		{
			bool first_edge(true);
			for (std::size_t a(0); a < var.alts.size(); ++a)
			{
				auto const &alt(var.alts[a]);
				if (sv_type::NONE != alt.type && sv_type::DEL != alt.type) continue;
				u64 const target_pos(ref_pos + var.ref.size());
				u64 edge_idx;
				if (sv_type::NONE == alt.type)
				{
					edge_idx = graph.add_edge(alt.alt);
					next_aligned_positions.emplace(target_pos, edge_destination{edge_idx, aln_pos + alt.alt.size()});
				}
				else
				{
					edge_idx = graph.add_edge();
					next_aligned_positions.emplace(target_pos, edge_destination{edge_idx, aln_pos});
				}
				edges_by_alt[a] = edge_idx;
				current_edge_targets.push_back(target_pos);
				if (first_edge) { min_edge = edge_idx; first_edge = false; }
				max_edge = edge_idx;
			}
		}

		// Column growth (:369-376).
		{
			u64 const ncol(graph.paths_by_edge_and_chrom_copy.number_of_columns());
			if (ncol && ncol <= max_edge)
			{
				u64 const multiplier(4 + ncol / column_allocation);
				graph.paths_by_edge_and_chrom_copy.resize_bits(graph.paths_by_edge_and_chrom_copy.number_of_rows() * multiplier * column_allocation);
			}
		}

		// Paths (:379-425).
		if (delegate.trace)
		{
			build_trace &tr(*delegate.trace);
			tr.copies = graph.ploidy_csum.back();
			tr.gt.resize(tr.gt.size() + tr.copies, 0);
			tr.edges_by_alt.insert(tr.edges_by_alt.end(), edges_by_alt.begin(), edges_by_alt.end());
			tr.alt_offsets.push_back(tr.edges_by_alt.size());
			tr.ref_pos.push_back(ref_pos);
			tr.edge_target_pos.insert(tr.edge_target_pos.end(), current_edge_targets.begin(), current_edge_targets.end());
		}
		for (auto const &sc : included)
		{
			auto const &gt(var.gt[sc.sample_vcf]);
			ORACLE_ASSERT(sc.copy_vcf < gt.size());
			u64 const base_idx(graph.ploidy_csum[sc.sample_out]);
			auto const alt_idx(gt[sc.copy_vcf]);
			if (delegate.trace)
				delegate.trace->gt[delegate.trace->gt.size() - delegate.trace->copies + base_idx + sc.copy_out] = alt_idx;
			if (0 == alt_idx) continue;
			if (NULL_ALLELE == alt_idx) continue;
			ORACLE_ASSERT(u64(alt_idx - 1) < edges_by_alt.size());
			u64 const edge_idx(edges_by_alt[alt_idx - 1]);
			if (EDGE_MAX == edge_idx) continue;
			u64 const row_idx(base_idx + sc.copy_out);
			if (ref_pos < target_ref_positions_by_chrom_copy[row_idx])
			{
				delegate.overlaps.push_back({var.lineno + vf.last_header_lineno, ref_pos, var.id, vf.sample_names[sc.sample_vcf], sc.copy_vcf, alt_idx});
				if (delegate.trace) { delegate.trace->overlaps.push_back(delegate.trace->ref_pos.size() - 1); delegate.trace->overlaps.push_back(row_idx); }
			}
			ORACLE_ASSERT(edge_idx - min_edge < current_edge_targets.size());
			target_ref_positions_by_chrom_copy[row_idx] = current_edge_targets[edge_idx - min_edge];
			graph.paths_by_edge_and_chrom_copy.set(row_idx, edge_idx);
		}
		prev_ref_pos = ref_pos;
	}

	// Sink node (:437-443).
	{
		u64 const ref_pos(ref_seq.size());
		add_target_nodes(ref_pos);
		graph.add_or_update_node(ref_pos, aln_pos + (ref_pos - prev_ref_pos));
	}

	// Shrink the column count to a multiple of 64 (:447-451) and transpose (:453).
	if (graph.paths_by_edge_and_chrom_copy.size())
	{
		u64 const ncol(row_col_divisor * ((graph.edge_count() + row_col_divisor - 1) / row_col_divisor));
		graph.paths_by_edge_and_chrom_copy.resize_bits(graph.paths_by_edge_and_chrom_copy.number_of_rows() * ncol);
	}
	graph.paths_by_chrom_copy_and_edge = transpose_matrix(graph.paths_by_edge_and_chrom_copy);
}


// ---------------------------------------------------------------------------------------
// output_sequence — libvcf2multialign/sequence_writer.cc:22-85
// ---------------------------------------------------------------------------------------
struct sequence_writing_delegate	// sequence_writer.hh:16-36
{
	u32 chromosome_copy_index{PLOIDY_MAX};
	sequence_writing_delegate() = default;
	explicit sequence_writing_delegate(u32 idx): chromosome_copy_index(idx) {}
	virtual ~sequence_writing_delegate() {}
	virtual void handle_node(variant_graph const &, u64) {}
};

// Byte sink abstraction so that the same walk can feed a std::string, a FILE or a counter.
// `faithful` mode (bench only) mimics the reference's per-character gap fill through a
// 4 KiB stream buffer (sequence_writer.cc:81; libbio file_handling.hh:26).
struct byte_sink
{
	virtual ~byte_sink() {}
	virtual void write(char const *p, std::size_t n) = 0;
	virtual void fill(char c, std::size_t n) = 0;
	void put(char c) { write(&c, 1); }
	void puts(std::string_view sv) { write(sv.data(), sv.size()); }
};

struct string_sink final : byte_sink
{
	std::string s;
	void write(char const *p, std::size_t n) override { s.append(p, n); }
	void fill(char c, std::size_t n) override { s.append(n, c); }
};

// Fixed 4 KiB buffer that is "flushed" by discarding; gap characters go in one at a time.
struct faithful_null_sink final : byte_sink
{
	char buf[4096]; std::size_t used{}; u64 total{}; volatile char guard{};
	void flush() { if (used) guard = buf[used - 1]; total += used; used = 0; }
	void write(char const *p, std::size_t n) override
	{
		while (n)
		{
			std::size_t const k(std::min(n, sizeof(buf) - used));
			std::memcpy(buf + used, p, k); used += k; p += k; n -= k;
			if (sizeof(buf) == used) flush();
		}
	}
	void fill(char c, std::size_t n) override
	{
		for (std::size_t i(0); i < n; ++i) { buf[used++] = c; if (sizeof(buf) == used) flush(); }
	}
};

// Bulk variant: row-sized buffer, memset gaps ("fair" CPU baseline).
struct bulk_null_sink final : byte_sink
{
	std::vector<char> buf; std::size_t used{}; u64 total{}; volatile char guard{};
	explicit bulk_null_sink(std::size_t cap): buf(cap) {}
	void flush() { if (used) guard = buf[used - 1]; total += used; used = 0; }
	void write(char const *p, std::size_t n) override
	{
		while (n)
		{
			std::size_t const k(std::min(n, buf.size() - used));
			std::memcpy(buf.data() + used, p, k); used += k; p += k; n -= k;
			if (buf.size() == used) flush();
		}
	}
	void fill(char c, std::size_t n) override
	{
		while (n)
		{
			std::size_t const k(std::min(n, buf.size() - used));
			std::memset(buf.data() + used, c, k); used += k; n -= k;
			if (buf.size() == used) flush();
		}
	}
};

inline void output_sequence(
	std::vector<char> const &ref_seq, variant_graph const &graph, byte_sink &stream,
	char const *fasta_identifier, bool const should_output_unaligned, sequence_writing_delegate &delegate)
{
	if (fasta_identifier) { stream.put('>'); stream.puts(fasta_identifier); stream.put('\n'); }

	u64 ref_pos(0), aln_pos(0), next_ref_pos(0), next_aln_pos(0), current_node(0);
	u64 const limit(graph.node_count() - 1);
	while (current_node < limit)
	{
		delegate.handle_node(graph, current_node);
		std::size_t label_size(0);
		bool followed_alt(false);
		if (PLOIDY_MAX != delegate.chromosome_copy_index)
		{
			u64 const lb(graph.alt_edge_count_csum[current_node]), rb(graph.alt_edge_count_csum[1 + current_node]);
			for (u64 e(lb); e < rb; ++e)
				if (graph.paths_by_chrom_copy_and_edge.get(e, delegate.chromosome_copy_index))
				{
					u64 const target(graph.alt_edge_targets[e]);
					auto const &label(graph.alt_edge_labels[e]);
					next_ref_pos = graph.reference_positions[target];
					next_aln_pos = graph.aligned_positions[target];
					ORACLE_ASSERT(label.size() <= next_aln_pos - aln_pos);
					stream.puts(label);
					current_node = target;
					label_size = label.size();
					followed_alt = true;
					break;
				}
		}
		if (!followed_alt)
		{
			next_ref_pos = graph.reference_positions[current_node + 1];
			next_aln_pos = graph.aligned_positions[current_node + 1];
			stream.write(ref_seq.data() + ref_pos, next_ref_pos - ref_pos);
			label_size = next_ref_pos - ref_pos;
			++current_node;
		}
		if (!should_output_unaligned)
			stream.fill('-', next_aln_pos - aln_pos - label_size);
		ref_pos = next_ref_pos;
		aln_pos = next_aln_pos;
	}
}


// ---------------------------------------------------------------------------------------
// Output drivers — output.hh:26-79, output.cc:18-76, haplotype_output.cc:22-132
// ---------------------------------------------------------------------------------------
struct output_delegate	// output.hh:26-38 (subprocess hooks omitted: no --pipe in the oracle)
{
	virtual ~output_delegate() {}
	virtual void will_handle_sample(std::string const &, u32, u32) {}
	virtual void will_handle_founder_sequence(u32) {}
	virtual void handled_sequences(u32) {}
	virtual void handled_node(u64) {}
};

// A "separate" output file as (name, bytes); output.cc:18-44 passes the file name as the
// FASTA identifier regardless of should_include_fasta_header.
using named_outputs = std::vector<std::pair<std::string, std::string>>;

struct output_base
{
	char const		*m_chromosome_id{};
	output_delegate	*m_delegate{};
	bool			m_should_output_reference{};
	bool			m_should_output_unaligned{};

	output_base(char const *chromosome_id, bool output_reference, bool output_unaligned, output_delegate &delegate):
		m_chromosome_id(chromosome_id), m_delegate(&delegate),
		m_should_output_reference(output_reference), m_should_output_unaligned(output_unaligned) {}

	void output_sequence_file(std::vector<char> const &ref_seq, variant_graph const &graph, std::string const &dst_name,
		sequence_writing_delegate &delegate, named_outputs &files) const
	{
		string_sink sink;
		output_sequence(ref_seq, graph, sink, dst_name.c_str(), m_should_output_unaligned, delegate);
		files.emplace_back(dst_name, std::move(sink.s));
	}

	std::string separate_name(std::string const &stem, bool const should_include_fasta_header) const
	{
		std::string n;
		if (m_chromosome_id) { n += m_chromosome_id; n += '.'; }
		n += stem;
		if (should_include_fasta_header) n += m_should_output_unaligned ? ".fa" : ".a2m";
		return n;
	}
};

struct haplotype_output final : output_base
{
	using output_base::output_base;

	// haplotype_output.cc:38-82
	void output_a2m(std::vector<char> const &ref_seq, variant_graph const &graph, byte_sink &stream) const
	{
		u32 seq_count(1);
		if (m_should_output_reference)
		{
			std::string id;
			if (m_chromosome_id) { id += m_chromosome_id; id += '\t'; }
			id += "REF";
			sequence_writing_delegate d;
			output_sequence(ref_seq, graph, stream, id.c_str(), m_should_output_unaligned, d);
			stream.put('\n');
			m_delegate->handled_sequences(seq_count);
		}
		for (u32 s(0); s < graph.sample_names.size(); ++s)
			for (u32 c(0), ploidy(graph.sample_ploidy(s)); c < ploidy; ++c)
			{
				m_delegate->will_handle_sample(graph.sample_names[s], s, c);
				std::string id;
				if (m_chromosome_id) { id += m_chromosome_id; id += '\t'; }
				id += graph.sample_names[s]; id += '-'; id += std::to_string(1 + c);
				sequence_writing_delegate d(graph.ploidy_csum[s] + c);
				output_sequence(ref_seq, graph, stream, id.c_str(), m_should_output_unaligned, d);
				stream.put('\n');
				++seq_count;
				m_delegate->handled_sequences(seq_count);
			}
	}

	// haplotype_output.cc:85-132
	void output_separate(std::vector<char> const &ref_seq, variant_graph const &graph, bool const should_include_fasta_header, named_outputs &files) const
	{
		if (m_should_output_reference)
		{
			sequence_writing_delegate d;
			output_sequence_file(ref_seq, graph, separate_name("REF", should_include_fasta_header), d, files);
		}
		for (u32 s(0); s < graph.sample_names.size(); ++s)
			for (u32 c(0), ploidy(graph.sample_ploidy(s)); c < ploidy; ++c)
			{
				m_delegate->will_handle_sample(graph.sample_names[s], s, c);
				sequence_writing_delegate d(graph.ploidy_csum[s] + c);
				output_sequence_file(ref_seq, graph, separate_name(graph.sample_names[s] + "." + std::to_string(1 + c), should_include_fasta_header), d, files);
			}
	}
};


// ---------------------------------------------------------------------------------------
// pbwt_context — include/vcf2multialign/pbwt.hh:22-145
// ---------------------------------------------------------------------------------------
struct pbwt_context
{
	static constexpr u64 DIVERGENCE_MAX = std::numeric_limits<u64>::max();

	// DIVERGENCE_MAX sorts first (pbwt.hh:44-45).
	struct div_less { bool operator()(u64 a, u64 b) const { return 1 + a < 1 + b; } };

	std::vector<u32>			permutation, prev_permutation;
	std::vector<u64>			divergence, prev_divergence;
	std::map<u64, u32, div_less>	divergence_value_counts;

	explicit pbwt_context(u32 const count): permutation(count), divergence(count, DIVERGENCE_MAX)
	{
		if (count)
		{
			divergence[0] = 0;
			divergence_value_counts[0] = 1;
			if (1 < count) divergence_value_counts[DIVERGENCE_MAX] = count - 1;
		}
		std::iota(permutation.begin(), permutation.end(), 0);
	}

	void swap_vectors()	// pbwt.hh:137-145
	{
		permutation.swap(prev_permutation);
		divergence.swap(prev_divergence);
		permutation.clear();
		divergence.clear();
	}

	// pbwt.hh:77-134; `column` = words of one edge column of paths_by_edge_and_chrom_copy.
	void update_divergence(u64 const *column, u64 const column_words, u64 const kk)
	{
		div_less const lt;
		u32 zero_idx(0), one_idx;
		{
			u32 ones(0);
			for (u64 w(0); w < column_words; ++w) ones += __builtin_popcountll(column[w]);
			one_idx = u32(prev_permutation.size()) - ones;
		}
		permutation.resize(prev_permutation.size());
		divergence.resize(prev_divergence.size());
		u64 pp(kk + 1), qq(kk + 1);
		for (u32 ii(0); ii < prev_permutation.size(); ++ii)
		{
			u32 const val_idx(prev_permutation[ii]);
			u64 const pd(prev_divergence[ii]);
			if (lt(pp, pd)) pp = pd;
			if (lt(qq, pd)) qq = pd;
			{
				auto const it(divergence_value_counts.find(pd));
				--it->second;
				if (0 == it->second) divergence_value_counts.erase(it);
			}
			if (0 == ((column[val_idx / 64] >> (val_idx % 64)) & 1))
			{
				++divergence_value_counts[pp];
				permutation[zero_idx] = val_idx; divergence[zero_idx] = pp; ++zero_idx; pp = 0;
			}
			else
			{
				++divergence_value_counts[qq];
				permutation[one_idx] = val_idx; divergence[one_idx] = qq; ++one_idx; qq = 0;
			}
		}
	}
};


// ---------------------------------------------------------------------------------------
// find_initial_cut_positions_lambda_min — libvcf2multialign/find_cut_positions.cc:93-211
// ---------------------------------------------------------------------------------------
inline u32 find_initial_cut_positions_lambda_min(
	variant_graph const &graph, u64 const min_distance, std::vector<u64> &out_cut_positions, output_delegate *delegate = nullptr)
{
	struct cut_position { u64 edge{}; u64 prev_edge{EDGE_MAX}; u64 node{}; u32 score{}; };
	auto update_if_needed([](cut_position &cur, u32 const eq_class_count, cut_position const &prev){	// :57-65
		u32 const cand(std::max(eq_class_count, prev.score));
		if (cand < cur.score) { cur.score = cand; cur.prev_edge = prev.edge; }
	});
	auto edge_lower_bound([](std::vector<cut_position> const &v, std::size_t const hi, u64 const edge){
		return std::size_t(std::lower_bound(v.begin(), v.begin() + hi, edge,
			[](cut_position const &c, u64 const e){ return c.edge < e; }) - v.begin());
	});

	out_cut_positions.clear();
	u32 const path_count(graph.total_chromosome_copies());
	u64 rightmost_target(0), edge_idx(0), prev_cut_pos_id(EDGE_MAX);
	pbwt_context pbwt(path_count);
	std::vector<cut_position> cuts;
	cuts.push_back({0, EDGE_MAX, 0, 0});
	u64 const colw(graph.paths_by_edge_and_chrom_copy.number_of_rows() / 64);

	for (u64 node(0); node < graph.node_count(); ++node)	// walker.advance() starts at node 0
	{
		if (rightmost_target <= node && prev_cut_pos_id != edge_idx)	// :126-129
		{
			cuts.push_back({edge_idx, EDGE_MAX, node, path_count});
			prev_cut_pos_id = edge_idx;
			std::size_t const cur(cuts.size() - 1);
			std::size_t rb(cuts.size());
			auto const &dvc(pbwt.divergence_value_counts);
			u32 eq_class_count(dvc.rbegin()->second);			// :137
			// All but the last (greatest) key, in descending order (:111-117, 138).
			if (dvc.size() > 1)
			{
				auto it(std::prev(dvc.end()));
				while (it != dvc.begin())
				{
					--it;
					std::size_t const pos(edge_lower_bound(cuts, rb, it->first));
					if (pos != rb)
					{
						rb = pos;
						if (min_distance <= graph.aligned_length(cuts[pos].node, node))
							update_if_needed(cuts[cur], eq_class_count, cuts[pos]);
					}
					eq_class_count += it->second;
				}
			}
			if (0 != rb)	// :161-165
			{
				--rb;
				update_if_needed(cuts[cur], eq_class_count, cuts[rb]);
			}
		}

		for (u64 e(graph.alt_edge_count_csum[node]); e < graph.alt_edge_count_csum[node + 1]; ++e)	// :170-176
		{
			pbwt.swap_vectors();
			pbwt.update_divergence(graph.paths_by_edge_and_chrom_copy.column_words(edge_idx), colw, edge_idx);
			++edge_idx;
			rightmost_target = std::max(rightmost_target, graph.alt_edge_targets[e]);
		}
		if (delegate) delegate->handled_node(node);
	}

	if (cuts.size() <= 1) return CUT_POSITION_SCORE_MAX;	// :182-183

	std::size_t it(cuts.size() - 1);
	u32 const retval(cuts[it].score);
	while (true)
	{
		out_cut_positions.push_back(cuts[it].node);
		u64 const prev_edge(cuts[it].prev_edge);
		if (EDGE_MAX == prev_edge) break;
		it = edge_lower_bound(cuts, it, prev_edge);
	}
	if (0 != out_cut_positions.back()) out_cut_positions.push_back(0);
	std::reverse(out_cut_positions.begin(), out_cut_positions.end());
	ORACLE_ASSERT(out_cut_positions.back() < graph.node_count());
	if (out_cut_positions.back() != graph.node_count() - 1)
		out_cut_positions.back() = graph.node_count() - 1;
	return retval;
}


// ---------------------------------------------------------------------------------------
// founder_sequence_greedy_output — libvcf2multialign/founder_sequence_greedy_output.cc
// ---------------------------------------------------------------------------------------
struct joined_path_eq_class	// founder_…cc:48-69
{
	u32 lhs_rep{}, rhs_rep{}, size{};
	bool operator<(joined_path_eq_class const &o) const { return size < o.size; }
};

// Per-cut class data as the reference derives it from (permutation, divergence)
// (founder_…cc:213-253). This struct is also the interface the GPU kernel K3 fills.
struct cut_class_data
{
	std::vector<u32>					rhs_eq_classes;			// representative per chromosome copy
	u32									rhs_distinct_eq_classes{};
	u32									rhs_first_path_eq_class{};
	bool								rhs_first_path_is_ref{true};	// over the segment *ending* at this cut
	std::vector<joined_path_eq_class>	joined;					// unsorted, in permutation order
};

// lb::matrix<u32>, column-major (lib/libbio/include/libbio/matrix/matrix.hh:63-68).
struct ploidy_matrix
{
	std::vector<u32> data; u64 rows{}, cols{};
	void resize(u64 r, u64 c) { rows = r; cols = c; data.assign(r * c, 0); }
	u32 &operator()(u64 y, u64 x) { return data[x * rows + y]; }
	u32 operator()(u64 y, u64 x) const { return data[x * rows + y]; }
};

// The greedy matcher proper (founder_…cc:255-448, 474-509), factored so that the class data
// can come either from the pBWT sweep below (oracle) or from elsewhere (cross-checks).
struct greedy_matcher
{
	u32 founder_count;
	bool keep_ref_edges;
	ploidy_matrix &assigned;
	std::multimap<u32, u32> assignments_by_eq_class;
	std::vector<char> reserved;
	std::vector<u32> arbitrarily_connected_rhs;

	greedy_matcher(u32 total_copies, u32 founders, bool keep, ploidy_matrix &out):
		founder_count(founders), keep_ref_edges(keep), assigned(out), reserved(total_copies, 0) {}

	// Called at cut index cut_pos_idx >= 1 with `joined` in permutation order.
	void handle_cut(
		u64 const cut_pos_idx, std::vector<joined_path_eq_class> &joined,
		u32 const lhs_distinct, u32 const rhs_distinct,
		bool const lhs_first_is_ref, bool const rhs_first_is_ref,
		u32 const lhs_first_class, u32 const rhs_first_class)
	{
		std::sort(joined.begin(), joined.end());	// :258
		if (!keep_ref_edges && lhs_first_is_ref && rhs_first_is_ref)	// :261-266
			std::erase_if(joined, [&](auto const &c){ return lhs_first_class == c.lhs_rep && rhs_first_class == c.rhs_rep; });

		if (1 == cut_pos_idx)	// :268-320
		{
			u32 remaining_founders(founder_count);
			u32 remaining_reserved(std::min(remaining_founders, lhs_distinct));
			remaining_founders -= remaining_reserved;
			u32 founder_idx(0);
			auto do_assign([&](joined_path_eq_class const &c){
				assignments_by_eq_class.emplace(c.lhs_rep, founder_idx);
				assigned(0, founder_idx) = c.lhs_rep;
				++founder_idx;
			});
			for (auto it(joined.rbegin()); it != joined.rend(); ++it)
			{
				if (reserved[it->lhs_rep])
				{
					if (remaining_founders) { --remaining_founders; do_assign(*it); }
				}
				else if (remaining_reserved)
				{
					--remaining_reserved;
					reserved[it->lhs_rep] = 1;
					do_assign(*it);
				}
			}
			// The reference loops forever here when `joined` is empty and founders remain (:309-319);
			// the oracle reports it instead of hanging.
			while (remaining_founders)
			{
				if (joined.empty()) throw oracle_error("find_matchings: reference would not terminate (empty joined classes)");
				for (auto it(joined.rbegin()); it != joined.rend(); ++it)
				{
					if (!remaining_founders) break;
					--remaining_founders;
					do_assign(*it);
				}
			}
		}

		// Subsequent assignment (:331-448).
		std::fill(reserved.begin(), reserved.end(), 0);
		arbitrarily_connected_rhs.clear();
		u32 remaining_founders(founder_count);
		u32 remaining_reserved(std::min(remaining_founders, rhs_distinct));
		remaining_founders -= remaining_reserved;

		auto try_assign([&](joined_path_eq_class const &c) -> bool {
			auto const it(assignments_by_eq_class.find(c.lhs_rep));
			if (assignments_by_eq_class.end() == it) return false;
			u32 const founder_idx(it->second);
			assignments_by_eq_class.erase(it);
			assigned(cut_pos_idx, founder_idx) = c.rhs_rep;
			return true;
		});
		auto assign_arbitrary([&](u32 const rhs_rep){
			ORACLE_ASSERT(!assignments_by_eq_class.empty());
			auto const it(assignments_by_eq_class.begin());
			u32 const founder_idx(it->second);
			assignments_by_eq_class.erase(it);
			assigned(cut_pos_idx, founder_idx) = rhs_rep;
		});

		{
			bool is_first(true), stop(false);
			while (!stop)
			{
				bool did_assign(false);
				for (auto it(joined.rbegin()); it != joined.rend(); ++it)
				{
					if (reserved[it->rhs_rep])
					{
						if (remaining_founders)
						{
							if (try_assign(*it)) { did_assign = true; --remaining_founders; }
						}
						else if (!is_first) { stop = true; break; }	// goto continue_subsequent_assignment
					}
					else if (remaining_reserved)
					{
						--remaining_reserved;
						if (try_assign(*it)) reserved[it->rhs_rep] = 1;
						else arbitrarily_connected_rhs.push_back(it->rhs_rep);
					}
				}
				if (stop) break;
				if (!remaining_founders) break;
				if (is_first) { is_first = false; continue; }
				if (!did_assign) break;
			}
		}

		for (u32 const rhs_rep : arbitrarily_connected_rhs)	// step 4
			if (!reserved[rhs_rep]) { assign_arbitrary(rhs_rep); reserved[rhs_rep] = 1; }

		while (!assignments_by_eq_class.empty())			// step 5
		{
			if (joined.empty()) throw oracle_error("find_matchings: reference would not terminate (step 5, empty joined classes)");
			for (auto it(joined.rbegin()); it != joined.rend(); ++it)
			{
				if (assignments_by_eq_class.empty()) break;
				assign_arbitrary(it->rhs_rep);
			}
		}

		assignments_by_eq_class.clear();					// finish_assignment
		for (u32 f(0); f < assigned.cols; ++f)
			assignments_by_eq_class.insert({assigned(cut_pos_idx, f), f});
	}
};

struct founder_sequence_greedy_output final : output_base
{
	std::vector<u64>	m_cut_positions;
	u64					m_min_distance{};
	u32					m_score{};
	ploidy_matrix		m_assigned_samples;
	bool				m_should_keep_ref_edges{};
	// Recorded per-cut class data (one entry per cut position index ≥ 0 reached), for cross-checking K3.
	std::vector<cut_class_data>	m_recorded_classes;
	bool				m_record_classes{};

	founder_sequence_greedy_output(char const *chromosome_id, bool output_reference, bool keep_ref_edges, bool output_unaligned, output_delegate &delegate):
		output_base(chromosome_id, output_reference, output_unaligned, delegate), m_should_keep_ref_edges(keep_ref_edges) {}

	bool find_cut_positions(variant_graph const &graph, u64 const min_dist)	// founder_…cc:139-151
	{
		u32 const score(find_initial_cut_positions_lambda_min(graph, min_dist, m_cut_positions, m_delegate));
		if (CUT_POSITION_SCORE_MAX == score) return false;
		m_min_distance = min_dist; m_score = score;
		return true;
	}

	// founder_…cc:154-512
	bool find_matchings(variant_graph const &graph, u32 const founder_count)
	{
		if (m_cut_positions.size() < 2) return false;
		u32 const H(graph.total_chromosome_copies());
		if (0 == H) return false;
		ORACLE_ASSERT(0 == m_cut_positions.front());

		m_assigned_samples.resize(m_cut_positions.size() - 1, founder_count);
		std::fill(m_assigned_samples.data.begin(), m_assigned_samples.data.end(), PLOIDY_MAX);
		m_recorded_classes.clear();

		greedy_matcher matcher(H, founder_count, m_should_keep_ref_edges, m_assigned_samples);

		u64 edge_idx(0), prev_cut_edge_idx(0), cut_pair_edge_idx(0);
		std::vector<u32> lhs_eq(H, PLOIDY_MAX), rhs_eq(H, PLOIDY_MAX);
		u32 lhs_distinct(0), rhs_distinct(0);
		std::vector<joined_path_eq_class> joined;
		bool lhs_first_is_ref(true), rhs_first_is_ref(true);
		u32 lhs_first_class(0), rhs_first_class(0);
		std::size_t cut_it(1);
		pbwt_context pbwt(H);
		u64 const colw(graph.paths_by_edge_and_chrom_copy.number_of_rows() / 64);
		u64 cut_pos_idx(0);

		for (u64 node(0); node < graph.node_count(); ++node)
		{
			ORACLE_ASSERT(cut_it < m_cut_positions.size());
			ORACLE_ASSERT(node <= m_cut_positions[cut_it]);
			if (node == m_cut_positions[cut_it])
			{
				lhs_eq.swap(rhs_eq);
				std::fill(rhs_eq.begin(), rhs_eq.end(), PLOIDY_MAX);
				lhs_distinct = rhs_distinct;
				lhs_first_class = rhs_first_class;
				rhs_distinct = 0;
				rhs_first_class = pbwt.permutation.front();

				u32 rep(PLOIDY_MAX);
				joined.clear();
				for (u32 i(0); i < H; ++i)
				{
					u32 const aa(pbwt.permutation[i]);
					u64 const dd(pbwt.divergence[i]);
					// u64 < divergence_value resolves to the built-in comparison through the implicit conversion
					// (pbwt.hh:46), not to divergence_value::operator<, so DIVERGENCE_MAX compares greatest here.
					if (prev_cut_edge_idx < dd) { rep = aa; ++rhs_distinct; }
					rhs_eq[aa] = rep;
					if (0 < cut_pos_idx)
					{
						if (cut_pair_edge_idx < dd) joined.push_back({lhs_eq[aa], rep, 0});
						ORACLE_ASSERT(!joined.empty());
						++joined.back().size;
					}
				}

				if (m_record_classes)
					m_recorded_classes.push_back({rhs_eq, rhs_distinct, rhs_first_class, rhs_first_is_ref, joined});

				if (0 < cut_pos_idx)
					matcher.handle_cut(cut_pos_idx, joined, lhs_distinct, rhs_distinct, lhs_first_is_ref, rhs_first_is_ref, lhs_first_class, rhs_first_class);

				++cut_pos_idx; ++cut_it;
				cut_pair_edge_idx = prev_cut_edge_idx;
				prev_cut_edge_idx = edge_idx;
				lhs_first_is_ref = rhs_first_is_ref;
				rhs_first_is_ref = true;
			}

			for (u64 e(graph.alt_edge_count_csum[node]); e < graph.alt_edge_count_csum[node + 1]; ++e)	// :460-469
			{
				pbwt.swap_vectors();
				pbwt.update_divergence(graph.paths_by_edge_and_chrom_copy.column_words(edge_idx), colw, edge_idx);
				rhs_first_is_ref &= !graph.paths_by_edge_and_chrom_copy.get(pbwt.permutation.front(), edge_idx);
				++edge_idx;
			}
			m_delegate->handled_node(node);
			if (cut_it == m_cut_positions.size()) break;	// sink reached: walker.advance() would end here too
		}

		if (1 == cut_pos_idx)	// trivial single-segment case (:474-509)
		{
			u32 rep(PLOIDY_MAX);
			joined.clear();
			for (u32 i(0); i < H; ++i)
			{
				u32 const aa(pbwt.permutation[i]);
				if (0 < pbwt.divergence[i]) { rep = aa; ++rhs_distinct; joined.push_back({PLOIDY_MAX, rep, 0}); }
				rhs_eq[aa] = rep;
				ORACLE_ASSERT(!joined.empty());
				++joined.back().size;
			}
			std::sort(joined.begin(), joined.end());
			if (!m_should_keep_ref_edges && rhs_first_is_ref)
				std::erase_if(joined, [&](auto const &c){ return rhs_first_class == c.rhs_rep; });
			u32 f(0);
			for (auto it(joined.rbegin()); it != joined.rend() && f < founder_count; ++it, ++f)
				m_assigned_samples(0, f) = it->rhs_rep;
		}
		return true;
	}

	// founder_…cc:78-115
	struct founder_delegate final : sequence_writing_delegate
	{
		ploidy_matrix const &assigned; u64 founder; std::vector<u64> const &cuts; u64 cut_pos_index{};
		founder_delegate(ploidy_matrix const &a, u64 f, std::vector<u64> const &c): assigned(a), founder(f), cuts(c)
		{
			ORACLE_ASSERT(!cuts.empty());
			ORACLE_ASSERT(0 == cuts.front());
		}
		void handle_node(variant_graph const &, u64 const node) override
		{
			ORACLE_ASSERT(node <= cuts[cut_pos_index]);
			if (node == cuts[cut_pos_index])
			{
				chromosome_copy_index = assigned(cut_pos_index, founder);
				++cut_pos_index;
			}
		}
	};

	// founder_…cc:515-550
	void output_a2m(std::vector<char> const &ref_seq, variant_graph const &graph, byte_sink &stream) const
	{
		if (m_should_output_reference)
		{
			std::string id;
			if (m_chromosome_id) { id += m_chromosome_id; id += '\t'; }
			id += "REF";
			sequence_writing_delegate d;
			output_sequence(ref_seq, graph, stream, id.c_str(), m_should_output_unaligned, d);
			stream.put('\n');
			m_delegate->handled_sequences(1);
		}
		for (u32 col(0); col < m_assigned_samples.cols; ++col)
		{
			m_delegate->will_handle_founder_sequence(col);
			std::string id;
			if (m_chromosome_id) { id += m_chromosome_id; id += '\t'; }
			id += std::to_string(1 + col);
			founder_delegate d(m_assigned_samples, col, m_cut_positions);
			output_sequence(ref_seq, graph, stream, id.c_str(), m_should_output_unaligned, d);
			stream.put('\n');
			m_delegate->handled_sequences(2 + col);
		}
	}

	// founder_…cc:553-597
	void output_separate(std::vector<char> const &ref_seq, variant_graph const &graph, bool const should_include_fasta_header, named_outputs &files) const
	{
		if (m_should_output_reference)
		{
			sequence_writing_delegate d;
			output_sequence_file(ref_seq, graph, separate_name("REF", should_include_fasta_header), d, files);
		}
		for (u32 col(0); col < m_assigned_samples.cols; ++col)
		{
			m_delegate->will_handle_founder_sequence(col);
			founder_delegate d(m_assigned_samples, col, m_cut_positions);
			output_sequence_file(ref_seq, graph, separate_name(std::to_string(1 + col), should_include_fasta_header), d, files);
		}
	}
};

} // namespace v2m_oracle
