// ORACLE — TEST INFRASTRUCTURE ONLY. C entry points over v2m_oracle.hh for ctypes
// (tests/, smoke(), bench.py cpu_baseline / --impl reference). Not part of the product.
#include "v2m_oracle.hh"

#include <atomic>
#include <chrono>
#include <memory>
#include <thread>

using namespace v2m_oracle;

namespace {

	thread_local std::string g_error;

	struct graph_box
	{
		variant_graph		graph;
		std::vector<char>	ref_seq;
		build_delegate		build;
		build_trace			trace;
		u64					handled_variants{};
		u64					chr_id_mismatches{};
		// flattened labels
		std::vector<u64>	label_offsets;
		std::string			label_chars;
		std::string			sample_names_joined;

		void flatten()
		{
			label_offsets.assign(1, 0);
			label_chars.clear();
			for (auto const &l : graph.alt_edge_labels) { label_chars += l; label_offsets.push_back(label_chars.size()); }
			sample_names_joined.clear();
			for (auto const &s : graph.sample_names) { sample_names_joined += s; sample_names_joined += '\n'; }
		}
	};

	struct founder_box
	{
		output_delegate delegate;
		std::string chrom;
		std::unique_ptr<founder_sequence_greedy_output> out;
	};

	template <typename t_fn>
	int guarded(t_fn &&fn)
	{
		try { fn(); return 0; }
		catch (std::exception const &e) { g_error = e.what(); return 1; }
		catch (...) { g_error = "unknown error"; return 1; }
	}

	long long copy_out(std::string const &s, char *dst, unsigned long long cap)
	{
		if (dst && s.size() <= cap) std::memcpy(dst, s.data(), s.size());
		return (long long) s.size();
	}
}

extern "C" {

char const *orc_last_error() { return g_error.c_str(); }

// ---- transpose ----
int orc_transpose(u64 const *src, u64 rows, u64 cols, u64 *dst)
{
	return guarded([&]{
		bit_matrix m(rows, cols);
		std::memcpy(m.words.data(), src, m.words.size() * 8);
		auto const t(transpose_matrix(m));
		std::memcpy(dst, t.words.data(), t.words.size() * 8);
	});
}

u64 orc_transpose8x8(u64 w, int terms) { return terms ? transpose8x8_terms(w) : transpose8x8(w); }

// ---- graph ----
void *orc_graph_from_files(char const *fasta_path, char const *vcf_path, char const *chr_id)
{
	auto box(std::make_unique<graph_box>());
	if (guarded([&]{
		read_single_fasta_sequence(fasta_path, box->ref_seq, nullptr);
		auto const vf(read_vcf(vcf_path));
		box->build.trace = &box->trace;
		build_variant_graph(box->ref_seq, vf, chr_id, box->graph, box->build, &box->handled_variants, &box->chr_id_mismatches);
		box->flatten();
	})) return nullptr;
	return box.release();
}

// The genotype block behind the graph's path bits (build_trace): dims6 = {records, copies, alts, edges, overlap pairs, 0}.
void orc_graph_trace_dims(void *h, u64 *dims6)
{
	auto const &t(static_cast<graph_box *>(h)->trace);
	dims6[0] = t.ref_pos.size(); dims6[1] = t.copies; dims6[2] = t.edges_by_alt.size();
	dims6[3] = t.edge_target_pos.size(); dims6[4] = t.overlaps.size() / 2; dims6[5] = 0;
}

void orc_graph_trace_copy(void *h, std::uint16_t *gt, u64 *alt_offsets, u64 *edges_by_alt, u64 *ref_pos, u64 *edge_target_pos, u64 *overlaps)
{
	auto const &t(static_cast<graph_box *>(h)->trace);
	std::copy(t.gt.begin(), t.gt.end(), gt);
	std::copy(t.alt_offsets.begin(), t.alt_offsets.end(), alt_offsets);
	std::copy(t.edges_by_alt.begin(), t.edges_by_alt.end(), edges_by_alt);
	std::copy(t.ref_pos.begin(), t.ref_pos.end(), ref_pos);
	std::copy(t.edge_target_pos.begin(), t.edge_target_pos.end(), edge_target_pos);
	std::copy(t.overlaps.begin(), t.overlaps.end(), overlaps);
}

// Graph from flat arrays (synthetic workloads). paths_ec = words of paths_by_edge_and_chrom_copy
// (rows = h_pad chromosome copies, cols = e_pad edges); the other orientation is derived with the oracle transpose
// unless paths_ce is given.
void *orc_graph_from_arrays(
	u64 node_count, u64 edge_count,
	u64 const *refpos, u64 const *alnpos, u64 const *targets, u64 const *csum,
	u64 const *label_offsets, char const *label_chars,
	u64 const *paths_ec, u64 const *paths_ce, u64 h_pad, u64 e_pad,
	u32 sample_count, u32 const *ploidy_csum, char const *sample_names_joined,
	char const *ref, u64 ref_len)
{
	auto box(std::make_unique<graph_box>());
	if (guarded([&]{
		auto &g(box->graph);
		g.reference_positions.assign(refpos, refpos + node_count);
		g.aligned_positions.assign(alnpos, alnpos + node_count);
		g.alt_edge_targets.assign(targets, targets + edge_count);
		g.alt_edge_count_csum.assign(csum, csum + node_count + 1);
		g.alt_edge_labels.resize(edge_count);
		for (u64 e(0); e < edge_count; ++e)
			g.alt_edge_labels[e].assign(label_chars + label_offsets[e], label_chars + label_offsets[e + 1]);
		g.paths_by_edge_and_chrom_copy = bit_matrix(h_pad, e_pad);
		std::memcpy(g.paths_by_edge_and_chrom_copy.words.data(), paths_ec, h_pad * e_pad / 8);
		if (paths_ce)
		{
			g.paths_by_chrom_copy_and_edge = bit_matrix(e_pad, h_pad);
			std::memcpy(g.paths_by_chrom_copy_and_edge.words.data(), paths_ce, h_pad * e_pad / 8);
		}
		else
			g.paths_by_chrom_copy_and_edge = transpose_matrix(g.paths_by_edge_and_chrom_copy);
		g.ploidy_csum.assign(ploidy_csum, ploidy_csum + sample_count + 1);
		{
			std::string_view names(sample_names_joined ? sample_names_joined : "");
			for (u32 s(0); s < sample_count; ++s)
			{
				auto const e(names.find('\n'));
				if (std::string_view::npos == e) { g.sample_names.emplace_back(names.empty() ? "S" + std::to_string(s) : std::string(names)); names = {}; }
				else { g.sample_names.emplace_back(names.substr(0, e)); names.remove_prefix(e + 1); }
			}
		}
		box->ref_seq.assign(ref, ref + ref_len);
		box->flatten();
	})) return nullptr;
	return box.release();
}

void orc_graph_free(void *h) { delete static_cast<graph_box *>(h); }

// dims: [0]=node_count [1]=edge_count [2]=h_pad (rows of edge×copy matrix) [3]=e_pad (its cols)
// [4]=total copies [5]=sample count [6]=label chars [7]=ref length [8]=overlap reports [9]=handled variants
// [10]=rows of copy×edge matrix [11]=cols of copy×edge matrix [12]=sample name bytes [13]=ref mismatches
void orc_graph_dims(void *h, u64 *dims)
{
	auto &b(*static_cast<graph_box *>(h));
	auto &g(b.graph);
	dims[0] = g.node_count(); dims[1] = g.edge_count();
	dims[2] = g.paths_by_edge_and_chrom_copy.number_of_rows(); dims[3] = g.paths_by_edge_and_chrom_copy.number_of_columns();
	dims[4] = g.total_chromosome_copies(); dims[5] = g.sample_names.size();
	dims[6] = b.label_chars.size(); dims[7] = b.ref_seq.size(); dims[8] = b.build.overlaps.size(); dims[9] = b.handled_variants;
	dims[10] = g.paths_by_chrom_copy_and_edge.number_of_rows(); dims[11] = g.paths_by_chrom_copy_and_edge.number_of_columns();
	dims[12] = b.sample_names_joined.size(); dims[13] = b.build.ref_mismatches.size();
}

void orc_graph_copy(
	void *h, u64 *refpos, u64 *alnpos, u64 *targets, u64 *csum, u64 *label_offsets, char *label_chars,
	u64 *paths_ec, u64 *paths_ce, u32 *ploidy_csum, char *sample_names_joined, char *ref)
{
	auto &b(*static_cast<graph_box *>(h));
	auto &g(b.graph);
	auto cp([](auto const &v, auto *dst){ if (dst && !v.empty()) std::memcpy(dst, v.data(), v.size() * sizeof(v[0])); });
	cp(g.reference_positions, refpos); cp(g.aligned_positions, alnpos); cp(g.alt_edge_targets, targets);
	cp(g.alt_edge_count_csum, csum); cp(b.label_offsets, label_offsets); cp(b.label_chars, label_chars);
	cp(g.paths_by_edge_and_chrom_copy.words, paths_ec); cp(g.paths_by_chrom_copy_and_edge.words, paths_ce);
	cp(g.ploidy_csum, ploidy_csum); cp(b.sample_names_joined, sample_names_joined); cp(b.ref_seq, ref);
}

// Overlap report i: out[0]=lineno out[1]=ref_pos out[2]=chrom copy out[3]=gt; text = "var_id\tsample".
long long orc_graph_overlap(void *h, u64 i, u64 *out, char *text, u64 cap)
{
	auto &b(*static_cast<graph_box *>(h));
	auto const &o(b.build.overlaps.at(i));
	out[0] = o.lineno; out[1] = o.ref_pos; out[2] = o.chrom_copy; out[3] = o.gt;
	return copy_out(o.var_id + "\t" + o.sample, text, cap);
}

// ---- rows ----
// One row via output_sequence (sequence_writer.cc:22-85). copy_idx = 0xffffffff → REF.
long long orc_output_sequence(void *h, u32 copy_idx, char const *fasta_id, int unaligned, char *dst, u64 cap)
{
	long long n(-1);
	guarded([&]{
		auto &b(*static_cast<graph_box *>(h));
		string_sink sink;
		sequence_writing_delegate d(copy_idx);
		output_sequence(b.ref_seq, b.graph, sink, fasta_id, unaligned, d);
		n = copy_out(sink.s, dst, cap);
	});
	return n;
}

// haplotype_output::output_a2m into memory (haplotype_output.cc:38-82).
long long orc_haplotype_output_a2m(void *h, char const *chrom_id, int output_ref, int unaligned, char *dst, u64 cap)
{
	long long n(-1);
	guarded([&]{
		auto &b(*static_cast<graph_box *>(h));
		output_delegate od;
		haplotype_output out(chrom_id, output_ref, unaligned, od);
		string_sink sink;
		out.output_a2m(b.ref_seq, b.graph, sink);
		n = copy_out(sink.s, dst, cap);
	});
	return n;
}

// Separate outputs serialised as: name '\0' u64 length (little endian) bytes, repeated.
static std::string pack_files(named_outputs const &files)
{
	std::string s;
	for (auto const &[name, bytes] : files)
	{
		s += name; s += '\0';
		u64 const len(bytes.size());
		s.append(reinterpret_cast<char const *>(&len), 8);
		s += bytes;
	}
	return s;
}

long long orc_haplotype_output_separate(void *h, char const *chrom_id, int output_ref, int unaligned, int include_header, char *dst, u64 cap)
{
	long long n(-1);
	guarded([&]{
		auto &b(*static_cast<graph_box *>(h));
		output_delegate od;
		haplotype_output out(chrom_id, output_ref, unaligned, od);
		named_outputs files;
		out.output_separate(b.ref_seq, b.graph, include_header, files);
		n = copy_out(pack_files(files), dst, cap);
	});
	return n;
}

// ---- founders ----
void *orc_founder_new(char const *chrom_id, int output_ref, int keep_ref_edges, int unaligned)
{
	auto fb(std::make_unique<founder_box>());
	if (chrom_id) fb->chrom = chrom_id;
	fb->out = std::make_unique<founder_sequence_greedy_output>(chrom_id ? fb->chrom.c_str() : nullptr, output_ref, keep_ref_edges, unaligned, fb->delegate);
	fb->out->m_record_classes = true;
	return fb.release();
}

void orc_founder_free(void *f) { delete static_cast<founder_box *>(f); }

// Returns 1 on success, 0 if the reference would return false, -1 on error.
int orc_founder_find_cut_positions(void *f, void *h, u64 min_distance)
{
	int res(-1);
	guarded([&]{ res = static_cast<founder_box *>(f)->out->find_cut_positions(static_cast<graph_box *>(h)->graph, min_distance) ? 1 : 0; });
	return res;
}

void orc_founder_set_cut_positions(void *f, u64 const *cuts, u64 n)
{
	static_cast<founder_box *>(f)->out->m_cut_positions.assign(cuts, cuts + n);
}

u64 orc_founder_cut_count(void *f) { return static_cast<founder_box *>(f)->out->m_cut_positions.size(); }
u32 orc_founder_score(void *f) { return static_cast<founder_box *>(f)->out->m_score; }
void orc_founder_cuts(void *f, u64 *dst)
{
	auto const &c(static_cast<founder_box *>(f)->out->m_cut_positions);
	std::memcpy(dst, c.data(), c.size() * 8);
}

int orc_founder_find_matchings(void *f, void *h, u32 founder_count)
{
	int res(-1);
	guarded([&]{ res = static_cast<founder_box *>(f)->out->find_matchings(static_cast<graph_box *>(h)->graph, founder_count) ? 1 : 0; });
	return res;
}

// Column-major segments × founders (lb::matrix layout).
void orc_founder_assigned(void *f, u64 *rows, u64 *cols, u32 *dst)
{
	auto const &m(static_cast<founder_box *>(f)->out->m_assigned_samples);
	*rows = m.rows; *cols = m.cols;
	if (dst) std::memcpy(dst, m.data.data(), m.data.size() * 4);
}

void orc_founder_set_assigned(void *f, u64 rows, u64 cols, u32 const *src)
{
	auto &m(static_cast<founder_box *>(f)->out->m_assigned_samples);
	m.resize(rows, cols);
	std::memcpy(m.data.data(), src, rows * cols * 4);
}

// Recorded class data at cut index i (0-based over reached cuts ≥ 1st non-zero cut):
// meta[0]=distinct meta[1]=first class meta[2]=first_is_ref (of the segment ending here) meta[3]=joined count
u64 orc_founder_recorded_cuts(void *f) { return static_cast<founder_box *>(f)->out->m_recorded_classes.size(); }
void orc_founder_recorded(void *f, u64 i, u32 *meta, u32 *rhs_eq, u32 *joined /* triples */)
{
	auto const &c(static_cast<founder_box *>(f)->out->m_recorded_classes.at(i));
	meta[0] = c.rhs_distinct_eq_classes; meta[1] = c.rhs_first_path_eq_class; meta[2] = c.rhs_first_path_is_ref; meta[3] = u32(c.joined.size());
	if (rhs_eq) std::memcpy(rhs_eq, c.rhs_eq_classes.data(), c.rhs_eq_classes.size() * 4);
	if (joined)
		for (std::size_t k(0); k < c.joined.size(); ++k)
		{
			joined[3 * k] = c.joined[k].lhs_rep; joined[3 * k + 1] = c.joined[k].rhs_rep; joined[3 * k + 2] = c.joined[k].size;
		}
}

long long orc_founder_output_a2m(void *f, void *h, char *dst, u64 cap)
{
	long long n(-1);
	guarded([&]{
		auto &b(*static_cast<graph_box *>(h));
		string_sink sink;
		static_cast<founder_box *>(f)->out->output_a2m(b.ref_seq, b.graph, sink);
		n = copy_out(sink.s, dst, cap);
	});
	return n;
}

long long orc_founder_output_separate(void *f, void *h, int include_header, char *dst, u64 cap)
{
	long long n(-1);
	guarded([&]{
		auto &b(*static_cast<graph_box *>(h));
		named_outputs files;
		static_cast<founder_box *>(f)->out->output_separate(b.ref_seq, b.graph, include_header, files);
		n = copy_out(pack_files(files), dst, cap);
	});
	return n;
}

// Greedy matcher alone, fed with externally computed class data (cross-check of K3 + host driver).
// Not used by the product; lets tests confirm that identical class data ⇒ identical assignments.

// ---- CPU baseline timing (bench.py only) ----
// Emits `n_rows` haplotype rows (copy indices first_copy, first_copy + 1, …; wraps) on `threads` threads into
// null sinks; mode 0 = faithful (4 KiB buffer, per-char gaps), 1 = bulk. Returns seconds; bases_out = rows × A.
double orc_time_rows(void *h, u32 first_copy, u32 n_rows, int threads, int mode, u64 *bases_out)
{
	auto &b(*static_cast<graph_box *>(h));
	auto const H(b.graph.total_chromosome_copies());
	u64 const A(b.graph.aligned_positions.back());
	std::atomic<u32> next(0);
	auto worker([&]{
		faithful_null_sink fs;
		bulk_null_sink bs(1u << 20);
		while (true)
		{
			u32 const i(next.fetch_add(1));
			if (i >= n_rows) break;
			sequence_writing_delegate d((first_copy + i) % H);
			std::string const id("S-" + std::to_string(i));
			if (0 == mode) { output_sequence(b.ref_seq, b.graph, fs, id.c_str(), false, d); fs.put('\n'); fs.flush(); }
			else { output_sequence(b.ref_seq, b.graph, bs, id.c_str(), false, d); bs.put('\n'); bs.flush(); }
		}
	});
	auto const t0(std::chrono::steady_clock::now());
	if (threads <= 1) worker();
	else
	{
		std::vector<std::thread> pool;
		for (int t(0); t < threads; ++t) pool.emplace_back(worker);
		for (auto &t : pool) t.join();
	}
	auto const t1(std::chrono::steady_clock::now());
	if (bases_out) *bases_out = u64(n_rows) * A;
	return std::chrono::duration<double>(t1 - t0).count();
}

double orc_time_transpose(u64 const *src, u64 rows, u64 cols, u64 *dst)
{
	bit_matrix m(rows, cols);
	std::memcpy(m.words.data(), src, m.words.size() * 8);
	auto const t0(std::chrono::steady_clock::now());
	auto const t(transpose_matrix(m));
	auto const t1(std::chrono::steady_clock::now());
	if (dst) std::memcpy(dst, t.words.data(), t.words.size() * 8);
	return std::chrono::duration<double>(t1 - t0).count();
}

// The "Paths" loop of build_variant_graph (variant_graph.cc:379-425, as restated in v2m_oracle.hh) alone, on a flat genotype
// block: the CPU baseline of the device's genotype ingestion. One thread, as the reference. Returns seconds.
double orc_time_paths_loop(std::uint16_t const *gt, u64 n_records, u32 copies, u64 const *alt_offsets, u64 const *edges_by_alt,
	u64 const *ref_pos, u64 const *edge_target_pos, u64 h_pad, u64 e_pad, u64 *dst_words, u64 *n_overlaps)
{
	bit_matrix m(h_pad, e_pad);
	std::vector<u64> target(copies, 0);
	u64 overlaps(0);
	auto const t0(std::chrono::steady_clock::now());
	for (u64 r(0); r < n_records; ++r)
	{
		u64 const a0(alt_offsets[r]), n_alts(alt_offsets[r + 1] - a0);
		std::uint16_t const *row(gt + r * copies);
		for (u32 h(0); h < copies; ++h)
		{
			auto const alt_idx(row[h]);
			if (0 == alt_idx) continue;
			if (NULL_ALLELE == alt_idx) continue;
			if (!(u64(alt_idx - 1) < n_alts)) continue;
			u64 const edge_idx(edges_by_alt[a0 + alt_idx - 1]);
			if (EDGE_MAX == edge_idx) continue;
			if (ref_pos[r] < target[h]) ++overlaps;
			target[h] = edge_target_pos[edge_idx];
			m.set(h, edge_idx);
		}
	}
	auto const t1(std::chrono::steady_clock::now());
	if (dst_words) std::memcpy(dst_words, m.words.data(), m.words.size() * 8);
	if (n_overlaps) *n_overlaps = overlaps;
	return std::chrono::duration<double>(t1 - t0).count();
}

} // extern "C"

// ---- measurement infrastructure: CPU mirror of the synthetic path-bit generator (vcf2multialign_b200/csrc/synth.cu,
// synth.py:allele_numpy) — not a restatement of the reference. bench.py's CPU legs use it to make the copy-major path
// lines of a few sampled chromosome copies of the big shapes without touching the GPU library.
namespace {
	inline u32 synth_hash32(u32 x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }
	inline u32 synth_mix(u32 seed, u32 a, u32 b) { return synth_hash32(synth_hash32(a ^ seed) + b); }
	inline u32 synth_allele(u32 seed, u32 h, u32 site, u32 n_alts, u32 thr)
	{
		u32 const off(synth_mix(seed, h, 23) & 8191u);
		u32 const block((site + off) >> 13);
		u32 const founder(synth_mix(seed, h * 0x9e3779b1u + block, 37) & 63u);
		u32 const fh(synth_mix(seed, site * 64u + founder, 41));
		bool const nonref((fh & 0xffffu) < thr);
		u32 alt(n_alts > 1 ? (fh >> 16) & 1u : 0u);
		bool const mut(0 == (synth_mix(seed, h * 0x85ebca6bu + site, 53) & 1023u));
		if (mut && !nonref) alt = 0;
		return (nonref != mut) ? alt + 1 : 0;
	}
}

extern "C" void orc_synth_path_lines(u32 seed, u32 first_copy, u32 n_copies, u64 E, u64 We, u32 const *edge_site, std::uint8_t const *edge_alt,
	std::uint8_t const *edge_nalts, u64 *dst /* n_copies x We words, copy-major lines */, int threads)
{
	std::fill(dst, dst + u64(n_copies) * We, u64(0));
	std::atomic<u32> next(0);
	auto worker([&]{
		while (true)
		{
			u32 const i(next.fetch_add(1));
			if (i >= n_copies) break;
			u32 const h(first_copy + i);
			u64 *line(dst + u64(i) * We);
			for (u64 e(0); e < E; ++e)
			{
				u32 const site(edge_site[e]);
				u32 const thr(65536u >> (2u + synth_mix(seed, site, 11) % 10u));
				if (1u + edge_alt[e] == synth_allele(seed, h, site, edge_nalts[e], thr)) line[e / 64] |= u64(1) << (e % 64);
			}
		}
	});
	std::vector<std::thread> pool;
	for (int t(1); t < threads; ++t) pool.emplace_back(worker);
	worker();
	for (auto &t : pool) t.join();
}
