// Parity test of the graph / cut-position interchange (vcf2multialign_b200/host/v2m_cereal.hh) against archives written
// by the real cereal headers (tests/golden/cereal/, made by oracle/cereal_golden.cc). Host only: no device is touched.
// TEST INFRASTRUCTURE: includes oracle/v2m_oracle.hh as the checker.
// Usage: cereal_test <fasta> <vcf> <min_distance> <golden.graph> <golden.cuts>
#include <fstream>
#include <iostream>
#include <sstream>

#include "../../oracle/v2m_oracle.hh"
#include "../../vcf2multialign_b200/host/v2m_cereal.hh"

namespace orc = v2m_oracle;
namespace b200 = v2m_b200;

namespace {
	int failures(0);
	void expect(bool ok, char const *what) { std::cout << (ok ? "ok   " : "FAIL ") << what << '\n'; if (!ok) ++failures; }

	std::string slurp(char const *path)
	{
		std::ifstream f(path, std::ios::binary);
		std::ostringstream ss; ss << f.rdbuf();
		return ss.str();
	}

	b200::bit_matrix convert(orc::bit_matrix const &m)
	{
		b200::bit_matrix out(m.bits ? m.number_of_rows() : 0, m.bits ? m.number_of_columns() : 0);
		out.values() = m.words;
		return out;
	}

	b200::variant_graph convert(orc::variant_graph const &g)
	{
		b200::variant_graph out;
		out.reference_positions = g.reference_positions; out.aligned_positions = g.aligned_positions;
		out.alt_edge_targets = g.alt_edge_targets; out.alt_edge_count_csum = g.alt_edge_count_csum;
		out.alt_edge_labels = g.alt_edge_labels; out.sample_names = g.sample_names; out.ploidy_csum = g.ploidy_csum;
		out.paths_by_edge_and_chrom_copy = convert(g.paths_by_edge_and_chrom_copy);
		out.paths_by_chrom_copy_and_edge = convert(g.paths_by_chrom_copy_and_edge);
		return out;
	}

	bool same(b200::variant_graph const &a, b200::variant_graph const &b)
	{
		return a.reference_positions == b.reference_positions && a.aligned_positions == b.aligned_positions && a.alt_edge_targets == b.alt_edge_targets
			&& a.alt_edge_count_csum == b.alt_edge_count_csum && a.alt_edge_labels == b.alt_edge_labels && a.sample_names == b.sample_names
			&& a.ploidy_csum == b.ploidy_csum && a.paths_by_edge_and_chrom_copy == b.paths_by_edge_and_chrom_copy
			&& a.paths_by_chrom_copy_and_edge == b.paths_by_chrom_copy_and_edge;
	}
}

int main(int argc, char **argv)
{
	if (argc != 6) { std::cerr << "usage: cereal_test fasta vcf min_distance golden.graph golden.cuts\n"; return 2; }
	try
	{
		std::vector<char> ref_seq;
		orc::variant_graph og;
		orc::build_delegate bd;
		orc::read_single_fasta_sequence(argv[1], ref_seq, nullptr);
		auto const vf(orc::read_vcf(argv[2]));
		orc::build_variant_graph(ref_seq, vf, "1", og, bd);
		b200::variant_graph const graph(convert(og));

		std::string const golden_graph(slurp(argv[4])), golden_cuts(slurp(argv[5]));
		expect(!golden_graph.empty() && !golden_cuts.empty(), "golden archives present");

		{
			std::ostringstream os;
			b200::save_variant_graph(os, graph);
			expect(os.str() == golden_graph, "save_variant_graph writes the bytes cereal writes");
			std::istringstream is(golden_graph);
			b200::variant_graph back;
			b200::load_variant_graph(is, back);
			expect(same(back, graph), "load_variant_graph reads cereal's archive back into the same graph");
		}
		{
			b200::cut_positions_file cuts;
			cuts.min_distance = std::stoull(argv[3]);
			cuts.score = orc::find_initial_cut_positions_lambda_min(og, cuts.min_distance, cuts.cut_positions);
			std::ostringstream os;
			b200::save_cut_positions(os, cuts);
			expect(os.str() == golden_cuts, "save_cut_positions writes the bytes cereal writes");
			std::istringstream is(golden_cuts);
			b200::cut_positions_file back;
			b200::load_cut_positions(is, back);
			expect(back.cut_positions == cuts.cut_positions && back.min_distance == cuts.min_distance && back.score == cuts.score, "load_cut_positions reads cereal's archive back");
		}
		{
			// a big-endian writer's archive: flag 0 and every value byte-swapped
			b200::cut_positions_file cuts{{0, 7, 300}, 50, 9};
			std::ostringstream os;
			b200::save_cut_positions(os, cuts);
			std::string le(os.str()), be(le);
			be[0] = 0;
			auto const swap_at([&](std::size_t off, std::size_t n) { for (std::size_t i(0); i < n / 2; ++i) std::swap(be[off + i], be[off + n - 1 - i]); });
			swap_at(1, 4); swap_at(5, 8); swap_at(13, 8);
			for (std::size_t i(0); i < 3; ++i) swap_at(21 + 8 * i, 8);
			swap_at(45, 4);
			std::istringstream is(be);
			b200::cut_positions_file back;
			b200::load_cut_positions(is, back);
			expect(back.cut_positions == cuts.cut_positions && 50 == back.min_distance && 9 == back.score, "archives of the other byte order are read");
			bool threw(false);
			try { std::istringstream cut(le.substr(0, le.size() - 3)); b200::load_cut_positions(cut, back); } catch (b200::archive_error const &) { threw = true; }
			expect(threw, "truncated archive is an error");
		}
	}
	catch (std::exception const &exc)
	{
		std::cout << "FAIL exception: " << exc.what() << std::endl;
		return 1;
	}
	std::cout << (failures ? "FAILED" : "PASSED") << std::endl;
	return failures ? 1 : 0;
}
