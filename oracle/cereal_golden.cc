// ORACLE / TEST INFRASTRUCTURE — writes golden cereal PortableBinary archives with the REAL cereal headers of the reference
// checkout (compiled where they lie: -I/root/reference/lib/cereal/include; nothing is copied into this repository).
// The serialize functions below restate, field by field, what the reference archives:
//   variant_graph::serialize                       include/vcf2multialign/variant_graph.hh:197-209 (versioned member)
//   cut_positions::serialize                       include/vcf2multialign/output.hh:133-139        (versioned member)
//   lb::int_vector serialize                       lib/libbio/include/libbio/int_vector/cereal_serialization.hh:16-22
//   lb::int_matrix save / load                     lib/libbio/include/libbio/int_matrix/cereal_serialization.hh:15-26
// (libbio's own headers cannot be included: they need Boost.)
// Usage: cereal_golden <fasta> <vcf> <min_distance> <out.graph> <out.cuts>
#include <cstdint>
#include <fstream>
#include <iostream>

#include <cereal/archives/portable_binary.hpp>
#include <cereal/types/string.hpp>
#include <cereal/types/vector.hpp>

#include "v2m_oracle.hh"

namespace {
	using namespace v2m_oracle;

	struct int_vector_mirror
	{
		std::vector<std::uint64_t>	m_values;
		std::size_t					m_size{};
	};
	template <typename t_archive> void serialize(t_archive &archive, int_vector_mirror &vec) { archive(vec.m_values, vec.m_size); }

	struct int_matrix_mirror
	{
		int_vector_mirror	m_data;
		std::size_t			m_stride{1};
	};
	template <typename t_archive> void save(t_archive &archive, int_matrix_mirror const &mat) { archive(mat.m_data, mat.m_stride); }
	template <typename t_archive> void load(t_archive &archive, int_matrix_mirror &mat) { archive(mat.m_data, mat.m_stride); }

	int_matrix_mirror mirror(bit_matrix const &m)
	{
		int_matrix_mirror out;
		out.m_data.m_values = m.words;
		out.m_data.m_size = m.bits;
		out.m_stride = m.stride;
		return out;
	}

	struct graph_mirror
	{
		std::vector<std::uint64_t>	reference_positions, aligned_positions, alt_edge_targets, alt_edge_count_csum;
		std::vector<std::string>	alt_edge_labels;
		int_matrix_mirror			paths_by_chrom_copy_and_edge, paths_by_edge_and_chrom_copy;
		std::vector<std::string>	sample_names;
		std::vector<std::uint32_t>	ploidy_csum;

		template <typename t_archive> void serialize(t_archive &ar, std::uint32_t const)
		{
			ar(reference_positions);
			ar(aligned_positions);
			ar(alt_edge_targets);
			ar(alt_edge_count_csum);
			ar(alt_edge_labels);
			ar(paths_by_chrom_copy_and_edge);
			ar(paths_by_edge_and_chrom_copy);
			ar(sample_names);
			ar(ploidy_csum);
		}
	};

	struct cuts_mirror
	{
		std::vector<std::uint64_t>	cut_positions;
		std::uint64_t				min_distance{};
		std::uint32_t				score{};

		template <typename t_archive> void serialize(t_archive &ar, std::uint32_t const)
		{
			ar(min_distance);
			ar(cut_positions);
			ar(score);
		}
	};
}

int main(int argc, char **argv)
{
	if (argc != 6) { std::cerr << "usage: cereal_golden <fasta> <vcf> <min_distance> <out.graph> <out.cuts>\n"; return 2; }
	try
	{
		std::vector<char> ref_seq;
		variant_graph g;
		build_delegate bd;
		read_single_fasta_sequence(argv[1], ref_seq, nullptr);
		auto const vf(read_vcf(argv[2]));
		build_variant_graph(ref_seq, vf, "1", g, bd);

		graph_mirror gm;
		gm.reference_positions = g.reference_positions; gm.aligned_positions = g.aligned_positions;
		gm.alt_edge_targets = g.alt_edge_targets; gm.alt_edge_count_csum = g.alt_edge_count_csum;
		gm.alt_edge_labels = g.alt_edge_labels;
		gm.paths_by_chrom_copy_and_edge = mirror(g.paths_by_chrom_copy_and_edge);
		gm.paths_by_edge_and_chrom_copy = mirror(g.paths_by_edge_and_chrom_copy);
		gm.sample_names = g.sample_names;
		gm.ploidy_csum.assign(g.ploidy_csum.begin(), g.ploidy_csum.end());
		{
			std::ofstream os(argv[4], std::ios::binary);
			cereal::PortableBinaryOutputArchive archive(os);
			archive(gm);
		}

		cuts_mirror cm;
		cm.min_distance = std::stoull(argv[3]);
		cm.score = find_initial_cut_positions_lambda_min(g, cm.min_distance, cm.cut_positions);
		{
			std::ofstream os(argv[5], std::ios::binary);
			cereal::PortableBinaryOutputArchive archive(os);
			archive(cm);
		}
		// Read both back through cereal as a self-check of the mirrors.
		{
			std::ifstream is(argv[4], std::ios::binary);
			cereal::PortableBinaryInputArchive archive(is);
			graph_mirror back;
			archive(back);
			if (back.alt_edge_labels != gm.alt_edge_labels || back.paths_by_edge_and_chrom_copy.m_data.m_values != gm.paths_by_edge_and_chrom_copy.m_data.m_values)
				throw std::runtime_error("cereal round trip differs");
		}
	}
	catch (std::exception const &exc)
	{
		std::cerr << "cereal_golden: " << exc.what() << '\n';
		return 1;
	}
	return 0;
}
