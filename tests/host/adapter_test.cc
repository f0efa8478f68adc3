// Parity test of the C++ host adapters (vcf2multialign_b200/host/v2m_b200.hh) against the CPU oracle.
// TEST INFRASTRUCTURE: includes oracle/v2m_oracle.hh as the checker. Usage: adapter_test <fasta> <vcf> <scratch dir> [founders]
// Prints one "ok …" line per check and exits non-zero on the first mismatch. Needs a CUDA device.
#include <cstdio>
#include <fstream>
#include <iostream>
#include <sstream>

#include <sys/stat.h>

#include "../../oracle/v2m_oracle.hh"
#include "../../vcf2multialign_b200/host/v2m_b200.hh"
#include "../../vcf2multialign_b200/host/v2m_cereal.hh"

namespace orc = v2m_oracle;
namespace b200 = v2m_b200;

namespace {

	int failures(0);

	void expect(bool const cond, std::string const &what)
	{
		std::cout << (cond ? "ok   " : "FAIL ") << what << std::endl;
		if (!cond) ++failures;
	}

	struct log_delegate_oracle final : orc::output_delegate
	{
		std::ostringstream log;
		void will_handle_sample(std::string const &s, orc::u32 i, orc::u32 c) override { log << "S " << s << ' ' << i << ' ' << c << '\n'; }
		void will_handle_founder_sequence(orc::u32 i) override { log << "F " << i << '\n'; }
		void handled_sequences(orc::u32 n) override { log << "H " << n << '\n'; }
	};

	struct log_delegate final : b200::output_delegate
	{
		std::ostringstream log;
		int exec_errors{}, exits{};
		void will_handle_sample(std::string const &s, sample_type i, ploidy_type c) override { log << "S " << s << ' ' << i << ' ' << c << '\n'; }
		void will_handle_founder_sequence(sample_type i) override { log << "F " << i << '\n'; }
		void handled_sequences(sequence_count_type n) override { log << "H " << n << '\n'; }
		void unable_to_execute_subprocess(int) override { ++exec_errors; }
		void exit_subprocess(pid_t pid) override { ++exits; ::waitpid(pid, nullptr, 0); }
	};

	// Within a block of rows the will_handle_* hooks fire together before the block's first handled_sequences; compare the
	// two kinds of events separately, in order.
	std::string only(std::string const &log, char kind)
	{
		std::istringstream in(log);
		std::string line, out;
		while (std::getline(in, line)) if (!line.empty() && (line[0] == kind || ('S' == kind && 'F' == line[0]))) out += line + '\n';
		return out;
	}

	std::string slurp(std::string const &path)
	{
		std::ifstream f(path, std::ios::binary);
		std::ostringstream ss; ss << f.rdbuf();
		return ss.str();
	}

	b200::variant_graph convert(orc::variant_graph const &g, bool const with_transposed)
	{
		b200::variant_graph out;
		out.reference_positions = g.reference_positions; out.aligned_positions = g.aligned_positions;
		out.alt_edge_targets = g.alt_edge_targets; out.alt_edge_count_csum = g.alt_edge_count_csum;
		out.alt_edge_labels = g.alt_edge_labels; out.sample_names = g.sample_names; out.ploidy_csum = g.ploidy_csum;
		out.paths_by_edge_and_chrom_copy = b200::bit_matrix(g.paths_by_edge_and_chrom_copy.number_of_rows(), g.paths_by_edge_and_chrom_copy.number_of_columns());
		out.paths_by_edge_and_chrom_copy.values() = g.paths_by_edge_and_chrom_copy.words;
		if (with_transposed)
		{
			out.paths_by_chrom_copy_and_edge = b200::bit_matrix(g.paths_by_chrom_copy_and_edge.number_of_rows(), g.paths_by_chrom_copy_and_edge.number_of_columns());
			out.paths_by_chrom_copy_and_edge.values() = g.paths_by_chrom_copy_and_edge.words;
		}
		return out;
	}
}


int main(int argc, char **argv)
{
	if (argc < 4) { std::cerr << "usage: adapter_test fasta vcf scratch_dir [founders]\n"; return 2; }
	std::string const scratch(argv[3]);
	unsigned const founders(argc > 4 ? unsigned(std::atoi(argv[4])) : 2);
	try
	{
		b200::check(v2m_cuda_init(0));

		std::vector<char> ref_seq;
		orc::variant_graph og;
		orc::build_delegate bd;
		orc::read_single_fasta_sequence(argv[1], ref_seq, nullptr);
		auto const vf(orc::read_vcf(argv[2]));
		orc::build_variant_graph(ref_seq, vf, "1", og, bd);
		b200::variant_graph const graph(convert(og, false));		// the adapter derives the transposed matrix on the device

		// transpose_matrix (transpose_matrix.hh:14)
		{
			auto const t(b200::transpose_matrix(graph.paths_by_edge_and_chrom_copy));
			expect(t.values() == og.paths_by_chrom_copy_and_edge.words && t.number_of_rows() == og.paths_by_chrom_copy_and_edge.number_of_rows(), "transpose_matrix");
			expect(0 == b200::transpose_matrix(b200::bit_matrix()).number_of_columns(), "transpose_matrix of an empty matrix");
		}

		// output_sequence (sequence_writer.hh:39-56)
		for (bool const unaligned : {false, true})
		{
			orc::u32 const copy(og.total_chromosome_copies() - 1);
			orc::string_sink sink; orc::sequence_writing_delegate od(copy);
			orc::output_sequence(ref_seq, og, sink, "some id", unaligned, od);
			std::ostringstream ss; b200::sequence_writing_delegate d(copy);
			b200::output_sequence(ref_seq, graph, ss, "some id", unaligned, d);
			expect(ss.str() == sink.s, std::string("output_sequence") + (unaligned ? " --unaligned" : ""));
		}

		// output_sequence, the lb::file_handle & overload (sequence_writer.hh:48-56) that output::output_sequence_file calls
		for (bool const unaligned : {false, true})
		{
			orc::u32 const copy(og.total_chromosome_copies() / 2);
			orc::string_sink sink; orc::sequence_writing_delegate od(copy);
			orc::output_sequence(ref_seq, og, sink, "row.a2m", unaligned, od);
			std::string const dst(scratch + "/one_row_" + std::to_string(unaligned));
			{
				b200::file_handle handle(::open(dst.c_str(), O_WRONLY | O_CREAT | O_EXCL, 0644));
				b200::sequence_writing_delegate d(copy);
				b200::output_sequence(ref_seq, graph, handle, "row.a2m", unaligned, d);
				b200::sequence_writing_delegate dref;						// PLOIDY_MAX: the REF row, appended to the same descriptor
				b200::output_sequence(ref_seq, graph, handle, nullptr, unaligned, dref);
			}
			orc::sequence_writing_delegate oref;
			orc::output_sequence(ref_seq, og, sink, nullptr, unaligned, oref);
			expect(slurp(dst) == sink.s, std::string("output_sequence(file_handle &)") + (unaligned ? " --unaligned" : ""));
		}
		b200::release_resident_graph();

		// haplotype_output::output_a2m(stream) with every option combination the CLI has for it
		for (int mask(0); mask < 8; ++mask)
		{
			char const *chrom((mask & 1) ? "chr7" : nullptr);
			bool const with_ref(mask & 2), unaligned(mask & 4);
			log_delegate_oracle odel; orc::haplotype_output oout(chrom, with_ref, unaligned, odel);
			orc::string_sink sink; oout.output_a2m(ref_seq, og, sink);
			log_delegate del; b200::haplotype_output out(nullptr, chrom, with_ref, unaligned, del);
			std::ostringstream ss; out.output_a2m(ref_seq, graph, ss);
			std::string const tag("haplotype_output::output_a2m chrom=" + std::string(chrom ? chrom : "-") + " ref=" + std::to_string(with_ref) + " unaligned=" + std::to_string(unaligned));
			expect(ss.str() == sink.s, tag + " bytes");
			expect(only(del.log.str(), 'H') == only(odel.log.str(), 'H') && only(del.log.str(), 'S') == only(odel.log.str(), 'S'), tag + " delegate calls");
		}

		// output_a2m(dst_name): exclusive create, then --pipe
		{
			log_delegate_oracle odel; orc::haplotype_output oout("c", true, false, odel);
			orc::string_sink sink; oout.output_a2m(ref_seq, og, sink);
			log_delegate del; b200::haplotype_output out(nullptr, "c", true, false, del);
			std::string const dst(scratch + "/all.a2m");
			out.output_a2m(ref_seq, graph, dst.c_str());
			expect(slurp(dst) == sink.s, "output_a2m(dst_name) file bytes");
			bool threw(false);
			try { out.output_a2m(ref_seq, graph, dst.c_str()); } catch (b200::error const &) { threw = true; }
			expect(threw, "output_a2m(dst_name) refuses to overwrite (O_EXCL)");

			std::string const script(scratch + "/sink.sh");
			{ std::ofstream f(script); f << "#!/bin/sh\ncat > \"$1\"\n"; }
			::chmod(script.c_str(), 0755);
			log_delegate pdel; b200::haplotype_output pout(script.c_str(), "c", true, false, pdel);
			std::string const pdst(scratch + "/piped.a2m");
			pout.output_a2m(ref_seq, graph, pdst.c_str());
			expect(slurp(pdst) == sink.s && 1 == pdel.exits, "output_a2m(dst_name) through --pipe");
			log_delegate bdel; b200::haplotype_output bout("/nonexistent/cmd", "c", true, false, bdel);
			bout.output_a2m(ref_seq, graph, (scratch + "/never.a2m").c_str());
			expect(1 == bdel.exec_errors && 0 == bdel.exits, "--pipe with a missing command reports unable_to_execute_subprocess");
		}

		// output_separate
		for (int mask(0); mask < 4; ++mask)
		{
			bool const unaligned(mask & 1), header(mask & 2);
			log_delegate_oracle odel; orc::haplotype_output oout("c", true, unaligned, odel);
			orc::named_outputs files; oout.output_separate(ref_seq, og, header, files);
			std::string const dir(scratch + "/sep" + std::to_string(mask));
			::mkdir(dir.c_str(), 0755);
			char cwd[4096]; if (!::getcwd(cwd, sizeof cwd)) return 2;
			if (::chdir(dir.c_str())) return 2;
			log_delegate del; b200::haplotype_output out(nullptr, "c", true, unaligned, del);
			out.output_separate(ref_seq, graph, header);
			bool same(true);
			for (auto const &[name, bytes] : files) same = same && slurp(name) == bytes;
			if (::chdir(cwd)) return 2;
			expect(same && !files.empty(), "haplotype_output::output_separate unaligned=" + std::to_string(unaligned) + " header=" + std::to_string(header));
			expect(only(del.log.str(), 'S') == only(odel.log.str(), 'S') && only(del.log.str(), 'H').empty(), "  delegate calls");
		}

		// output_separate through --pipe: one subprocess per row (output.cc:26-38)
		{
			log_delegate_oracle odel; orc::haplotype_output oout(nullptr, true, false, odel);
			orc::named_outputs files; oout.output_separate(ref_seq, og, true, files);
			std::string const dir(scratch + "/sep_pipe");
			::mkdir(dir.c_str(), 0755);
			std::string const script(scratch + "/sink_sep.sh");
			{ std::ofstream f(script); f << "#!/bin/sh\ncat > \"$1\"\n"; }
			::chmod(script.c_str(), 0755);
			char cwd[4096]; if (!::getcwd(cwd, sizeof cwd)) return 2;
			if (::chdir(dir.c_str())) return 2;
			log_delegate del; b200::haplotype_output out(script.c_str(), nullptr, true, false, del);
			out.output_separate(ref_seq, graph, true);
			bool same(true);
			for (auto const &[name, bytes] : files) same = same && slurp(name) == bytes;
			if (::chdir(cwd)) return 2;
			expect(same && !files.empty() && std::size_t(del.exits) == files.size(), "haplotype_output::output_separate through --pipe");
		}

		// founder_sequence_greedy_output: find_matchings + output_a2m + output_separate
		{
			log_delegate_oracle odel;
			orc::founder_sequence_greedy_output oout(nullptr, true, false, false, odel);
			bool const have_cuts(oout.find_cut_positions(og, 0));
			expect(have_cuts, "oracle find_cut_positions");
			if (have_cuts && oout.find_matchings(og, founders))
			{
				log_delegate del;
				b200::founder_sequence_greedy_output out(nullptr, nullptr, true, false, false, del);
				expect(out.find_cut_positions(ref_seq, graph, 0), "founder_sequence_greedy_output::find_cut_positions returns true");
				expect(out.cut_positions() == oout.m_cut_positions, "  cut_positions identical");
				expect(out.max_segmentation_height() == oout.m_score, "  max_segmentation_height identical");
				expect(out.find_matchings(ref_seq, graph, founders), "founder_sequence_greedy_output::find_matchings returns true");
				expect(out.assigned_samples().data == oout.m_assigned_samples.data && out.assigned_samples().rows == oout.m_assigned_samples.rows, "  assigned_samples identical");
				odel.log.str("");
				orc::string_sink sink; oout.output_a2m(ref_seq, og, sink);
				std::ostringstream ss; out.output_a2m(ref_seq, graph, ss);
				expect(ss.str() == sink.s, "founder_sequence_greedy_output::output_a2m bytes");
				expect(only(del.log.str(), 'H') == only(odel.log.str(), 'H') && only(del.log.str(), 'S') == only(odel.log.str(), 'S'), "  delegate calls");
				std::cout << ss.str();
			}
			// The reference's own parameter lists (output.hh:121-126): no reference sequence for the two sweeps, the graph stays
			// resident from find_cut_positions to output_a2m; cut positions through their cereal file.
			if (have_cuts && oout.m_assigned_samples.data.size())
			{
				log_delegate del;
				b200::founder_sequence_greedy_output out(nullptr, "chrQ", true, false, false, del);
				expect(out.find_cut_positions(graph, 0), "find_cut_positions(graph, minimum_distance)");
				expect(out.cut_positions() == oout.m_cut_positions && out.max_segmentation_height() == oout.m_score, "  cut_positions / score identical");
				std::string const cut_file(scratch + "/cuts.cereal");
				out.output_cut_positions(cut_file.c_str());
				b200::founder_sequence_greedy_output other(nullptr, "chrQ", true, false, false, del);
				other.load_cut_positions(cut_file.c_str());
				expect(other.cut_positions() == out.cut_positions() && other.max_segmentation_height() == out.max_segmentation_height()
					&& other.minimum_distance() == 0, "output_cut_positions(path) / load_cut_positions(path) round trip");
				expect(other.find_matchings(graph, founders), "find_matchings(graph, founder_count) on loaded cut positions");
				expect(other.assigned_samples().data == oout.m_assigned_samples.data, "  assigned_samples identical");
				log_delegate_oracle odel2;
				orc::founder_sequence_greedy_output oout2("chrQ", true, false, false, odel2);
				oout2.m_cut_positions = oout.m_cut_positions; oout2.m_assigned_samples = oout.m_assigned_samples;
				orc::string_sink sink; oout2.output_a2m(ref_seq, og, sink);
				std::ostringstream ss; other.output_a2m(ref_seq, graph, ss);		// the reference sequence arrives only now
				expect(ss.str() == sink.s, "  output_a2m after the reference-less sweeps");
			}
			b200::founder_sequence_greedy_output none(nullptr, nullptr, true, false, false, *new log_delegate);
			expect(!none.find_matchings(ref_seq, graph, founders), "find_matchings without cut positions returns false");
		}
	}
	catch (std::exception const &exc)
	{
		std::cout << "FAIL exception: " << exc.what() << std::endl;
		return 1;
	}
	std::cout << (failures ? "FAILED" : "ALL OK") << std::endl;
	return failures ? 1 : 0;
}
