// Graph and cut-position interchange with the reference: the files vcf2multialign writes with
// --output-graph / --output-cut-positions and reads with --input-graph / --input-cut-positions are cereal
// PortableBinary archives (vcf2multialign/main.cc:393-401, 418-426; founder_sequence_greedy_output.cc:121-136) of
//   variant_graph::serialize                      include/vcf2multialign/variant_graph.hh:197-209
//   founder_sequence_greedy_output::cut_positions include/vcf2multialign/output.hh:133-139
//   lb::int_vector / lb::int_matrix               lib/libbio/include/libbio/int_vector/cereal_serialization.hh:16-22,
//                                                 lib/libbio/include/libbio/int_matrix/cereal_serialization.hh:15-26
// This header reads and writes exactly that byte stream without cereal or libbio, so a graph built by a stock
// vcf2multialign elsewhere can be materialised here and vice versa.
//
// The encoding (cereal 1.3, archives/portable_binary.hpp): one byte — 1 if the writer was little-endian; every arithmetic
// value raw in the writer's byte order; every container preceded by its element count as a 64-bit cereal::size_type;
// std::string = count + bytes; a class with a *versioned* serialize member is preceded, the first time its type occurs in
// the archive, by its 32-bit class version (0: the reference registers none). int_vector<1> = (vector of 64-bit words,
// element count); int_matrix = (int_vector, stride = number of rows; 1 for a default-constructed matrix).
// The golden files under tests/golden/cereal/ were written by the real cereal headers (tests/golden/make_cereal_golden.sh).
#pragma once

#include <bit>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <istream>
#include <ostream>
#include <stdexcept>
#include <string>
#include <vector>

#include "v2m_b200.hh"

namespace v2m_b200 {

struct archive_error : std::runtime_error { using std::runtime_error::runtime_error; };

namespace portable_binary {

class writer
{
	std::ostream	&m_os;
public:
	explicit writer(std::ostream &os): m_os(os) { put<std::uint8_t>(std::endian::native == std::endian::little ? 1 : 0); }

	template <typename T> void put(T const v) { m_os.write(reinterpret_cast<char const *>(&v), sizeof(T)); }
	void class_version(std::uint32_t const v = 0) { put<std::uint32_t>(v); }
	template <typename T> void put(std::vector<T> const &vec)
	{
		put<std::uint64_t>(vec.size());
		if (!vec.empty()) m_os.write(reinterpret_cast<char const *>(vec.data()), std::streamsize(vec.size() * sizeof(T)));
	}
	void put(std::string const &s) { put<std::uint64_t>(s.size()); m_os.write(s.data(), std::streamsize(s.size())); }
	void put(std::vector<std::string> const &vec) { put<std::uint64_t>(vec.size()); for (auto const &s : vec) put(s); }
	void put(bit_matrix const &m)
	{
		put(m.values());																		// int_vector::m_values
		put<std::uint64_t>(m.number_of_rows() * m.number_of_columns());							// int_vector::m_size
		put<std::uint64_t>(m.number_of_rows() ? m.number_of_rows() : 1);						// int_matrix::m_stride
	}
	void finish() { m_os.flush(); if (!m_os) throw archive_error("portable binary archive: write failed"); }
};


class reader
{
	std::istream	&m_is;
	bool			m_swap{};

	void raw(void *dst, std::size_t n)
	{
		m_is.read(static_cast<char *>(dst), std::streamsize(n));
		if (std::size_t(m_is.gcount()) != n) throw archive_error("portable binary archive: truncated input");
	}

	template <typename T> static T swapped(T v)
	{
		if constexpr (1 == sizeof(T)) return v;
		else
		{
			unsigned char b[sizeof(T)];
			std::memcpy(b, &v, sizeof(T));
			for (std::size_t i(0); i < sizeof(T) / 2; ++i) std::swap(b[i], b[sizeof(T) - 1 - i]);
			std::memcpy(&v, b, sizeof(T));
			return v;
		}
	}

	std::uint64_t count(std::size_t const element_bytes)
	{
		auto const n(get<std::uint64_t>());
		// A count cannot exceed what is left in a sane file; refuse obviously broken sizes before allocating.
		if (n > (std::uint64_t(1) << 40) / (element_bytes ? element_bytes : 1)) throw archive_error("portable binary archive: implausible element count");
		return n;
	}

public:
	explicit reader(std::istream &is): m_is(is)
	{
		std::uint8_t flag(0);
		raw(&flag, 1);
		if (flag > 1) throw archive_error("portable binary archive: bad endianness flag");
		m_swap = (1 == flag) != (std::endian::native == std::endian::little);
	}

	template <typename T> T get() { T v; raw(&v, sizeof(T)); return m_swap ? swapped(v) : v; }
	std::uint32_t class_version() { return get<std::uint32_t>(); }
	template <typename T> void get(std::vector<T> &vec)
	{
		vec.resize(count(sizeof(T)));
		if (!vec.empty()) raw(vec.data(), vec.size() * sizeof(T));
		if (m_swap) for (auto &v : vec) v = swapped(v);
	}
	void get(std::string &s) { s.resize(count(1)); if (!s.empty()) raw(s.data(), s.size()); }
	void get(std::vector<std::string> &vec) { vec.resize(count(8)); for (auto &s : vec) get(s); }
	void get(bit_matrix &m)
	{
		std::vector<std::uint64_t> words;
		get(words);
		auto const bits(get<std::uint64_t>()), stride(get<std::uint64_t>());
		if (0 == stride || 0 != bits % stride || words.size() < (bits + 63) / 64) throw archive_error("portable binary archive: inconsistent bit matrix");
		m = bit_matrix(bits ? stride : 0, bits ? bits / stride : 0);
		words.resize((bits + 63) / 64);
		m.values() = std::move(words);
	}
};

} // namespace portable_binary


// archive(graph) — main.cc:398-399 / :423-424 with variant_graph::serialize (variant_graph.hh:197-209).
inline void save_variant_graph(std::ostream &os, variant_graph const &g)
{
	portable_binary::writer ar(os);
	ar.class_version();
	ar.put(g.reference_positions);
	ar.put(g.aligned_positions);
	ar.put(g.alt_edge_targets);
	ar.put(g.alt_edge_count_csum);
	ar.put(g.alt_edge_labels);
	ar.put(g.paths_by_chrom_copy_and_edge);
	ar.put(g.paths_by_edge_and_chrom_copy);
	ar.put(g.sample_names);
	ar.put(g.ploidy_csum);
	ar.finish();
}

inline void load_variant_graph(std::istream &is, variant_graph &g)
{
	portable_binary::reader ar(is);
	(void) ar.class_version();
	ar.get(g.reference_positions);
	ar.get(g.aligned_positions);
	ar.get(g.alt_edge_targets);
	ar.get(g.alt_edge_count_csum);
	ar.get(g.alt_edge_labels);
	ar.get(g.paths_by_chrom_copy_and_edge);
	ar.get(g.paths_by_edge_and_chrom_copy);
	ar.get(g.sample_names);
	ar.get(g.ploidy_csum);
	if (g.aligned_positions.size() != g.reference_positions.size() || g.alt_edge_count_csum.size() != g.reference_positions.size() + 1
		|| g.alt_edge_labels.size() != g.alt_edge_targets.size() || g.ploidy_csum.size() != g.sample_names.size() + 1)
		throw archive_error("variant graph archive: array sizes do not agree");
}

// founder_sequence_greedy_output::cut_positions (output.hh:89-97, serialize :133-139).
struct cut_positions_file
{
	std::vector<std::uint64_t>	cut_positions;
	std::uint64_t				min_distance{};
	std::uint32_t				score{};
};

inline void save_cut_positions(std::ostream &os, cut_positions_file const &c)
{
	portable_binary::writer ar(os);
	ar.class_version();
	ar.put<std::uint64_t>(c.min_distance);
	ar.put(c.cut_positions);
	ar.put<std::uint32_t>(c.score);
	ar.finish();
}

inline void load_cut_positions(std::istream &is, cut_positions_file &c)
{
	portable_binary::reader ar(is);
	(void) ar.class_version();
	c.min_distance = ar.get<std::uint64_t>();
	ar.get(c.cut_positions);
	c.score = ar.get<std::uint32_t>();
}

// File-name forms, as main.cc and founder_sequence_greedy_output.cc:121-136 use them (writing fails if the file exists).
inline void load_variant_graph(char const *path, variant_graph &g)
{
	std::ifstream is(path, std::ios::binary);
	if (!is) throw archive_error(std::string("unable to open ") + path);
	load_variant_graph(is, g);
}

inline void save_variant_graph(char const *path, variant_graph const &g)
{
	int const fd(::open(path, O_WRONLY | O_CREAT | O_EXCL, 0644));					// lb::writing_open_mode::CREATE
	if (-1 == fd) throw archive_error(std::string("unable to create ") + path);
	::close(fd);
	std::ofstream os(path, std::ios::binary | std::ios::trunc);
	save_variant_graph(os, g);
}


// founder_sequence_greedy_output::load_cut_positions / output_cut_positions (output.hh:121-122, founder_…cc:121-136).
template <typename t_graph>
void load_cut_positions(founder_sequence_greedy_output_t<t_graph> &out, char const *path)
{
	std::ifstream is(path, std::ios::binary);
	if (!is) throw archive_error(std::string("unable to open ") + path);
	cut_positions_file c;
	load_cut_positions(is, c);
	out.set_cut_positions(std::move(c.cut_positions), c.min_distance, c.score);
}

template <typename t_graph>
void output_cut_positions(founder_sequence_greedy_output_t<t_graph> const &out, char const *path)
{
	int const fd(::open(path, O_WRONLY | O_CREAT | O_EXCL, 0644));					// lb::writing_open_mode::CREATE
	if (-1 == fd) throw archive_error(std::string("unable to create ") + path);
	::close(fd);
	std::ofstream os(path, std::ios::binary | std::ios::trunc);
	save_cut_positions(os, cut_positions_file{out.cut_positions(), out.minimum_distance(), out.max_segmentation_height()});
}

// The members with the reference's names (output.hh:121-122), declared in v2m_b200.hh.
template <typename t_graph>
void founder_sequence_greedy_output_t<t_graph>::load_cut_positions(char const *path) { v2m_b200::load_cut_positions(*this, path); }

template <typename t_graph>
void founder_sequence_greedy_output_t<t_graph>::output_cut_positions(char const *path) const { v2m_b200::output_cut_positions(*this, path); }

} // namespace v2m_b200
