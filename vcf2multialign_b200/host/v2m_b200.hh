// Host-side C++ adapters: the reference's interfaces for the materialisation path, implemented on libv2m_cuda.so
// (include/v2m_cuda.h). A maintainer of vcf2multialign links these in place of libvcf2multialign's
//     transpose_matrix.cc, sequence_writer.cc, haplotype_output.cc, founder_sequence_greedy_output.cc (find_matchings + output)
// (see INTEGRATION.md). Names, argument meaning, delegate call order and error behaviour follow
//     include/vcf2multialign/{transpose_matrix,sequence_writer,output}.hh and libvcf2multialign/output.cc.
//
// The adapters are templates over the graph type: anything shaped like v2m::variant_graph
// (include/vcf2multialign/variant_graph.hh:36-79 — public members reference_positions, aligned_positions,
// alt_edge_targets, alt_edge_count_csum, alt_edge_labels, paths_by_edge_and_chrom_copy, paths_by_chrom_copy_and_edge,
// sample_names, ploidy_csum) whose bit matrices offer values().data(), number_of_rows(), number_of_columns() as
// lb::bit_matrix does (lib/libbio/include/libbio/int_matrix/int_matrix.hh). libbio cannot be compiled in this image
// (Boost is absent), so `v2m_b200::variant_graph` / `v2m_b200::bit_matrix` below are minimal stand-ins with the same
// member names and memory layout, used by the tests.
//
// There is no CPU fallback: every call ends in libv2m_cuda.so and throws v2m_b200::error when the library reports one.
#pragma once

#include <cstdint>
#include <concepts>
#include <cstring>
#include <exception>
#include <limits>
#include <memory>
#include <ostream>
#include <stdexcept>
#include <string>
#include <vector>

#include <fcntl.h>
#include <sys/wait.h>
#include <unistd.h>

#include "../../include/v2m_cuda.h"

namespace v2m_b200 {

struct error : std::runtime_error { using std::runtime_error::runtime_error; };

inline void check(int const status)
{
	if (V2M_OK != status) throw error(std::string("libv2m_cuda: ") + v2m_cuda_last_error());
}


// ---- stand-ins for the libbio containers (layout: int_matrix.hh:144-175, int_vector.hh:253-261) -------------------------
class bit_matrix
{
	std::vector<std::uint64_t>	m_values;
	std::uint64_t				m_rows{}, m_cols{};

public:
	bit_matrix() = default;
	bit_matrix(std::uint64_t rows, std::uint64_t cols): m_values((rows * cols + 63) / 64, 0), m_rows(rows), m_cols(cols) {}
	std::uint64_t number_of_rows() const { return m_rows; }
	std::uint64_t number_of_columns() const { return m_cols; }
	std::vector<std::uint64_t> &values() { return m_values; }
	std::vector<std::uint64_t> const &values() const { return m_values; }
	bool operator()(std::uint64_t y, std::uint64_t x) const { auto const i(x * m_rows + y); return (m_values[i / 64] >> (i % 64)) & 1; }
	void set(std::uint64_t y, std::uint64_t x) { auto const i(x * m_rows + y); m_values[i / 64] |= std::uint64_t(1) << (i % 64); }
	bool operator==(bit_matrix const &o) const { return m_rows == o.m_rows && m_cols == o.m_cols && m_values == o.m_values; }
};

struct variant_graph		// member names of include/vcf2multialign/variant_graph.hh:57-66
{
	typedef std::uint64_t	position_type;
	typedef std::uint64_t	node_type;
	typedef std::uint64_t	edge_type;
	typedef std::uint32_t	ploidy_type;
	constexpr static ploidy_type PLOIDY_MAX{std::numeric_limits<ploidy_type>::max()};

	std::vector<position_type>	reference_positions;
	std::vector<position_type>	aligned_positions;
	std::vector<node_type>		alt_edge_targets;
	std::vector<edge_type>		alt_edge_count_csum;
	std::vector<std::string>	alt_edge_labels;
	bit_matrix					paths_by_chrom_copy_and_edge;
	bit_matrix					paths_by_edge_and_chrom_copy;
	std::vector<std::string>	sample_names;
	std::vector<ploidy_type>	ploidy_csum;

	node_type node_count() const { return reference_positions.size(); }
	edge_type edge_count() const { return alt_edge_targets.size(); }
	ploidy_type total_chromosome_copies() const { return ploidy_csum.empty() ? 0 : ploidy_csum.back(); }
};

typedef std::vector<char> sequence_type;		// variant_graph.hh:33


// ---- v2m::transpose_matrix (include/vcf2multialign/transpose_matrix.hh:14) ------------------------------------------------
template <typename t_matrix = bit_matrix>
[[nodiscard]] t_matrix transpose_matrix(t_matrix const &src)
{
	if (0 == src.number_of_columns()) return t_matrix();									// transpose_matrix.cc:48-49
	t_matrix dst(src.number_of_columns(), src.number_of_rows());
	// rows / cols not multiples of 64: the library reports V2M_ERR_ARGUMENT where the reference asserts (:53-54)
	check(v2m_cuda_transpose_bits(src.values().data(), src.number_of_rows(), src.number_of_columns(), dst.values().data()));
	return dst;
}


// ---- device residency ---------------------------------------------------------------------------------------------------
class device_graph
{
	v2m_graph_t m_handle{};

	template <typename t_graph>
	void upload(char const *ref_seq, std::uint64_t const ref_len, t_graph const &graph)
	{
		std::vector<std::uint64_t> label_offsets(graph.alt_edge_labels.size() + 1, 0);
		std::string label_chars;
		for (std::size_t e(0); e < graph.alt_edge_labels.size(); ++e)
		{
			label_chars += graph.alt_edge_labels[e];
			label_offsets[e + 1] = label_chars.size();
		}
		v2m_graph_view view{};
		view.node_count = graph.reference_positions.size();
		view.edge_count = graph.alt_edge_targets.size();
		view.reference_positions = graph.reference_positions.data();
		view.aligned_positions = graph.aligned_positions.data();
		view.alt_edge_targets = graph.alt_edge_targets.data();
		view.alt_edge_count_csum = graph.alt_edge_count_csum.data();
		view.label_offsets = label_offsets.data();
		view.label_chars = label_chars.data();
		auto const &ec(graph.paths_by_edge_and_chrom_copy);
		auto const &ce(graph.paths_by_chrom_copy_and_edge);
		view.h_pad = ec.number_of_rows();
		view.e_pad = ec.number_of_columns();
		view.paths_by_edge_and_chrom_copy = ec.values().data();
		// When the host graph has not been transposed yet (or was built without variant_graph.cc:453), K1 does it on the device.
		view.paths_by_chrom_copy_and_edge = (ce.number_of_rows() == view.e_pad && ce.number_of_columns() == view.h_pad && !ce.values().empty()) ? ce.values().data() : nullptr;
		view.total_chromosome_copies = graph.ploidy_csum.empty() ? 0 : graph.ploidy_csum.back();
		view.ref_seq = ref_seq;
		view.ref_len = ref_len;
		check(v2m_cuda_graph_upload(&view, &m_handle));
	}

public:
	device_graph() = default;
	device_graph(device_graph const &) = delete;
	device_graph &operator=(device_graph const &) = delete;
	device_graph(device_graph &&o) noexcept: m_handle(o.m_handle) { o.m_handle = nullptr; }
	~device_graph() { if (m_handle) v2m_cuda_graph_free(m_handle); }

	template <typename t_graph>
	device_graph(sequence_type const &ref_seq, t_graph const &graph) { upload(ref_seq.data(), ref_seq.size(), graph); }

	// Without the reference sequence: enough for find_cut_positions / find_matchings, which take none (output.hh:123, 126).
	template <typename t_graph>
	explicit device_graph(t_graph const &graph) { upload(nullptr, 0, graph); }

	// Supplies the reference later (a no-op if the graph has it): rows can be written from then on.
	void attach_reference(sequence_type const &ref_seq) const { check(v2m_cuda_graph_attach_reference(m_handle, ref_seq.data(), ref_seq.size())); }

	v2m_graph_t get() const { return m_handle; }
};

namespace detail {

	// The reference hands the same `variant_graph const &` to find_cut_positions, find_matchings and output_* one after the
	// other (main.cc:489-549) and to output_sequence once per row (output.cc:34, 42). The graph is uploaded on the first of
	// these calls and stays resident while the caller keeps passing the same, unchanged graph: identity is the address and
	// size of its node, edge and path arrays (a graph edited in place between calls must be released by hand).
	struct graph_key
	{
		void const *nodes{}, *targets{}, *paths{};
		std::size_t n_nodes{}, n_edges{}, n_path_words{};
		bool operator==(graph_key const &) const = default;
	};

	template <typename t_graph>
	graph_key key_of(t_graph const &g)
	{
		return {g.reference_positions.data(), g.alt_edge_targets.data(), g.paths_by_edge_and_chrom_copy.values().data(),
			g.reference_positions.size(), g.alt_edge_targets.size(), g.paths_by_edge_and_chrom_copy.values().size()};
	}

	class resident_graph
	{
		std::unique_ptr<device_graph>	m_graph;
		graph_key						m_key;

	public:
		template <typename t_graph>
		device_graph const &get(t_graph const &graph, sequence_type const *ref_seq)
		{
			graph_key const key(key_of(graph));
			if (!m_graph || !(key == m_key))
			{
				m_graph.reset();
				m_graph = ref_seq ? std::make_unique<device_graph>(*ref_seq, graph) : std::make_unique<device_graph>(graph);
				m_key = key;
			}
			else if (ref_seq)
				m_graph->attach_reference(*ref_seq);
			return *m_graph;
		}

		void release() { m_graph.reset(); m_key = graph_key{}; }
	};
}

class device_plan
{
	v2m_plan_t m_handle{};

public:
	device_plan() = default;
	explicit device_plan(v2m_plan_t h): m_handle(h) {}
	device_plan(device_plan const &) = delete;
	device_plan &operator=(device_plan const &) = delete;
	device_plan(device_plan &&o) noexcept: m_handle(o.m_handle) { o.m_handle = nullptr; }
	~device_plan() { if (m_handle) v2m_cuda_plan_free(m_handle); }
	v2m_plan_t get() const { return m_handle; }
};



// ---- sequence_writer.hh:16-56 --------------------------------------------------------------------------------------------
// The reference calls handle_node once per visited node so that a delegate may switch chromosome_copy_index on the way
// (only founder_sequence_writing_delegate does, founder_sequence_greedy_output.cc:106-114). On the device the walk of a
// row is resolved from a plan, so the two delegate kinds the reference has are the two plan kinds here: a fixed
// chromosome copy (this struct; handle_node is the no-op of haplotype_output.cc:24) and founder rows (below).
struct sequence_writing_delegate
{
	typedef variant_graph::ploidy_type ploidy_type;
	constexpr static ploidy_type PLOIDY_MAX{variant_graph::PLOIDY_MAX};
	ploidy_type chromosome_copy_index{PLOIDY_MAX};
	sequence_writing_delegate() = default;
	sequence_writing_delegate(ploidy_type const idx): chromosome_copy_index(idx) {}
};

namespace detail {

	inline int ostream_sink(const void *bytes, std::uint64_t n, void *ctx)
	{
		auto &stream(*static_cast<std::ostream *>(ctx));
		stream.write(static_cast<char const *>(bytes), std::streamsize(n));
		return stream ? 0 : 1;
	}

	inline int fd_sink(const void *bytes, std::uint64_t n, void *ctx)
	{
		int const fd(*static_cast<int *>(ctx));
		auto const *p(static_cast<char const *>(bytes));
		while (n)
		{
			ssize_t const w(::write(fd, p, n));
			if (w < 0) { if (EINTR == errno) continue; return 1; }
			p += w; n -= std::uint64_t(w);
		}
		return 0;
	}

	// One row of a single-row plan as '>' id '\n' bytes (no trailing newline: sequence_writer.cc:35-36, 83-85), through the
	// library's pinned block pipeline.
	struct row_target { v2m_sink sink; void *sink_ctx; char const *fasta_identifier; };

	inline int row_to_sink(std::uint32_t, const void *bytes, std::uint64_t n, void *ctx)
	{
		auto const &t(*static_cast<row_target const *>(ctx));
		if (t.fasta_identifier)
		{
			std::string header(">");
			header += t.fasta_identifier; header += '\n';
			if (t.sink(header.data(), header.size(), t.sink_ctx)) return 1;
		}
		return n ? t.sink(bytes, n, t.sink_ctx) : 0;
	}

	inline void write_row(v2m_plan_t plan, char const *fasta_identifier, bool unaligned, v2m_sink sink, void *sink_ctx)
	{
		row_target target{sink, sink_ctx, fasta_identifier};
		check(v2m_cuda_stream_rows(plan, unaligned ? 1 : 0, &row_to_sink, &target));
	}

	inline resident_graph &thread_resident_graph() { thread_local resident_graph rg; return rg; }
}

// Frees the graph the free functions below keep resident for the calling thread.
inline void release_resident_graph() { detail::thread_resident_graph().release(); }

// v2m::output_sequence(ref_seq, graph, stream, fasta_identifier, should_output_unaligned, delegate) — sequence_writer.cc:22-85.
// The reference calls this once per row (output.cc:34, 42): the graph is uploaded by the first call and stays resident for
// the calling thread while the same graph is passed (detail::resident_graph; release_resident_graph() drops it).
template <typename t_graph>
void output_sequence(sequence_type const &ref_seq, t_graph const &graph, std::ostream &stream, char const *fasta_identifier,
	bool const should_output_unaligned, sequence_writing_delegate &delegate)
{
	device_graph const &dg(detail::thread_resident_graph().get(graph, &ref_seq));
	v2m_plan_t p{};
	std::uint32_t const idx(delegate.chromosome_copy_index);
	check(v2m_cuda_plan_haplotypes(dg.get(), &idx, 1, &p));
	device_plan const plan(p);
	detail::write_row(plan.get(), fasta_identifier, should_output_unaligned, detail::ostream_sink, &stream);
}

// The lb::file_handle & overload (sequence_writer.hh:48-56, sequence_writer.cc:88-100) — the one output::output_sequence_file
// calls (output.cc:34, 42). Anything with get() returning the descriptor will do; the bytes go out with write(2) and the
// descriptor stays open, as with the reference's never-closing lb::file_ostream.
struct file_handle					// stand-in for lb::file_handle (lib/libbio/include/libbio/file_handle.hh): owns the descriptor
{
	int fd{-1};
	explicit file_handle(int const fd_): fd(fd_) {}
	file_handle(file_handle const &) = delete;
	file_handle &operator=(file_handle const &) = delete;
	~file_handle() { if (fd >= 0) ::close(fd); }
	int get() const { return fd; }
};

template <typename t_graph, typename t_handle>
requires requires(t_handle &h) { { h.get() } -> std::convertible_to<int>; }
void output_sequence(sequence_type const &ref_seq, t_graph const &graph, t_handle &handle, char const *fasta_identifier,
	bool const should_output_unaligned, sequence_writing_delegate &delegate)
{
	device_graph const &dg(detail::thread_resident_graph().get(graph, &ref_seq));
	v2m_plan_t p{};
	std::uint32_t const idx(delegate.chromosome_copy_index);
	check(v2m_cuda_plan_haplotypes(dg.get(), &idx, 1, &p));
	device_plan const plan(p);
	int fd(handle.get());
	detail::write_row(plan.get(), fasta_identifier, should_output_unaligned, detail::fd_sink, &fd);
}


// ---- output.hh:26-68, output.cc --------------------------------------------------------------------------------------------
struct output_delegate
{
	typedef std::uint32_t sample_type;
	typedef variant_graph::ploidy_type ploidy_type;
	typedef std::uint32_t sequence_count_type;
	virtual ~output_delegate() {}
	virtual void will_handle_sample(std::string const &sample, sample_type const sample_idx, ploidy_type const chr_copy_idx) = 0;
	virtual void will_handle_founder_sequence(sample_type const idx) = 0;
	virtual void handled_sequences(sequence_count_type const sequence_count) = 0;
	// lb::subprocess_status / subprocess_type do not exist without libbio: errno of the failed exec, and the child's pid.
	virtual void unable_to_execute_subprocess(int const error_number) = 0;
	virtual void exit_subprocess(pid_t const pid) = 0;		// must reap the child (the reference's delegate calls proc.close())
};

template <typename t_graph>
class output_t
{
protected:
	char const		*m_pipe_cmd{};
	char const		*m_chromosome_id{};
	output_delegate	*m_delegate{};
	bool			m_should_output_reference{};
	bool			m_should_output_unaligned{};
	detail::resident_graph	m_resident;			// the graph on the device, kept across find_* / output_* calls on the same graph

	struct row_desc { std::string a2m_id; std::string file_stem; };

	// The rows this object writes, in order, as a device plan + per-row hooks.
	virtual device_plan make_plan(device_graph const &dg, t_graph const &graph, std::vector<row_desc> &rows) = 0;
	virtual void before_row(t_graph const &graph, std::uint32_t row) = 0;
	virtual void after_row(std::uint32_t row, bool a2m_stream) = 0;

	std::string a2m_id(std::string const &tail) const
	{
		std::string id;
		if (m_chromosome_id) { id += m_chromosome_id; id += '\t'; }
		return id + tail;
	}

	std::string file_name(std::string const &stem, bool const should_include_fasta_header) const		// haplotype_output.cc:92-102, 116-126
	{
		std::string name;
		if (m_chromosome_id) { name += m_chromosome_id; name += '.'; }
		name += stem;
		if (should_include_fasta_header) name += m_should_output_unaligned ? ".fa" : ".a2m";
		return name;
	}

	// Opens dst_name exclusively (lb::writing_open_mode::CREATE, lb/src/file_handling.cc:81-94) or starts `pipe_cmd dst_name`
	// with its stdin piped (output.cc:26-38); returns the fd to write to, -1 if the delegate was told the exec failed.
	int open_destination(char const *dst_name, pid_t &child)
	{
		child = -1;
		if (!m_pipe_cmd)
		{
			int const fd(::open(dst_name, O_RDWR | O_CREAT | O_EXCL | O_CLOEXEC, 0644));
			if (fd < 0) throw error(std::string("unable to create ") + dst_name + ": " + std::strerror(errno));
			return fd;
		}
		int in[2], status[2];
		if (::pipe(in) || ::pipe2(status, O_CLOEXEC)) throw error("pipe(2) failed");
		pid_t const pid(::fork());
		if (pid < 0) throw error("fork(2) failed");
		if (0 == pid)
		{
			::dup2(in[0], STDIN_FILENO);
			::close(in[0]); ::close(in[1]); ::close(status[0]);
			char *const argv[] = {const_cast<char *>(m_pipe_cmd), const_cast<char *>(dst_name), nullptr};
			::execvp(m_pipe_cmd, argv);
			int const err(errno);
			(void) !::write(status[1], &err, sizeof(err));
			::_exit(127);
		}
		::close(in[0]); ::close(status[1]);
		int err(0);
		ssize_t const got(::read(status[0], &err, sizeof(err)));
		::close(status[0]);
		if (got > 0)
		{
			::close(in[1]);
			::waitpid(pid, nullptr, 0);
			m_delegate->unable_to_execute_subprocess(err);
			return -1;
		}
		child = pid;
		return in[1];
	}

public:
	output_t(char const *pipe_cmd, char const *chromosome_id, bool const should_output_reference, bool const should_output_unaligned, output_delegate &delegate):
		m_pipe_cmd(pipe_cmd), m_chromosome_id(chromosome_id), m_delegate(&delegate),
		m_should_output_reference(should_output_reference), m_should_output_unaligned(should_output_unaligned) {}
	virtual ~output_t() {}

	// output_a2m(ref_seq, graph, stream): '>' id '\n' row '\n' per row (haplotype_output.cc:38-82, founder_…cc:515-550).
	// Rows are produced block-wise on the device; will_handle_* of a block's rows fire when its first row is written,
	// handled_sequences after each row, all from the calling thread and in row order.
	void output_a2m(sequence_type const &ref_seq, t_graph const &graph, std::ostream &stream)
	{
		stream_a2m(ref_seq, graph, detail::ostream_sink, &stream);
	}

	void output_a2m(sequence_type const &ref_seq, t_graph const &graph, char const *const dst_name)			// output.cc:47-76
	{
		pid_t child;
		int fd(open_destination(dst_name, child));
		if (fd < 0) return;
		// A regular file (opened read-write above) is filled through a shared mapping by several threads inside the library;
		// the stdin of a --pipe subprocess takes sequential write(2).
		try { stream_a2m(ref_seq, graph, detail::fd_sink, &fd, child > 0 ? -1 : fd); }
		catch (...) { ::close(fd); if (child > 0) ::waitpid(child, nullptr, 0); throw; }
		::close(fd);
		if (child > 0) m_delegate->exit_subprocess(child);
	}

	// One file (or --pipe subprocess) per row; the file name doubles as the FASTA identifier and there is no trailing
	// newline (output.cc:18-44). The rows come block-wise out of the same pinned two-stream pipeline as the A2M stream
	// (v2m_cuda_stream_rows); each row's file is opened, written and closed from the calling thread, in row order.
	void output_separate(sequence_type const &ref_seq, t_graph const &graph, bool const should_include_fasta_header)
	{
		device_graph const &dg(m_resident.get(graph, &ref_seq));
		std::vector<row_desc> rows;
		device_plan const plan(make_plan(dg, graph, rows));
		separate_ctx ctx{this, &graph, &rows, should_include_fasta_header, nullptr};
		int const status(v2m_cuda_stream_rows(plan.get(), m_should_output_unaligned ? 1 : 0, &separate_row, &ctx));
		if (ctx.failure) std::rethrow_exception(ctx.failure);
		check(status);
	}

	// Drops the device copy of the graph (it is also dropped with the object).
	void release_device_graph() { m_resident.release(); }

private:
	struct separate_ctx { output_t *self; t_graph const *graph; std::vector<row_desc> const *rows; bool fasta_header; std::exception_ptr failure; };

	static int separate_row(std::uint32_t const r, const void *bytes, std::uint64_t const n, void *ctx_)
	{
		auto &ctx(*static_cast<separate_ctx *>(ctx_));
		output_t &self(*ctx.self);
		try
		{
			self.before_row(*ctx.graph, r);
			std::string const name(self.file_name((*ctx.rows)[r].file_stem, ctx.fasta_header));
			pid_t child;
			int fd(self.open_destination(name.c_str(), child));
			if (fd < 0) return 0;										// the delegate has been told (unable_to_execute_subprocess)
			std::string const header('>' + name + '\n');				// output.cc:36, 42: the file name is the FASTA identifier
			bool const ok(0 == detail::fd_sink(header.data(), header.size(), &fd) && (0 == n || 0 == detail::fd_sink(bytes, n, &fd)));
			::close(fd);
			if (!ok) { if (child > 0) ::waitpid(child, nullptr, 0); throw error("output_separate: write to " + name + " failed"); }
			if (child > 0) self.m_delegate->exit_subprocess(child);
			self.after_row(r, false);
			return 0;
		}
		catch (...) { ctx.failure = std::current_exception(); return 1; }
	}

	struct hook_ctx { output_t *self; t_graph const *graph; std::uint32_t next_before{}; std::uint32_t n_rows{}; };

	static void row_done(std::uint32_t row, void *ctx)
	{
		auto &h(*static_cast<hook_ctx *>(ctx));
		for (; h.next_before <= row && h.next_before < h.n_rows; ++h.next_before) h.self->before_row(*h.graph, h.next_before);
		h.self->after_row(row, true);
	}

	void stream_a2m(sequence_type const &ref_seq, t_graph const &graph, v2m_sink sink, void *sink_ctx, int const file_fd = -1)
	{
		device_graph const &dg(m_resident.get(graph, &ref_seq));
		std::vector<row_desc> rows;
		device_plan const plan(make_plan(dg, graph, rows));
		std::string ids;
		for (auto const &r : rows) { ids += r.a2m_id; ids += '\0'; }
		hook_ctx ctx{this, &graph, 0, std::uint32_t(rows.size())};
		if (file_fd >= 0) check(v2m_cuda_write_a2m(plan.get(), ids.data(), m_should_output_unaligned ? 1 : 0, file_fd, &row_done, &ctx));
		else check(v2m_cuda_stream_a2m(plan.get(), ids.data(), m_should_output_unaligned ? 1 : 0, sink, sink_ctx, &row_done, &ctx));
	}
};


// ---- haplotype_output (output.hh:71-78, haplotype_output.cc:38-132) ---------------------------------------------------
template <typename t_graph>
class haplotype_output_t final : public output_t<t_graph>
{
	typedef output_t<t_graph> base;
	using typename base::row_desc;
	struct row_origin { std::uint32_t sample; std::uint32_t copy; bool is_ref; };
	std::vector<row_origin>	m_origin;
	std::uint32_t			m_seq_count{1};			// haplotype_output.cc:46: starts at 1 even with --omit-reference

	device_plan make_plan(device_graph const &dg, t_graph const &graph, std::vector<row_desc> &rows) override
	{
		std::vector<std::uint32_t> copy_idx;
		m_origin.clear();
		m_seq_count = 1;
		if (this->m_should_output_reference)
		{
			copy_idx.push_back(V2M_PLOIDY_MAX);
			rows.push_back({this->a2m_id("REF"), "REF"});
			m_origin.push_back({0, 0, true});
		}
		for (std::uint32_t s(0); s < graph.sample_names.size(); ++s)
		{
			std::uint32_t const ploidy(graph.ploidy_csum[s + 1] - graph.ploidy_csum[s]);
			for (std::uint32_t c(0); c < ploidy; ++c)
			{
				copy_idx.push_back(graph.ploidy_csum[s] + c);													// haplotype_output.cc:28-30
				std::string const &name(graph.sample_names[s]);
				rows.push_back({this->a2m_id(name + '-' + std::to_string(1 + c)), name + '.' + std::to_string(1 + c)});	// :69-72, :116-126
				m_origin.push_back({s, c, false});
			}
		}
		v2m_plan_t p{};
		check(v2m_cuda_plan_haplotypes(dg.get(), copy_idx.data(), std::uint32_t(copy_idx.size()), &p));
		return device_plan(p);
	}

	void before_row(t_graph const &graph, std::uint32_t const row) override
	{
		auto const &o(m_origin[row]);
		if (!o.is_ref) this->m_delegate->will_handle_sample(graph.sample_names[o.sample], o.sample, o.copy);		// :67
	}

	void after_row(std::uint32_t const row, bool const a2m_stream) override
	{
		if (!a2m_stream) return;									// output_separate reports nothing (haplotype_output.cc:85-132)
		if (m_origin[row].is_ref) this->m_delegate->handled_sequences(1);		// :58
		else this->m_delegate->handled_sequences(++m_seq_count);				// :78-79
	}

public:
	using base::base;
};


// ---- founder_sequence_greedy_output (output.hh:81-130, founder_sequence_greedy_output.cc) -----------------------------------
template <typename t_graph>
class founder_sequence_greedy_output_t final : public output_t<t_graph>
{
public:
	typedef variant_graph::ploidy_type ploidy_type;
	constexpr static ploidy_type PLOIDY_MAX{variant_graph::PLOIDY_MAX};
	typedef std::vector<std::uint64_t> cut_position_vector;
	typedef ploidy_type cut_position_score_type;																	// find_cut_positions.hh:16

	// lb::matrix<ploidy_type> (column-major, lib/libbio/include/libbio/matrix/matrix.hh:63-68): rows = segments, columns = founders.
	struct ploidy_matrix
	{
		std::vector<ploidy_type> data; std::uint64_t rows{}, cols{};
		ploidy_type operator()(std::uint64_t y, std::uint64_t x) const { return data[x * rows + y]; }
		std::uint64_t number_of_rows() const { return rows; }
		std::uint64_t number_of_columns() const { return cols; }
	};

private:
	typedef output_t<t_graph> base;
	using typename base::row_desc;
	cut_position_vector	m_cut_positions;
	ploidy_matrix		m_assigned_samples;
	std::uint64_t		m_min_distance{};
	cut_position_score_type	m_score{};
	bool				m_should_keep_ref_edges{};

	device_plan make_plan(device_graph const &dg, t_graph const &, std::vector<row_desc> &rows) override
	{
		// A founder column assigned PLOIDY_MAX in every segment follows REF edges only: that is the REF row (:519-531).
		std::uint64_t const segments(m_assigned_samples.rows);
		// The matrix is 54 MB on config 3: it is only copied when the REF column has to be put in front of it.
		std::vector<std::uint32_t> with_ref;
		std::uint32_t const *assigned(m_assigned_samples.data.data());
		std::uint32_t n_rows(0);
		if (this->m_should_output_reference)
		{
			with_ref.reserve(segments * (1 + m_assigned_samples.cols));
			with_ref.assign(segments, V2M_PLOIDY_MAX);
			with_ref.insert(with_ref.end(), m_assigned_samples.data.begin(), m_assigned_samples.data.end());
			assigned = with_ref.data();
			rows.push_back({this->a2m_id("REF"), "REF"});
			++n_rows;
		}
		for (std::uint32_t f(0); f < m_assigned_samples.cols; ++f, ++n_rows)
			rows.push_back({this->a2m_id(std::to_string(1 + f)), std::to_string(1 + f)});								// :539-542
		v2m_plan_t p{};
		check(v2m_cuda_plan_founders(dg.get(), m_cut_positions.data(), segments, assigned, n_rows, &p));
		return device_plan(p);
	}

	void before_row(t_graph const &, std::uint32_t const row) override
	{
		std::uint32_t const first(this->m_should_output_reference ? 1 : 0);
		if (row >= first) this->m_delegate->will_handle_founder_sequence(row - first);									// :535
	}

	void after_row(std::uint32_t const row, bool const a2m_stream) override
	{
		if (!a2m_stream) return;
		std::uint32_t const first(this->m_should_output_reference ? 1 : 0);
		if (row < first) this->m_delegate->handled_sequences(1);														// :530
		else this->m_delegate->handled_sequences(2 + (row - first));													// :548
	}

public:
	founder_sequence_greedy_output_t(char const *pipe_cmd, char const *chromosome_id, bool const should_output_reference,
		bool const should_keep_ref_edges, bool const should_output_unaligned, output_delegate &delegate):
		base(pipe_cmd, chromosome_id, should_output_reference, should_output_unaligned, delegate), m_should_keep_ref_edges(should_keep_ref_edges) {}

	[[nodiscard]] cut_position_vector const &cut_positions() const { return m_cut_positions; }
	[[nodiscard]] ploidy_matrix const &assigned_samples() const { return m_assigned_samples; }
	[[nodiscard]] cut_position_score_type max_segmentation_height() const { return m_score; }								// output.hh:124
	// What load_cut_positions (--input-cut-positions) provides in the reference.
	void set_cut_positions(cut_position_vector cuts) { m_cut_positions = std::move(cuts); }
	void set_cut_positions(cut_position_vector cuts, std::uint64_t const min_distance, cut_position_score_type const score)
	{
		m_cut_positions = std::move(cuts); m_min_distance = min_distance; m_score = score;
	}
	[[nodiscard]] std::uint64_t minimum_distance() const { return m_min_distance; }

	// founder_sequence_greedy_output::load_cut_positions / output_cut_positions (output.hh:121-122, founder_…cc:121-136): the
	// reference's cereal PortableBinary file, byte for byte. Defined in host/v2m_cereal.hh (include it to use them).
	void load_cut_positions(char const *path);
	void output_cut_positions(char const *path) const;

	// output.hh:123, founder_sequence_greedy_output.cc:139-152 → find_initial_cut_positions_lambda_min (find_cut_positions.cc:93-211).
	// The reference's parameter list: the sweep needs no reference sequence, so the graph goes up without one and stays
	// resident for find_matchings (which then starts from the pBWT states this call leaves behind) and for output_*.
	[[nodiscard]] bool find_cut_positions(t_graph const &graph, std::uint64_t const minimum_distance)
	{
		return find_cut_positions(this->m_resident.get(graph, nullptr), graph, minimum_distance);
	}

	[[nodiscard]] bool find_cut_positions(sequence_type const &ref_seq, t_graph const &graph, std::uint64_t const minimum_distance)
	{
		return find_cut_positions(this->m_resident.get(graph, &ref_seq), graph, minimum_distance);
	}

	[[nodiscard]] bool find_cut_positions(device_graph const &dg, t_graph const &graph, std::uint64_t const minimum_distance)
	{
		std::vector<std::uint64_t> cuts(std::max<std::size_t>(graph.reference_positions.size(), 1));
		std::uint64_t n(0);
		std::uint32_t score(0);
		int found(0);
		check(v2m_cuda_find_cut_positions(dg.get(), minimum_distance, cuts.data(), cuts.size(), &n, &score, &found));
		if (!found) return false;
		cuts.resize(n);
		m_cut_positions.assign(cuts.begin(), cuts.end());
		m_min_distance = minimum_distance;
		m_score = score;
		return true;
	}

	// output.hh:126, founder_sequence_greedy_output.cc:154-512.
	[[nodiscard]] bool find_matchings(t_graph const &graph, ploidy_type const founder_count)
	{
		return find_matchings(this->m_resident.get(graph, nullptr), founder_count);
	}

	[[nodiscard]] bool find_matchings(sequence_type const &ref_seq, t_graph const &graph, ploidy_type const founder_count)
	{
		return find_matchings(this->m_resident.get(graph, &ref_seq), founder_count);
	}

	[[nodiscard]] bool find_matchings(device_graph const &dg, ploidy_type const founder_count)						// :154-512
	{
		std::uint64_t const segments(m_cut_positions.empty() ? 0 : m_cut_positions.size() - 1);
		m_assigned_samples.rows = segments; m_assigned_samples.cols = founder_count;
		m_assigned_samples.data.assign(segments * founder_count, PLOIDY_MAX);
		int found(0);
		check(v2m_cuda_find_matchings(dg.get(), m_cut_positions.data(), m_cut_positions.size(), founder_count, m_should_keep_ref_edges ? 1 : 0,
			m_assigned_samples.data.data(), &found));
		return 0 != found;
	}
};

typedef output_t<variant_graph>							output;
typedef haplotype_output_t<variant_graph>				haplotype_output;
typedef founder_sequence_greedy_output_t<variant_graph>	founder_sequence_greedy_output;

} // namespace v2m_b200
