// C ABI of libv2m_cuda.so (include/v2m_cuda.h): graph residency, plans, row emission, transpose.
// Host-side preparation here is O(N + E) array flattening only; all per-row / per-bit work runs in the kernels.
#include <algorithm>
#include <atomic>
#include <cerrno>
#include <chrono>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <thread>
#include <unistd.h>

#include "../../include/v2m_cuda.h"
#include "common.cuh"
#include "device_graph.hh"

namespace v2m {

	std::string &last_error()
	{
		thread_local std::string msg;
		return msg;
	}

	int fail(int status, std::string msg)
	{
		last_error() = std::move(msg);
		return status;
	}

	static std::atomic<u64> g_launches{0};
	void count_launch(unsigned n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

	template <typename T>
	int dev_alloc(T *&ptr, u64 count)
	{
		ptr = nullptr;
		V2M_CUDA_TRY(cudaMalloc(reinterpret_cast<void **>(&ptr), std::max<u64>(count, 1) * sizeof(T)));
		return 0;
	}

	template <typename T>
	int dev_upload(T *&ptr, T const *src, u64 count, u64 alloc_count = 0, bool src_on_device = false)
	{
		alloc_count = std::max(alloc_count, count);
		if (int const st = dev_alloc(ptr, alloc_count)) return st;
		if (alloc_count > count)
			V2M_CUDA_TRY(cudaMemset(ptr + count, 0, (alloc_count - count) * sizeof(T)));
		if (count)
			V2M_CUDA_TRY(cudaMemcpy(ptr, src, count * sizeof(T), src_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice));
		return 0;
	}

	struct plan_holder
	{
		device_plan				p;
		graph_holder const		*owner{};
		std::vector<u64>		unaligned_lengths;	// host copy of p.lengths once computed

		~plan_holder() { cudaFree(p.taken); cudaFree(p.prefix); cudaFree(p.lengths); cudaFree(p.line_idx); cudaFree(p.own_lines); }
	};
}

namespace v2m {

	// Large host arrays in pageable memory (the path matrices: 4 GB on c4) go up through two pinned 64 MiB buffers that are
	// filled by a few threads and sent with asynchronous copies, block k + 1 being staged while block k is on the link: 16 – 20
	// GB/s where a plain cudaMemcpy from pageable memory reaches 11. The pinned pair is page-locked once per process.
	class staged_uploader
	{
		static constexpr u64 block_bytes = u64(64) << 20;
		std::mutex		m_mutex;		// one upload at a time
		u8				*m_pinned[2] = {nullptr, nullptr};

	public:
		~staged_uploader() { for (auto p : m_pinned) if (p) cudaFreeHost(p); }

		// Returns an empty string on success, the error text otherwise (the caller may be a worker thread).
		std::string upload(void *d_dst, void const *h_src, u64 const bytes)
		{
			std::lock_guard<std::mutex> const lock(m_mutex);
			for (auto &p : m_pinned)
				if (!p && cudaSuccess != cudaMallocHost(reinterpret_cast<void **>(&p), block_bytes)) { p = nullptr; return "cudaMallocHost failed"; }
			cudaStream_t streams[2] = {nullptr, nullptr};
			for (auto &st : streams) if (cudaSuccess != cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking)) return "cudaStreamCreate failed";
			std::string error;
			unsigned const n_threads(std::max(1u, std::min(8u, std::thread::hardware_concurrency())));
			int k(0);
			for (u64 off(0); off < bytes && error.empty(); off += block_bytes, k ^= 1)
			{
				u64 const n(std::min(block_bytes, bytes - off));
				if (cudaSuccess != cudaStreamSynchronize(streams[k])) { error = "staged upload: stream error"; break; }		// buffer k is free again
				u64 const piece(((n + n_threads - 1) / n_threads + 4095) & ~u64(4095));
				std::vector<std::thread> pool;
				u8 *const dst(m_pinned[k]);
				u8 const *const src(static_cast<u8 const *>(h_src) + off);
				for (unsigned t(1); t < n_threads; ++t)
					if (u64(t) * piece < n)
						pool.emplace_back([=]() { std::memcpy(dst + t * piece, src + t * piece, std::min(piece, n - t * piece)); });
				std::memcpy(dst, src, std::min(piece, n));
				for (auto &th : pool) th.join();
				if (cudaSuccess != cudaMemcpyAsync(static_cast<u8 *>(d_dst) + off, dst, n, cudaMemcpyHostToDevice, streams[k])) error = "staged upload: H2D copy failed";
			}
			for (auto &st : streams) { if (cudaSuccess != cudaStreamSynchronize(st) && error.empty()) error = "staged upload: stream error"; cudaStreamDestroy(st); }
			return error;
		}
	};

	static staged_uploader &uploader() { static staged_uploader u; return u; }
}

using namespace v2m;

struct v2m_plan_s : v2m::plan_holder {};

// The reference sequence and the REF-aligned template built from it (the bytes of the REF row).
static int attach_reference(graph_holder &holder, char const *ref_seq, u64 const ref_len)
{
	device_graph &g(holder.g);
	g.L = ref_len;
	if (int const st = dev_upload(g.ref, reinterpret_cast<u8 const *>(ref_seq), g.L, g.L + 16)) return st;
	if (int const st = dev_alloc(g.tmpl, g.A_alloc)) return st;
	return launch_build_template(g, nullptr);
}


extern "C" {

const char *v2m_cuda_last_error(void) { return last_error().c_str(); }

int v2m_cuda_device_count(int *count)
{
	if (!count) return fail(V2M_ERR_ARGUMENT, "count is NULL");
	V2M_CUDA_TRY(cudaGetDeviceCount(count));
	return V2M_OK;
}

int v2m_cuda_init(int device)
{
	V2M_CUDA_TRY(cudaSetDevice(device));
	V2M_CUDA_TRY(cudaFree(nullptr));
	return V2M_OK;
}

uint64_t v2m_cuda_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

void *v2m_cuda_host_alloc(uint64_t bytes)
{
	void *p(nullptr);
	if (cudaSuccess != cudaMallocHost(&p, bytes)) { fail(V2M_ERR_CUDA, "cudaMallocHost failed"); return nullptr; }
	return p;
}

void v2m_cuda_host_free(void *p) { if (p) cudaFreeHost(p); }

void *v2m_cuda_device_alloc(uint64_t bytes)
{
	void *p(nullptr);
	if (cudaSuccess != cudaMalloc(&p, bytes)) { fail(V2M_ERR_CUDA, "cudaMalloc failed"); return nullptr; }
	return p;
}

void v2m_cuda_device_free(void *p) { if (p) cudaFree(p); }

int v2m_cuda_memcpy_d2h(void *dst, const void *d_src, uint64_t bytes)
{
	V2M_CUDA_TRY(cudaMemcpy(dst, d_src, bytes, cudaMemcpyDeviceToHost));
	return V2M_OK;
}

int v2m_cuda_memcpy_h2d(void *d_dst, const void *src, uint64_t bytes)
{
	V2M_CUDA_TRY(cudaMemcpy(d_dst, src, bytes, cudaMemcpyHostToDevice));
	return V2M_OK;
}

int v2m_cuda_synchronize(void)
{
	V2M_CUDA_TRY(cudaDeviceSynchronize());
	return V2M_OK;
}


// ---- K1 ---------------------------------------------------------------------------------------------------
int v2m_cuda_transpose_bits_device(const void *d_src, uint64_t rows, uint64_t cols, void *d_dst, void *stream)
{
	if (0 != rows % 64 || 0 != cols % 64)
		return fail(V2M_ERR_ARGUMENT, "transpose_bits: rows and cols must be multiples of 64 (transpose_matrix.cc:53-54)");
	if (0 == cols || 0 == rows) return V2M_OK;
	if (!d_src || !d_dst) return fail(V2M_ERR_ARGUMENT, "transpose_bits: NULL buffer");
	return launch_transpose_bits(d_src, rows, cols, d_dst, static_cast<cudaStream_t>(stream));
}

int v2m_cuda_transpose_bits(const uint64_t *src, uint64_t rows, uint64_t cols, uint64_t *dst)
{
	if (0 != rows % 64 || 0 != cols % 64)
		return fail(V2M_ERR_ARGUMENT, "transpose_bits: rows and cols must be multiples of 64 (transpose_matrix.cc:53-54)");
	if (0 == cols || 0 == rows) return V2M_OK;
	if (!src || !dst) return fail(V2M_ERR_ARGUMENT, "transpose_bits: NULL buffer");
	u64 const bytes(rows / 8 * cols);
	u64 *d_src(nullptr), *d_dst(nullptr);
	if (int const st = dev_alloc(d_src, bytes / 8)) return st;
	if (int const st = dev_alloc(d_dst, bytes / 8)) { cudaFree(d_src); return st; }
	int st(V2M_OK);
	do
	{
		if (cudaSuccess != cudaMemcpy(d_src, src, bytes, cudaMemcpyHostToDevice)) { st = fail(V2M_ERR_CUDA, "H2D copy failed"); break; }
		if ((st = launch_transpose_bits(d_src, rows, cols, d_dst, nullptr))) break;
		if (cudaSuccess != cudaMemcpy(dst, d_dst, bytes, cudaMemcpyDeviceToHost)) { st = fail(V2M_ERR_CUDA, "D2H copy failed"); break; }
	} while (false);
	cudaFree(d_src); cudaFree(d_dst);
	return st;
}


// ---- graph ------------------------------------------------------------------------------------------------
int v2m_cuda_graph_upload(const v2m_graph_view *view, v2m_graph_t *out)
{
	if (!view || !out) return fail(V2M_ERR_ARGUMENT, "graph_upload: NULL argument");
	*out = nullptr;
	u64 const N(view->node_count), E(view->edge_count);
	if (N < 1) return fail(V2M_ERR_ARGUMENT, "graph_upload: empty graph");
	if (!view->reference_positions || !view->aligned_positions || !view->alt_edge_count_csum)
		return fail(V2M_ERR_ARGUMENT, "graph_upload: NULL node arrays");
	if (E && (!view->alt_edge_targets || !view->label_offsets))
		return fail(V2M_ERR_ARGUMENT, "graph_upload: NULL edge arrays");
	if (0 != view->h_pad % 64 || 0 != view->e_pad % 64 || view->e_pad < E || view->h_pad < view->total_chromosome_copies)
		return fail(V2M_ERR_ARGUMENT, "graph_upload: path matrix dimensions must be multiples of 64 covering all edges and chromosome copies");
	if (view->h_pad && view->e_pad && !view->paths_by_edge_and_chrom_copy && !view->paths_by_chrom_copy_and_edge)
		return fail(V2M_ERR_ARGUMENT, "graph_upload: no path matrix given");
	u64 const A(view->aligned_positions[N - 1]);
	if (A >= (u64(1) << 32) - 2 * K2_TCOL || view->ref_len >= (u64(1) << 32) || view->reference_positions[N - 1] >= (u64(1) << 32))
		return fail(V2M_ERR_ARGUMENT, "graph_upload: aligned length or reference length does not fit 32-bit device positions");
	if (view->ref_seq && view->reference_positions[N - 1] > view->ref_len)
		return fail(V2M_ERR_ARGUMENT, "graph_upload: node positions exceed the reference length");

	auto holder(std::make_unique<v2m_graph_s>());
	device_graph &g(holder->g);
	V2M_CUDA_TRY(cudaGetDevice(&g.device));
	bool const trace(nullptr != std::getenv("V2M_TRACE"));		// host-side phase times on stderr
	auto mark([&, t_last = std::chrono::steady_clock::now()](char const *what) mutable {
		if (!trace) return;
		cudaDeviceSynchronize();
		auto const now(std::chrono::steady_clock::now());
		std::fprintf(stderr, "[graph_upload] %-32s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
		t_last = now;
	});
	g.N = N; g.E = E; g.L = view->ref_seq ? view->ref_len : 0; g.A = A;
	g.h_pad = view->h_pad; g.e_pad = view->e_pad; g.Wh = g.h_pad / 64; g.We = g.e_pad / 64;
	g.H = view->total_chromosome_copies;
	g.n_tiles = (A + K2_TCOL - 1) / K2_TCOL;
	g.A_alloc = std::max<u64>(g.n_tiles, 1) * K2_TCOL;
	// Path matrices in host memory start their way to the device now, on a worker thread, while this thread flattens the node
	// and edge arrays (0.2 s of host work on c4 against 0.36 s for a plain upload of the 4 GB matrix).
	struct background_upload
	{
		std::thread	worker;
		std::string	error;
		~background_upload() { if (worker.joinable()) worker.join(); }
	} paths_upload;
	u64 const path_words(g.Wh * g.e_pad);
	bool const paths_in_background(path_words * 8 >= (u64(64) << 20) && !view->paths_on_device);
	if (paths_in_background)
	{
		if (view->paths_by_edge_and_chrom_copy) if (int const st = dev_alloc(g.paths_eh, path_words)) return st;
		if (view->paths_by_chrom_copy_and_edge) if (int const st = dev_alloc(g.paths_he, path_words)) return st;
		int const device(g.device);
		u64 *const d_eh(g.paths_eh), *const d_he(g.paths_he);
		u64 const *const h_eh(view->paths_by_edge_and_chrom_copy), *const h_he(view->paths_by_chrom_copy_and_edge);
		paths_upload.worker = std::thread([&paths_upload, device, d_eh, d_he, h_eh, h_he, path_words]() {
			if (cudaSuccess != cudaSetDevice(device)) { paths_upload.error = "cudaSetDevice failed"; return; }
			if (h_eh) paths_upload.error = uploader().upload(d_eh, h_eh, path_words * 8);
			if (h_he && paths_upload.error.empty()) paths_upload.error = uploader().upload(d_he, h_he, path_words * 8);
		});
	}
	holder->csum.assign(view->alt_edge_count_csum, view->alt_edge_count_csum + N + 1);
	holder->alnpos.assign(view->aligned_positions, view->aligned_positions + N);
	if (E) holder->targets.assign(view->alt_edge_targets, view->alt_edge_targets + E);
	if (holder->csum[N] != E)
		return fail(V2M_ERR_ARGUMENT, "graph_upload: alt_edge_count_csum does not match edge_count");

	// Flatten to 32-bit device arrays.
	std::vector<u32> refpos(N), alnpos(N), edge_src(g.e_pad, 0), edge_tgt(g.e_pad, 0);
	for (u64 n(0); n < N; ++n)
	{
		refpos[n] = u32(view->reference_positions[n]);
		alnpos[n] = u32(view->aligned_positions[n]);
		if (n && (view->aligned_positions[n] <= view->aligned_positions[n - 1] || view->reference_positions[n] <= view->reference_positions[n - 1]))
			return fail(V2M_ERR_ARGUMENT, "graph_upload: node positions must be strictly increasing");
	}
	std::vector<edge_rec> edges(g.e_pad, edge_rec{0, 0, 0, 0});
	std::vector<uint2> edge_ref(g.e_pad, make_uint2(0, 0));
	// K2a's edge classes (emit_rows.cu): simple / sibling / complex
	std::vector<u64> edge_simple(g.We, ~u64(0)), edge_sibling(g.We, 0);
	u64 reach(0);									// max target node over the edges of earlier nodes
	for (u64 n(0); n < N; ++n)
	{
		u64 const lb(holder->csum[n]), rb(holder->csum[n + 1]);
		if (lb > rb || rb > E) return fail(V2M_ERR_ARGUMENT, "graph_upload: alt_edge_count_csum is not a cumulative sum");
		for (u64 e(lb); e < rb; ++e)
		{
			u64 const t(view->alt_edge_targets[e]);
			if (t <= n || t >= N) return fail(V2M_ERR_ARGUMENT, "graph_upload: ALT edge target out of range");
			u64 const lo(view->label_offsets[e]), hi(view->label_offsets[e + 1]);
			if (hi < lo || hi - lo > alnpos[t] - alnpos[n])
				return fail(V2M_ERR_ARGUMENT, "graph_upload: ALT label longer than its aligned span (sequence_writer.cc:61)");
			if (hi >= (u64(1) << 32)) return fail(V2M_ERR_ARGUMENT, "graph_upload: label store too large");
			edge_src[e] = u32(n); edge_tgt[e] = u32(t);
			u64 const k(e - lb);					// 0: the node's first ALT edge
			if (reach <= n && k >= 1 && k <= 3 && e % 64 >= k) edge_sibling[e / 64] |= u64(1) << (e % 64);
			else if (reach > n || k >= 1) edge_simple[e / 64] &= ~(u64(1) << (e % 64));
			edges[e] = edge_rec{alnpos[n], alnpos[t], u32(lo), u32(hi - lo)};
			edge_ref[e] = make_uint2(refpos[n], refpos[t]);
		}
		for (u64 e(lb); e < rb; ++e) reach = std::max<u64>(reach, view->alt_edge_targets[e]);
	}
	g.label_bytes = E ? view->label_offsets[E] : 0;
	mark("host copies + node / edge arrays");

	// Per tile of K2_TCOL positions (aligned columns / reference positions): the edges whose span may intersect it.
	auto const tile_edge_ranges([E](u64 const n_tiles, auto const &span /* e → (begin, end) */) {
		std::vector<uint2> out(std::max<u64>(n_tiles, 1), make_uint2(0, 0));
		u64 e_lo(0), e_hi(0), e_back(0);
		u32 reach(0);			// max span end over edges [0, e_back]
		for (u64 b(0); b < n_tiles; ++b)
		{
			u64 const t0(b * K2_TCOL), t1(t0 + K2_TCOL);
			while (e_lo < E && span(e_lo).x < t0) ++e_lo;
			while (e_hi < E && span(e_hi).x < t1) ++e_hi;
			// first edge whose running maximum of span ends exceeds t0
			while (e_back < e_lo)
			{
				reach = std::max(reach, span(e_back).y);
				if (reach > t0) break;
				++e_back;
			}
			out[b] = make_uint2(unsigned(std::min(e_back, e_lo)), unsigned(e_hi));
		}
		return out;
	});
	std::vector<uint2> const tile_edges(tile_edge_ranges(g.n_tiles, [&edges](u64 e){ return make_uint2(edges[e].a0, edges[e].a1); }));
	std::vector<uint2> tile_labels(tile_edges.size(), make_uint2(0, 0));
	for (u64 b(0); b < g.n_tiles; ++b)
		if (tile_edges[b].x < tile_edges[b].y)
			tile_labels[b] = make_uint2(edges[tile_edges[b].x].label_off, edges[tile_edges[b].y - 1].label_off + edges[tile_edges[b].y - 1].label_len);
	g.ref_end = refpos[N - 1];
	g.n_ref_tiles = (u64(g.ref_end) + K2_TCOL - 1) / K2_TCOL;
	std::vector<uint2> const ref_tile_edges(tile_edge_ranges(g.n_ref_tiles, [&edge_ref](u64 e){ return edge_ref[e]; }));

	if (int const st = dev_upload(g.refpos, refpos.data(), N)) return st;
	if (int const st = dev_upload(g.alnpos, alnpos.data(), N)) return st;
	if (int const st = dev_upload(g.edge_src, edge_src.data(), g.e_pad)) return st;
	if (int const st = dev_upload(g.edge_tgt, edge_tgt.data(), g.e_pad)) return st;
	if (int const st = dev_upload(g.edge_simple, edge_simple.data(), g.We)) return st;
	if (int const st = dev_upload(g.edge_sibling, edge_sibling.data(), g.We)) return st;
	if (int const st = dev_upload(g.edges, edges.data(), g.e_pad)) return st;
	if (int const st = dev_upload(g.label_chars, reinterpret_cast<u8 const *>(view->label_chars), g.label_bytes, g.label_bytes + 32)) return st;
	if (int const st = dev_upload(g.tile_edges, tile_edges.data(), tile_edges.size())) return st;
	if (int const st = dev_upload(g.tile_labels, tile_labels.data(), tile_labels.size())) return st;
	if (int const st = dev_upload(g.edge_ref, edge_ref.data(), g.e_pad)) return st;
	mark("tile tables");
	if (int const st = dev_upload(g.ref_tile_edges, ref_tile_edges.data(), ref_tile_edges.size())) return st;
	mark("uploads of the small arrays");

	if (path_words)
	{
		if (paths_in_background)
		{
			paths_upload.worker.join();
			if (!paths_upload.error.empty()) return fail(V2M_ERR_CUDA, "graph_upload: " + paths_upload.error);
		}
		if (view->paths_by_edge_and_chrom_copy)
		{
			if (!paths_in_background)
				if (int const st = dev_upload(g.paths_eh, view->paths_by_edge_and_chrom_copy, path_words, 0, 0 != view->paths_on_device)) return st;
		}
		else if (int const st = dev_alloc(g.paths_eh, path_words)) return st;

		if (view->paths_by_chrom_copy_and_edge)
		{
			if (!paths_in_background)
				if (int const st = dev_upload(g.paths_he, view->paths_by_chrom_copy_and_edge, path_words, 0, 0 != view->paths_on_device)) return st;
		}
		else
		{
			// variant_graph.cc:453: paths_by_chrom_copy_and_edge = transpose_matrix(paths_by_edge_and_chrom_copy)
			if (int const st = dev_alloc(g.paths_he, path_words)) return st;
			if (int const st = launch_transpose_bits(g.paths_eh, g.h_pad, g.e_pad, g.paths_he, nullptr)) return st;
		}
		if (!view->paths_by_edge_and_chrom_copy)
			if (int const st = launch_transpose_bits(g.paths_he, g.e_pad, g.h_pad, g.paths_eh, nullptr)) return st;
	}

	mark("path matrices (upload / K1)");
	// Without a reference the graph serves the sweeps (find_cut_positions, find_matchings, segment_classes) and plans; rows
	// need v2m_cuda_graph_attach_reference first.
	if (view->ref_seq)
		if (int const st = attach_reference(*holder, view->ref_seq, view->ref_len)) return st;
	V2M_CUDA_TRY(cudaDeviceSynchronize());
	mark("reference + template");
	*out = holder.release();
	return V2M_OK;
}

int v2m_cuda_graph_attach_reference(v2m_graph_t g, const char *ref_seq, uint64_t ref_len)
{
	if (!g || !ref_seq) return fail(V2M_ERR_ARGUMENT, "graph_attach_reference: NULL argument");
	if (g->g.tmpl) return V2M_OK;		// already there: the reference of a graph does not change
	if (ref_len >= (u64(1) << 32) || g->g.ref_end > ref_len)
		return fail(V2M_ERR_ARGUMENT, "graph_attach_reference: node positions exceed the reference length (or it does not fit 32-bit device positions)");
	if (int const st = attach_reference(*g, ref_seq, ref_len)) return st;
	V2M_CUDA_TRY(cudaDeviceSynchronize());
	return V2M_OK;
}

int v2m_cuda_graph_free(v2m_graph_t g)
{
	delete g;
	return V2M_OK;
}

int v2m_cuda_graph_aligned_length(v2m_graph_t g, uint64_t *out)
{
	if (!g || !out) return fail(V2M_ERR_ARGUMENT, "NULL argument");
	*out = g->g.A;
	return V2M_OK;
}

int v2m_cuda_graph_download_paths_by_chrom_copy_and_edge(v2m_graph_t g, uint64_t *dst)
{
	if (!g || !dst) return fail(V2M_ERR_ARGUMENT, "NULL argument");
	V2M_CUDA_TRY(cudaMemcpy(dst, g->g.paths_he, g->g.Wh * g->g.e_pad * 8, cudaMemcpyDeviceToHost));
	return V2M_OK;
}

int v2m_cuda_graph_download_path_lines(v2m_graph_t g, uint32_t first_copy, uint32_t n_copies, uint64_t *dst)
{
	if (!g || (!dst && n_copies)) return fail(V2M_ERR_ARGUMENT, "NULL argument");
	if (u64(first_copy) + n_copies > g->g.h_pad) return fail(V2M_ERR_ARGUMENT, "graph_download_path_lines: chromosome copy range out of bounds");
	if (n_copies && g->g.We)
		V2M_CUDA_TRY(cudaMemcpy(dst, g->g.paths_he + u64(first_copy) * g->g.We, u64(n_copies) * g->g.We * 8, cudaMemcpyDeviceToHost));
	return V2M_OK;
}

int v2m_cuda_graph_download_template(v2m_graph_t g, uint8_t *dst)
{
	if (!g || !dst) return fail(V2M_ERR_ARGUMENT, "NULL argument");
	if (!g->g.tmpl) return fail(V2M_ERR_STATE, "graph_download_template: the graph was uploaded without a reference sequence");
	V2M_CUDA_TRY(cudaMemcpy(dst, g->g.tmpl, g->g.A, cudaMemcpyDeviceToHost));
	return V2M_OK;
}


// ---- plans ------------------------------------------------------------------------------------------------
int v2m_cuda_plan_haplotypes(v2m_graph_t g, const uint32_t *copy_idx, uint32_t n_rows, v2m_plan_t *out)
{
	if (!g || !out || (n_rows && !copy_idx)) return fail(V2M_ERR_ARGUMENT, "plan_haplotypes: NULL argument");
	*out = nullptr;
	for (u32 i(0); i < n_rows; ++i)
		if (V2M_PLOIDY_MAX != copy_idx[i] && copy_idx[i] >= g->g.H)
			return fail(V2M_ERR_ARGUMENT, "plan_haplotypes: chromosome copy index out of range");
	auto holder(std::make_unique<v2m_plan_s>());
	holder->owner = g;
	holder->p.graph = &g->g;
	holder->p.n_rows = n_rows;
	if (int const st = dev_alloc(holder->p.taken, u64(n_rows) * g->g.We)) return st;
	if (int const st = dev_upload(holder->p.line_idx, copy_idx, n_rows)) return st;
	holder->p.lines = g->g.paths_he;
	int const st(launch_resolve_rows(g->g, holder->p.lines, holder->p.line_idx, n_rows, holder->p.taken, nullptr));
	cudaError_t const err(cudaDeviceSynchronize());
	if (st) return st;
	if (cudaSuccess != err) return fail(V2M_ERR_CUDA, std::string("resolve_rows: ") + cudaGetErrorString(err));
	*out = holder.release();
	return V2M_OK;
}

int v2m_cuda_plan_founders(v2m_graph_t g, const uint64_t *cut_positions, uint64_t n_segments,
	const uint32_t *assigned, uint32_t n_founders, v2m_plan_t *out)
{
	if (!g || !out || !cut_positions || (n_founders && n_segments && !assigned)) return fail(V2M_ERR_ARGUMENT, "plan_founders: NULL argument");
	*out = nullptr;
	if (n_segments < 1) return fail(V2M_ERR_ARGUMENT, "plan_founders: need at least one segment");
	if (0 != cut_positions[0]) return fail(V2M_ERR_ARGUMENT, "plan_founders: first cut position must be node 0 (founder_sequence_greedy_output.cc:102)");
	std::vector<u64> cut_edges(n_segments + 1);
	for (u64 s(0); s <= n_segments; ++s)
	{
		u64 const node(cut_positions[s]);
		if (node >= g->g.N || (s && node <= cut_positions[s - 1])) return fail(V2M_ERR_ARGUMENT, "plan_founders: cut positions must be increasing node ids");
		cut_edges[s] = g->csum[node];
	}
	// The last segment runs to the end of the edge list whatever the final cut node is.
	cut_edges[n_segments] = g->g.E;

	bool const trace(nullptr != std::getenv("V2M_TRACE"));		// host-side phase times on stderr
	auto mark([&, t_last = std::chrono::steady_clock::now()](char const *what) mutable {
		if (!trace) return;
		cudaDeviceSynchronize();
		auto const now(std::chrono::steady_clock::now());
		std::fprintf(stderr, "[plan_founders] %-32s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
		t_last = now;
	});
	auto holder(std::make_unique<v2m_plan_s>());
	holder->owner = g;
	holder->p.graph = &g->g;
	holder->p.n_rows = n_founders;
	if (int const st = dev_alloc(holder->p.taken, u64(n_founders) * g->g.We)) return st;
	mark("cut edges + taken buffer");
	u64 *d_cut_edges(nullptr), *d_lines(nullptr);
	u32 *d_assigned(nullptr), *d_bad(nullptr);
	int st(V2M_OK);
	do
	{
		if ((st = dev_upload(d_cut_edges, cut_edges.data(), cut_edges.size()))) break;
		// `assigned` is 54 MB on c3 and lives in pageable memory: this copy (4 – 6 ms) is most of the call. Sending it through the
		// pinned staging pair in 8 MiB pieces measured the same 3.9 – 6.2 ms (the host-side copy into the pinned pieces is the
		// bound) and cost 60 ms on a process's first call, so it stays a plain copy; callers that keep `assigned` in memory
		// from v2m_cuda_host_alloc get link speed (~1 ms).
		if ((st = dev_upload(d_assigned, assigned, n_segments * n_founders))) break;
		mark("assigned upload");
		if ((st = dev_alloc(d_lines, u64(n_founders) * g->g.We))) break;
		// the range check of `assigned` runs on the device (13.5 M entries on c3: a host loop cost more than the whole plan)
		if ((st = dev_alloc(d_bad, 1))) break;
		if (cudaSuccess != cudaMemsetAsync(d_bad, 0, sizeof(u32), nullptr)) { st = fail(V2M_ERR_CUDA, "plan_founders: memset failed"); break; }
		if ((st = launch_check_assigned(d_assigned, n_segments * n_founders, g->g.H, d_bad, nullptr))) break;
		if ((st = launch_compose_founders(g->g, d_cut_edges, n_segments, d_assigned, n_founders, d_lines, nullptr))) break;
		if ((st = launch_resolve_rows(g->g, d_lines, nullptr, n_founders, holder->p.taken, nullptr))) break;
		u32 bad(0);
		cudaError_t const err(cudaMemcpy(&bad, d_bad, sizeof(u32), cudaMemcpyDeviceToHost));		// synchronises with the default stream
		if (cudaSuccess != err) st = fail(V2M_ERR_CUDA, std::string("plan_founders: ") + cudaGetErrorString(err));
		else if (bad) st = fail(V2M_ERR_ARGUMENT, "plan_founders: assigned chromosome copy out of range");
		mark("check + compose + resolve");
	} while (false);
	cudaFree(d_cut_edges); cudaFree(d_assigned); cudaFree(d_bad);
	if (st) { cudaFree(d_lines); return st; }
	holder->p.own_lines = d_lines;
	holder->p.lines = d_lines;
	*out = holder.release();
	return V2M_OK;
}

int v2m_cuda_plan_resolve(v2m_plan_t p, void *stream)
{
	if (!p) return fail(V2M_ERR_ARGUMENT, "plan_resolve: NULL plan");
	return launch_resolve_rows(*p->p.graph, p->p.lines, p->p.line_idx, p->p.n_rows, p->p.taken, static_cast<cudaStream_t>(stream));
}

int v2m_cuda_plan_free(v2m_plan_t p)
{
	delete p;
	return V2M_OK;
}

int v2m_cuda_plan_row_count(v2m_plan_t p, uint32_t *out)
{
	if (!p || !out) return fail(V2M_ERR_ARGUMENT, "NULL argument");
	*out = p->p.n_rows;
	return V2M_OK;
}

int v2m_cuda_plan_download_taken(v2m_plan_t p, uint64_t *dst)
{
	if (!p || !dst) return fail(V2M_ERR_ARGUMENT, "NULL argument");
	V2M_CUDA_TRY(cudaMemcpy(dst, p->p.taken, u64(p->p.n_rows) * p->p.graph->We * 8, cudaMemcpyDeviceToHost));
	return V2M_OK;
}


// ---- rows -------------------------------------------------------------------------------------------------
static int check_emit_args(v2m_plan_t p, uint32_t row_begin, uint32_t row_end, uint64_t col_begin, uint64_t col_end)
{
	if (!p) return fail(V2M_ERR_ARGUMENT, "emit_rows: NULL plan");
	if (!p->p.graph->tmpl) return fail(V2M_ERR_STATE, "emit_rows: the graph was uploaded without a reference sequence (v2m_cuda_graph_attach_reference)");
	if (row_begin > row_end || row_end > p->p.n_rows) return fail(V2M_ERR_ARGUMENT, "emit_rows: row range out of bounds");
	if (col_begin > col_end || col_end > p->p.graph->A) return fail(V2M_ERR_ARGUMENT, "emit_rows: column range out of bounds");
	return V2M_OK;
}

int v2m_cuda_emit_rows_device(v2m_plan_t p, uint32_t row_begin, uint32_t row_end, uint64_t col_begin, uint64_t col_end,
	void *d_out, uint64_t pitch, void *stream)
{
	if (int const st = check_emit_args(p, row_begin, row_end, col_begin, col_end)) return st;
	if (row_begin == row_end || col_begin == col_end) return V2M_OK;
	if (!d_out) return fail(V2M_ERR_ARGUMENT, "emit_rows_device: NULL output");
	if (0 != col_begin % 16 || 0 != pitch % 16 || 0 != reinterpret_cast<std::uintptr_t>(d_out) % 16 || pitch < col_end - col_begin)
		return fail(V2M_ERR_ARGUMENT, "emit_rows_device: output, pitch and col_begin must be 16-byte aligned and pitch >= columns");
	return launch_emit_rows(*p->p.graph, p->p.taken, row_begin, row_end, col_begin, col_end, d_out, pitch, static_cast<cudaStream_t>(stream));
}

namespace {
	// Two device tiles + two streams of the calling host thread, grown on demand, released with v2m_cuda_release_staging()
	// or when the thread ends.
	struct staging_area
	{
		u8				*d_buf[2] = {nullptr, nullptr};
		cudaStream_t	streams[2] = {nullptr, nullptr};
		u64				capacity{};
		int				device{-1};

		void release()
		{
			for (int i(0); i < 2; ++i)
			{
				if (d_buf[i]) cudaFree(d_buf[i]);
				if (streams[i]) cudaStreamDestroy(streams[i]);
				d_buf[i] = nullptr; streams[i] = nullptr;
			}
			capacity = 0; device = -1;
		}

		int reserve(u64 const bytes)
		{
			int dev(0);
			V2M_CUDA_TRY(cudaGetDevice(&dev));
			if (dev != device) { release(); device = dev; }
			for (int i(0); i < 2; ++i)
				if (!streams[i] && cudaSuccess != cudaStreamCreateWithFlags(&streams[i], cudaStreamNonBlocking))
					return fail(V2M_ERR_CUDA, "cudaStreamCreate failed");
			if (bytes <= capacity) return V2M_OK;
			for (int i(0); i < 2; ++i)
			{
				if (d_buf[i]) { cudaFree(d_buf[i]); d_buf[i] = nullptr; }
				capacity = 0;
				if (int const st = dev_alloc(d_buf[i], bytes)) return st;
			}
			capacity = bytes;
			return V2M_OK;
		}

		~staging_area() { release(); }
	};

	staging_area &thread_staging() { thread_local staging_area sa; return sa; }
}

// Row pitch of the device buffers: rows start on a 128-byte line (whole-line bulk stores: +2 % on emit_rows against a
// 16-byte pitch), but not at a multiple of 4 KiB, which puts the same columns of all rows on the same DRAM channels (−13 %).
static u64 device_row_pitch(u64 const row_bytes)
{
	static u64 const align([] { char const *env(std::getenv("V2M_ROW_PITCH_ALIGN")); u64 const v(env ? std::strtoull(env, nullptr, 10) : 0); return (v >= 16 && 0 == (v & (v - 1))) ? v : u64(128); }());	// experiments
	u64 pitch((std::max<u64>(row_bytes, 1) + align - 1) & ~(align - 1));
	if (0 == pitch % 4096) pitch += align;
	return pitch;
}

void v2m_cuda_release_staging(void) { thread_staging().release(); }

uint64_t v2m_cuda_row_pitch(uint64_t row_bytes) { return device_row_pitch(row_bytes); }

int v2m_cuda_emit_rows(v2m_plan_t p, uint32_t row_begin, uint32_t row_end, uint64_t col_begin, uint64_t col_end,
	uint8_t *out, uint64_t pitch)
{
	if (int const st = check_emit_args(p, row_begin, row_end, col_begin, col_end)) return st;
	if (row_begin == row_end || col_begin == col_end) return V2M_OK;
	if (!out || pitch < col_end - col_begin) return fail(V2M_ERR_ARGUMENT, "emit_rows: NULL output or pitch < columns");

	// Device tiles start at a 16-byte column boundary; the first (col_begin % 16) bytes of each tile row are skipped on the way out.
	u64 const cb(col_begin & ~u64(15));
	u64 const skip(col_begin - cb);
	u64 const width(col_end - col_begin);
	u64 const d_pitch(device_row_pitch(col_end - cb));
	static u64 const budget([] { char const *env(std::getenv("V2M_STAGE_MIB")); u64 const v(env ? std::strtoull(env, nullptr, 10) : 0); return (v ? v : u64(256)) << 20; }());	// bytes per staging buffer (experiments: V2M_STAGE_MIB)
	u32 const chunk_rows(u32(std::max<u64>(1, std::min<u64>(row_end - row_begin, budget / d_pitch))));

	// Staging tiles and streams are kept per host thread: a caller streaming a plan block by block (bench.py's end-to-end leg,
	// the C++ adapters) should not pay two cudaMalloc / cudaFree pairs — device-wide synchronisations — per call.
	staging_area &sa(thread_staging());
	int st(sa.reserve(u64(chunk_rows) * d_pitch));
	if (st) return st;
	int k(0);
	for (u32 r0(row_begin); r0 < row_end && !st; r0 += chunk_rows, k ^= 1)
	{
		u32 const r1(std::min<u64>(row_end, u64(r0) + chunk_rows));
		// Work on a stream is ordered, so buffer k is free again once its previous copy has been issued on the same stream.
		if ((st = launch_emit_rows(*p->p.graph, p->p.taken, r0, r1, cb, col_end, sa.d_buf[k], d_pitch, sa.streams[k]))) break;
		// Rows at the device pitch on the host side too: one 1-D copy per block (56.1 GB/s on this pool's boxes against 53.9 for
		// the 2-D copy of the same rows, profiles/r02_d2h_ceiling_n1.txt); the bytes between a row's end and the next row's
		// start are then written as well (the header says so). Otherwise a 2-D copy.
		cudaError_t err;
		if (pitch == d_pitch && 0 == skip)
		{
			// The copy engine is ~4 % slower on a copy whose length is not a multiple of its burst size (an odd row width):
			// the bulk of the block goes in one copy of a multiple of 4 KiB, the ragged tail in a second, small one.
			u8 *const dst(out + u64(r0 - row_begin) * pitch);
			u64 const total(u64(r1 - r0 - 1) * pitch + width), bulk(total & ~u64(4095));
			err = bulk ? cudaMemcpyAsync(dst, sa.d_buf[k], bulk, cudaMemcpyDeviceToHost, sa.streams[k]) : cudaSuccess;
			if (cudaSuccess == err && total > bulk) err = cudaMemcpyAsync(dst + bulk, sa.d_buf[k] + bulk, total - bulk, cudaMemcpyDeviceToHost, sa.streams[k]);
		}
		else
			err = cudaMemcpy2DAsync(out + u64(r0 - row_begin) * pitch, pitch, sa.d_buf[k] + skip, d_pitch, width, r1 - r0, cudaMemcpyDeviceToHost, sa.streams[k]);
		if (cudaSuccess != err)
			st = fail(V2M_ERR_CUDA, "D2H copy of rows failed");
	}
	for (int i(0); i < 2; ++i)
		if (cudaSuccess != cudaStreamSynchronize(sa.streams[i]) && !st) st = fail(V2M_ERR_CUDA, std::string("emit_rows: ") + cudaGetErrorString(cudaGetLastError()));
	return st;
}

// ---- --unaligned ------------------------------------------------------------------------------------------------
static int ensure_unaligned_prefix(v2m_plan_t p)
{
	if (p->p.prefix) return V2M_OK;
	device_graph const &g(*p->p.graph);
	u32 const n_rows(p->p.n_rows);
	int *d_prefix(nullptr);
	u64 *d_lengths(nullptr);
	if (int const st = dev_alloc(d_prefix, u64(n_rows) * (g.We + 1))) return st;
	if (int const st = dev_alloc(d_lengths, n_rows)) { cudaFree(d_prefix); return st; }
	int st(launch_unaligned_prefix(g, p->p.taken, n_rows, d_prefix, d_lengths, nullptr));
	p->unaligned_lengths.assign(n_rows, 0);
	if (!st && n_rows && cudaSuccess != cudaMemcpy(p->unaligned_lengths.data(), d_lengths, u64(n_rows) * 8, cudaMemcpyDeviceToHost))
		st = fail(V2M_ERR_CUDA, std::string("unaligned_prefix: ") + cudaGetErrorString(cudaGetLastError()));
	if (st) { cudaFree(d_prefix); cudaFree(d_lengths); return st; }
	p->p.prefix = d_prefix;
	p->p.lengths = d_lengths;
	return V2M_OK;
}

int v2m_cuda_plan_row_lengths(v2m_plan_t p, int unaligned, uint64_t *out)
{
	if (!p || (!out && p->p.n_rows)) return fail(V2M_ERR_ARGUMENT, "plan_row_lengths: NULL argument");
	if (!unaligned)
	{
		std::fill(out, out + p->p.n_rows, p->p.graph->A);
		return V2M_OK;
	}
	if (int const st = ensure_unaligned_prefix(p)) return st;
	std::copy(p->unaligned_lengths.begin(), p->unaligned_lengths.end(), out);
	return V2M_OK;
}

int v2m_cuda_emit_rows_unaligned(v2m_plan_t p, uint32_t row_begin, uint32_t row_end, const uint64_t *offsets, uint8_t *out)
{
	if (!p) return fail(V2M_ERR_ARGUMENT, "emit_rows_unaligned: NULL plan");
	if (!p->p.graph->ref) return fail(V2M_ERR_STATE, "emit_rows_unaligned: the graph was uploaded without a reference sequence (v2m_cuda_graph_attach_reference)");
	if (row_begin > row_end || row_end > p->p.n_rows) return fail(V2M_ERR_ARGUMENT, "emit_rows_unaligned: row range out of bounds");
	if (row_begin == row_end) return V2M_OK;
	if (!offsets || !out) return fail(V2M_ERR_ARGUMENT, "emit_rows_unaligned: NULL argument");
	if (int const st = ensure_unaligned_prefix(p)) return st;
	u32 const n(row_end - row_begin);
	u64 extent(0);
	for (u32 i(0); i < n; ++i) extent = std::max(extent, offsets[i] + p->unaligned_lengths[row_begin + i]);
	if (0 == extent) return V2M_OK;
	u8 *d_out(nullptr);
	u64 *d_offsets(nullptr);
	int st(V2M_OK);
	do
	{
		if ((st = dev_alloc(d_out, extent))) break;
		if ((st = dev_upload(d_offsets, offsets, n))) break;
		if ((st = launch_emit_unaligned(*p->p.graph, p->p.taken, p->p.prefix, row_begin, row_end, d_offsets, d_out, nullptr))) break;
		// Only the rows' own bytes are copied back; contiguous rows go in one copy.
		u32 i(0);
		while (i < n && !st)
		{
			u64 const b(offsets[i]);
			u64 e(b + p->unaligned_lengths[row_begin + i]);
			u32 j(i + 1);
			while (j < n && offsets[j] == e) { e += p->unaligned_lengths[row_begin + j]; ++j; }
			if (e > b && cudaSuccess != cudaMemcpyAsync(out + b, d_out + b, e - b, cudaMemcpyDeviceToHost, nullptr))
				st = fail(V2M_ERR_CUDA, "emit_rows_unaligned: D2H copy failed");
			i = j;
		}
		cudaError_t const err(cudaDeviceSynchronize());
		if (!st && cudaSuccess != err) st = fail(V2M_ERR_CUDA, std::string("emit_rows_unaligned: ") + cudaGetErrorString(err));
	} while (false);
	cudaFree(d_out); cudaFree(d_offsets);
	return st;
}

// Plain D2H copies, no kernel: the host link's ceiling for the end-to-end roofline (bench.py e2e.peak, scripts/microbench).
int v2m_cuda_time_d2h_copies(void *h_dst, const void *d_src, uint64_t bytes, uint64_t block_bytes, int iters, double *seconds)
{
	if (!h_dst || !d_src || !seconds || 0 == bytes || 0 == block_bytes || iters < 1) return fail(V2M_ERR_ARGUMENT, "time_d2h_copies: bad argument");
	staging_area &sa(thread_staging());
	if (int const st = sa.reserve(0)) return st;
	V2M_CUDA_TRY(cudaDeviceSynchronize());
	auto const t0(std::chrono::steady_clock::now());
	int k(0);
	for (int it(0); it < iters; ++it)
		for (u64 off(0); off < bytes; off += block_bytes, k ^= 1)
			V2M_CUDA_TRY(cudaMemcpyAsync(static_cast<u8 *>(h_dst) + off, static_cast<u8 const *>(d_src) + off, std::min(block_bytes, bytes - off), cudaMemcpyDeviceToHost, sa.streams[k]));
	for (int i(0); i < 2; ++i) V2M_CUDA_TRY(cudaStreamSynchronize(sa.streams[i]));
	*seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
	return V2M_OK;
}

int v2m_cuda_time_emit_rows(v2m_plan_t p, uint32_t row_begin, uint32_t row_end, void *d_out, uint64_t pitch, int iters, float *ms_mean)
{
	if (!p || !ms_mean || iters < 1) return fail(V2M_ERR_ARGUMENT, "time_emit_rows: bad argument");
	cudaEvent_t e0, e1;
	V2M_CUDA_TRY(cudaEventCreate(&e0));
	V2M_CUDA_TRY(cudaEventCreate(&e1));
	int st(V2M_OK);
	cudaEventRecord(e0, nullptr);
	for (int i(0); i < iters && !st; ++i)
		st = v2m_cuda_emit_rows_device(p, row_begin, row_end, 0, p->p.graph->A, d_out, pitch, nullptr);
	cudaEventRecord(e1, nullptr);
	cudaError_t const err(cudaEventSynchronize(e1));
	float ms(0);
	cudaEventElapsedTime(&ms, e0, e1);
	cudaEventDestroy(e0); cudaEventDestroy(e1);
	if (st) return st;
	if (cudaSuccess != err) return fail(V2M_ERR_CUDA, std::string("time_emit_rows: ") + cudaGetErrorString(err));
	*ms_mean = ms / iters;
	return V2M_OK;
}

} // extern "C"


// ---- whole-stream driver ------------------------------------------------------------------------------------
namespace {

	int write_all(int fd, void const *buf, size_t n)
	{
		auto const *p(static_cast<char const *>(buf));
		while (n)
		{
			ssize_t const w(::write(fd, p, n));
			if (w < 0)
			{
				if (EINTR == errno) continue;
				return fail(V2M_ERR_STATE, std::string("write(2) failed: ") + std::strerror(errno));
			}
			p += w; n -= size_t(w);
		}
		return V2M_OK;
	}
}

static int fd_sink(const void *bytes, uint64_t n, void *ctx) { return write_all(*static_cast<int *>(ctx), bytes, size_t(n)); }

namespace {
	// Seekable destinations opened for reading and writing are filled through a shared mapping: write(2) / pwrite(2) into the
	// page cache is one copy at a time per file (the inode's write lock; ~3 GB/s on tmpfs whatever the number of threads), far
	// below what the device delivers, whereas copies into a mapping fault and fill pages on as many threads as there are
	// pieces. Pipes, sockets, O_APPEND and write-only descriptors (output.cc:26-38 --pipe, …) stay on sequential write(2).
	struct copy_piece { void *dst; void const *src; u64 n; };

	struct file_mapping
	{
		int		fd{-1};
		u8		*base{};		// mapping of [map_offset, map_offset + map_bytes) of the file
		u64		map_offset{}, map_bytes{};
		u64		position{};		// file offset of the next byte of the stream
		u64		grown_from{};	// size of the file before it was extended for the stream
		bool	grown{};

		bool open(int const fd_, u64 const stream_bytes)
		{
			struct stat sb;
			if (0 != ::fstat(fd_, &sb) || !S_ISREG(sb.st_mode)) return false;
			int const flags(::fcntl(fd_, F_GETFL));
			if (-1 == flags || (flags & O_APPEND) || O_RDWR != (flags & O_ACCMODE)) return false;
			off_t const pos(::lseek(fd_, 0, SEEK_CUR));
			if (pos < 0 || 0 == stream_bytes) return false;
			u64 const page(u64(::sysconf(_SC_PAGESIZE)));
			u64 const end(u64(pos) + stream_bytes);
			// The blocks of the extended range are allocated up front: a store into a hole of a mapped file on a full disk (or
			// over quota) raises SIGBUS, where write(2) — the caller's fallback when this returns false — reports ENOSPC.
			if (u64(sb.st_size) < end)
			{
				if (0 != ::posix_fallocate(fd_, sb.st_size, off_t(end) - sb.st_size)) return false;
				grown_from = u64(sb.st_size); grown = true;
			}
			map_offset = u64(pos) / page * page;
			map_bytes = end - map_offset;
			void *m(::mmap(nullptr, map_bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd_, off_t(map_offset)));
			if (MAP_FAILED == m)
			{
				if (grown) (void) ::ftruncate(fd_, off_t(grown_from));
				grown = false;
				return false;
			}
			fd = fd_; base = static_cast<u8 *>(m); position = u64(pos);
			return true;
		}

		u8 *at(u64 const file_offset) const { return base + (file_offset - map_offset); }

		int close(bool const ok)
		{
			if (!base) return V2M_OK;
			::munmap(base, map_bytes);
			base = nullptr;
			if (!ok && grown) (void) ::ftruncate(fd, off_t(std::max(grown_from, position)));		// keep what was written, drop the rest
			if (ok && ::lseek(fd, off_t(position), SEEK_SET) < 0) return fail(V2M_ERR_STATE, std::string("lseek failed: ") + std::strerror(errno));
			return V2M_OK;
		}
	};

	void copy_parallel(std::vector<copy_piece> const &pieces, unsigned threads)
	{
		threads = std::max(1u, std::min<unsigned>(threads, unsigned(pieces.size())));
		std::atomic<std::size_t> next{0};
		auto const work([&]() {
			for (;;)
			{
				std::size_t const i(next.fetch_add(1, std::memory_order_relaxed));
				if (i >= pieces.size()) return;
				std::memcpy(pieces[i].dst, pieces[i].src, pieces[i].n);
			}
		});
		std::vector<std::thread> pool;
		for (unsigned t(1); t < threads; ++t) pool.emplace_back(work);
		work();
		for (auto &th : pool) th.join();
	}
}

static int stream_a2m_impl(v2m_plan_t p, const char *ids, int unaligned, v2m_sink sink, void *sink_ctx, v2m_row_callback row_done, void *ctx, int positional_fd,
	v2m_row_sink row_sink = nullptr);


extern "C" int v2m_cuda_write_a2m(v2m_plan_t p, const char *ids, int unaligned, int fd, v2m_row_callback row_done, void *ctx)
{
	return stream_a2m_impl(p, ids, unaligned, fd_sink, &fd, row_done, ctx, fd);
}

extern "C" int v2m_cuda_stream_a2m(v2m_plan_t p, const char *ids, int unaligned, v2m_sink sink, void *sink_ctx, v2m_row_callback row_done, void *ctx)
{
	return stream_a2m_impl(p, ids, unaligned, sink, sink_ctx, row_done, ctx, -1);
}

extern "C" int v2m_cuda_stream_rows(v2m_plan_t p, int unaligned, v2m_row_sink sink, void *sink_ctx)
{
	if (!sink) return fail(V2M_ERR_ARGUMENT, "stream_rows: NULL sink");
	return stream_a2m_impl(p, nullptr, unaligned, nullptr, sink_ctx, nullptr, nullptr, -1, sink);
}

static int stream_a2m_impl(v2m_plan_t p, const char *ids, int unaligned, v2m_sink sink, void *sink_ctx, v2m_row_callback row_done, void *ctx, int const positional_fd,
	v2m_row_sink const row_sink)
{
	if (!p || (!sink && !row_sink) || (!row_sink && !ids && p->p.n_rows)) return fail(V2M_ERR_ARGUMENT, "stream_a2m: NULL argument");
	if (!p->p.graph->tmpl) return fail(V2M_ERR_STATE, "stream_a2m: the graph was uploaded without a reference sequence (v2m_cuda_graph_attach_reference)");
	device_graph const &g(*p->p.graph);
	u32 const n_rows(p->p.n_rows);
	if (0 == n_rows) return V2M_OK;

	std::vector<u64> lengths(n_rows, g.A);
	if (unaligned)
		if (int const st = v2m_cuda_plan_row_lengths(p, 1, lengths.data())) return st;

	// Row blocks of up to ~128 MiB go through two pinned buffers: while block k is being written to fd,
	// block k + 1 is produced and copied. Rows keep a 128-byte aligned pitch on the device; headers are interleaved here.
	u64 const max_len(*std::max_element(lengths.begin(), lengths.end()));
	u64 const pitch(device_row_pitch(max_len));
	u32 const block_rows(u32(std::max<u64>(1, std::min<u64>(n_rows, (u64(256) << 20) / pitch))));
	staging_area &thread_streams(thread_staging());
	if (int const rc = thread_streams.reserve(0)) return rc;
	file_mapping mapping;
	bool positional(false);
	if (positional_fd >= 0)
	{
		u64 total(0);
		char const *id(ids);
		for (u32 r(0); r < n_rows; ++r) { u64 const l(std::strlen(id)); total += 1 + l + 1 + lengths[r] + 1; id += l + 1; }
		positional = mapping.open(positional_fd, total);
	}
	unsigned const writer_threads(std::max(1u, std::min(16u, std::thread::hardware_concurrency())));
	std::vector<copy_piece> pieces;
	u8 *h_buf[2] = {nullptr, nullptr};
	u8 *d_buf[2] = {nullptr, nullptr};
	u64 *d_row_offsets(nullptr);				// --unaligned: i · pitch for the rows of a block
	cudaStream_t streams[2] = {nullptr, nullptr};
	int st(V2M_OK);

	auto produce([&](u32 r0, int k) -> int {
		u32 const r1(std::min<u64>(n_rows, u64(r0) + block_rows));
		if (unaligned)
		{
			// Rows at pitch offsets like the aligned ones (the bytes between a row's end and the next pitch are not used).
			if (int const s = launch_emit_unaligned(g, p->p.taken, p->p.prefix, r0, r1, d_row_offsets, d_buf[k], streams[k])) return s;
		}
		else if (int const s = launch_emit_rows(g, p->p.taken, r0, r1, 0, g.A, d_buf[k], pitch, streams[k])) return s;
		if (cudaSuccess != cudaMemcpyAsync(h_buf[k], d_buf[k], u64(r1 - r0) * pitch, cudaMemcpyDeviceToHost, streams[k]))
			return fail(V2M_ERR_CUDA, "write_a2m: D2H copy failed");
		return V2M_OK;
	});

	do
	{
		for (int i(0); i < 2 && !st; ++i)
		{
			if (cudaSuccess != cudaMallocHost(reinterpret_cast<void **>(&h_buf[i]), u64(block_rows) * pitch)) { st = fail(V2M_ERR_CUDA, "write_a2m: cudaMallocHost failed"); break; }
			if ((st = dev_alloc(d_buf[i], u64(block_rows) * pitch))) break;
			streams[i] = thread_streams.streams[i];		// the calling thread's two streams: no stream (and no unit ticket for it) is created per call
		}
		if (st) break;
		if (unaligned)
		{
			if ((st = ensure_unaligned_prefix(p))) break;
			std::vector<u64> offsets(block_rows);
			for (u32 i(0); i < block_rows; ++i) offsets[i] = u64(i) * pitch;
			if ((st = dev_upload(d_row_offsets, offsets.data(), offsets.size()))) break;
			if (cudaSuccess != cudaDeviceSynchronize()) { st = fail(V2M_ERR_CUDA, "write_a2m: upload of the row offsets failed"); break; }	// the streams below do not wait for the default stream
		}

		char const *id(ids);
		if ((st = produce(0, 0))) break;
		int k(0);
		for (u32 r0(0); r0 < n_rows && !st; r0 += block_rows, k ^= 1)
		{
			u32 const r1(std::min<u64>(n_rows, u64(r0) + block_rows));
			if (r1 < n_rows && (st = produce(r1, k ^ 1))) break;
			if (cudaSuccess != cudaStreamSynchronize(streams[k])) { st = fail(V2M_ERR_CUDA, std::string("write_a2m: ") + cudaGetErrorString(cudaGetLastError())); break; }
			if (positional)
			{
				// The same bytes at the same file offsets as the sequential form below, copied into the mapping by several threads.
				constexpr u64 piece_bytes(u64(4) << 20);
				pieces.clear();
				for (u32 r(r0); r < r1; ++r)
				{
					u64 const id_len(std::strlen(id));
					u8 *dst(mapping.at(mapping.position));
					dst[0] = '>'; std::memcpy(dst + 1, id, id_len); dst[1 + id_len] = '\n';
					id += id_len + 1;
					mapping.position += id_len + 2;
					u8 const *row(h_buf[k] + u64(r - r0) * pitch);
					for (u64 off(0); off < lengths[r]; off += piece_bytes)
						pieces.push_back({mapping.at(mapping.position + off), row + off, std::min(piece_bytes, lengths[r] - off)});
					mapping.position += lengths[r];
					*mapping.at(mapping.position) = '\n';
					mapping.position += 1;
				}
				copy_parallel(pieces, writer_threads);
				if (row_done) for (u32 r(r0); r < r1; ++r) row_done(r, ctx);
				continue;
			}
			if (row_sink)
			{
				// row bodies only, one call per row, in row order: the caller adds whatever surrounds a row
				for (u32 r(r0); r < r1 && !st; ++r)
					if (row_sink(r, h_buf[k] + u64(r - r0) * pitch, lengths[r], sink_ctx)) st = fail(V2M_ERR_STATE, "stream_rows: the sink reported an error");
				continue;
			}
			for (u32 r(r0); r < r1 && !st; ++r)
			{
				// '>' id '\n' bytes '\n' (sequence_writer.cc:35-36; haplotype_output.cc:56-57, 75-76)
				std::string header(">");
				header += id; header += '\n';
				id += std::strlen(id) + 1;
				auto const put([&](void const *bytes, u64 n) -> int {
					int const rc(sink(bytes, n, sink_ctx));
					if (rc && sink != fd_sink) fail(V2M_ERR_STATE, "stream_a2m: the sink reported an error");
					return rc ? V2M_ERR_STATE : V2M_OK;
				});
				if ((st = put(header.data(), header.size()))) break;
				if ((st = put(h_buf[k] + u64(r - r0) * pitch, lengths[r]))) break;
				if ((st = put("\n", 1))) break;
				if (row_done) row_done(r, ctx);
			}
		}
	} while (false);

	for (int i(0); i < 2; ++i)
	{
		if (streams[i]) cudaStreamSynchronize(streams[i]);
		cudaFree(d_buf[i]);
		if (h_buf[i]) cudaFreeHost(h_buf[i]);
	}
	cudaFree(d_row_offsets);
	// Leave the descriptor where sequential writes would have left it.
	if (positional) { int const cst(mapping.close(V2M_OK == st)); if (!st) st = cst; }
	return st;
}

extern "C" int v2m_cuda_time_transpose_bits(const void *d_src, uint64_t rows, uint64_t cols, void *d_dst, int iters, float *ms_mean)
{
	if (!ms_mean || iters < 1) return fail(V2M_ERR_ARGUMENT, "time_transpose_bits: bad argument");
	cudaEvent_t e0, e1;
	V2M_CUDA_TRY(cudaEventCreate(&e0));
	V2M_CUDA_TRY(cudaEventCreate(&e1));
	int st(V2M_OK);
	cudaEventRecord(e0, nullptr);
	for (int i(0); i < iters && !st; ++i)
		st = v2m_cuda_transpose_bits_device(d_src, rows, cols, d_dst, nullptr);
	cudaEventRecord(e1, nullptr);
	cudaError_t const err(cudaEventSynchronize(e1));
	float ms(0);
	cudaEventElapsedTime(&ms, e0, e1);
	cudaEventDestroy(e0); cudaEventDestroy(e1);
	if (st) return st;
	if (cudaSuccess != err) return fail(V2M_ERR_CUDA, std::string("time_transpose_bits: ") + cudaGetErrorString(err));
	*ms_mean = ms / iters;
	return V2M_OK;
}

// ---- CUDA events for callers without CUDA headers (bench.py times the launching stream with these) ------------------
extern "C" void *v2m_cuda_event_create(void)
{
	cudaEvent_t ev(nullptr);
	if (cudaSuccess != cudaEventCreate(&ev)) { fail(V2M_ERR_CUDA, "cudaEventCreate failed"); return nullptr; }
	return ev;
}

extern "C" void v2m_cuda_event_destroy(void *ev) { if (ev) cudaEventDestroy(static_cast<cudaEvent_t>(ev)); }

extern "C" int v2m_cuda_event_record(void *ev, void *stream)
{
	V2M_CUDA_TRY(cudaEventRecord(static_cast<cudaEvent_t>(ev), static_cast<cudaStream_t>(stream)));
	return V2M_OK;
}

extern "C" int v2m_cuda_event_elapsed_ms(void *e0, void *e1, float *ms)
{
	V2M_CUDA_TRY(cudaEventSynchronize(static_cast<cudaEvent_t>(e1)));
	V2M_CUDA_TRY(cudaEventElapsedTime(ms, static_cast<cudaEvent_t>(e0), static_cast<cudaEvent_t>(e1)));
	return V2M_OK;
}
