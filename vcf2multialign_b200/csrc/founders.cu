// K3 — per-cut haplotype classes for founder_sequence_greedy_output::find_matchings
// (reference: libvcf2multialign/founder_sequence_greedy_output.cc:201-253 and 460-469, which drive
// pbwt_context::update_divergence, include/vcf2multialign/pbwt.hh:77-134, once per ALT edge column).
//
// What the reference computes, restated: after the edge columns [0, e) have been
// processed, `permutation` is the stable sort of the chromosome copies by their bits on [0, e) read co-lexicographically
// (column e-1 most significant, 0 < 1) and `divergence[i]` is the first column from which permutation[i] and
// permutation[i-1] agree up to e (e itself for i = 0). At the cut that ends segment [s, e) (previous cut at s, the one
// before at s'), runs of positions with divergence <= s are the classes of equal bit strings on [s, e); the first member of
// a run is its representative; runs with divergence <= s' are the "joined" classes over [s', e).
//
// The reference's sweep is one sequential chain of E column steps (E·H std::map updates). Here the chain is cut into P
// chunks (P = CTAs resident on the GPU) at cut positions and run as a three-pass scan:
//   pass A  k3_local_kernel     P CTAs. Chunk j sweeps its own columns starting from the identity order: the result is the
//                               order of the chromosome copies by the chunk's bits alone → dense class ids (copies equal on
//                               the whole chunk share an id; ids ascend in co-lex order) and, per class, the divergence
//                               against the class before it.
//   pass B  k3_compose_kernel   1 CTA. The true (permutation, divergence) at every chunk boundary, by running the same
//                               stable-partition step on the ≤ 13 bit planes of each chunk's class ids (a pBWT over those
//                               artificial columns gives the same order and propagates the incoming divergences exactly;
//                               pairs that the chunk separates take the class divergence from pass A).
//   pass C  k3_emit_kernel      P CTAs. Chunk j repeats its sweep from the true state and, at every cut it owns, derives the
//                               class representatives, the distinct count, permutation.front(), rhs_first_path_is_ref and the
//                               joined (lhs_rep, rhs_rep, size) runs in permutation order.
// One column step is a CTA-wide stable partition of H <= 16,384 positions kept in shared memory (permutation u16,
// divergence u32): ballots give the zero/one masks, the divergence maxima "since the previous zero / one" are two
// warp-segmented max-scans (skipped for words that are all zeros or all ones, the common case for rare alleles), one
// cross-warp exchange through shared memory, then an in-place scatter. Columns arrive through bulk async copies
// (16 columns per copy, double buffered).
//
// Roofline: this kernel is NOT HBM-bound. It reads the edge-major path matrix twice (2 · E_pad · H_pad / 8 bytes, 1.4 GB on
// c3) but each chunk is a dependent chain of column steps (two CTA barriers each); the bound is issue/latency of that
// chain, E / P steps long. DESIGN.md §4 gives the measured step time.
#include "pbwt_sweep.cuh"

namespace v2m {

// ---------------------------------------------------------------------------------------------------------
// Pass C: class data at every cut
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ u32 k3_widen(u32 const rep) { return 0xffffu == rep ? PLOIDY_MAX : rep; }

// founder_sequence_greedy_output.cc:213-253 at cut c (segment [s, e), previous segment starting at sp) on the state in
// shared memory. Writes rep_cur[h] for every chromosome copy; with t_emit also the cut's outputs.
template <int KW, bool t_emit>
__device__ __forceinline__ void k3_cut(k3_params const &P, k3_smem const &s, u32 const c, u32 const s_thr, u32 const sp_thr,
	unsigned short const *rep_prev, unsigned short *rep_cur, bool const is_ref, u64 &cursor, u64 const cap_end)
{
	unsigned const lane(threadIdx.x & 31), warp(threadIdx.x >> 5);
	u32 const base(warp * KW * 32);
	u32 const below(lanes_below(lane));
	u32 *ex(s.ex + 128);
	constexpr u32 NONE(0xffffffffu);

	u32 p[KW], h1[KW], h2[KW];
	u32 cnt1(0), cnt2(0), last1(0xffffu), first2(NONE);
#pragma unroll
	for (int k = 0; k < KW; ++k)
	{
		u32 const i(base + k * 32 + lane);
		p[k] = s.perm[i];
		u32 const d(s.div[i]);
		bool const valid(i < P.H);
		h1[k] = __ballot_sync(FULL, valid && d > s_thr);
		h2[k] = __ballot_sync(FULL, valid && d > sp_thr);
		cnt1 += __popc(h1[k]); cnt2 += __popc(h2[k]);
		u32 const v(__shfl_sync(FULL, p[k], h1[k] ? 31 - __clz(h1[k]) : 0));
		if (h1[k]) last1 = v;
		if (NONE == first2 && h2[k]) first2 = base + k * 32 + __ffs(h2[k]) - 1;
	}
	if (0 == lane) { ex[warp] = cnt1; ex[32 + warp] = cnt2; ex[64 + warp] = last1; ex[96 + warp] = first2; }
	__syncthreads();

	u32 const cnt1_u(ex[lane]), cnt2_u(ex[32 + lane]), last1_u(ex[64 + lane]), first2_u(ex[96 + lane]);
	u32 const tot1(__reduce_add_sync(FULL, cnt1_u)), tot2(__reduce_add_sync(FULL, cnt2_u));
	u32 const pre2(__reduce_add_sync(FULL, lane < warp ? cnt2_u : 0u));
	u32 carry;
	{
		u32 const m(__ballot_sync(FULL, 0 != cnt1_u) & lanes_below(warp));
		u32 const v(__shfl_sync(FULL, last1_u, m ? 31 - __clz(m) : 0));
		carry = m ? v : 0xffffu;
	}
	u32 next_after[KW];
	{
		u32 const m(__ballot_sync(FULL, 0 != cnt2_u) & ~lanes_upto(warp));
		u32 const v(__shfl_sync(FULL, first2_u, m ? __ffs(m) - 1 : 0));
		u32 na(m ? v : P.H);
#pragma unroll
		for (int k = KW - 1; k >= 0; --k)
		{
			next_after[k] = na;
			if (h2[k]) na = base + k * 32 + __ffs(h2[k]) - 1;
		}
	}

	u32 run2(pre2);
#pragma unroll
	for (int k = 0; k < KW; ++k)
	{
		u32 const i(base + k * 32 + lane);
		u32 const m(h1[k] & lanes_upto(lane));
		u32 const v(__shfl_sync(FULL, p[k], m ? 31 - __clz(m) : 0));
		u32 const rep(m ? v : carry);
		u32 const w(__shfl_sync(FULL, p[k], h1[k] ? 31 - __clz(h1[k]) : 0));
		if (h1[k]) carry = w;
		if (i < P.H)
		{
			rep_cur[p[k]] = (unsigned short) rep;
			if (t_emit && P.rhs_eq) P.rhs_eq[u64(c - 1) * P.H + p[k]] = k3_widen(rep);
		}
		if (t_emit && (h2[k] >> lane & 1))
		{
			u32 const after(h2[k] & ~lanes_upto(lane));
			u32 const next(after ? base + k * 32 + __ffs(after) - 1 : next_after[k]);
			u64 const off(cursor + run2 + __popc(h2[k] & below));
			if (off < cap_end)
			{
				u32 *t(P.joined + 3 * off);
				t[0] = (1 == c) ? PLOIDY_MAX : k3_widen(rep_prev[p[k]]);
				t[1] = k3_widen(rep);
				t[2] = next - i;
			}
		}
		run2 += __popc(h2[k]);
	}
	if (t_emit)
	{
		if (0 == threadIdx.x)
		{
			P.distinct[c - 1] = tot1;
			P.first_class[c - 1] = p[0];
			P.first_is_ref[c - 1] = is_ref ? 1 : 0;
			P.joined_count[c - 1] = tot2;
		}
		cursor += tot2;
	}
}

template <int KW>
__global__ void __launch_bounds__(K3_THREADS, 1)
k3_emit_kernel(k3_params const P)
{
	extern __shared__ __align__(128) u8 k3_raw[];
	k3_smem const s(k3_carve(k3_raw, P.Hp, P.Wh));
	k3_state const st(k3_make_state(s, P.H));
	u32 const j(blockIdx.x);
	u32 const c0(P.chunk_cut[j]), c1(P.chunk_cut[j + 1]);
	u32 const eb(P.cut_edges[c0]), ee(P.cut_edges[c1]);

	{
		unsigned short const *bperm(P.bperm + u64(j) * P.Hp);
		u32 const *bdiv(P.bdiv + u64(j) * P.Hp);
		for (u32 i(threadIdx.x); i < P.Hp; i += K3_THREADS)
		{
			s.perm[i] = bperm[i]; s.div[i] = bdiv[i];
			s.rep[0][i] = 0xffffu; s.rep[1][i] = 0xffffu;
		}
	}
	__syncthreads();

	u64 cursor(P.joined_base[j]);
	u64 const cap_end(P.joined_base[j + 1]);
	u32 cur(0);					// rep[cur] receives the next cut's representatives, rep[cur ^ 1] holds the previous cut's
	if (c0 >= 1)
	{
		// The cut this chunk starts at belongs to the chunk before; its representatives are the lhs of this chunk's first cut.
		k3_cut<KW, false>(P, s, c0, P.cut_edges[c0 - 1], 0, s.rep[1], s.rep[0], true, cursor, cap_end);
		cur = 1;
	}
	k3_cols_begin(P, s, eb, ee);

	u32 c(c0 + 1);
	u32 next_cut_edge(P.cut_edges[c]);
	bool is_ref(true);
	auto const do_cut([&]() {
		u32 const s_thr(P.cut_edges[c - 1]);
		u32 const sp_thr(c >= 2 ? P.cut_edges[c - 2] : s_thr);
		k3_cut<KW, true>(P, s, c, s_thr, sp_thr, s.rep[cur ^ 1], s.rep[cur], is_ref, cursor, cap_end);
		cur ^= 1; ++c; is_ref = true;
		next_cut_edge = (c <= c1) ? P.cut_edges[c] : 0xffffffffu;
	});
	k3_sweep<KW>(P, s, st, eb, ee,
		[&](u32 const k) { if (k == next_cut_edge) do_cut(); },
		// founder_sequence_greedy_output.cc:466: permutation.front() carries a zero iff the column has one
		[&](u32, u32 const zeros) { is_ref = is_ref && 0 != zeros; });
	if (c <= c1) do_cut();		// the cut at the chunk's end
}

} // namespace v2m


// ---------------------------------------------------------------------------------------------------------
// Host driver
// ---------------------------------------------------------------------------------------------------------
namespace v2m {

namespace {

	template <int KW>
	int k3_launch(int const pass, k3_params const &P, u32 const n_ctas, u32 const smem, cudaStream_t stream)
	{
		switch (pass)
		{
			case 0:
				V2M_CUDA_TRY(cudaFuncSetAttribute(k3_local_kernel<KW>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
				k3_local_kernel<KW><<<n_ctas, K3_THREADS, smem, stream>>>(P);
				V2M_CHECK_LAUNCH("k3_local_kernel");
				break;
			case 1:
				V2M_CUDA_TRY(cudaFuncSetAttribute(k3_compose_kernel<KW>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
				k3_compose_kernel<KW><<<1, K3_THREADS, smem, stream>>>(P);
				V2M_CHECK_LAUNCH("k3_compose_kernel");
				break;
			case 3:
				V2M_CUDA_TRY(cudaFuncSetAttribute(k3_seed_kernel<KW>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
				k3_seed_kernel<KW><<<n_ctas, K3_THREADS, smem, stream>>>(P);
				V2M_CHECK_LAUNCH("k3_seed_kernel");
				break;
			default:
				V2M_CUDA_TRY(cudaFuncSetAttribute(k3_emit_kernel<KW>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
				k3_emit_kernel<KW><<<n_ctas, K3_THREADS, smem, stream>>>(P);
				V2M_CHECK_LAUNCH("k3_emit_kernel");
				break;
		}
		return 0;
	}

	int k3_dispatch(u32 const KW, int const pass, k3_params const &P, u32 const n_ctas, u32 const smem, cudaStream_t stream)
	{
		switch (KW)
		{
			case 1: return k3_launch<1>(pass, P, n_ctas, smem, stream);
			case 2: return k3_launch<2>(pass, P, n_ctas, smem, stream);
			case 3: return k3_launch<3>(pass, P, n_ctas, smem, stream);
			case 4: return k3_launch<4>(pass, P, n_ctas, smem, stream);
			case 5: return k3_launch<5>(pass, P, n_ctas, smem, stream);
			case 6: return k3_launch<6>(pass, P, n_ctas, smem, stream);
			case 7: return k3_launch<7>(pass, P, n_ctas, smem, stream);
			case 8: return k3_launch<8>(pass, P, n_ctas, smem, stream);
			case 10: return k3_launch<10>(pass, P, n_ctas, smem, stream);
			case 12: return k3_launch<12>(pass, P, n_ctas, smem, stream);
			case 14: return k3_launch<14>(pass, P, n_ctas, smem, stream);
			case 16: return k3_launch<16>(pass, P, n_ctas, smem, stream);
			default: return fail(V2M_ERR_ARGUMENT, "segment_classes: more than 16384 chromosome copies are not supported");
		}
	}
}

struct k3_timing { float ms_local{}, ms_compose{}, ms_emit{}; u64 n_chunks{}, n_cuts{}, n_edges{}, joined_triples{}; };
k3_timing &last_k3_timing() { thread_local k3_timing t; return t; }

// Results of K3 on the host, in the layout of v2m_cuda_segment_classes.
int segment_classes_host(graph_holder const &gh, u64 const *cut_positions, u64 const n_cuts, bool const want_rhs_eq, int const max_chunks,
	std::vector<u32> &rhs_eq, std::vector<u32> &distinct, std::vector<u32> &first_class, std::vector<u8> &first_is_ref,
	std::vector<u64> &joined_offsets, std::vector<u32> &joined)
{
	device_graph const &g(gh.g);
	if (!cut_positions || n_cuts < 1) return fail(V2M_ERR_ARGUMENT, "segment_classes: no cut positions");
	if (0 != cut_positions[0]) return fail(V2M_ERR_ARGUMENT, "segment_classes: first cut position must be node 0 (founder_sequence_greedy_output.cc:168)");
	if (0 == g.H) return fail(V2M_ERR_ARGUMENT, "segment_classes: no chromosome copies");
	u32 const KW(k3_words_per_warp(g.H));
	if (0 == KW) return fail(V2M_ERR_ARGUMENT, "segment_classes: more than 16384 chromosome copies are not supported");
	u32 const Hp(KW * K3_WARPS * 32);
	u64 const n_seg(n_cuts - 1);

	distinct.assign(n_seg, 0); first_class.assign(n_seg, 0); first_is_ref.assign(n_seg, 0);
	joined_offsets.assign(n_cuts, 0); joined.clear(); rhs_eq.clear();
	if (0 == n_seg) return V2M_OK;

	std::vector<u32> cut_edges(n_cuts);
	for (u64 c(0); c < n_cuts; ++c)
	{
		u64 const node(cut_positions[c]);
		if (node >= g.N || (c && node <= cut_positions[c - 1])) return fail(V2M_ERR_ARGUMENT, "segment_classes: cut positions must be increasing node ids");
		cut_edges[c] = u32(gh.csum[node]);
	}
	cut_edges[n_seg] = u32(g.E);		// the sweep runs to the sink whatever the last cut node is
	for (u64 c(1); c < n_cuts; ++c)
		if (cut_edges[c] <= cut_edges[c - 1])
			return fail(V2M_ERR_ARGUMENT, "segment_classes: every segment must contain at least one ALT edge (find_cut_positions.cc:129 guarantees it for computed cuts)");

	// Chunks: cut-aligned, balanced by columns + 2 × cuts (a cut costs about as much as two column steps).
	int dev(0), sms(0);
	V2M_CUDA_TRY(cudaGetDevice(&dev));
	V2M_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
	u32 const smem(k3_smem_bytes(Hp, u32(g.Wh)));
	if (int const rc = k3_check_smem(smem, "segment_classes")) return rc;
	u64 P(max_chunks > 0 ? u64(max_chunks) : u64(sms));
	if (char const *env = (max_chunks > 0 ? nullptr : std::getenv("V2M_K3_CHUNKS")))		// tests force small / odd chunk counts
		if (int const v = std::atoi(env); v > 0) P = u64(v);
	P = std::max<u64>(1, std::min<u64>(P, n_seg));
	std::vector<u32> chunk_cut{0};
	{
		auto const weight([&](u64 c) { return u64(cut_edges[c]) + 2 * c; });
		u64 const W(weight(n_seg));
		u64 c(0);
		for (u64 j(1); j < P; ++j)
		{
			u64 const target(W * j / P);
			while (c < n_seg && weight(c) < target) ++c;
			if (c > chunk_cut.back() && c < n_seg) chunk_cut.push_back(u32(c));
		}
		chunk_cut.push_back(u32(n_seg));
	}
	// States left behind by find_cut_positions on this graph: chunks then start at the first cut at or after each of them, and
	// a short sweep from the known state to that cut replaces passes A and B (on c3 cuts are a few columns apart).
	std::vector<u32> seed_src, seed_edge;
	dev_buf d_seed_perm, d_seed_div, d_seed_src, d_seed_edge;
	{
		std::lock_guard<std::mutex> const lock(gh.sweep_mutex);
		bool const wanted(max_chunks <= 0 && !std::getenv("V2M_K3_CHUNKS") && !std::getenv("V2M_K3_NO_SEED"));
		if (wanted && gh.sweep.Hp == Hp && gh.sweep.edges.size() > 1)
		{
			std::vector<u32> cc{0}, src{0}, edge{0};
			for (u32 j(1); j < gh.sweep.edges.size(); ++j)
			{
				u32 const e(gh.sweep.edges[j]);
				u64 const c(std::lower_bound(cut_edges.begin(), cut_edges.end(), e) - cut_edges.begin());
				if (c >= n_seg) break;
				if (c == cc.back()) { if (cc.size() > 1) { src.back() = j; edge.back() = e; } continue; }	// a closer state for the same cut
				cc.push_back(u32(c)); src.push_back(j); edge.push_back(e);
			}
			if (cc.size() > 1)
			{
				cc.push_back(u32(n_seg));
				chunk_cut = cc; seed_src = src; seed_edge = edge;
				// private copies: another thread may run find_cut_positions on the graph while this call is in flight
				u64 const n(gh.sweep.edges.size());
				if (d_seed_perm.alloc(n * Hp * 2) || d_seed_div.alloc(n * Hp * 4)) return V2M_ERR_CUDA;
				V2M_CUDA_TRY(cudaMemcpy(d_seed_perm.p, gh.sweep.perm, n * Hp * 2, cudaMemcpyDeviceToDevice));
				V2M_CUDA_TRY(cudaMemcpy(d_seed_div.p, gh.sweep.div, n * Hp * 4, cudaMemcpyDeviceToDevice));
			}
		}
	}
	bool const seeded(!seed_src.empty());
	u32 const n_chunks(u32(chunk_cut.size() - 1));

	dev_buf d_chunk_cut, d_cut_edges, d_cls, d_classdiv, d_ncls, d_bperm, d_bdiv, d_rhs, d_distinct, d_first, d_isref, d_jcount, d_joined, d_jbase;
	if (d_chunk_cut.alloc(chunk_cut.size() * 4) || d_cut_edges.alloc(n_cuts * 4) || d_cls.alloc(u64(n_chunks) * Hp * 2) || d_classdiv.alloc(u64(n_chunks) * Hp * 4)
		|| d_ncls.alloc(n_chunks * 4) || d_bperm.alloc(u64(n_chunks) * Hp * 2) || d_bdiv.alloc(u64(n_chunks) * Hp * 4)
		|| d_distinct.alloc(n_seg * 4) || d_first.alloc(n_seg * 4) || d_isref.alloc(n_seg) || d_jcount.alloc(n_seg * 4) || d_jbase.alloc((n_chunks + 1) * 8))
		return V2M_ERR_CUDA;
	if (want_rhs_eq && d_rhs.alloc(n_seg * g.H * 4)) return V2M_ERR_CUDA;
	V2M_CUDA_TRY(cudaMemcpy(d_chunk_cut.p, chunk_cut.data(), chunk_cut.size() * 4, cudaMemcpyHostToDevice));
	V2M_CUDA_TRY(cudaMemcpy(d_cut_edges.p, cut_edges.data(), n_cuts * 4, cudaMemcpyHostToDevice));

	k3_params kp{};
	kp.paths_eh = g.paths_eh; kp.Wh = u32(g.Wh); kp.H = g.H; kp.Hp = Hp; kp.n_chunks = n_chunks;
	kp.chunk_cut = d_chunk_cut.as<u32>(); kp.cut_edges = d_cut_edges.as<u32>(); kp.n_cuts = u32(n_cuts);
	kp.cls = d_cls.as<unsigned short>(); kp.classdiv = d_classdiv.as<u32>(); kp.n_classes = d_ncls.as<u32>();
	kp.bperm = d_bperm.as<unsigned short>(); kp.bdiv = d_bdiv.as<u32>();
	kp.rhs_eq = want_rhs_eq ? d_rhs.as<u32>() : nullptr;
	kp.distinct = d_distinct.as<u32>(); kp.first_class = d_first.as<u32>(); kp.first_is_ref = d_isref.as<u8>(); kp.joined_count = d_jcount.as<u32>();
	kp.joined_base = d_jbase.as<u64>();

	cudaEvent_t ev[4];
	for (auto &e : ev) V2M_CUDA_TRY(cudaEventCreate(&e));
	struct ev_guard { cudaEvent_t *e; ~ev_guard() { for (int i(0); i < 4; ++i) cudaEventDestroy(e[i]); } } const guard{ev};
	if (seeded)
	{
		if (d_seed_src.alloc(seed_src.size() * 4) || d_seed_edge.alloc(seed_edge.size() * 4)) return V2M_ERR_CUDA;
		V2M_CUDA_TRY(cudaMemcpy(d_seed_src.p, seed_src.data(), seed_src.size() * 4, cudaMemcpyHostToDevice));
		V2M_CUDA_TRY(cudaMemcpy(d_seed_edge.p, seed_edge.data(), seed_edge.size() * 4, cudaMemcpyHostToDevice));
		kp.seed_perm = d_seed_perm.as<unsigned short>(); kp.seed_div = d_seed_div.as<u32>();
		kp.seed_src = d_seed_src.as<u32>(); kp.seed_edge = d_seed_edge.as<u32>();
	}
	cudaEventRecord(ev[0], nullptr);
	if (seeded)
	{
		if (int const st = k3_dispatch(KW, 3, kp, n_chunks, smem, nullptr)) return st;
		cudaEventRecord(ev[1], nullptr);
	}
	else
	{
		if (n_chunks > 1)
			if (int const st = k3_dispatch(KW, 0, kp, n_chunks - 1, smem, nullptr)) return st;
		cudaEventRecord(ev[1], nullptr);
		if (int const st = k3_dispatch(KW, 1, kp, 1, smem, nullptr)) return st;
	}
	cudaEventRecord(ev[2], nullptr);

	// Joined triples go to per-chunk regions of a pool; a first guess of 48 per cut + H per chunk, exact sizes on a retry.
	std::vector<u64> jbase(n_chunks + 1, 0);
	for (u32 j(0); j < n_chunks; ++j)
		jbase[j + 1] = jbase[j] + u64(chunk_cut[j + 1] - chunk_cut[j]) * 48 + g.H;
	std::vector<u32> jcount(n_seg);
	for (int attempt(0); attempt < 2; ++attempt)
	{
		if (d_joined.alloc(jbase[n_chunks] * 12)) return V2M_ERR_CUDA;
		kp.joined = d_joined.as<u32>();
		V2M_CUDA_TRY(cudaMemcpy(d_jbase.p, jbase.data(), jbase.size() * 8, cudaMemcpyHostToDevice));
		if (attempt) cudaEventRecord(ev[2], nullptr);
		if (int const st = k3_dispatch(KW, 2, kp, n_chunks, smem, nullptr)) return st;
		cudaEventRecord(ev[3], nullptr);
		V2M_CUDA_TRY(cudaMemcpy(jcount.data(), d_jcount.p, n_seg * 4, cudaMemcpyDeviceToHost));
		bool fits(true);
		std::vector<u64> need(n_chunks + 1, 0);
		for (u32 j(0); j < n_chunks; ++j)
		{
			u64 sum(0);
			for (u32 c(chunk_cut[j]); c < chunk_cut[j + 1]; ++c) sum += jcount[c];
			need[j + 1] = need[j] + sum;
			if (sum > jbase[j + 1] - jbase[j]) fits = false;
		}
		if (fits) break;
		if (attempt) return fail(V2M_ERR_STATE, "segment_classes: joined pool overflow after resizing");
		jbase = need;
	}

	V2M_CUDA_TRY(cudaMemcpy(distinct.data(), d_distinct.p, n_seg * 4, cudaMemcpyDeviceToHost));
	V2M_CUDA_TRY(cudaMemcpy(first_class.data(), d_first.p, n_seg * 4, cudaMemcpyDeviceToHost));
	V2M_CUDA_TRY(cudaMemcpy(first_is_ref.data(), d_isref.p, n_seg, cudaMemcpyDeviceToHost));
	for (u64 c(0); c < n_seg; ++c) joined_offsets[c + 1] = joined_offsets[c] + jcount[c];
	joined.resize(3 * joined_offsets[n_seg]);
	for (u32 j(0); j < n_chunks; ++j)
	{
		u64 const lo(joined_offsets[chunk_cut[j]]), hi(joined_offsets[chunk_cut[j + 1]]);
		if (hi > lo)
			V2M_CUDA_TRY(cudaMemcpy(joined.data() + 3 * lo, d_joined.as<u32>() + 3 * jbase[j], (hi - lo) * 12, cudaMemcpyDeviceToHost));
	}
	if (want_rhs_eq)
	{
		rhs_eq.resize(n_seg * g.H);
		V2M_CUDA_TRY(cudaMemcpy(rhs_eq.data(), d_rhs.p, n_seg * g.H * 4, cudaMemcpyDeviceToHost));
	}
	V2M_CUDA_TRY(cudaDeviceSynchronize());
	{
		k3_timing &t(last_k3_timing());
		t = k3_timing{};
		cudaEventElapsedTime(&t.ms_local, ev[0], ev[1]);
		cudaEventElapsedTime(&t.ms_compose, ev[1], ev[2]);
		cudaEventElapsedTime(&t.ms_emit, ev[2], ev[3]);
		t.n_chunks = n_chunks; t.n_cuts = n_cuts; t.n_edges = g.E; t.joined_triples = joined_offsets[n_seg];
	}
	return V2M_OK;
}

} // namespace v2m


extern "C" int v2m_cuda_last_segment_classes_timing(float *ms3, uint64_t *counts4)
{
	v2m::k3_timing const &t(v2m::last_k3_timing());
	if (ms3) { ms3[0] = t.ms_local; ms3[1] = t.ms_compose; ms3[2] = t.ms_emit; }
	if (counts4) { counts4[0] = t.n_chunks; counts4[1] = t.n_cuts; counts4[2] = t.n_edges; counts4[3] = t.joined_triples; }
	return V2M_OK;
}

extern "C" int v2m_cuda_segment_classes(v2m_graph_t g, const uint64_t *cut_positions, uint64_t n_cuts,
	uint32_t *rhs_eq, uint32_t *distinct, uint32_t *first_class, uint8_t *first_is_ref,
	uint64_t *joined_offsets, uint32_t *joined, uint64_t joined_cap)
{
	using namespace v2m;
	if (!g || !joined_offsets) return fail(V2M_ERR_ARGUMENT, "segment_classes: NULL argument");
	std::vector<u32> v_rhs, v_distinct, v_first, v_joined;
	std::vector<u8> v_isref;
	std::vector<u64> v_off;
	if (int const st = segment_classes_host(*g, cut_positions, n_cuts, nullptr != rhs_eq, 0, v_rhs, v_distinct, v_first, v_isref, v_off, v_joined)) return st;
	std::copy(v_off.begin(), v_off.end(), joined_offsets);
	if (distinct) std::copy(v_distinct.begin(), v_distinct.end(), distinct);
	if (first_class) std::copy(v_first.begin(), v_first.end(), first_class);
	if (first_is_ref) std::copy(v_isref.begin(), v_isref.end(), first_is_ref);
	if (rhs_eq) std::copy(v_rhs.begin(), v_rhs.end(), rhs_eq);
	if (v_off.back() > joined_cap || (!joined && v_off.back()))
		return fail(V2M_ERR_ARGUMENT, "segment_classes: joined capacity too small (needed count is in joined_offsets[n_cuts - 1])");
	std::copy(v_joined.begin(), v_joined.end(), joined);
	return V2M_OK;
}
