// K4 — cut positions: founder_sequence_greedy_output::find_cut_positions (include/vcf2multialign/output.hh:123,
// libvcf2multialign/founder_sequence_greedy_output.cc:139-152) = find_initial_cut_positions_lambda_min
// (libvcf2multialign/find_cut_positions.cc:93-211).
//
// What the reference computes, restated. Candidate cut positions are the nodes no ALT edge spans
// (rightmost_seen_alt_edge_target <= node, :126) that come with a new edge index (:129); they depend on the graph only and
// are found on the host in one pass over the nodes. `cut_positions` (the reference's local vector) holds a sentinel
// {edge 0, node 0, score 0} followed by the candidates; below "index" means an index into that vector.
// At a candidate `cur` with edge index e, after the pBWT sweep (pbwt.hh:77-134) has consumed the columns [0, e), the
// reference walks the distinct divergence values in descending order, skipping the greatest (which is always e, the value
// the first zero / first one of the last column received), and keeps eq = the number of positions whose value is
// greater than the current one. Every value d maps to the index M(d) of the first entry with edge >= d (:143); each index
// is considered once, when it is first reached (:144-146), with eq = #{positions whose value maps to a greater index, or
// equals e}. The candidate's score is the minimum over the reached indices q that are at least min_distance aligned
// columns away (:150) of max(eq_q, score[q]), first minimum in descending q (:57-65: strict <); its initial value is the
// number of chromosome copies, which also makes the final step (:161-165, eq = all positions) a no-op.
//
// Since max(eq_q, score[q]) >= eq_q and eq_q grows as q descends, only the reached indices with eq_q < score matter, and
// scores are small (the segmentation height). The device therefore emits, for every candidate, the reached indices in
// descending order with their eq, cut off at eq >= cap; the host runs the (sequential, cheap) recurrence over these lists
// exactly as the reference does and asks for a larger cap if a list was cut off before the recurrence could stop.
//
// Device side: the same chunked three-pass sweep as K3 (pbwt_sweep.cuh) with the candidates' edge indices as chunk
// boundaries; pass C here is k4_lists_kernel. At a candidate, positions whose index is at most K4_W - 1 entries back
// ("young", the common case: neighbours in the pBWT order diverged recently) are counted in K4_W shared-memory bins, one
// CTA-wide scan turns the bins into (index, eq) pairs; only if those cover fewer than cap positions are older indices
// extracted one at a time by CTA-wide max reductions.
//
// Roofline: like K3 this is a latency chain of E / P column steps plus one list extraction per candidate, not an
// HBM-bound kernel; it reads the edge-major path matrix twice (pass A and pass C).
#include <chrono>
#include <cstdio>
#include <limits>
#include <memory>

#include "pbwt_sweep.cuh"

namespace v2m {

constexpr u32 K4_W = K3_THREADS;				// age bins: one per thread

struct k4_params
{
	k3_params		k;				// cut_edges = the candidates' edge indices (+ E when the last candidate is not at the sink's edge index)
	u32 const		*mtab;			// [E + 1] M(d): index of the first cut_positions entry with edge >= d
	u32				n_cand;			// candidates c = 0 … n_cand-1 ↔ index c + 1
	u32				cap;
	u32				*list_q;		// [n_cand][cap] reached indices, descending
	u32				*list_eq;		// [n_cand][cap] eq at each of them
	u32				*list_n;		// [n_cand] number of pairs
	u8				*list_complete;	// [n_cand] 1: the list holds every reached index
};

// CTA-wide reductions through 32 words of shared memory; one barrier each. Consecutive calls must alternate `slot`.
__device__ __forceinline__ u32 k4_block_add(u32 v, u32 *slot)
{
	unsigned const lane(threadIdx.x & 31), warp(threadIdx.x >> 5);
	v = __reduce_add_sync(FULL, v);
	if (0 == lane) slot[warp] = v;
	__syncthreads();
	return __reduce_add_sync(FULL, slot[lane]);
}

__device__ __forceinline__ u32 k4_block_max(u32 v, u32 *slot)
{
	unsigned const lane(threadIdx.x & 31), warp(threadIdx.x >> 5);
	v = __reduce_max_sync(FULL, v);
	if (0 == lane) slot[warp] = v;
	__syncthreads();
	return __reduce_max_sync(FULL, slot[lane]);
}

// The (index, eq) list of candidate c from the divergence array in shared memory. All threads of the CTA call it.
template <int KW>
__device__ __forceinline__ void k4_candidate(k4_params const &P, k3_smem const &s, u32 *bins, u32 const c)
{
	unsigned const lane(threadIdx.x & 31), warp(threadIdx.x >> 5);
	u32 const base(warp * KW * 32);
	u32 *ex(s.ex + 128);							// four 32-word slots
	u32 const H(P.k.H), cap(P.cap);
	u32 const cur(c + 1), e_t(P.k.cut_edges[c]);
	u32 *out_q(P.list_q + u64(c) * cap), *out_eq(P.list_eq + u64(c) * cap);

	u32 q[KW];
	u32 oldmask(0), n_top(0), n_young(0);
#pragma unroll
	for (int k = 0; k < KW; ++k)
	{
		u32 const i(base + k * 32 + lane);
		bool const valid(i < H);
		u32 const d(valid ? s.div[i] : 0);
		bool const top(valid && d == e_t);
		u32 const qq((valid && !top) ? P.mtab[d] : 0);
		u32 const age(cur - qq);
		bool const young(valid && !top && age < K4_W);
		if (young) atomicAdd(&bins[age], 1u);
		q[k] = qq;
		if (valid && !top && !young) oldmask |= 1u << k;
		n_top += top ? 1u : 0u;
		n_young += young ? 1u : 0u;
	}
	u32 const tot(k4_block_add(n_top | (n_young << 16), ex));				// H <= 16,384: both halves fit 16 bits
	u32 const top_total(tot & 0xffffu), young_total(tot >> 16);

	// Bins → pairs: thread t owns age t, i.e. index cur - t; descending index = ascending age.
	u32 const cnt(bins[threadIdx.x]);
	u32 const v(cnt | ((cnt ? 1u : 0u) << 16));
	u32 incl(v);
#pragma unroll
	for (unsigned d = 1; d < 32; d <<= 1)
	{
		u32 const t(__shfl_up_sync(FULL, incl, d));
		if (lane >= d) incl += t;
	}
	if (31 == lane) ex[32 + warp] = incl;
	__syncthreads();
	u32 const wt(ex[32 + lane]);
	u32 const excl(__reduce_add_sync(FULL, lane < warp ? wt : 0u) + incl - v);
	u32 const nz_total(__reduce_add_sync(FULL, wt) >> 16);
	u32 const eq_here(top_total + (excl & 0xffffu));
	bool const emit(0 != cnt && eq_here < cap);
	if (emit)
	{
		u32 const idx(excl >> 16);			// <= eq_here - top_total < cap
		out_q[idx] = cur - threadIdx.x;
		out_eq[idx] = eq_here;
	}
	if (cnt) bins[threadIdx.x] = 0;
	u32 n_entries(__syncthreads_count(emit));

	u32 eq(top_total + young_total);			// positions accounted for so far
	bool complete(n_entries == nz_total && eq == H);
	if (eq < cap && eq < H)
	{
		// Every young pair was emitted (their eq is below cap); continue with the older indices, greatest first.
		u32 last(0xffffffffu);
		for (u32 it(0); ; ++it)
		{
			u32 m(0);
#pragma unroll
			for (int k = 0; k < KW; ++k)
				if ((oldmask >> k & 1u) && q[k] < last) m = max(m, q[k] + 1);
			u32 const best(k4_block_max(m, ex + 64));
			if (0 == best) break;
			u32 mine(0);
#pragma unroll
			for (int k = 0; k < KW; ++k)
				if ((oldmask >> k & 1u) && q[k] + 1 == best) ++mine;
			u32 const count(k4_block_add(mine, ex + 96));
			if (0 == threadIdx.x) { out_q[n_entries] = best - 1; out_eq[n_entries] = eq; }
			++n_entries;
			eq += count;
			last = best - 1;
			if (eq >= cap) break;
		}
		complete = (eq == H);
	}
	if (0 == threadIdx.x)
	{
		P.list_n[c] = n_entries;
		P.list_complete[c] = complete ? 1 : 0;
	}
}


template <int KW>
__global__ void __launch_bounds__(K3_THREADS, 1)
k4_lists_kernel(k4_params const P)
{
	extern __shared__ __align__(128) u8 k3_raw[];
	k3_smem const s(k3_carve(k3_raw, P.k.Hp, P.k.Wh));
	k3_state const st(k3_make_state(s, P.k.H));
	u32 *bins(reinterpret_cast<u32 *>(s.rep[0]));			// rep[0] + rep[1] = 4 · Hp bytes >= 4 · K4_W
	u32 const j(blockIdx.x);
	u32 const c0(P.k.chunk_cut[j]), c1(P.k.chunk_cut[j + 1]);
	u32 const eb(P.k.cut_edges[c0]), ee(P.k.cut_edges[c1]);

	{
		unsigned short const *bperm(P.k.bperm + u64(j) * P.k.Hp);
		u32 const *bdiv(P.k.bdiv + u64(j) * P.k.Hp);
		for (u32 i(threadIdx.x); i < P.k.Hp; i += K3_THREADS) { s.perm[i] = bperm[i]; s.div[i] = bdiv[i]; }
		bins[threadIdx.x] = 0;
	}
	__syncthreads();
	k3_cols_begin(P.k, s, eb, ee);

	// The candidate at the chunk's first column belongs to this chunk (its state is the chunk's initial state); candidate 0
	// (edge index 0: nothing swept yet) is the host's.
	u32 c(c0);
	u32 next_edge(eb);
	k3_sweep<KW>(P.k, s, st, eb, ee,
		[&](u32 const k) {
			if (k != next_edge) return;
			if (0 != c) k4_candidate<KW>(P, s, bins, c);
			++c;
			next_edge = (c < c1) ? P.k.cut_edges[c] : 0xffffffffu;
		},
		[](u32, u32) {});
	// The last chunk also owns the candidate at the end of the sweep, if the last cut is a candidate.
	if (j + 1 == P.k.n_chunks && c1 < P.n_cand && 0 != c1) k4_candidate<KW>(P, s, bins, c1);
}


// The lists are short (the recurrence stops at eq >= score) while their slots are cap wide: pack them before they cross
// the host link. One warp per candidate; offsets = exclusive sum of list_n (host).
__global__ void __launch_bounds__(256)
k4_pack_lists_kernel(u32 const *__restrict__ list_q, u32 const *__restrict__ list_eq, u32 const *__restrict__ list_n,
	u64 const *__restrict__ offsets, u32 const cap, u32 const n_cand, u32 *__restrict__ out_q, unsigned short *__restrict__ out_eq)
{
	u32 const lane(threadIdx.x & 31), n_warps(gridDim.x * (blockDim.x >> 5));
	for (u32 c(blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)); c < n_cand; c += n_warps)
	{
		u32 const n(list_n[c]);
		u64 const src(u64(c) * cap), dst(offsets[c]);
		for (u32 i(lane); i < n; i += 32)
		{
			out_q[dst + i] = list_q[src + i];
			out_eq[dst + i] = (unsigned short) list_eq[src + i];		// eq < cap <= H + 1 <= 16,385
		}
	}
}

} // namespace v2m


// ---------------------------------------------------------------------------------------------------------
// Host driver
// ---------------------------------------------------------------------------------------------------------
namespace v2m {

namespace {

	template <int KW>
	int k4_launch(k4_params const &P, u32 const n_ctas, u32 const smem, cudaStream_t stream)
	{
		V2M_CUDA_TRY(cudaFuncSetAttribute(k4_lists_kernel<KW>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
		k4_lists_kernel<KW><<<n_ctas, K3_THREADS, smem, stream>>>(P);
		V2M_CHECK_LAUNCH("k4_lists_kernel");
		return 0;
	}

	// pass 0 / 1: the shared passes A / B; pass 2: k4_lists_kernel.
	int k4_dispatch(u32 const KW, int const pass, k4_params const &P, u32 const n_ctas, u32 const smem, cudaStream_t stream)
	{
#define V2M_K4_CASE(kw) case kw: return pass < 2 ? k3_launch_ab<kw>(pass, P.k, n_ctas, smem, stream) : k4_launch<kw>(P, n_ctas, smem, stream)
		switch (KW)
		{
			V2M_K4_CASE(1); V2M_K4_CASE(2); V2M_K4_CASE(3); V2M_K4_CASE(4);
			V2M_K4_CASE(5); V2M_K4_CASE(6); V2M_K4_CASE(7); V2M_K4_CASE(8);
			V2M_K4_CASE(10); V2M_K4_CASE(12); V2M_K4_CASE(14); V2M_K4_CASE(16);
			default: return fail(V2M_ERR_ARGUMENT, "find_cut_positions: more than 16384 chromosome copies are not supported");
		}
#undef V2M_K4_CASE
	}

	constexpr u64 EDGE_MAX = std::numeric_limits<u64>::max();					// variant_graph.hh:53
	constexpr u32 CUT_POSITION_SCORE_MAX = std::numeric_limits<u32>::max();	// find_cut_positions.hh:17

	// find_cut_positions.cc:30-42
	struct cut_position { u64 edge{}; u64 prev_edge{EDGE_MAX}; u64 node{}; u32 score{}; u32 prev_index{}; };		// prev_index: where prev_edge was found
}

struct k4_timing { float ms_local{}, ms_compose{}, ms_lists{}, ms_host{}; u64 n_chunks{}, n_candidates{}, n_edges{}, list_entries{}, cap{}, attempts{}; };
k4_timing &last_k4_timing() { thread_local k4_timing t; return t; }


int find_cut_positions_host(graph_holder const &gh, u64 const min_distance, std::vector<u64> &out_cut_positions, u32 &out_score)
{
	device_graph const &g(gh.g);
	out_cut_positions.clear();
	out_score = CUT_POSITION_SCORE_MAX;
	if (0 == g.H) return fail(V2M_ERR_ARGUMENT, "find_cut_positions: no chromosome copies (the reference reads an empty divergence map, find_cut_positions.cc:137)");
	u32 const KW(k3_words_per_warp(g.H));
	if (0 == KW) return fail(V2M_ERR_ARGUMENT, "find_cut_positions: more than 16384 chromosome copies are not supported");
	u32 const Hp(KW * K3_WARPS * 32);
	if (g.E >= 0xfffffff0u) return fail(V2M_ERR_ARGUMENT, "find_cut_positions: edge indices do not fit 32 bits");
	auto const t_begin(std::chrono::steady_clock::now());
	bool const trace(nullptr != std::getenv("V2M_TRACE"));		// host-side phase times on stderr
	auto mark([&, t_last = t_begin](char const *what) mutable {
		if (!trace) return;
		cudaDeviceSynchronize();
		auto const now(std::chrono::steady_clock::now());
		std::fprintf(stderr, "[find_cut_positions] %-28s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
		t_last = now;
	});

	// Candidates (find_cut_positions.cc:126-132, 170-176): a function of the graph alone.
	std::vector<cut_position> cuts;
	cuts.push_back(cut_position{0, EDGE_MAX, 0, 0});
	{
		u64 rightmost(0), prev_id(EDGE_MAX);
		for (u64 node(0); node < g.N; ++node)
		{
			u64 const edge_idx(gh.csum[node]);
			if (rightmost <= node && prev_id != edge_idx)
			{
				cuts.push_back(cut_position{edge_idx, EDGE_MAX, node, g.H});
				prev_id = edge_idx;
			}
			for (u64 e(edge_idx); e < gh.csum[node + 1]; ++e) rightmost = std::max(rightmost, gh.targets[e]);
		}
	}
	u64 const n_cand(cuts.size() - 1);
	mark("candidates");
	k4_timing timing{};
	timing.n_candidates = n_cand; timing.n_edges = g.E;

	if (n_cand >= 2)		// candidate 0 sits at edge index 0 and never improves on its initial score (see the header comment)
	{
		// cut_edges for the sweep: the candidates' edge indices, closed by E.
		std::vector<u32> cut_edges(n_cand);
		for (u64 c(0); c < n_cand; ++c) cut_edges[c] = u32(cuts[c + 1].edge);
		if (cut_edges.back() != u32(g.E)) cut_edges.push_back(u32(g.E));
		u64 const n_cuts(cut_edges.size()), n_seg(n_cuts - 1);

		// M(d) for d in [0, E]: M(0) = 0 (the sentinel); otherwise 1 + #{candidates with edge < d}.
		std::vector<u32> mtab(g.E + 1);
		{
			u64 c(0);
			mtab[0] = 0;
			for (u64 d(1); d <= g.E; ++d)
			{
				while (c < n_cand && cuts[c + 1].edge < d) ++c;
				mtab[d] = u32(1 + c);
			}
		}

		int dev(0), sms(0);
		V2M_CUDA_TRY(cudaGetDevice(&dev));
		V2M_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
		u32 const smem(k3_smem_bytes(Hp, u32(g.Wh)));
		if (int const rc = k3_check_smem(smem, "find_cut_positions")) return rc;
		u64 P(sms);
		if (char const *env = std::getenv("V2M_K3_CHUNKS"))		// tests force small / odd chunk counts
			if (int const v = std::atoi(env); v > 0) P = u64(v);
		std::vector<u32> const chunk_cut(k3_make_chunks(cut_edges, n_seg, P, 2));
		u32 const n_chunks(u32(chunk_cut.size() - 1));
		timing.n_chunks = n_chunks;

		dev_buf d_chunk_cut, d_cut_edges, d_cls, d_classdiv, d_ncls, d_bperm, d_bdiv, d_mtab, d_q, d_eq, d_n, d_complete;
		if (d_chunk_cut.alloc(chunk_cut.size() * 4) || d_cut_edges.alloc(n_cuts * 4) || d_cls.alloc(u64(n_chunks) * Hp * 2) || d_classdiv.alloc(u64(n_chunks) * Hp * 4)
			|| d_ncls.alloc(n_chunks * 4) || d_bperm.alloc(u64(n_chunks) * Hp * 2) || d_bdiv.alloc(u64(n_chunks) * Hp * 4)
			|| d_mtab.alloc(mtab.size() * 4) || d_n.alloc(n_cand * 4) || d_complete.alloc(n_cand))
			return V2M_ERR_CUDA;
		V2M_CUDA_TRY(cudaMemcpy(d_chunk_cut.p, chunk_cut.data(), chunk_cut.size() * 4, cudaMemcpyHostToDevice));
		V2M_CUDA_TRY(cudaMemcpy(d_cut_edges.p, cut_edges.data(), n_cuts * 4, cudaMemcpyHostToDevice));
		V2M_CUDA_TRY(cudaMemcpy(d_mtab.p, mtab.data(), mtab.size() * 4, cudaMemcpyHostToDevice));
		mark("tables, chunks, uploads");

		k4_params kp{};
		kp.k.paths_eh = g.paths_eh; kp.k.Wh = u32(g.Wh); kp.k.H = g.H; kp.k.Hp = Hp; kp.k.n_chunks = n_chunks;
		kp.k.chunk_cut = d_chunk_cut.as<u32>(); kp.k.cut_edges = d_cut_edges.as<u32>(); kp.k.n_cuts = u32(n_cuts);
		kp.k.cls = d_cls.as<unsigned short>(); kp.k.classdiv = d_classdiv.as<u32>(); kp.k.n_classes = d_ncls.as<u32>();
		kp.k.bperm = d_bperm.as<unsigned short>(); kp.k.bdiv = d_bdiv.as<u32>();
		kp.mtab = d_mtab.as<u32>(); kp.n_cand = u32(n_cand);
		kp.list_n = d_n.as<u32>(); kp.list_complete = d_complete.as<u8>();

		cudaEvent_t ev[4];
		for (auto &e : ev) V2M_CUDA_TRY(cudaEventCreate(&e));
		struct ev_guard { cudaEvent_t *e; ~ev_guard() { for (int i(0); i < 4; ++i) cudaEventDestroy(e[i]); } } const guard{ev};
		cudaEventRecord(ev[0], nullptr);
		if (n_chunks > 1)
			if (int const st = k4_dispatch(KW, 0, kp, n_chunks - 1, smem, nullptr)) return st;
		cudaEventRecord(ev[1], nullptr);
		if (int const st = k4_dispatch(KW, 1, kp, 1, smem, nullptr)) return st;
		cudaEventRecord(ev[2], nullptr);
		mark("passes A, B");

		std::vector<u32> list_n(n_cand);
		std::vector<u64> list_off(n_cand + 1);
		std::unique_ptr<u32[]> list_q;
		std::unique_ptr<unsigned short[]> list_eq;
		std::vector<u8> list_complete(n_cand);
		dev_buf d_off, d_pq, d_peq;
		u32 cap(std::min<u32>(g.H + 1, 64));
		if (char const *env = std::getenv("V2M_K4_CAP"))			// tests force tiny caps to exercise the retry
			if (int const v = std::atoi(env); v > 0) cap = std::min<u32>(g.H + 1, u32(v));
		for (;;)
		{
			++timing.attempts;
			if (d_q.alloc(n_cand * cap * 4) || d_eq.alloc(n_cand * cap * 4)) return V2M_ERR_CUDA;
			kp.cap = cap; kp.list_q = d_q.as<u32>(); kp.list_eq = d_eq.as<u32>();
			V2M_CUDA_TRY(cudaMemset(d_n.p, 0, n_cand * 4));
			V2M_CUDA_TRY(cudaMemset(d_complete.p, 1, n_cand));
			cudaEventRecord(ev[2], nullptr);
			if (int const st = k4_dispatch(KW, 2, kp, n_chunks, smem, nullptr)) return st;
			cudaEventRecord(ev[3], nullptr);
			mark("lists");
			V2M_CUDA_TRY(cudaMemcpy(list_n.data(), d_n.p, n_cand * 4, cudaMemcpyDeviceToHost));
			V2M_CUDA_TRY(cudaMemcpy(list_complete.data(), d_complete.p, n_cand, cudaMemcpyDeviceToHost));
			list_off[0] = 0;
			for (u64 c(0); c < n_cand; ++c) list_off[c + 1] = list_off[c] + list_n[c];
			u64 const n_pairs(list_off[n_cand]);
			if (d_off.alloc((n_cand + 1) * 8) || d_pq.alloc(std::max<u64>(1, n_pairs) * 4) || d_peq.alloc(std::max<u64>(1, n_pairs) * 2)) return V2M_ERR_CUDA;
			V2M_CUDA_TRY(cudaMemcpy(d_off.p, list_off.data(), (n_cand + 1) * 8, cudaMemcpyHostToDevice));
			k4_pack_lists_kernel<<<u32(sms) * 8, 256>>>(d_q.as<u32>(), d_eq.as<u32>(), d_n.as<u32>(), d_off.as<u64>(), cap, u32(n_cand), d_pq.as<u32>(), d_peq.as<unsigned short>());
			V2M_CHECK_LAUNCH("k4_pack_lists_kernel");
			list_q.reset(new u32[std::max<u64>(1, n_pairs)]);				// not value-initialised: the copies fill them
			list_eq.reset(new unsigned short[std::max<u64>(1, n_pairs)]);
			V2M_CUDA_TRY(cudaMemcpy(list_q.get(), d_pq.p, n_pairs * 4, cudaMemcpyDeviceToHost));
			V2M_CUDA_TRY(cudaMemcpy(list_eq.get(), d_peq.p, n_pairs * 2, cudaMemcpyDeviceToHost));
			mark("pack + copy lists");

			// The recurrence (find_cut_positions.cc:131-166) over the lists.
			bool need_more(false);
			u64 entries(0);
			for (u64 c(1); c < n_cand; ++c)
			{
				cut_position &cur(cuts[c + 1]);
				cur.score = g.H; cur.prev_edge = EDGE_MAX;
				u32 const n(list_n[c]);
				u32 const *lq(list_q.get() + list_off[c]);
				unsigned short const *le(list_eq.get() + list_off[c]);
				entries += n;
				bool stopped(false);
				for (u32 i(0); i < n; ++i)
				{
					u32 const eq(le[i]);
					if (eq >= cur.score) { stopped = true; break; }		// max(eq, ·) >= eq, and eq only grows from here on
					cut_position const &prev(cuts[lq[i]]);
					if (min_distance <= gh.alnpos[cur.node] - gh.alnpos[prev.node])
					{
						u32 const cand(std::max(eq, prev.score));			// cut_position::update_if_needed, :57-65
						if (cand < cur.score) { cur.score = cand; cur.prev_edge = prev.edge; cur.prev_index = lq[i]; }
					}
				}
				// Pairs left out have eq >= cap: harmless iff the score is already at most cap.
				if (!stopped && !list_complete[c] && cur.score > cap) { need_more = true; break; }
			}
			timing.list_entries = entries; timing.cap = cap;
			mark("recurrence");
			if (!need_more) break;
			if (cap > g.H) return fail(V2M_ERR_STATE, "find_cut_positions: incomplete candidate list at full capacity");
			cap = (cap < 512 && g.H + 1 > 512) ? 512 : g.H + 1;
		}
		V2M_CUDA_TRY(cudaDeviceSynchronize());
		{
			// The states at the chunk boundaries stay with the graph for find_matchings (device_graph.hh: sweep_states).
			std::lock_guard<std::mutex> const lock(gh.sweep_mutex);
			gh.sweep.clear();
			gh.sweep.edges.resize(n_chunks);
			for (u32 j(0); j < n_chunks; ++j) gh.sweep.edges[j] = cut_edges[chunk_cut[j]];
			gh.sweep.perm = static_cast<unsigned short *>(d_bperm.release());
			gh.sweep.div = static_cast<u32 *>(d_bdiv.release());
			gh.sweep.Hp = Hp;
		}
		cudaEventElapsedTime(&timing.ms_local, ev[0], ev[1]);
		cudaEventElapsedTime(&timing.ms_compose, ev[1], ev[2]);
		cudaEventElapsedTime(&timing.ms_lists, ev[2], ev[3]);
	}

	// find_cut_positions.cc:181-210: trace the solution back from the last candidate.
	if (cuts.size() <= 1) { last_k4_timing() = timing; return V2M_OK; }
	{
		std::size_t it(cuts.size() - 1);
		out_score = cuts[it].score;
		for (;;)
		{
			out_cut_positions.push_back(cuts[it].node);
			if (EDGE_MAX == cuts[it].prev_edge) break;
			it = cuts[it].prev_index;		// the reference finds prev_edge by lower_bound (:190-196); the entries M() reaches have distinct edges, so this is that entry
		}
		if (0 != out_cut_positions.back()) out_cut_positions.push_back(0);
		std::reverse(out_cut_positions.begin(), out_cut_positions.end());
		if (out_cut_positions.back() != g.N - 1) out_cut_positions.back() = g.N - 1;
	}
	mark("trace-back");
	timing.ms_host = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
	last_k4_timing() = timing;
	return V2M_OK;
}

} // namespace v2m


extern "C" int v2m_cuda_find_cut_positions(v2m_graph_t g, uint64_t min_distance, uint64_t *cut_positions, uint64_t cap,
	uint64_t *n_cuts, uint32_t *score, int *found)
{
	using namespace v2m;
	if (!g || !n_cuts || !found) return fail(V2M_ERR_ARGUMENT, "find_cut_positions: NULL argument");
	*found = 0; *n_cuts = 0;
	std::vector<u64> cuts;
	u32 sc(0);
	if (int const st = find_cut_positions_host(*g, min_distance, cuts, sc)) return st;
	if (score) *score = sc;
	*n_cuts = cuts.size();
	if (std::numeric_limits<u32>::max() == sc) return V2M_OK;		// the reference returns false (founder_…cc:145-146)
	if (cuts.size() > cap || (!cut_positions && !cuts.empty()))
		return fail(V2M_ERR_ARGUMENT, "find_cut_positions: cut position capacity too small (needed count is in *n_cuts)");
	std::copy(cuts.begin(), cuts.end(), cut_positions);
	*found = 1;
	return V2M_OK;
}

extern "C" int v2m_cuda_last_cut_positions_timing(float *ms4, uint64_t *counts6)
{
	v2m::k4_timing const &t(v2m::last_k4_timing());
	if (ms4) { ms4[0] = t.ms_local; ms4[1] = t.ms_compose; ms4[2] = t.ms_lists; ms4[3] = t.ms_host; }
	if (counts6) { counts6[0] = t.n_chunks; counts6[1] = t.n_candidates; counts6[2] = t.n_edges; counts6[3] = t.list_entries; counts6[4] = t.cap; counts6[5] = t.attempts; }
	return V2M_OK;
}
