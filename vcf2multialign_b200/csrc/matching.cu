// Host half of founder_sequence_greedy_output::find_matchings (libvcf2multialign/founder_sequence_greedy_output.cc:154-512):
// the per-cut class data comes from K3 (founders.cu) and the greedy assignment of founders to class representatives runs
// here, on the host, with the same std::sort the reference uses — the assignment is a short sequential pass per cut whose tie
// behaviour depends on std::sort's introsort and on std::multimap's ordering of equal keys (SURVEY.md §7 "hard parts"), so
// it is kept as host C++ by design. The reference's std::multimap<class, founder> holds at most founder_count entries and is
// rebuilt at every cut (27 M node allocations on c3); it is replaced by a sorted vector with the same observable behaviour.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <utility>
#include <vector>

#include "../../include/v2m_cuda.h"
#include "common.cuh"
#include "device_graph.hh"

namespace v2m {

int segment_classes_host(graph_holder const &gh, u64 const *cut_positions, u64 const n_cuts, bool const want_rhs_eq, int const max_chunks,
	std::vector<u32> &rhs_eq, std::vector<u32> &distinct, std::vector<u32> &first_class, std::vector<u8> &first_is_ref,
	std::vector<u64> &joined_offsets, std::vector<u32> &joined);

namespace {

	// joined_path_eq_class (founder_sequence_greedy_output.cc:48-69): ordered by size only.
	struct joined_class
	{
		u32 lhs, rhs, size;
		bool operator<(joined_class const &other) const { return size < other.size; }
	};

	// std::multimap<u32, u32> (class representative → founder) as the reference uses it (founder_…cc:174, 283, 340-360,
	// 441-447), without a tree: the keys are chromosome copy indices, so every key owns a FIFO of founders threaded through
	// m_next and found by direct indexing. What the reference's container guarantees and the matching depends on:
	// insert places a pair after the pairs with an equal key (libstdc++ _M_get_insert_equal_pos descends to the right on
	// equality) → FIFO per key; find returns the first pair with the key → the FIFO's head; begin is the first pair of the
	// smallest key → the head of the smallest key with a non-empty FIFO (keys sorted on first use, then a cursor that only
	// moves forward, since between two rebuilds pairs are only removed).
	class founders_by_class
	{
		static constexpr u32 NONE = 0xffffffffu;
		std::vector<u32>	m_stamp, m_head, m_tail;		// by key
		std::vector<u32>	m_next;							// by founder
		std::vector<u32>	m_keys;							// distinct keys of this epoch
		u32					m_epoch{1}, m_count{0}, m_cursor{0};
		bool				m_sorted{false};
		// A founder that has not been given a class yet is filed under PLOIDY_MAX (:170-172, 441-447), the greatest key:
		// it gets the slot after the last chromosome copy.
		u32 index_of(u32 const key) const { return std::min<u32>(key, u32(m_stamp.size() - 1)); }
	public:
		founders_by_class(u32 const n_copies, u32 const n_founders): m_stamp(n_copies + 1, 0), m_head(n_copies + 1), m_tail(n_copies + 1), m_next(n_founders, NONE) {}
		bool empty() const { return 0 == m_count; }
		void clear()
		{
			if (0 == ++m_epoch) { std::fill(m_stamp.begin(), m_stamp.end(), 0); m_epoch = 1; }
			m_keys.clear(); m_count = 0; m_cursor = 0; m_sorted = false;
		}
		void emplace(u32 const class_key, u32 const founder)
		{
			u32 const key(index_of(class_key));
			m_next[founder] = NONE;
			if (m_stamp[key] != m_epoch)
			{
				m_stamp[key] = m_epoch; m_head[key] = founder; m_tail[key] = founder;
				m_keys.push_back(key); m_sorted = false; m_cursor = 0;
			}
			else if (NONE == m_head[key]) { m_head[key] = founder; m_tail[key] = founder; }
			else { m_next[m_tail[key]] = founder; m_tail[key] = founder; }
			++m_count;
		}
		// find(key) + erase: the first founder filed under key.
		bool pop(u32 const class_key, u32 &founder)
		{
			u32 const key(index_of(class_key));
			if (m_stamp[key] != m_epoch || NONE == m_head[key]) return false;
			founder = m_head[key];
			m_head[key] = m_next[founder];
			--m_count;
			return true;
		}
		// begin() + erase.
		bool pop_first(u32 &founder)
		{
			if (0 == m_count) return false;
			if (!m_sorted) { std::sort(m_keys.begin(), m_keys.end()); m_sorted = true; m_cursor = 0; }
			while (NONE == m_head[m_keys[m_cursor]]) ++m_cursor;
			return pop(m_keys[m_cursor], founder);
		}
	};

	// The reference's "reserved" bit vector (founder_…cc:175, 333), cleared per cut: a stamp per chromosome copy instead of an
	// O(H) fill per cut.
	class stamp_set
	{
		std::vector<u32>	m_stamp;
		u32					m_epoch{1};
	public:
		explicit stamp_set(u32 const n): m_stamp(n, 0) {}
		void clear() { if (0 == ++m_epoch) { std::fill(m_stamp.begin(), m_stamp.end(), 0); m_epoch = 1; } }
		bool test(u32 const i) const { return m_stamp[i] == m_epoch; }
		void set(u32 const i) { m_stamp[i] = m_epoch; }
	};

	// State carried from cut to cut (founder_sequence_greedy_output.cc:174-176).
	class founder_matcher
	{
		u32							m_founders;
		u64							m_segments;
		u32							*m_assigned;			// column-major: (segment, founder) → m_assigned[founder * segments + segment]
		founders_by_class			m_by_class;				// class representative on the current lhs → founder
		stamp_set					m_reserved;				// by chromosome copy
		std::vector<u32>			m_arbitrary_rhs;

		// The row of the segment being matched is kept here and leaves for m_assigned in blocks of BLOCK segments, so that the
		// column-major matrix is written in runs instead of one word per founder per cut.
		static constexpr u64		BLOCK = 512;
		std::vector<u32>			m_row;					// [founder]
		std::vector<u32>			m_block;				// [segment - m_block_first][founder]
		u64							m_block_first{0}, m_block_fill{0};

		void flush_block()
		{
			for (u32 f(0); f < m_founders; ++f)
			{
				u32 *dst(m_assigned + u64(f) * m_segments + m_block_first);
				for (u64 i(0); i < m_block_fill; ++i) dst[i] = m_block[i * m_founders + f];
			}
			m_block_first += m_block_fill;
			m_block_fill = 0;
		}

		// Segments are finished in order, each exactly once.
		void finish_row()
		{
			std::copy(m_row.begin(), m_row.end(), m_block.begin() + m_block_fill * m_founders);
			if (BLOCK == ++m_block_fill) flush_block();
		}

	public:
		founder_matcher(u32 copies, u32 founders, u64 segments, u32 *assigned):
			m_founders(founders), m_segments(segments), m_assigned(assigned), m_by_class(copies, founders), m_reserved(copies),
			m_row(founders, V2M_PLOIDY_MAX), m_block(BLOCK * founders) {}

		void finish() { flush_block(); }

		// :268-320 — founders of the first segment, largest joined classes first; every lhs class gets one founder
		// before any class gets a second one.
		int seed_first_segment(std::vector<joined_class> const &by_size, u32 const lhs_distinct)
		{
			u32 free_founders(m_founders);
			u32 to_reserve(std::min(free_founders, lhs_distinct));
			free_founders -= to_reserve;
			u32 next_founder(0);
			auto const place([&](joined_class const &jc) {
				m_by_class.emplace(jc.lhs, next_founder);
				m_row[next_founder] = jc.lhs;
				++next_founder;
			});
			for (auto it(by_size.rbegin()); it != by_size.rend(); ++it)
			{
				if (m_reserved.test(it->lhs))
				{
					if (free_founders) { --free_founders; place(*it); }
				}
				else if (to_reserve)
				{
					--to_reserve;
					m_reserved.set(it->lhs);
					place(*it);
				}
			}
			// :309-319 cycles over the classes until every founder is placed; with no class left it would spin forever.
			if (free_founders && by_size.empty())
				return fail(V2M_ERR_STATE, "find_matchings: no joined class left for the remaining founders (the reference does not terminate on this input, founder_sequence_greedy_output.cc:309-319)");
			while (free_founders)
				for (auto it(by_size.rbegin()); it != by_size.rend() && free_founders; ++it)
				{
					--free_founders;
					place(*it);
				}
			finish_row();
			return V2M_OK;
		}

		// :331-447 — carry the founders across the cut that ends segment `segment - 1` (cut index `segment`).
		int extend(u64 const segment, std::vector<joined_class> const &by_size, u32 const rhs_distinct)
		{
			m_reserved.clear();
			m_arbitrary_rhs.clear();
			std::fill(m_row.begin(), m_row.end(), V2M_PLOIDY_MAX);		// :170-172
			u32 free_founders(m_founders);
			u32 to_reserve(std::min(free_founders, rhs_distinct));
			free_founders -= to_reserve;

			auto const follow([&](joined_class const &jc) -> bool {			// :340-352
				u32 founder;
				if (!m_by_class.pop(jc.lhs, founder)) return false;
				m_row[founder] = jc.rhs;
				return true;
			});
			auto const take_any([&](u32 const rhs) -> bool {					// :354-360
				u32 founder;
				if (!m_by_class.pop_first(founder)) return false;
				m_row[founder] = rhs;
				return true;
			});

			// Steps 1–3 (:363-413).
			bool first_round(true), done(false);
			while (!done)
			{
				bool progressed(false);
				for (auto it(by_size.rbegin()); it != by_size.rend(); ++it)
				{
					if (m_reserved.test(it->rhs))
					{
						if (free_founders)
						{
							if (follow(*it)) { progressed = true; --free_founders; }
						}
						else if (!first_round) { done = true; break; }			// :383-386
					}
					else if (to_reserve)
					{
						--to_reserve;
						if (follow(*it)) m_reserved.set(it->rhs);
						else m_arbitrary_rhs.push_back(it->rhs);
					}
				}
				if (done || !free_founders) break;
				if (first_round) { first_round = false; continue; }
				if (!progressed) break;
			}

			// Step 4 (:417-426).
			for (u32 const rhs : m_arbitrary_rhs)
				if (!m_reserved.test(rhs))
				{
					if (!take_any(rhs))
						return fail(V2M_ERR_STATE, "find_matchings: no founder left to connect arbitrarily (the reference asserts, founder_sequence_greedy_output.cc:355)");
					m_reserved.set(rhs);
				}

			// Step 5 (:429-438).
			if (!m_by_class.empty() && by_size.empty())
				return fail(V2M_ERR_STATE, "find_matchings: no joined class left for the unconnected founders (the reference does not terminate on this input, founder_sequence_greedy_output.cc:429-438)");
			while (!m_by_class.empty())
				for (auto it(by_size.rbegin()); it != by_size.rend() && !m_by_class.empty(); ++it)
					take_any(it->rhs);

			// :441-447
			m_by_class.clear();
			for (u32 f(0); f < m_founders; ++f)
				m_by_class.emplace(m_row[f], f);
			finish_row();
			return V2M_OK;
		}
	};
}

} // namespace v2m


extern "C" int v2m_cuda_find_matchings(v2m_graph_t g, const uint64_t *cut_positions, uint64_t n_cuts, uint32_t founder_count,
	int keep_ref_edges, uint32_t *assigned, int *found)
{
	using namespace v2m;
	if (!g || !found) return fail(V2M_ERR_ARGUMENT, "find_matchings: NULL argument");
	*found = 0;
	if (n_cuts < 2 || 0 == g->g.H) return V2M_OK;		// :162-166
	if (!cut_positions || (founder_count && !assigned)) return fail(V2M_ERR_ARGUMENT, "find_matchings: NULL argument");
	u64 const segments(n_cuts - 1);
	std::fill(assigned, assigned + segments * founder_count, V2M_PLOIDY_MAX);		// :170-172

	bool const trace(nullptr != std::getenv("V2M_TRACE"));		// host-side phase times on stderr
	auto t_last(std::chrono::steady_clock::now());
	auto mark([&](char const *what) {
		if (!trace) return;
		auto const now(std::chrono::steady_clock::now());
		std::fprintf(stderr, "[find_matchings] %-28s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
		t_last = now;
	});
	std::vector<u32> rhs_eq, distinct, first_class, joined;
	std::vector<u8> first_is_ref;
	std::vector<u64> offsets;
	if (int const st = segment_classes_host(*g, cut_positions, n_cuts, false, 0, rhs_eq, distinct, first_class, first_is_ref, offsets, joined)) return st;
	mark("segment classes (K3 + copies)");

	std::vector<joined_class> by_size;
	auto const classes_of_cut([&](u64 const c) {		// cut index c >= 1 → its joined classes, in permutation order
		by_size.clear();
		for (u64 i(offsets[c - 1]); i < offsets[c]; ++i) by_size.push_back({joined[3 * i], joined[3 * i + 1], joined[3 * i + 2]});
	});

	if (2 == n_cuts)
	{
		// The trivial case (:474-509): one segment, founders take the largest classes.
		classes_of_cut(1);
		std::sort(by_size.begin(), by_size.end());
		// rhs_first_path_is_ref was reset to true when the only cut was handled (:456-457) and no edge follows the sink,
		// so the test at :500 reduces to !m_should_keep_ref_edges.
		if (!keep_ref_edges)
			std::erase_if(by_size, [&](joined_class const &jc) { return first_class[0] == jc.rhs; });
		u32 f(0);
		for (auto it(by_size.rbegin()); it != by_size.rend() && f < founder_count; ++it, ++f) assigned[f] = it->rhs;
		*found = 1;
		return V2M_OK;
	}

	founder_matcher matcher(g->g.H, founder_count, segments, assigned);
	for (u64 c(2); c < n_cuts; ++c)
	{
		// In the reference's loop this is cut_pos_idx = c - 1: lhs = segment c - 2, rhs = segment c - 1.
		classes_of_cut(c);
		std::sort(by_size.begin(), by_size.end());														// :258
		if (!keep_ref_edges && first_is_ref[c - 2] && first_is_ref[c - 1])								// :261-266
			std::erase_if(by_size, [&](joined_class const &jc) { return first_class[c - 2] == jc.lhs && first_class[c - 1] == jc.rhs; });
		if (2 == c)
			if (int const st = matcher.seed_first_segment(by_size, distinct[0])) return st;
		if (int const st = matcher.extend(c - 1, by_size, distinct[c - 1])) return st;
	}
	matcher.finish();
	mark("greedy assignment");
	*found = 1;
	return V2M_OK;
}
