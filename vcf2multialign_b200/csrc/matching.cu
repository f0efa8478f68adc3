// Host half of founder_sequence_greedy_output::find_matchings (libvcf2multialign/founder_sequence_greedy_output.cc:154-512):
// the per-cut class data comes from K3 (founders.cu) and the greedy assignment of founders to class representatives runs
// here, on the host, with the same std::sort the reference uses — the assignment is a short sequential pass per cut whose tie
// behaviour depends on std::sort's introsort and on std::multimap's ordering of equal keys (SURVEY.md §7 "hard parts"), so
// it is kept as host C++ by design. The reference's std::multimap<class, founder> holds at most founder_count entries and is
// rebuilt at every cut (27 M node allocations on c3); it is replaced by a sorted vector with the same observable behaviour.
#include <algorithm>
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

	// std::multimap<u32, u32> as the reference uses it (founder_…cc:174, 283, 340-360, 441-447), on a sorted vector:
	// insert places the pair after the pairs with an equal key (libstdc++ _M_get_insert_equal_pos descends to the right on
	// equality), find returns the first pair with the key (lower bound), begin is the first pair of the smallest key.
	class flat_multimap
	{
		std::vector<std::pair<u32, u32>>	m_items;
	public:
		typedef std::vector<std::pair<u32, u32>>::iterator iterator;
		iterator begin() { return m_items.begin(); }
		iterator end() { return m_items.end(); }
		bool empty() const { return m_items.empty(); }
		void clear() { m_items.clear(); }
		void erase(iterator it) { m_items.erase(it); }
		void emplace(u32 const key, u32 const value)
		{
			auto const it(std::upper_bound(m_items.begin(), m_items.end(), key, [](u32 const k, std::pair<u32, u32> const &p) { return k < p.first; }));
			m_items.insert(it, {key, value});
		}
		iterator find(u32 const key)
		{
			auto const it(std::lower_bound(m_items.begin(), m_items.end(), key, [](std::pair<u32, u32> const &p, u32 const k) { return p.first < k; }));
			return (it != m_items.end() && it->first == key) ? it : m_items.end();
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
		flat_multimap				m_by_class;				// class representative on the current lhs → founder
		stamp_set					m_reserved;				// by chromosome copy
		std::vector<u32>			m_arbitrary_rhs;

		u32 &slot(u64 const segment, u32 const founder) { return m_assigned[u64(founder) * m_segments + segment]; }

	public:
		founder_matcher(u32 copies, u32 founders, u64 segments, u32 *assigned):
			m_founders(founders), m_segments(segments), m_assigned(assigned), m_reserved(copies) {}

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
				slot(0, next_founder) = jc.lhs;
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
			return V2M_OK;
		}

		// :331-447 — carry the founders across the cut that ends segment `segment - 1` (cut index `segment`).
		int extend(u64 const segment, std::vector<joined_class> const &by_size, u32 const rhs_distinct)
		{
			m_reserved.clear();
			m_arbitrary_rhs.clear();
			u32 free_founders(m_founders);
			u32 to_reserve(std::min(free_founders, rhs_distinct));
			free_founders -= to_reserve;

			auto const follow([&](joined_class const &jc) -> bool {			// :340-352
				auto const it(m_by_class.find(jc.lhs));
				if (m_by_class.end() == it) return false;
				u32 const founder(it->second);
				m_by_class.erase(it);
				slot(segment, founder) = jc.rhs;
				return true;
			});
			auto const take_any([&](u32 const rhs) -> bool {					// :354-360
				if (m_by_class.empty()) return false;
				auto const it(m_by_class.begin());
				u32 const founder(it->second);
				m_by_class.erase(it);
				slot(segment, founder) = rhs;
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
				m_by_class.emplace(slot(segment, f), f);
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

	std::vector<u32> rhs_eq, distinct, first_class, joined;
	std::vector<u8> first_is_ref;
	std::vector<u64> offsets;
	if (int const st = segment_classes_host(*g, cut_positions, n_cuts, false, 0, rhs_eq, distinct, first_class, first_is_ref, offsets, joined)) return st;

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
	*found = 1;
	return V2M_OK;
}
