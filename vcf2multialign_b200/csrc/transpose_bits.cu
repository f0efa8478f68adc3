// K1 transpose_bits — the device side of v2m::transpose_matrix
// (reference: libvcf2multialign/transpose_matrix.cc:41-109; called once per graph at variant_graph.cc:453).
//
// dst(x, y) = src(y, x) on libbio's column-major, LSB-first packed bit matrices. Seen as arrays of 32-bit words
// (little-endian halves of the 64-bit words), the source is C "lines" (its columns) of R/32 words and the destination
// R lines of C/32 words.
//
// One CTA moves a tile of TC = 512 source lines × 1024 bits (32 words):
//   load    512 threads × 16 passes of 8-byte loads (lines are only 8-byte aligned in general: R/8 bytes per line,
//           R % 64 == 0), into shared memory with the 8-byte pair index XOR-swizzled by the line group, so that the
//           transposing reads below hit 32 distinct banks;
//   compute every thread owns one 32×32 sub-tile: 32 conflict-free LDS, five block-swap stages in registers
//           (common.cuh:transpose32) — ~12 instructions per 128 B per warp, an order of magnitude fewer than 32
//           ballots per sub-tile, which would leave the kernel issue-bound instead of HBM-bound;
//   store   lanes are laid out 16 line groups × 2 word columns, so each STG.32 of a warp writes two full 64-byte
//           runs of two destination lines; CTAs of consecutive line tiles (blockIdx.y) extend the same destination lines, so
//           the 126 MB L2 merges them into full lines before write-back.
//   grid    row bands vary fastest. A 128-byte piece of a source line starts at an 8-byte boundary only, so it shares its
//           first and last sectors with the pieces of the neighbouring bands; with the bands of one line tile resident
//           together those sectors leave DRAM once (ncu, c5: DRAM reads 3.81 GB for 3.79 GB algorithmic; with line tiles
//           fastest it was 7.41 GB). A variant that read whole lines through one bulk copy per CTA and stored 16-byte runs
//           was measured and dropped: half-sector stores cost read-modify-write traffic in the L2 (5.0 GB written, 4.1 ms).
//           Also measured and dropped on c5: one persistent CTA per SM with two cp.async-filled tile buffers (2.08 ms: 16
//           warps per SM do not cover the transpose's issue time) and 256-line tiles with 4 CTAs per SM (1.96 ms: 32-byte
//           store runs); 256 threads with two sub-tiles each, cp.async loads and 3 CTAs per SM (1.63 ms: occupancy is not
//           the limit); 1024-line × 512-bit tiles whose stores are single 128-byte runs (transpose_bits_tall_kernel below:
//           1.62 ms — what the longer runs save on the misaligned destination lines, the 64-byte source pieces lose);
//           this kernel runs 1.61 ms there.
//   shapes  the destination line is C / 8 bytes long. When that is not a multiple of 32 (c5: 750,000 bytes) every other
//           destination line starts in the middle of a sector, its 64-byte runs touch three sectors instead of two, and the
//           kernel reaches 72 % of the copy peak; with C = 6,000,128 (or the c2 shape, or 12,000,000) it reaches 83 %
//           (scripts/bench_transpose.py --cols …). The layout is the reference's, so the lines cannot be padded.
//   round 2 what bounds it (profiles/r02_k1_ablation.txt, r02_k1_variants.txt, r02_k1_wide_ncu_summary.txt): nothing is saturated —
//           DRAM 57 % active, L2 41 – 47 %, ALU pipe 51 %, issue slots 42 %. The transposes alone are issue-bound at 0.72 ms on c5
//           (~850 instructions per sub-tile, a third of them addresses and predicates), loads hide under them (0.87 ms together)
//           but stores do not (1.41 ms; 0.96 on a sector-aligned shape), and loads and stores together cost more than the sum of
//           what each adds alone. Tried on top of the kernels below, none faster on c5: half-size tiles at 4 CTAs per SM (1.62 ms),
//           a persistent kernel with cp.async prefetch (1.65), two sub-tiles per thread with 8-byte stores in 256-byte runs (the
//           wide kernel, 1.64; 2.86 when held to 3 CTAs per SM — the L1 that is left no longer covers the loads), ld.global.cg /
//           L1::no_allocate / .lu / L2 evict-first loads (1.62 / 1.94 / 1.92 / 1.92). The wide kernel wins where destination lines
//           are 8 mod 16 bytes long and is the launcher's choice there. Its own ablations: transposes 0.63 ms, + stores 0.82 (the
//           256-byte runs cost next to nothing, on every shape), + loads 1.23 (its 32-byte source pieces straddle two sectors three
//           times out of four; L2 → L1 traffic 10.5 GB for 3.79), everything 1.64. Loads and stores with the transposes left out
//           take as long as the whole kernel for the 512-line tiles (1.61 ms on c5, 1.41 aligned): the integer work is hidden, the
//           memory side of this tiling is the bound. A cluster of two CTAs that shares a 2048-line × 512-row tile through
//           distributed shared memory (64-byte pieces AND 256-byte runs) was built and measured: bit-exact, 6.3 ms — 64 scalar
//           ld.shared::cluster per thread alone take 5 ms; removed.
//
// Algorithmic bytes: 2 · R · C / 8 (each bit read once, written once). Roofline: HBM read + write.
#include <algorithm>
#include <atomic>
#include <cstdlib>

#include "common.cuh"

namespace v2m {

// Cache policies (make VARIANT=… EXTRA=-DV2M_K1_LOAD_POLICY=1 …; 0 = plain __ldg / plain stores). Measured on c5 / 5056 ×
// 1,155,008 (profiles/r01_transpose_cache_policy.txt): plain 1.609 / 0.331 ms; evict-last stores 1.567 / 0.324 ms (the
// default: partially written destination lines stay in the L2 until the neighbouring tiles complete them); st.global.cs
// 1.605 / 0.333 ms; evict-first, L1-no-allocate loads 2.018 / 0.400 ms (the bands' shared boundary sectors are lost).
#ifndef V2M_K1_LOAD_POLICY
#define V2M_K1_LOAD_POLICY 0		// 1: L2 evict-first on the source (read once)
#endif
#ifndef V2M_K1_STORE_POLICY
#define V2M_K1_STORE_POLICY 1		// 1: L2 evict-last on the destination (lines wait for the neighbouring tiles), 2: st.global.cs
#endif

__device__ __forceinline__ u64 k1_policy_evict_last()
{
	u64 policy;
	asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(policy));
	return policy;
}

__device__ __forceinline__ u64 k1_load(u64 const *p, u64 const policy)
{
#if V2M_K1_LOAD_POLICY == 1
	u64 v;
	asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.b64 %0, [%1], %2;" : "=l"(v) : "l"(p), "l"(policy));
	return v;
#else
	return __ldg(p);
#endif
}

__device__ __forceinline__ void k1_store(u32 *p, u32 const v, u64 const policy)
{
#if V2M_K1_STORE_POLICY == 1
	asm volatile("st.global.L2::cache_hint.b32 [%0], %1, %2;" :: "l"(p), "r"(v), "l"(policy) : "memory");
#elif V2M_K1_STORE_POLICY == 2
	__stcs(p, v);
#else
	*p = v;
#endif
}

constexpr int K1_TC = 512;			// source lines per tile
constexpr int K1_TW = 32;			// 32-bit words per line per tile (1024 bits)
constexpr int K1_THREADS = 512;

__global__ void __launch_bounds__(K1_THREADS, 2)
transpose_bits_kernel(u64 const * __restrict__ src, u32 * __restrict__ dst, u64 const R64, u64 const C, u64 const dst_pitch32)
{
	// R64 = source rows / 64 = 64-bit words per source line; C = number of source lines; dst_pitch32 = 32-bit words between
	// destination lines (C / 32 for a whole matrix, more when the lines being written are a column range of longer ones).
	extern __shared__ __align__(16) u64 tile[];					// K1_TC * K1_TW / 2 words = 64 KiB (dynamic: above the 48 KiB static limit)

	// Row bands vary fastest (blockIdx.x): the CTAs that share the boundary sectors of a 128-byte piece of a source line run
	// together, so those sectors come out of DRAM once and out of the L2 afterwards.
	u64 const c0(u64(blockIdx.y) * K1_TC);
	u64 const p0(u64(blockIdx.x) * (K1_TW / 2));				// first 64-bit word (pair) of the row band
	unsigned const tid(threadIdx.x);
	[[maybe_unused]] u64 const ld_policy(V2M_K1_LOAD_POLICY ? l2_policy_evict_first() : 0), st_policy(1 == V2M_K1_STORE_POLICY ? k1_policy_evict_last() : 0);

	// ---- load -------------------------------------------------------------------------------------------
	{
		unsigned const q(tid & 15);								// pair within the 128-byte piece of a line
		unsigned const l_in_pass(tid >> 4);						// 32 lines per pass
		bool const pair_ok(p0 + q < R64);
		u64 vals[K1_TC / 32];
#pragma unroll
		for (int p = 0; p < K1_TC / 32; ++p)
		{
			u64 const line(c0 + 32 * p + l_in_pass);
			vals[p] = (pair_ok && line < C) ? k1_load(src + line * R64 + p0 + q, ld_policy) : 0;
		}
#pragma unroll
		for (int p = 0; p < K1_TC / 32; ++p)
		{
			unsigned const l(32 * p + l_in_pass);
			tile[l * 16 + (q ^ ((l >> 5) & 15))] = vals[p];		// line group = l >> 5 = p
		}
	}
	__syncthreads();

	// ---- transpose 32×32 sub-tiles in registers ------------------------------------------------------------
	unsigned const lane(tid & 31), warp(tid >> 5);
	unsigned const xl(lane & 15);								// line group (32 lines) within the tile
	unsigned const yl(lane >> 4);								// low bit of the word column
	unsigned const yw(2 * warp + yl);							// 32-bit word column within the tile, 0 … 31
	u32 const *tile32(reinterpret_cast<u32 const *>(tile));
	u32 a[32];
	{
		unsigned const phys(((warp ^ xl) << 1) | yl);			// swizzled word index inside the line
#pragma unroll
		for (int i = 0; i < 32; ++i)
			a[i] = tile32[(32 * xl + i) * 32 + phys];
	}
	transpose32(a);

	// ---- store ---------------------------------------------------------------------------------------------
	u64 const C32(C / 32);
	u64 const xw(c0 / 32 + xl);									// destination word within the line
	u64 const r0(64 * p0 + 32 * yw);							// first destination line of this sub-tile
	if (xw < C32 && r0 < 64 * R64)								// rows are a multiple of 64, sub-tiles of 32: all-or-nothing
	{
		u32 *out(dst + r0 * dst_pitch32 + xw);
#pragma unroll
		for (int j = 0; j < 32; ++j)
			k1_store(out + u64(j) * dst_pitch32, a[j], st_policy);
	}
}



// The same transpose on tiles of 1024 source lines × 512 bits (16 words): a warp's store is one 128-byte run of one
// destination line instead of two 64-byte runs, which halves the partial-sector stores when destination lines are not
// a multiple of 32 bytes long. Lane ↔ line group (32 of them), warp ↔ word column (16). Shared layout, in 32-bit words:
// line l at l · 16 + 16 · (l >> 9), its word c at position c ^ ((l >> 5) & 15): the 32 lanes of a transposing read (lines
// 32 · lane + i, one column) hit 32 distinct banks, and the loading warps' 8-byte stores stay conflict-free.
constexpr int K1T_TC = 1024;
constexpr int K1T_TW = 16;

template <int MODE>
__global__ void __launch_bounds__(K1_THREADS, 2)
transpose_bits_tall_kernel(u64 const * __restrict__ src, u32 * __restrict__ dst, u64 const R64, u64 const C, u64 const dst_pitch32)
{
	extern __shared__ __align__(16) u64 tile[];					// K1T_TC · K1T_TW words + 16 words of padding
	u64 const c0(u64(blockIdx.y) * K1T_TC);
	u64 const p0(u64(blockIdx.x) * (K1T_TW / 2));				// first 64-bit word (pair) of the row band
	unsigned const tid(threadIdx.x);
	[[maybe_unused]] u64 const ld_policy(V2M_K1_LOAD_POLICY ? l2_policy_evict_first() : 0), st_policy(1 == V2M_K1_STORE_POLICY ? k1_policy_evict_last() : 0);

	{
		unsigned const q(tid & 7);								// pair within the 64-byte piece of a line
		unsigned const l_in_pass(tid >> 3);						// 64 lines per pass
		bool const pair_ok(p0 + q < R64);
		u64 vals[K1T_TC / 64];
#pragma unroll
		for (int p = 0; p < K1T_TC / 64; ++p)
		{
			u64 const line(c0 + 64 * p + l_in_pass);
			vals[p] = (MODE & 2) ? u64(tid) * 0x9e3779b97f4a7c15ull + p : (pair_ok && line < C) ? k1_load(src + line * R64 + p0 + q, ld_policy) : 0;
		}
#pragma unroll
		for (int p = 0; p < K1T_TC / 64; ++p)
		{
			unsigned const l(64 * p + l_in_pass);
			unsigned const sw((l >> 5) & 15);
			u64 const v(vals[p]);
			tile[l * 8 + 8 * (l >> 9) + (q ^ (sw >> 1))] = (sw & 1) ? (v >> 32 | v << 32) : v;
		}
	}
	__syncthreads();

	unsigned const lane(tid & 31), yw(tid >> 5);				// lane = line group, yw = word column
	u32 const *tile32(reinterpret_cast<u32 const *>(tile));
	u32 a[32];
	{
		unsigned const base(32 * lane * 16 + 16 * (lane >> 4) + (yw ^ (lane & 15)));
#pragma unroll
		for (int i = 0; i < 32; ++i)
			a[i] = tile32[base + i * 16];
	}
	if (!(MODE & 4)) transpose32(a);

	u64 const C32(C / 32);
	u64 const xw(c0 / 32 + lane);
	u64 const r0(64 * p0 + 32 * yw);
	if (xw < C32 && r0 < 64 * R64)
	{
		u32 *out(dst + r0 * dst_pitch32 + xw);
#pragma unroll
		for (int j = 0; j < 32; ++j)
			if (!(MODE & 1) || 0x12345678u == a[j]) k1_store(out + u64(j) * dst_pitch32, a[j], st_policy);
	}
}


// The sheared form of the same transpose, for the matrices the path actually hands over (thousands of rows, 10^5 … 10^7
// source lines). What bounds the two tile kernels above is neither DRAM nor the bit work but the shape of their stores
// (scripts/microbench/k1_patterns.cu, profiles/r02_k1_patterns.txt: their store pattern alone, with no loads and no bit work,
// takes 1.55 ms on c5 — the whole kernel's time; their reads alone 0.53 ms): 32-bit stores in 64-byte runs reach 3.1 TB/s
// write-only when the runs start on sector boundaries and 2.0 – 2.4 TB/s when destination lines start mid-sector, whereas
// 64-bit stores in 128-byte runs that start on a sector boundary reach 5.4 – 5.6 TB/s on every shape. Hence:
//   (1) tall tiles (1024 lines × 512 bits: a warp's 32 sub-tiles lie side by side along one destination line) and a word swap
//       between neighbouring lanes per pair of rows, so that a warp's store is two 128-byte runs with 8 bytes per lane (staging
//       the transposed tile in shared memory for the same store shape cost more than it gained: 1.90 ms on c5, three barriers
//       and ~850 instructions per warp and tile);
//   (2) every destination line's runs start on 32-byte sector boundaries. Line r starts at byte r · C / 8, i.e. s(r) = 0, 8,
//       16 or 24 bytes into a sector, so its run in tile x is bytes [128 x − s, 128 x − s + 128) of the line: source lines
//       [1024 x − 8 s, 1024 x − 8 s + 1024). Rows of different s sit side by side in every 32 × 32 sub-tile, so the sub-tile
//       is *sheared* on the way into registers: word i of a thread's sub-tile is assembled from the same word of up to four
//       source lines, 64 · k lines apart, each masked to the rows whose shift is 8 k bytes (the row masks are the same for
//       every sub-tile, since sub-tiles start at multiples of 32 rows). One transpose32 then yields every row's own window.
//       The tile loads a halo of 192 source lines (L2 hits: the neighbouring tile's CTA reads them as its own at the same
//       time) and the sub-tile read costs up to four LDS + three LOP3 per word instead of one LDS;
//   RESULT (profiles/r02_k1_sheared.txt): correct on every shape (tests/test_gpu_parity.py runs all three kernels against the
//   CPU restatement) and indifferent to the destination lines' phase, as intended — but 1.85 ms on c5 against the tile kernel's 1.57,
//   2.19 against 1.73 (tall) on the c4 shape, 1.74 against 1.43 on a sector-aligned shape. ncu: the ALU pipe is 71 % busy and
//   `math pipe throttle` is the top stall (1.16 G warp instructions against the tile kernel's 0.77 G): with the stores out of
//   the way the transpose is bound by its own integer work — 64 PRMT + 144 LOP3 / SHF for the bit stages plus index
//   arithmetic, selects for the swizzle and the lane swap — and the shear and the halo add to exactly that. It therefore stays
//   opt-in (V2M_K1_TALL=2) as the record of the experiment; the launcher's default is unchanged. The same shear on the 512-line
//   × 1024-bit tile kernel (128-byte source pieces, the cheapest tiling) ran 2.02 ms on c5 with 32-bit stores and 1.92 ms with
//   the lane swap, 2.26 ms on the c4 shape: the 37 % halo and the shear cost more than the aligned 64-byte runs give back.
//   (3) tiles are independent, two CTAs per SM, row bands fastest in the grid, so that at any moment the CTAs in flight write
//       ~30 consecutive runs (3.8 KB) of every destination line. A variant in which persistent CTAs walked along the lines and
//       carried each line's last bytes from tile to tile in shared memory (no halo, no shear) ran 2.50 ms on c5 against the
//       tile kernel's 1.57: its CTAs write 128-byte pieces that are ~26 KB apart in every destination line, and DRAM pages
//       see one piece at a time.
constexpr int K1H_LINES = 1024;								// a tile's own source lines = 128 bytes of every destination line
constexpr int K1H_ROWS = 512;								// rows per band: 8 64-bit words of every source line
constexpr int K1H_HALO = 192;								// source lines before the tile's own: 24 bytes per destination line
constexpr int K1H_SRC_LINES = K1H_LINES + K1H_HALO;
constexpr int K1H_PASSES = K1H_SRC_LINES / 64;
constexpr int K1H_HALO_PASSES = K1H_HALO / 64;
constexpr int K1H_TILE_WORDS = K1H_SRC_LINES * 16 + 16 * ((K1H_SRC_LINES - 1) >> 9) + 16;	// 32-bit words, tall-tile swizzle (>= K1H_ROWS * 32)
constexpr int K1H_SMEM = K1H_TILE_WORDS * 4;

struct k1h_shifts { u32 rows[4]; };							// rows[k]: bit j set iff line 32 m + j starts 8 k bytes into a sector

// EDGE = false: a tile whose own lines, halo and row band all lie inside the matrix — no bounds checks, addresses by constant
// strides (the kernel is bound by the ALU pipe: 71 % busy in the first version, of which a third was address arithmetic,
// predicates and selects; profiles/r02_k1_sheared_ncu_summary.txt).
template <bool EDGE>
__device__ __forceinline__ void k1h_tile(u64 const * __restrict__ src, u8 * __restrict__ dst, u64 const R64, u64 const C, k1h_shifts const &sh,
	u64 *tile, u32 const band, u32 const x)
{
	u32 const *tile32(reinterpret_cast<u32 const *>(tile));
	unsigned const tid(threadIdx.x), lane(tid & 31), warp(tid >> 5);
	u64 const LB(C / 8);											// bytes per destination line
	u64 const p0(u64(band) * (K1H_ROWS / 64));
	bool const sheared(0 != (sh.rows[1] | sh.rows[2] | sh.rows[3]));

	// ---- source tile (halo first) → shared memory, the tall tile's swizzled layout ------------------------------------------
	{
		unsigned const q(tid & 7), lp(tid >> 3);					// 64-bit word within the 64-byte piece of a line; 64 lines per pass
		u64 vals[K1H_PASSES];
		if (EDGE)
		{
			bool const pair_ok(p0 + q < R64);
			long long const line0((long long) x * K1H_LINES - K1H_HALO);
#pragma unroll
			for (int p = 0; p < K1H_PASSES; ++p)
			{
				long long const line(line0 + 64 * p + lp);
				bool const wanted(p >= K1H_HALO_PASSES || sheared);
				vals[p] = (wanted && pair_ok && line >= 0 && u64(line) < C) ? __ldg(src + u64(line) * R64 + p0 + q) : 0;
			}
		}
		else
		{
			u64 const *first(src + (u64(x) * K1H_LINES - K1H_HALO + lp) * R64 + p0 + q);
			u64 const stride(64 * R64);
#pragma unroll
			for (int p = 0; p < K1H_PASSES; ++p)
				vals[p] = (p >= K1H_HALO_PASSES || sheared) ? __ldg(first + u64(p) * stride) : 0;
		}
#pragma unroll
		for (int p = 0; p < K1H_PASSES; ++p)
		{
			unsigned const l(64 * p + lp);
			unsigned const sw((l >> 5) & 15);
			u64 const v(vals[p]);
			tile[l * 8 + 8 * (l >> 9) + (q ^ (sw >> 1))] = (sw & 1) ? (v >> 32 | v << 32) : v;
		}
	}
	__syncthreads();

	// ---- sheared 32 × 32 sub-tile: lane = line group of the tile's own lines, warp = word column ----------------------------
	// Rows that start 8 k bytes into a sector take their bits from the source lines 64 k earlier, i.e. from line group G0 - 2 k.
	u32 a[32];
	{
		unsigned const G0(lane + K1H_HALO / 32);					// local line group of this lane's unshifted window
		auto const base_of([&](unsigned const G) { return 512 * G + 16 * (G >> 4) + (warp ^ (G & 15)); });
		unsigned const b0(base_of(G0));
		if (!sheared)
		{
#pragma unroll
			for (int i = 0; i < 32; ++i) a[i] = tile32[b0 + i * 16];
		}
		else if (0 == sh.rows[1])
		{
			// destination lines are 16 mod 32 bytes long: shifts 0 and 16
			unsigned const b2(base_of(G0 - 4));
			u32 const m0(sh.rows[0]);
#pragma unroll
			for (int i = 0; i < 32; ++i) a[i] = (tile32[b0 + i * 16] & m0) | (tile32[b2 + i * 16] & ~m0);
		}
		else
		{
			unsigned const b1(base_of(G0 - 2)), b2(base_of(G0 - 4)), b3(base_of(G0 - 6));
			u32 const m0(sh.rows[0]), m2(sh.rows[2]), m01(sh.rows[0] | sh.rows[1]);
#pragma unroll
			for (int i = 0; i < 32; ++i)
			{
				u32 const lo((tile32[b0 + i * 16] & m0) | (tile32[b1 + i * 16] & ~m0));
				u32 const hi((tile32[b2 + i * 16] & m2) | (tile32[b3 + i * 16] & ~m2));
				a[i] = (lo & m01) | (hi & ~m01);
			}
		}
	}
	transpose32(a);

	// ---- stores: neighbouring lanes swap one word per pair of rows, so that a lane holds 8 bytes of one row -------------------
	// a[j] is word `lane` of row 32 · warp + j. Even lanes send a[j + 1] and receive the odd neighbour's a[j] (word lane + 1 of
	// row j); odd lanes the other way round: a warp's store is two sector-aligned 128-byte runs (rows j and j + 1), 8 bytes a lane.
	u64 const r0(u64(band) * K1H_ROWS + 32 * warp);
	if (EDGE && r0 >= 64 * R64) return;									// rows are a multiple of 64, sub-tiles of 32: all or nothing
	unsigned const odd(lane & 1);
	u64 const X(u64(x) * (K1H_LINES / 8) + 4 * (lane & ~1u));			// byte offset of this lane's 8 bytes in an unshifted run
	u64 const lineA((r0 + odd) * LB), lineB(lineA + 2 * LB);			// this lane's first two destination lines; 4 · LB is a multiple of 32,
	u8 *const pA(dst + lineA + X - (lineA & 31));						// so its lines alternate between two sector phases: runs start
	u8 *const pB(dst + lineB + X - (lineB & 31));						// (line start & 31) bytes early
	u64 const step(4 * LB);
#pragma unroll
	for (int jj = 0; jj < 16; ++jj)
	{
		u32 const got(__shfl_xor_sync(0xffffffffu, odd ? a[2 * jj] : a[2 * jj + 1], 1));
		u32 const lo(odd ? got : a[2 * jj]), hi(odd ? a[2 * jj + 1] : got);
		u8 *const p(((jj & 1) ? pB : pA) + u64(jj >> 1) * step);
		if (EDGE)
		{
			u64 const line((jj & 1 ? lineB : lineA) + u64(jj >> 1) * step);
			long long const o((long long) X - (long long) (line & 31));
			if (o < 0 || u64(o) + 8 > LB) continue;
		}
		asm volatile("st.global.v2.u32 [%0], {%1, %2};" :: "l"(p), "r"(lo), "r"(hi) : "memory");
	}
}

__global__ void __launch_bounds__(K1_THREADS, 2)
transpose_bits_sheared_kernel(u64 const * __restrict__ src, u8 * __restrict__ dst, u64 const R64, u64 const C, k1h_shifts const sh)
{
	extern __shared__ __align__(16) u64 tile[];
	u32 const band(blockIdx.x), x(blockIdx.y);
	bool const edge(0 == x || (u64(x) + 1) * K1H_LINES > C || (u64(band) + 1) * (K1H_ROWS / 64) > R64);
	if (edge) k1h_tile<true>(src, dst, R64, C, sh, tile, band, x);
	else k1h_tile<false>(src, dst, R64, C, sh, tile, band, x);
}



// Half-size tiles: 512 source lines × 512 rows (64-byte source pieces), 256 threads, 4 CTAs per SM — the same lanes ↔ sub-tiles
// map and the same 64-byte store runs as transpose_bits_kernel, but four CTAs in four different phases (load / transpose /
// store) per SM instead of two. Shared layout in 32-bit words: word c of line l at l · 16 + 16 · (l >> 5) + (((c >> 1) ^ ((l >> 6)
// & 7)) << 1 | (c & 1)): the transposing read of a warp (16 line groups × 2 word columns) hits 32 distinct banks.
constexpr int K1S_TC = 512;
constexpr int K1S_TW = 16;
constexpr int K1S_THREADS = 256;
constexpr int K1S_SMEM = (K1S_TC * K1S_TW + 16 * (K1S_TC / 32)) * 4;

// MODE (ablations for profiles/r02_k1_ablation.txt, V2M_K1_MODE): 0 the kernel; 1 no stores (kept alive by a test that never
// holds); 2 no loads (the tile is made of thread indices); 3 neither; 4 loads and stores but no transposes (the memory side alone).
template <bool EDGE, int MODE = 0>
__device__ __forceinline__ void k1s_tile(u64 const * __restrict__ src, u32 * __restrict__ dst, u64 const R64, u64 const C, u64 const dst_pitch32, u64 *tile)
{
	u64 const c0(u64(blockIdx.y) * K1S_TC);
	u64 const p0(u64(blockIdx.x) * (K1S_TW / 2));
	unsigned const tid(threadIdx.x);
	[[maybe_unused]] u64 const st_policy(1 == V2M_K1_STORE_POLICY ? k1_policy_evict_last() : 0);
	{
		unsigned const q(tid & 7), lp(tid >> 3);				// 32 lines per pass
		u64 const *first(src + (c0 + lp) * R64 + p0 + q);
		u64 const stride(32 * R64);
		[[maybe_unused]] bool const pair_ok(p0 + q < R64);
		u64 vals[K1S_TC / 32];
#pragma unroll
		for (int p = 0; p < K1S_TC / 32; ++p)
			vals[p] = (MODE & 2) ? u64(tid) * 0x9e3779b97f4a7c15ull + p : (!EDGE || (pair_ok && c0 + 32 * p + lp < C)) ? __ldg(first + u64(p) * stride) : 0;
#pragma unroll
		for (int p = 0; p < K1S_TC / 32; ++p)
			tile[(32 * p + lp) * 8 + 8 * p + (q ^ ((p >> 1) & 7))] = vals[p];
	}
	__syncthreads();

	unsigned const lane(tid & 31), warp(tid >> 5);
	unsigned const xl(lane & 15), yl(lane >> 4);
	unsigned const yw(2 * warp + yl);
	u32 const *tile32(reinterpret_cast<u32 const *>(tile));
	u32 a[32];
	{
		unsigned const base((32 * xl) * 16 + 16 * xl + (((warp ^ (xl >> 1)) << 1) | yl));
#pragma unroll
		for (int i = 0; i < 32; ++i) a[i] = tile32[base + 16 * i];
	}
	if (!(MODE & 4)) transpose32(a);

	u64 const xw(c0 / 32 + xl);
	u64 const r0(64 * p0 + 32 * yw);
	if (!EDGE || (xw < C / 32 && r0 < 64 * R64))
	{
		u32 *out(dst + r0 * dst_pitch32 + xw);
#pragma unroll
		for (int j = 0; j < 32; ++j)
			if (!(MODE & 1) || 0x12345678u == a[j]) k1_store(out + u64(j) * dst_pitch32, a[j], st_policy);
	}
}

template <int MODE>
__global__ void __launch_bounds__(K1S_THREADS, 4)
transpose_bits_small_kernel(u64 const * __restrict__ src, u32 * __restrict__ dst, u64 const R64, u64 const C, u64 const dst_pitch32)
{
	extern __shared__ __align__(16) u64 tile[];
	bool const edge((u64(blockIdx.y) + 1) * K1S_TC > C || (u64(blockIdx.x) + 1) * (K1S_TW / 2) > R64);
	if (edge) k1s_tile<true, MODE>(src, dst, R64, C, dst_pitch32, tile);
	else k1s_tile<false, MODE>(src, dst, R64, C, dst_pitch32, tile);
}


// The tall tile with bulk stores (V2M_K1_TALL=9): the transposed sub-tiles go back into the (by then dead) shared tile in
// destination order — 128 bytes per destination line — and every thread hands one line's 128 bytes to the copy engine
// (cp.async.bulk shared → global). Needs 16-byte aligned runs, i.e. destination lines that are a multiple of 16 bytes long
// (c5: 750,000; not c4 / c2, whose odd lines start 8 bytes into a 16-byte unit). Per thread 32 STS with immediate offsets and
// one bulk copy replace 32 STG with a 64-bit address each.
// RESULT (profiles/r02_k1_ablation.txt): bit-exact, and slower — 1.84 ms on c5 against 1.62 (1.86 against 1.42 on the sector-
// aligned shape): transposes + stores take the same 1.15 ms as with the tall kernel's STG (29.6 M bulk copies of 128 bytes do
// not reach the L2 in a better shape than 128-byte runs of 4-byte stores), and the third barrier and the wait for the copy
// engine to have read the tile lengthen every CTA's tail. Opt-in, as the record of the experiment.
template <int MODE>
__global__ void __launch_bounds__(K1_THREADS, 2)
transpose_bits_tall_bulk_kernel(u64 const * __restrict__ src, u8 * __restrict__ dst, u64 const R64, u64 const C, u64 const dst_pitch_bytes)
{
	extern __shared__ __align__(16) u64 tile[];
	u64 const c0(u64(blockIdx.y) * K1T_TC);
	u64 const p0(u64(blockIdx.x) * (K1T_TW / 2));
	unsigned const tid(threadIdx.x);
	{
		unsigned const q(tid & 7), l_in_pass(tid >> 3);
		bool const pair_ok(p0 + q < R64);
		u64 vals[K1T_TC / 64];
#pragma unroll
		for (int p = 0; p < K1T_TC / 64; ++p)
		{
			u64 const line(c0 + 64 * p + l_in_pass);
			vals[p] = (MODE & 2) ? u64(tid) * 0x9e3779b97f4a7c15ull + p : (pair_ok && line < C) ? __ldg(src + line * R64 + p0 + q) : 0;
		}
#pragma unroll
		for (int p = 0; p < K1T_TC / 64; ++p)
		{
			unsigned const l(64 * p + l_in_pass);
			unsigned const sw((l >> 5) & 15);
			u64 const v(vals[p]);
			tile[l * 8 + 8 * (l >> 9) + (q ^ (sw >> 1))] = (sw & 1) ? (v >> 32 | v << 32) : v;
		}
	}
	__syncthreads();

	unsigned const lane(tid & 31), yw(tid >> 5);
	u32 *tile32(reinterpret_cast<u32 *>(tile));
	u32 a[32];
	{
		unsigned const base(32 * lane * 16 + 16 * (lane >> 4) + (yw ^ (lane & 15)));
#pragma unroll
		for (int i = 0; i < 32; ++i) a[i] = tile32[base + i * 16];
	}
	__syncthreads();											// the source tile is dead: its place takes the tile in destination order
	transpose32(a);
	{
		u32 *stage(tile32 + (32 * yw) * 32 + lane);				// line 32 · yw + j of the band, word `lane` of its 128 bytes
#pragma unroll
		for (int j = 0; j < 32; ++j) stage[j * 32] = a[j];
	}
	fence_proxy_async_smem();
	__syncthreads();

	u64 const LB(C / 8);
	u64 const x_bytes(u64(blockIdx.y) * (K1T_TC / 8));
	u64 const row(64 * p0 + tid);
	if (!(MODE & 1) && row < 64 * R64 && x_bytes < LB)
	{
		u32 const bytes(u32(min(u64(K1T_TC / 8), LB - x_bytes)));		// a multiple of 16: LB is
		bulk_s2g(dst + row * dst_pitch_bytes + x_bytes, tile32 + tid * 32, bytes);
	}
	bulk_commit();
	bulk_wait_read<0>();
}


// The tall tile as a stream (opt-in, V2M_K1_TALL=8; the record of an experiment): persistent CTAs, two per SM, draw tiles from
// a counter, row bands fastest, and the source pieces of the next tile are on their way into shared memory (8-byte cp.async: no
// registers, no waiting warps) while the current one is transposed and stored. One buffer is enough: the copies are issued
// right after every thread has read its sub-tiles into registers. 256 threads; every thread takes the two sub-tiles that share
// its 32 source lines and lie in the word columns 2 w and 2 w + 1 (w = warp): one LDS.64 per source line feeds both, so the
// swizzle can stay at 64-bit granularity. Stores as in the tall kernel: 128 bytes of one destination line per warp instruction.
// Why it was tried — the ablations (profiles/r02_k1_ablation.txt, c5): the transposes alone take 0.72 ms (issue-bound, IPC 3.7 of
// 4), loads + transposes 0.87, transposes + stores 1.41 (tile kernel's store shape; 1.10 with the tall kernel's), everything 1.60
// – 1.65: the load phase of a CTA stretches when stores are in flight, and with two (or four: the half-size-tile kernel, same
// time) CTAs per SM nothing covers it.
// What came out: (1) tiles must be handed out by a counter — with a fixed stride the CTAs that run ahead spread the front of
// partially written sectors over more of the destination than the L2 holds (stores alone: 2.17 ms on c5, 1.15 – 1.20 with the
// counter, 0.85 on a sector-aligned shape either way); (2) cp.async in 4- or 8-byte pieces is itself slow — loads + transposes
// 1.37 ms (4-byte) and 1.25 ms (8-byte) against 0.87 for LDG.64 + STS — so the kernel ends at 1.65 ms on c5, 1.89 on c4, 1.54
// aligned: no better than the kernels it was meant to beat. 16-byte cp.async and bulk copies need 16-byte aligned source lines,
// which the layout does not give (R / 8 bytes per line, 8 mod 16 for 5,056 rows).
// Shared layout in 64-bit words: word q of line l at l · 8 + 8 · (l >> 8) + (q ^ ((l >> 5) & 7)): a half-warp's LDS.64 (lines
// 32 · lane + i, one q) covers 16 distinct bank pairs.
constexpr int K1P_THREADS = 256;
constexpr int K1P_SMEM = (K1T_TC * 8 + 8 * (K1T_TC / 256)) * 8;

template <int MODE>
__global__ void __launch_bounds__(K1P_THREADS, 2)
transpose_bits_tall_pairs_kernel(u64 const * __restrict__ src, u32 * __restrict__ dst, u64 const R64, u64 const C, u64 const dst_pitch32,
	u32 const n_bands, u32 const n_tiles, unsigned *ticket)
{
	extern __shared__ __align__(16) u64 tile[];
	__shared__ u32 s_next;
	unsigned const tid(threadIdx.x);
	[[maybe_unused]] u64 const st_policy(1 == V2M_K1_STORE_POLICY ? k1_policy_evict_last() : 0);
	u32 const sbase(smem_u32(tile));
	unsigned const q(tid & 7), lp(tid >> 3);					// load side: 64-bit word within the 64-byte piece, 32 lines per pass
	unsigned const lane(tid & 31), qw(tid >> 5);				// transpose side: lane = line group, qw = pair of word columns
	unsigned const rbase(256 * lane + 8 * (lane >> 3) + (qw ^ (lane & 7)));
	u64 const C32(C / 32);

	auto const issue([&](u32 const t) {
		u64 const c0(u64(t / n_bands) * K1T_TC);
		u64 const p0(u64(t % n_bands) * (K1T_TW / 2));
		bool const pair_ok(p0 + q < R64);
		u64 const *first(src + (c0 + lp) * R64 + p0 + q);
		u64 const stride(32 * R64);
		u32 const d0(sbase + 8 * (lp * 8));
#pragma unroll
		for (int p = 0; p < K1T_TC / 32; ++p)
		{
			bool const ok(pair_ok && c0 + 32 * p + lp < C);
			u64 const *g(ok ? first + u64(p) * stride : src);
			u32 const n(ok ? 8u : 0u);
			u32 const d(d0 + 8 * (32 * p * 8 + 8 * (p >> 3)) + 8 * (q ^ (p & 7)));
			asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" :: "r"(d), "l"(g), "r"(n) : "memory");
		}
		asm volatile("cp.async.commit_group;" ::: "memory");
	});

	u32 const total(n_bands * n_tiles);
	if (0 == tid) s_next = atomicAdd(ticket, 1u);
	__syncthreads();
	u32 t(s_next);
	__syncthreads();
	if (t < total && !(MODE & 2)) issue(t);
	while (t < total)
	{
		if (0 == tid) s_next = atomicAdd(ticket, 1u);
		asm volatile("cp.async.wait_group 0;" ::: "memory");
		__syncthreads();
		u32 const next(s_next);
		u32 a[32], b[32];
#pragma unroll
		for (int i = 0; i < 32; ++i)
		{
			u64 const v(tile[rbase + 8 * i]);
			a[i] = u32(v); b[i] = u32(v >> 32);
		}
		__syncthreads();
		if (next < total && !(MODE & 2)) issue(next);
		transpose32(a);
		transpose32(b);

		u64 const xw(u64(t / n_bands) * (K1T_TC / 32) + lane);
		u64 const r0(64 * u64(t % n_bands) * (K1T_TW / 2) + 64 * qw);
		if (xw < C32 && r0 < 64 * R64)							// rows are a multiple of 64: both sub-tiles or neither
		{
			u32 *out(dst + r0 * dst_pitch32 + xw);
#pragma unroll
			for (int j = 0; j < 32; ++j)
				if (!(MODE & 1) || 0x12345678u == a[j]) k1_store(out + u64(j) * dst_pitch32, a[j], st_policy);
			out += 32 * dst_pitch32;
#pragma unroll
			for (int j = 0; j < 32; ++j)
				if (!(MODE & 1) || 0x12345678u == b[j]) k1_store(out + u64(j) * dst_pitch32, b[j], st_policy);
		}
		t = next;
	}
}


// Quarter-height tall tiles (V2M_K1_TALL=12): 1024 source lines × 256 rows (32-byte source pieces), 256 threads, one sub-tile per
// thread, 4 CTAs per SM — the tall kernel's stores (lane = line group: 128 bytes of one destination line per warp store) at the
// occupancy at which the half-size-tile kernel's loads cost next to nothing. Separates the two things the wide kernel's loads
// could be suffering from: the 32-byte pieces or the 16 warps per SM.
// RESULT (profiles/r02_k1_ablation.txt): 1.575 ms on c5, 1.73 on c4, 1.59 aligned, 0.327 on c2 — the best of all kernels on c5
// and c4 by 1 – 2 %, which is nothing; opt-in. Transposes 0.75 ms, with loads 1.03, with stores 0.98, loads + stores without the
// transposes 1.51: reads and writes do not overlap in DRAM time, and that sum is what every kernel in this file ends up at.
// Shared layout in 32-bit words: word c of line l at l · 8 + 8 · (l >> 8) + (c ^ ((l >> 5) & 7)).
constexpr int K1Q_TC = 1024;
constexpr int K1Q_THREADS = 256;
constexpr int K1Q_SMEM = (K1Q_TC * 8 + 8 * (K1Q_TC / 256)) * 4;

template <bool EDGE, int MODE>
__device__ __forceinline__ void k1q_tile(u64 const * __restrict__ src, u32 * __restrict__ dst, u64 const R64, u64 const C, u64 const dst_pitch32, u64 *tile)
{
	u64 const c0(u64(blockIdx.y) * K1Q_TC);
	u64 const p0(u64(blockIdx.x) * 4);
	unsigned const tid(threadIdx.x);
	[[maybe_unused]] u64 const st_policy(1 == V2M_K1_STORE_POLICY ? k1_policy_evict_last() : 0);
	{
		unsigned const q(tid & 3), lp(tid >> 2);					// 64 lines per pass
		u64 const *first(src + (c0 + lp) * R64 + p0 + q);
		u64 const stride(64 * R64);
		[[maybe_unused]] bool const pair_ok(p0 + q < R64);
		bool const swap(0 != ((lp >> 5) & 1));						// (l >> 5) & 1
		u64 vals[K1Q_TC / 64];
#pragma unroll
		for (int p = 0; p < K1Q_TC / 64; ++p)
			vals[p] = (MODE & 2) ? u64(tid) * 0x9e3779b97f4a7c15ull + p : (!EDGE || (pair_ok && c0 + 64 * p + lp < C)) ? __ldg(first + u64(p) * stride) : 0;
#pragma unroll
		for (int p = 0; p < K1Q_TC / 64; ++p)
		{
			u64 const v(vals[p]);
			tile[(64 * p + lp) * 4 + 4 * (p >> 2) + (q ^ (p & 3))] = swap ? (v >> 32 | v << 32) : v;
		}
	}
	__syncthreads();

	unsigned const lane(tid & 31), w(tid >> 5);
	u32 const *tile32(reinterpret_cast<u32 const *>(tile));
	u32 a[32];
	{
		unsigned const base(256 * lane + 8 * (lane >> 3) + (w ^ (lane & 7)));
#pragma unroll
		for (int i = 0; i < 32; ++i) a[i] = tile32[base + 8 * i];
	}
	if (!(MODE & 4)) transpose32(a);

	u64 const xw(c0 / 32 + lane);
	u64 const r0(64 * p0 + 32 * w);
	if (!EDGE || (xw < C / 32 && r0 < 64 * R64))
	{
		u32 *out(dst + r0 * dst_pitch32 + xw);
#pragma unroll
		for (int j = 0; j < 32; ++j)
			if (!(MODE & 1) || 0x12345678u == a[j]) k1_store(out + u64(j) * dst_pitch32, a[j], st_policy);
	}
}

template <int MODE>
__global__ void __launch_bounds__(K1Q_THREADS, 4)
transpose_bits_quarter_kernel(u64 const * __restrict__ src, u32 * __restrict__ dst, u64 const R64, u64 const C, u64 const dst_pitch32)
{
	extern __shared__ __align__(16) u64 tile[];
	bool const edge((u64(blockIdx.y) + 1) * K1Q_TC > C || (u64(blockIdx.x) + 1) * 4 > R64);
	if (edge) k1q_tile<true, MODE>(src, dst, R64, C, dst_pitch32, tile);
	else k1q_tile<false, MODE>(src, dst, R64, C, dst_pitch32, tile);
}


// The wide form: tiles of 2048 source lines × 256 rows, 256 threads, every thread transposes TWO sub-tiles that lie side by
// side along the destination lines (source lines 64 · lane … + 31 and … + 32 … + 63, word column = warp), so that it holds 8
// consecutive bytes of each of its 32 destination lines and a warp's store is one 256-byte run of one line, 8 bytes per lane,
// with no exchange between lanes. Why: the store shape is what bounds the tile kernels (profiles/r02_k1_patterns.txt — the L2
// takes 64-byte runs of 4-byte pieces at 2.0 – 2.4 TB/s when lines start mid-sector, 256-byte runs of 8-byte pieces at 3.9 –
// 4.3 TB/s, and a run's two partial sectors weigh 2 in 9 instead of 2 in 3), and the shear that removes the partial sectors
// altogether costs more integer work than it saves (above). The price is on the read side: a tile takes 32-byte pieces of the
// source lines (4 words), which straddle two sectors three times out of four; the neighbouring row band's CTA reads the other
// halves at the same time (row bands vary fastest in the grid), so they come out of the L2.
// Shared layout in 32-bit words: word c of line l at l · 8 + 8 · (l >> 6) + (c ^ ((l >> 8) & 7)) — the 32 lanes of a
// transposing read (lines 64 · lane + i, one word column) hit banks 8 · (lane & 3) + (c ^ (lane >> 2)): all distinct.
constexpr int K1W_LINES = 2048;
constexpr int K1W_ROWS = 256;								// 4 64-bit words of every source line
constexpr int K1W_THREADS = 256;
constexpr int K1W_TILE_WORDS = K1W_LINES * 8 + 8 * (K1W_LINES / 64);
constexpr int K1W_SMEM = K1W_TILE_WORDS * 4;

template <int MODE>
__global__ void __launch_bounds__(K1W_THREADS, 2)
transpose_bits_wide_kernel(u64 const * __restrict__ src, u32 * __restrict__ dst, u64 const R64, u64 const C, u64 const dst_pitch32)
{
	extern __shared__ __align__(16) u64 tile[];
	u64 const c0(u64(blockIdx.y) * K1W_LINES);
	u64 const p0(u64(blockIdx.x) * (K1W_ROWS / 64));			// first 64-bit word of the row band
	unsigned const tid(threadIdx.x);
	[[maybe_unused]] u64 const st_policy(1 == V2M_K1_STORE_POLICY ? k1_policy_evict_last() : 0);

	// ---- load: 4 lanes per 32-byte piece, 64 lines per pass ----------------------------------------------------------------
	{
		unsigned const q(tid & 3), lp(tid >> 2);
		bool const pair_ok(p0 + q < R64);
		u64 const *first(src + (c0 + lp) * R64 + p0 + q);
		u64 const stride(64 * R64);
#pragma unroll
		for (int half = 0; half < 2; ++half)
		{
			u64 vals[K1W_LINES / 128];
#pragma unroll
			for (int p = 0; p < K1W_LINES / 128; ++p)
			{
				int const pass(half * (K1W_LINES / 128) + p);
				vals[p] = (MODE & 2) ? u64(tid) * 0x9e3779b97f4a7c15ull + p : (pair_ok && c0 + 64 * pass + lp < C) ? __ldg(first + u64(pass) * stride) : 0;
			}
#pragma unroll
			for (int p = 0; p < K1W_LINES / 128; ++p)
			{
				int const pass(half * (K1W_LINES / 128) + p);
				unsigned const l(64 * pass + lp);
				unsigned const sw((l >> 8) & 7);				// = (pass >> 2) & 7: known at compile time
				u64 const v(vals[p]);
				tile[l * 4 + 4 * pass + (q ^ (sw >> 1))] = (sw & 1) ? (v >> 32 | v << 32) : v;
			}
		}
	}
	__syncthreads();

	// ---- two 32 × 32 sub-tiles per thread ---------------------------------------------------------------------------------------
	unsigned const lane(tid & 31), yw(tid >> 5);					// lane = 64 source lines, yw = word column (32 rows)
	u32 const *tile32(reinterpret_cast<u32 const *>(tile));
	u32 a[32], b[32];
	{
		unsigned const base(520 * lane + (yw ^ ((lane >> 2) & 7)));
#pragma unroll
		for (int i = 0; i < 32; ++i) a[i] = tile32[base + 8 * i];
#pragma unroll
		for (int i = 0; i < 32; ++i) b[i] = tile32[base + 256 + 8 * i];
	}
	if (!(MODE & 4))
	{
		transpose32(a);
		transpose32(b);
	}

	// ---- store: 8 bytes per lane, 256 bytes of one destination line per warp -------------------------------------------------------
	u64 const C32(C / 32);
	u64 const xw(c0 / 32 + 2 * lane);								// C is a multiple of 64: a pair of words is inside the line or not
	u64 const r0(64 * p0 + 32 * yw);
	if (xw < C32 && r0 < 64 * R64)
	{
		u32 *out(dst + r0 * dst_pitch32 + xw);
#pragma unroll
		for (int j = 0; j < 32; ++j)
		{
			u32 *const p(out + u64(j) * dst_pitch32);
			if ((MODE & 1) && 0x12345678u != a[j]) continue;
#if V2M_K1_STORE_POLICY == 1
			asm volatile("st.global.L2::cache_hint.v2.b32 [%0], {%1, %2}, %3;" :: "l"(p), "r"(a[j]), "r"(b[j]), "l"(st_policy) : "memory");
#else
			asm volatile("st.global.v2.b32 [%0], {%1, %2};" :: "l"(p), "r"(a[j]), "r"(b[j]) : "memory");
#endif
		}
	}
}


// Tile counters of the streaming kernels: a ring of 64 slots in device memory, one per launch in flight, zeroed on the launch's
// stream before the kernel starts.
static __device__ unsigned k1_ticket_ring[64];

static int k1_ticket(unsigned **out, cudaStream_t stream)
{
	static std::atomic<unsigned> next{0};
	void *base(nullptr);
	V2M_CUDA_TRY(cudaGetSymbolAddress(&base, k1_ticket_ring));
	unsigned *const slot(static_cast<unsigned *>(base) + next.fetch_add(1, std::memory_order_relaxed) % 64);
	V2M_CUDA_TRY(cudaMemsetAsync(slot, 0, sizeof(unsigned), stream));
	*out = slot;
	return 0;
}

// dst_line_bits >= cols: the destination lines written are the first cols bits of lines that are dst_line_bits long (K5 fills
// paths_by_chrom_copy_and_edge a range of edge columns at a time; d_dst then points at the range's first word of line 0).
int launch_transpose_bits_into(void const *d_src, u64 rows, u64 cols, void *d_dst, u64 dst_line_bits, cudaStream_t stream)
{
	if (0 == cols || 0 == rows) return 0;
	u64 const R64(rows / 64);
	u64 const dst_pitch32(dst_line_bits / 32);
	// Experiments and tests (V2M_K1_TALL): 0 tile kernel, 1 tall, 2 sheared, 3 wide, 6 half-size tiles, 8 streaming, 9 bulk stores, 12 quarter-height tall tiles; otherwise the
	// choice below. V2M_K1_MODE: the ablations of kernels 1, 3, 6, 8 and 9.
	int const force([] { char const *env(std::getenv("V2M_K1_TALL")); return env ? std::atoi(env) : -1; }());
	int const mode([] { char const *env(std::getenv("V2M_K1_MODE")); return env ? std::atoi(env) : 0; }());
	{
		// The path matrix as build_variant_graph hands it over (rows = chromosome copies, one source line per edge) and anything
		// of that kind: at least two full row bands and enough lines for the halo not to matter.
		u64 const n_bands((rows + K1H_ROWS - 1) / K1H_ROWS);
		u64 const LB(cols / 8);
		k1h_shifts sh{{0, 0, 0, 0}};
		for (unsigned j(0); j < 32; ++j) sh.rows[((j * LB) % 32) / 8] |= 1u << j;
		bool const any_shift(0 != (sh.rows[1] | sh.rows[2] | sh.rows[3]));
		u64 const n_tiles((LB + (any_shift ? 24 : 0) + 127) / 128);
		bool const fits(n_tiles <= 65535u && n_bands <= 0x7fffffffu);
		if (fits && 2 == force && dst_line_bits == cols)		// opt-in: measured slower than the tile kernels on every shape (see the kernel's header)
		{
			V2M_CUDA_TRY(cudaFuncSetAttribute(transpose_bits_sheared_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, K1H_SMEM));
			transpose_bits_sheared_kernel<<<dim3(unsigned(n_bands), unsigned(n_tiles)), K1_THREADS, K1H_SMEM, stream>>>(
				static_cast<u64 const *>(d_src), static_cast<u8 *>(d_dst), R64, cols, sh);
			V2M_CHECK_LAUNCH("transpose_bits_sheared_kernel");
			return 0;
		}
	}
	if (9 == force && 0 == (cols / 8) % 16 && 0 == (dst_line_bits / 8) % 16 && 0 == reinterpret_cast<u64>(d_dst) % 16)
	{
		dim3 const grid(unsigned((R64 + K1T_TW / 2 - 1) / (K1T_TW / 2)), unsigned((cols + K1T_TC - 1) / K1T_TC));
		if (grid.y > 65535u) return fail(1, "transpose_bits: too many source lines in one launch");
		constexpr int smem_bytes((K1T_TC * K1T_TW + 16) * sizeof(u32));
		auto const launch([&](auto kernel) -> int {
			V2M_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
			kernel<<<grid, K1_THREADS, smem_bytes, stream>>>(static_cast<u64 const *>(d_src), static_cast<u8 *>(d_dst), R64, cols, dst_line_bits / 8);
			V2M_CHECK_LAUNCH("transpose_bits_tall_bulk_kernel");
			return 0;
		});
		return 1 == mode ? launch(transpose_bits_tall_bulk_kernel<1>) : 2 == mode ? launch(transpose_bits_tall_bulk_kernel<2>) : 3 == mode ? launch(transpose_bits_tall_bulk_kernel<3>) : launch(transpose_bits_tall_bulk_kernel<0>);
	}
	if (12 == force)
	{
		dim3 const grid(unsigned((R64 + 3) / 4), unsigned((cols + K1Q_TC - 1) / K1Q_TC));
		if (grid.y > 65535u) return fail(1, "transpose_bits: too many source lines in one launch");
		auto const launch([&](auto kernel) -> int {
			V2M_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, K1Q_SMEM));
			kernel<<<grid, K1Q_THREADS, K1Q_SMEM, stream>>>(static_cast<u64 const *>(d_src), static_cast<u32 *>(d_dst), R64, cols, dst_pitch32);
			V2M_CHECK_LAUNCH("transpose_bits_quarter_kernel");
			return 0;
		});
		return 1 == mode ? launch(transpose_bits_quarter_kernel<1>) : 2 == mode ? launch(transpose_bits_quarter_kernel<2>) : 3 == mode ? launch(transpose_bits_quarter_kernel<3>) : 4 == mode ? launch(transpose_bits_quarter_kernel<4>) : launch(transpose_bits_quarter_kernel<0>);
	}
	if (8 == force)
	{
		u64 const n_bands((R64 + K1T_TW / 2 - 1) / (K1T_TW / 2)), n_tiles((cols + K1T_TC - 1) / K1T_TC);
		if (n_bands * n_tiles > 0xffffffffull) return fail(1, "transpose_bits: too many tiles in one launch");
		int dev(0), sms(0);
		V2M_CUDA_TRY(cudaGetDevice(&dev));
		V2M_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
		unsigned const ctas(unsigned(std::min<u64>(n_bands * n_tiles, u64(2) * sms)));
		unsigned *ticket(nullptr);
		if (int const rc = k1_ticket(&ticket, stream)) return rc;
		auto const launch([&](auto kernel) -> int {
			V2M_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, K1P_SMEM));
			kernel<<<ctas, K1P_THREADS, K1P_SMEM, stream>>>(static_cast<u64 const *>(d_src), static_cast<u32 *>(d_dst), R64, cols, dst_pitch32, u32(n_bands), u32(n_tiles), ticket);
			V2M_CHECK_LAUNCH("transpose_bits_tall_pairs_kernel");
			return 0;
		});
		return 1 == mode ? launch(transpose_bits_tall_pairs_kernel<1>) : 2 == mode ? launch(transpose_bits_tall_pairs_kernel<2>) : 3 == mode ? launch(transpose_bits_tall_pairs_kernel<3>) : launch(transpose_bits_tall_pairs_kernel<0>);
	}
	if (6 == force)
	{
		dim3 const grid(unsigned((R64 + K1S_TW / 2 - 1) / (K1S_TW / 2)), unsigned((cols + K1S_TC - 1) / K1S_TC));
		if (grid.y > 65535u) return fail(1, "transpose_bits: too many source lines in one launch");
		auto const launch([&](auto kernel) -> int {
			V2M_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, K1S_SMEM));
			kernel<<<grid, K1S_THREADS, K1S_SMEM, stream>>>(static_cast<u64 const *>(d_src), static_cast<u32 *>(d_dst), R64, cols, dst_pitch32);
			V2M_CHECK_LAUNCH("transpose_bits_small_kernel");
			return 0;
		});
		return 1 == mode ? launch(transpose_bits_small_kernel<1>) : 2 == mode ? launch(transpose_bits_small_kernel<2>) : 3 == mode ? launch(transpose_bits_small_kernel<3>) : 4 == mode ? launch(transpose_bits_small_kernel<4>) : launch(transpose_bits_small_kernel<0>);
	}
	// The choice. Destination lines whose length is a multiple of 16 bytes (c5: 750,000) take the 512 × 1024 tile kernel; others —
	// the path matrices of c4 (787,560-byte lines) and c2 (144,376), and the way back from copy-major to edge-major lines (632
	// bytes) — the wide kernel when there are enough rows to fill its 256-row bands, else the tall one
	// (profiles/r02_k1_variants.txt: c4 1.78 → 1.73 – 1.75 ms, c2 0.336 – 0.351 → 0.327 – 0.340, c4 back 3.50 – 3.64 → 2.58).
	bool const misaligned(0 != (cols / 8) % 16);
	bool const wide_ok(0 == dst_line_bits % 64 && 0 == reinterpret_cast<u64>(d_dst) % 8);
	bool const wide(wide_ok && (force >= 0 ? 3 == force : (misaligned && rows > 512)));
	if (wide)
	{
		dim3 const grid(unsigned((R64 + K1W_ROWS / 64 - 1) / (K1W_ROWS / 64)), unsigned((cols + K1W_LINES - 1) / K1W_LINES));
		if (grid.y > 65535u) return fail(1, "transpose_bits: too many source lines in one launch");
		auto const launch([&](auto kernel) -> int {
			V2M_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, K1W_SMEM));
			kernel<<<grid, K1W_THREADS, K1W_SMEM, stream>>>(static_cast<u64 const *>(d_src), static_cast<u32 *>(d_dst), R64, cols, dst_pitch32);
			V2M_CHECK_LAUNCH("transpose_bits_wide_kernel");
			return 0;
		});
		return 1 == mode ? launch(transpose_bits_wide_kernel<1>) : 2 == mode ? launch(transpose_bits_wide_kernel<2>) : 3 == mode ? launch(transpose_bits_wide_kernel<3>) : 4 == mode ? launch(transpose_bits_wide_kernel<4>) : launch(transpose_bits_wide_kernel<0>);
	}
	bool const tall(force >= 0 ? 1 == force : (rows <= 512 || misaligned));
	if (tall)
	{
		dim3 const grid(unsigned((R64 + K1T_TW / 2 - 1) / (K1T_TW / 2)), unsigned((cols + K1T_TC - 1) / K1T_TC));
		if (grid.y > 65535u) return fail(1, "transpose_bits: too many source lines in one launch");
		constexpr int smem_bytes((K1T_TC * K1T_TW + 16) * sizeof(u32));
		auto const launch([&](auto kernel) -> int {
			V2M_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
			kernel<<<grid, K1_THREADS, smem_bytes, stream>>>(static_cast<u64 const *>(d_src), static_cast<u32 *>(d_dst), R64, cols, dst_pitch32);
			V2M_CHECK_LAUNCH("transpose_bits_tall_kernel");
			return 0;
		});
		return 1 == mode ? launch(transpose_bits_tall_kernel<1>) : 2 == mode ? launch(transpose_bits_tall_kernel<2>) : 3 == mode ? launch(transpose_bits_tall_kernel<3>) : 4 == mode ? launch(transpose_bits_tall_kernel<4>) : launch(transpose_bits_tall_kernel<0>);
	}
	dim3 const grid(unsigned((R64 + K1_TW / 2 - 1) / (K1_TW / 2)), unsigned((cols + K1_TC - 1) / K1_TC));
	if (grid.y > 65535u) return fail(1, "transpose_bits: more than 33,553,920 source lines in one launch");
	constexpr int smem_bytes(K1_TC * (K1_TW / 2) * sizeof(u64));
	V2M_CUDA_TRY(cudaFuncSetAttribute(transpose_bits_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));	// per device
	transpose_bits_kernel<<<grid, K1_THREADS, smem_bytes, stream>>>(static_cast<u64 const *>(d_src), static_cast<u32 *>(d_dst), R64, cols, dst_pitch32);
	V2M_CHECK_LAUNCH("transpose_bits_kernel");
	return 0;
}

int launch_transpose_bits(void const *d_src, u64 rows, u64 cols, void *d_dst, cudaStream_t stream)
{
	return launch_transpose_bits_into(d_src, rows, cols, d_dst, cols, stream);
}

} // namespace v2m
