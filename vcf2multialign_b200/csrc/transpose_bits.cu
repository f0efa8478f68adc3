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
//
// Algorithmic bytes: 2 · R · C / 8 (each bit read once, written once). Roofline: HBM read + write.
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
transpose_bits_kernel(u64 const * __restrict__ src, u32 * __restrict__ dst, u64 const R64, u64 const C)
{
	// R64 = source rows / 64 = 64-bit words per source line; C = number of source lines.
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
		u32 *out(dst + r0 * C32 + xw);
#pragma unroll
		for (int j = 0; j < 32; ++j)
			k1_store(out + u64(j) * C32, a[j], st_policy);
	}
}



// The same transpose on tiles of 1024 source lines × 512 bits (16 words): a warp's store is one 128-byte run of one
// destination line instead of two 64-byte runs, which halves the partial-sector stores when destination lines are not
// a multiple of 32 bytes long. Lane ↔ line group (32 of them), warp ↔ word column (16). Shared layout, in 32-bit words:
// line l at l · 16 + 16 · (l >> 9), its word c at position c ^ ((l >> 5) & 15): the 32 lanes of a transposing read (lines
// 32 · lane + i, one column) hit 32 distinct banks, and the loading warps' 8-byte stores stay conflict-free.
constexpr int K1T_TC = 1024;
constexpr int K1T_TW = 16;

__global__ void __launch_bounds__(K1_THREADS, 2)
transpose_bits_tall_kernel(u64 const * __restrict__ src, u32 * __restrict__ dst, u64 const R64, u64 const C)
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
			vals[p] = (pair_ok && line < C) ? k1_load(src + line * R64 + p0 + q, ld_policy) : 0;
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
	transpose32(a);

	u64 const C32(C / 32);
	u64 const xw(c0 / 32 + lane);
	u64 const r0(64 * p0 + 32 * yw);
	if (xw < C32 && r0 < 64 * R64)
	{
		u32 *out(dst + r0 * C32 + xw);
#pragma unroll
		for (int j = 0; j < 32; ++j)
			k1_store(out + u64(j) * C32, a[j], st_policy);
	}
}


// The streaming form of the same transpose, for long matrices (many source lines, few enough row bands to keep every band of a
// stretch of lines resident at once). What bounds the two tile kernels above is not DRAM and not the bit work but the shape
// of their stores (scripts/microbench/k1_patterns.cu, profiles/r02_k1_patterns.txt: the store pattern alone, with no loads
// and no bit work, runs 1.55 ms on c5 — the whole kernel's time; reads alone 0.53 ms): 32-bit stores in 64-byte runs reach
// 3.1 TB/s write-only when the runs start on sector boundaries and 2.0 – 2.4 TB/s when destination lines start mid-sector;
// 64-bit stores in 128-byte runs that start on a sector boundary reach 5.4 – 5.6 TB/s on every shape.
// Hence: (1) the transposed tile goes back to shared memory in destination order, so the stores need not follow the threads'
// sub-tiles: a half warp writes one 128-byte run of one destination line with 64-bit stores; (2) run starts are moved back to
// the 32-byte sector boundary below them (a destination line starts at row · C / 8, any multiple of 8), which takes the last
// <= 24 bytes of the line's previous run: a CTA therefore walks consecutive line tiles of ONE row band and carries those
// bytes from tile to tile in shared memory. Only the first run of a walk starts mid-sector and only its end flushes a partial
// sector. (3) The source tile (1024 lines × 512 bits, the tall tile's swizzled layout) and the destination-order tile share
// one 64 KiB buffer — the sub-tiles are in registers in between — so two CTAs fit an SM and one CTA's loads (issued for the
// next tile before the stores of the current one) overlap the other's stores.
// Grid: bands × walks, band fastest, at most two CTAs per SM so that all of it is resident; the bands of a walk then read
// neighbouring 64-byte pieces of the same source lines at about the same time and share their boundary sectors in the L2.
constexpr int K1S_LINES = 1024;								// source lines per tile
constexpr int K1S_ROWS = 512;								// rows per band: 8 64-bit words of every source line
constexpr int K1S_TILE_WORDS = K1S_LINES * 16 + 16;			// 32-bit words of the swizzled source tile (>= K1S_ROWS * 32)
constexpr int K1S_SMEM = K1S_TILE_WORDS * 4 + K1S_ROWS * 32;	// + 32 carried bytes per row

__device__ __forceinline__ void k1s_store(u8 *p, uint2 const v)
{
	asm volatile("st.global.v2.u32 [%0], {%1, %2};" :: "l"(p), "r"(v.x), "r"(v.y) : "memory");
}

__global__ void __launch_bounds__(K1_THREADS, 2)
transpose_bits_stream_kernel(u64 const * __restrict__ src, u8 * __restrict__ dst, u64 const R64, u64 const C,
	u32 const n_bands, u32 const walk_tiles, u32 const n_tiles)
{
	extern __shared__ __align__(16) u64 tile[];
	u32 *tile32(reinterpret_cast<u32 *>(tile));
	u8 *stage(reinterpret_cast<u8 *>(tile));					// [K1S_ROWS][128] once the sub-tiles are in registers
	u8 *carry(stage + K1S_TILE_WORDS * 4);					// [K1S_ROWS][32]: the last 32 bytes of each row's previous run

	unsigned const tid(threadIdx.x), lane(tid & 31), warp(tid >> 5);
	u32 const band(blockIdx.x % n_bands), walk(blockIdx.x / n_bands);
	u32 const x_begin(walk * walk_tiles), x_end(min(n_tiles, x_begin + walk_tiles));
	u64 const LB(C / 8);										// bytes per destination line
	u64 const R(64 * R64);
	u64 const p0(u64(band) * (K1S_ROWS / 64));

	unsigned const q(tid & 7), lp(tid >> 3);					// 64-bit word within the 64-byte piece of a line; 64 lines per pass
	bool const pair_ok(p0 + q < R64);
	u64 vals[K1S_LINES / 64];
	auto const load_tile([&](u32 const x) {
		u64 const c0(u64(x) * K1S_LINES);
#pragma unroll
		for (int p = 0; p < K1S_LINES / 64; ++p)
		{
			u64 const line(c0 + 64 * p + lp);
			vals[p] = (pair_ok && line < C) ? __ldg(src + line * R64 + p0 + q) : 0;
		}
	});

	unsigned const half(lane >> 4), l8(8 * (lane & 15));		// emission: a half warp per destination line, 8 bytes per lane
	load_tile(x_begin);
#pragma unroll 1
	for (u32 x(x_begin); x < x_end; ++x)
	{
		// source tile → shared memory (transpose_bits_tall_kernel's layout)
#pragma unroll
		for (int p = 0; p < K1S_LINES / 64; ++p)
		{
			unsigned const l(64 * p + lp);
			unsigned const sw((l >> 5) & 15);
			u64 const v(vals[p]);
			tile[l * 8 + 8 * (l >> 9) + (q ^ (sw >> 1))] = (sw & 1) ? (v >> 32 | v << 32) : v;
		}
		__syncthreads();
		u32 a[32];
		{
			unsigned const base(32 * lane * 16 + 16 * (lane >> 4) + (warp ^ (lane & 15)));	// lane = line group, warp = word column
#pragma unroll
			for (int i = 0; i < 32; ++i)
				a[i] = tile32[base + i * 16];
		}
		__syncthreads();											// every sub-tile is in registers: the buffer changes layout
		transpose32(a);
#pragma unroll
		for (int j = 0; j < 32; ++j)
			tile32[(32 * warp + j) * 32 + lane] = a[j];				// row 32 · warp + j of the band, word `lane` of its 128-byte run
		__syncthreads();
		if (x + 1 < x_end) load_tile(x + 1);						// in flight while this tile is stored

		bool const first(x == x_begin), last(x + 1 == x_end);
		u64 const X(u64(x) * (K1S_LINES / 8));						// byte offset of this tile's run in a destination line
#pragma unroll 4
		for (int it = 0; it < K1S_ROWS / 32; ++it)
		{
			unsigned const rl(32 * it + 2 * warp + half);
			u64 const r(u64(band) * K1S_ROWS + rl);
			u64 const line0(r * LB);
			unsigned const s(unsigned(line0) & 31u);				// the line starts s bytes into a sector: its runs start s bytes early
			bool const from_carry(l8 < s);
			u8 const *from(from_carry ? carry + rl * 32 + (32 - s + l8) : stage + rl * 128 + (l8 - s));
			uint2 const v(*reinterpret_cast<uint2 const *>(from));
			u64 const o(X + l8 - s);								// wraps only where from_carry && first, which is not stored
			if (r < R && !(from_carry && first) && o + 8 <= LB) k1s_store(dst + line0 + o, v);
			if (last && from_carry)
			{
				// end of the walk: the bytes held back for the next run go out now
				u64 const of(X + 128 - s + l8);
				if (r < R && of + 8 <= LB) k1s_store(dst + line0 + of, *reinterpret_cast<uint2 const *>(stage + rl * 128 + (128 - s + l8)));
			}
			uint2 keep(v);
			if (l8 >= 96) keep = *reinterpret_cast<uint2 const *>(stage + rl * 128 + l8);
			__syncwarp();
			if (l8 >= 96) *reinterpret_cast<uint2 *>(carry + rl * 32 + (l8 - 96)) = keep;
		}
		__syncthreads();											// the buffer is free for the next source tile
	}
}


int launch_transpose_bits(void const *d_src, u64 rows, u64 cols, void *d_dst, cudaStream_t stream)
{
	if (0 == cols || 0 == rows) return 0;
	u64 const R64(rows / 64);
	static int const force([] { char const *env(std::getenv("V2M_K1_TALL")); return env ? std::atoi(env) : -1; }());	// experiments: 0 / 1 / 2 (2: streaming kernel wherever its grid allows)
	{
		int dev(0), sms(0);
		V2M_CUDA_TRY(cudaGetDevice(&dev));
		V2M_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
		u64 const n_bands((rows + K1S_ROWS - 1) / K1S_ROWS), n_tiles((cols + K1S_LINES - 1) / K1S_LINES);
		u64 const slots(2 * u64(sms));
		// Long matrices with few row bands (the path matrix as build_variant_graph hands it over: rows = chromosome copies, one
		// source line per edge): every band of a stretch of lines resident at once, >= 8 tiles per walk.
		bool const streaming(force >= 0 ? 2 == force && n_bands <= slots : (n_bands <= slots && rows >= 1024 && n_tiles >= 8 * (slots / n_bands)));
		if (streaming)
		{
			u64 const walks_wanted(std::max<u64>(1, slots / n_bands));
			u32 const walk_tiles(u32((n_tiles + walks_wanted - 1) / walks_wanted));
			u32 const n_walks(u32((n_tiles + walk_tiles - 1) / walk_tiles));
			V2M_CUDA_TRY(cudaFuncSetAttribute(transpose_bits_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, K1S_SMEM));
			transpose_bits_stream_kernel<<<unsigned(n_bands * n_walks), K1_THREADS, K1S_SMEM, stream>>>(
				static_cast<u64 const *>(d_src), static_cast<u8 *>(d_dst), R64, cols, u32(n_bands), walk_tiles, u32(n_tiles));
			V2M_CHECK_LAUNCH("transpose_bits_stream_kernel");
			return 0;
		}
	}
	// Measured (scripts/bench_transpose.py, V2M_K1_TALL=0/1): the tall tiles win when the matrix has at most 512 rows (their
	// 512-bit band wastes less of a short line: 64 × 4,096,000 runs 2.6 × faster) and when destination lines are an odd
	// multiple of 8 bytes long (5056 × 3,000,000: 68 % against 61 % of the copy peak); they lose 6 points on sector-aligned
	// destination lines (their 64-byte source pieces straddle more sectors) and tie on c5.
	bool const tall(force >= 0 ? 1 == force : (rows <= 512 || 0 != (cols / 8) % 16));
	if (tall)
	{
		dim3 const grid(unsigned((R64 + K1T_TW / 2 - 1) / (K1T_TW / 2)), unsigned((cols + K1T_TC - 1) / K1T_TC));
		if (grid.y > 65535u) return fail(1, "transpose_bits: too many source lines in one launch");
		constexpr int smem_bytes((K1T_TC * K1T_TW + 16) * sizeof(u32));
		V2M_CUDA_TRY(cudaFuncSetAttribute(transpose_bits_tall_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
		transpose_bits_tall_kernel<<<grid, K1_THREADS, smem_bytes, stream>>>(static_cast<u64 const *>(d_src), static_cast<u32 *>(d_dst), R64, cols);
		V2M_CHECK_LAUNCH("transpose_bits_tall_kernel");
		return 0;
	}
	dim3 const grid(unsigned((R64 + K1_TW / 2 - 1) / (K1_TW / 2)), unsigned((cols + K1_TC - 1) / K1_TC));
	if (grid.y > 65535u) return fail(1, "transpose_bits: more than 33,553,920 source lines in one launch");
	constexpr int smem_bytes(K1_TC * (K1_TW / 2) * sizeof(u64));
	V2M_CUDA_TRY(cudaFuncSetAttribute(transpose_bits_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));	// per device
	transpose_bits_kernel<<<grid, K1_THREADS, smem_bytes, stream>>>(static_cast<u64 const *>(d_src), static_cast<u32 *>(d_dst), R64, cols);
	V2M_CHECK_LAUNCH("transpose_bits_kernel");
	return 0;
}

} // namespace v2m
