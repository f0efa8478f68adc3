"""CPU model of the shared-memory layouts of the K1 kernels (vcf2multialign_b200/csrc/transpose_bits.cu): for every kernel the
address a loading thread writes word c of line l to is the address the transposing thread reads it from (the formulas are
restated here from the kernels' comments), no two words share an address, the tile fits its allocation, and the 32 lanes of a
transposing read hit 32 distinct banks (16 distinct bank pairs per half-warp for the 64-bit reads). The kernels themselves are
checked bit for bit on the GPU (test_transpose_every_kernel_on_ragged_shapes); this file pins the layout arithmetic they rely on."""
import numpy as np


def _banks_ok(addr_by_lane):
    return len({int(a) % 32 for a in addr_by_lane}) == 32


def test_tile_kernel_layout():
    """512 lines x 32 words; 64-bit word q of line l at l * 16 + (q ^ ((l >> 5) & 15)); lanes = 16 line groups x 2 word columns."""
    lines, words = 512, 32
    addr = np.empty((lines, words), np.int64)
    for l in range(lines):
        for c in range(words):
            addr[l, c] = 2 * (l * 16 + ((c >> 1) ^ ((l >> 5) & 15))) + (c & 1)
    assert len(np.unique(addr)) == lines * words and addr.max() < lines * words
    for warp in range(16):
        for i in (0, 1, 17, 31):
            read = []
            for lane in range(32):
                xl, yl = lane & 15, lane >> 4
                phys = ((warp ^ xl) << 1) | yl
                a = (32 * xl + i) * 32 + phys
                assert a == addr[32 * xl + i, 2 * warp + yl]
                read.append(a)
            assert _banks_ok(read)


def test_tall_kernel_layout():
    """1024 lines x 16 words; word c of line l at l * 16 + 16 * (l >> 9) + (c ^ ((l >> 5) & 15)); lane = line group, warp = word column."""
    lines, words = 1024, 16
    addr = np.empty((lines, words), np.int64)
    for l in range(lines):
        sw = (l >> 5) & 15
        for q in range(8):
            base = 2 * (l * 8 + 8 * (l >> 9) + (q ^ (sw >> 1)))
            lo, hi = (base + 1, base) if sw & 1 else (base, base + 1)      # halves exchanged in every other line group
            addr[l, 2 * q], addr[l, 2 * q + 1] = lo, hi
            assert lo == l * 16 + 16 * (l >> 9) + ((2 * q) ^ sw) and hi == l * 16 + 16 * (l >> 9) + ((2 * q + 1) ^ sw)
    assert len(np.unique(addr)) == lines * words and addr.max() < lines * words + 16
    for yw in range(16):
        for i in (0, 5, 31):
            read = [32 * lane * 16 + 16 * (lane >> 4) + (yw ^ (lane & 15)) + i * 16 for lane in range(32)]
            assert all(read[lane] == addr[32 * lane + i, yw] for lane in range(32))
            assert _banks_ok(read)


def test_wide_kernel_layout():
    """2048 lines x 8 words; word c of line l at l * 8 + 8 * (l >> 6) + (c ^ ((l >> 8) & 7)); lane = 64 lines, warp = word column."""
    lines, words = 2048, 8
    addr = np.empty((lines, words), np.int64)
    for l in range(lines):
        sw = (l >> 8) & 7
        for c in range(words):
            addr[l, c] = l * 8 + 8 * (l >> 6) + (c ^ sw)
        for q in range(4):                                                  # the loader's 64-bit store
            base = 2 * (l * 4 + 4 * (l >> 6) + (q ^ (sw >> 1)))
            lo, hi = (base + 1, base) if sw & 1 else (base, base + 1)
            assert (lo, hi) == (addr[l, 2 * q], addr[l, 2 * q + 1])
    assert len(np.unique(addr)) == lines * words and addr.max() < lines * 8 + 8 * (lines // 64)
    for yw in range(8):
        for i in (0, 9, 31):
            base = [520 * lane + (yw ^ ((lane >> 2) & 7)) for lane in range(32)]
            a_read = [b + 8 * i for b in base]
            b_read = [b + 256 + 8 * i for b in base]
            assert all(a_read[lane] == addr[64 * lane + i, yw] and b_read[lane] == addr[64 * lane + 32 + i, yw] for lane in range(32))
            assert _banks_ok(a_read) and _banks_ok(b_read)


def test_half_size_and_quarter_height_layouts():
    """512 x 16 words (V2M_K1_TALL=6) and 1024 x 8 words (12)."""
    addr = np.empty((512, 16), np.int64)
    for l in range(512):
        for c in range(16):
            addr[l, c] = l * 16 + 16 * (l >> 5) + ((((c >> 1) ^ ((l >> 6) & 7)) << 1) | (c & 1))
    assert len(np.unique(addr)) == addr.size and addr.max() < 512 * 16 + 16 * 16
    for warp in range(8):
        for i in (0, 31):
            read = []
            for lane in range(32):
                xl, yl = lane & 15, lane >> 4
                a = (32 * xl) * 16 + 16 * xl + (((warp ^ (xl >> 1)) << 1) | yl) + 16 * i
                assert a == addr[32 * xl + i, 2 * warp + yl]
                read.append(a)
            assert _banks_ok(read)
    addr = np.empty((1024, 8), np.int64)
    for l in range(1024):
        for c in range(8):
            addr[l, c] = l * 8 + 8 * (l >> 8) + (c ^ ((l >> 5) & 7))
    assert len(np.unique(addr)) == addr.size and addr.max() < 1024 * 8 + 8 * 4
    for w in range(8):
        for i in (0, 31):
            read = [256 * lane + 8 * (lane >> 3) + (w ^ (lane & 7)) + 8 * i for lane in range(32)]
            assert all(read[lane] == addr[32 * lane + i, w] for lane in range(32))
            assert _banks_ok(read)


def test_streaming_kernel_layout_is_conflict_free_for_64_bit_reads():
    """V2M_K1_TALL=8: 64-bit word q of line l at l * 8 + 8 * (l >> 8) + (q ^ ((l >> 5) & 7)); a half-warp's LDS.64 must cover 16
    distinct bank pairs."""
    addr = np.empty((1024, 8), np.int64)
    for l in range(1024):
        for q in range(8):
            addr[l, q] = l * 8 + 8 * (l >> 8) + (q ^ ((l >> 5) & 7))
    assert len(np.unique(addr)) == addr.size and addr.max() < 1024 * 8 + 8 * 4
    for qw in range(8):
        for i in (0, 13, 31):
            read = [256 * lane + 8 * (lane >> 3) + (qw ^ (lane & 7)) + 8 * i for lane in range(32)]
            assert all(read[lane] == addr[32 * lane + i, qw] for lane in range(32))
            for half in (read[:16], read[16:]):
                assert len({a % 16 for a in half}) == 16
