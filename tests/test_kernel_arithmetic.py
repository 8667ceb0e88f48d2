"""CPU: integer identities the kernels rely on, restated in Python and checked exhaustively or on random inputs (the
GPU parity tests cover the kernels themselves; these pin WHY the tricks are exact).

  * warp_transpose_bits (csrc/common.cuh): five block-swap stages transpose the 32 x 32 bit matrix held one row per lane;
  * scatter_kernel (csrc/binning.cu): t // w == (t * ceil(2^16 / w)) >> 16 for the rect sizes it is used on, and the
    5-step search over the inclusive scan of the lanes' tile counts finds the owner of every flattened instance;
  * tile_scan_kernel (csrc/binning.cu): the split of the tiles over the cluster's CTAs covers every tile exactly once;
  * render_fwd_kernel's strip-mask scan (csrc/render_fwd.cu): one aligned word per lane and four ballots queue the
    entries in list order.
"""
import random


def _transpose_stages(rows):
    x = list(rows)
    for j, m in ((16, 0x0000FFFF), (8, 0x00FF00FF), (4, 0x0F0F0F0F), (2, 0x33333333), (1, 0x55555555)):
        y = [x[lane ^ j] for lane in range(32)]                      # __shfl_xor_sync(x, j)
        x = [(((y[l] >> j) & m) | (x[l] & ~m & 0xFFFFFFFF)) if (l & j) else ((x[l] & m) | ((y[l] & m) << j)) & 0xFFFFFFFF
             for l in range(32)]
    return x


def test_warp_transpose_bits_is_a_bit_matrix_transpose():
    rng = random.Random(7)
    cases = [[rng.getrandbits(32) for _ in range(32)] for _ in range(200)]
    cases += [[0] * 32, [0xFFFFFFFF] * 32, [1 << i for i in range(32)], [0xFFFFFFFF if i == 5 else 0 for i in range(32)]]
    for rows in cases:
        t = _transpose_stages(rows)
        for lane in range(32):
            want = 0
            for k in range(32):
                want |= ((rows[k] >> lane) & 1) << k
            assert t[lane] == want
        assert _transpose_stages(t) == list(rows)                   # an involution


def test_scatter_division_by_reciprocal_multiply_is_exact():
    # scatter_kernel: rects of at most 16 tiles, so t < 16 and w <= 16; checked well beyond that range
    for w in range(1, 65):
        m = 65535 // w + 1                                          # ceil(2^16 / w), as the kernel computes it
        assert m == -(-65536 // w)
        for t in range(0, 1024):
            assert (t * m) >> 16 == t // w, (t, w)


def test_scatter_owner_search_finds_every_instance_once():
    rng = random.Random(11)
    for _ in range(300):
        n = [rng.choice([0, 0, 1, 1, 2, 3, 4, 6, 9, 16]) for _ in range(32)]
        if rng.random() < 0.1:
            n = [0] * 32
        incl, run = [], 0
        for v in n:
            run += v
            incl.append(run)
        total = incl[31]
        seen = []
        for j in range(total):
            s = 0
            for step in (16, 8, 4, 2, 1):                           # first lane whose inclusive count exceeds j
                if incl[s + step - 1] <= j:
                    s += step
            s = min(s, 31)
            first = incl[s] - n[s]
            assert n[s] > 0 and first <= j < incl[s]
            seen.append((s, j - first))
        assert seen == [(s, t) for s in range(32) for t in range(n[s])]     # surfel-major, each instance once


def test_tile_scan_cluster_split_covers_every_tile_once():
    THREADS, CLUSTER = 1024, 8
    for num_tiles in [1, 63, 64, 1023, 1024, 1025, 4056, 8487, 8192, 8193, 32400, 130000]:
        chunks = (num_tiles + THREADS - 1) // THREADS
        per = ((chunks + CLUSTER - 1) // CLUSTER) * THREADS
        covered = []
        for rank in range(CLUSTER):
            begin = min(rank * per, num_tiles)
            end = min(begin + per, num_tiles)
            covered += list(range(begin, end))
        assert covered == list(range(num_tiles))


def test_strip_mask_word_scan_queues_entries_in_list_order():
    rng = random.Random(3)
    CHUNK, QUEUE = 32, 256
    for _ in range(200):
        range_x = rng.randrange(0, 5000)
        total = rng.randrange(0, 700)
        range_y = range_x + total
        strip_bit = 1 << rng.randrange(8)
        density = rng.choice([0.02, 0.3, 0.9])
        mem = bytearray(rng.getrandbits(8) for _ in range(range_y + 8))
        for i in range(len(mem)):
            if rng.random() > density:
                mem[i] &= ~strip_bit & 0xFF
        want = [o + 1 for o in range(total) if mem[range_x + o] & strip_bit]
        queue, scan, qcount = [], 0, 0
        while scan < total:                                         # the chunk loop: fill(), then CHUNK entries are consumed
            while qcount < CHUNK and scan < total:                  # fill(): rounds of 128 entries
                first = range_x + scan
                rnd = []                                            # (rank, value) pairs written this round
                lower = 0                                           # hits of lower lanes = the four masked popcounts
                for lane in range(32):
                    wpos = (first & ~3) + 4 * lane
                    at = lower
                    for b in range(4):
                        pos = wpos + b
                        if wpos < range_y and first <= pos < range_y and mem[pos] & strip_bit:
                            rnd.append((at, pos - range_x + 1))
                            at += 1
                    lower = at
                assert [r for r, _ in rnd] == list(range(len(rnd)))  # dense ranks, ascending with (lane, byte)
                queue += [v for _, v in rnd]
                qcount += len(rnd)
                assert qcount <= CHUNK - 1 + 128 <= QUEUE           # the ring never wraps onto unread entries
                scan = (first & ~3) + 128 - range_x
            qcount = max(0, qcount - CHUNK)
        assert queue == want
