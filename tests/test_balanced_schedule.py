"""Host-side restatement of the balanced persistent grid's schedule (csrc/wkv6_chunk_bwd.cu: the top of
wkv6_bwd_chunk_kernel and unit_at()) and the properties the kernel relies on: every (sequence, chunk) unit is visited exactly
once, chunks of a sequence are visited in order, a sequence is cut at most once, the CTA that owns the first part (piece A)
has the lower index and runs it FIRST, the CTA that owns the rest (piece B) runs it LAST, and the reciprocal divisions are
exact in the range pick_balanced() admits.  No GPU needed."""
import random


def _cta_units(j, ncta, BH, n):
    """Units (stage, sequence, chunk) of CTA j in visiting order -- the kernel's closed form, including its magic divisions."""
    U = BH * n
    gs, ge = j * U // ncta, (j + 1) * U // ncta
    gsu = min((gs + n - 1) // n * n, ge)
    ged = max(ge // n * n, gsu)
    bhA, lenA = ged // n, ge - ged
    bhW, nW = gsu // n, (ged - gsu) // n
    bhB, cB0 = gs // n, gs - gs // n * n
    lenW = nW * n
    magic = (1 << 31) // n + 1
    out = []
    for it in range(ge - gs):
        x = it - lenA
        iw = (max(x, 0) * magic) >> 31
        if x < 0:
            out.append(("A", bhA, it))
        elif x < lenW:
            out.append(("W", bhW + iw, x - iw * n))
        else:
            out.append(("B", bhB, cB0 + (x - lenW)))
    return out


def _check(BH, n, ncta):
    seen = {}
    for j in range(ncta):
        for it, (st, bh, c) in enumerate(_cta_units(j, ncta, BH, n)):
            assert 0 <= bh < BH and 0 <= c < n, (BH, n, ncta, j, it, st, bh, c)
            assert (bh, c) not in seen
            seen[(bh, c)] = (j, it, st)
    assert len(seen) == BH * n
    for (bh, c), (j, it, st) in seen.items():
        if c == 0:
            assert st in ("A", "W")                       # a sequence starts as a whole one or as a piece A
            continue
        pj, pit, pst = seen[(bh, c - 1)]
        if st == "A":
            assert (pj, pit, pst) == (j, it - 1, "A")
        elif st == "W":
            assert (pj, pit, pst) == (j, it - 1, "W")
        elif pst == "B":
            assert (pj, pit) == (j, it - 1)
        else:                                             # first chunk of a piece B: handed over by the previous CTA's piece A,
            assert pst == "A" and pj == j - 1 and pit < it   # which that CTA ran first (earlier in its own timeline)
    for j in range(ncta):
        stages = [st for st, _, _ in _cta_units(j, ncta, BH, n)]
        order = "".join(s for i, s in enumerate(stages) if i == 0 or stages[i - 1] != s)
        assert order in ("A", "W", "B", "AW", "AB", "WB", "AWB"), order    # A first, B last


def test_bench_shape_and_corner_cases():
    for BH, n, ncta in [(128, 12, 74), (256, 12, 148), (75, 3, 74), (1000, 1, 148), (149, 40, 148), (74, 5, 74),
                        (200, 7, 66), (148, 12, 148), (296, 12, 148), (152, 4, 148)]:
        _check(BH, n, ncta)


def test_random_shapes():
    rnd = random.Random(1)
    for _ in range(200):
        ncta = rnd.choice([74, 148, 66, 132, 37])
        _check(rnd.randint(ncta, 5 * ncta), rnd.randint(1, 40), ncta)


def test_reciprocal_division_is_exact_in_the_admitted_range():
    """x / d == (x * (2^31 / d + 1)) >> 31 for x * d < 2^31 (pick_balanced() rejects larger launches)."""
    rnd = random.Random(2)
    for d in list(range(1, 70)) + [rnd.randint(70, 40000) for _ in range(200)]:
        magic = (1 << 31) // d + 1
        lim = ((1 << 31) - 1) // d
        for x in [0, 1, d - 1, d, d + 1, lim // 2, lim - 1, lim] + [rnd.randint(0, lim) for _ in range(50)]:
            if x < 0:
                continue
            assert (x * magic) >> 31 == x // d, (x, d)
