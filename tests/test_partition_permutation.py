"""The reference BLAS builder partitions a node's primitives with an in-place swap loop
(rayca-soft/src/bvh/blas.rs:283-291).  The device builder (rayca_b200/csrc/gpu_fast_build.cu, ref_tables /
ref_scatter) applies the same reordering as a closed-form permutation computed from the left/right flags and a
prefix sum; this checks the closed form against the loop, exhaustively for short ranges and on random ones."""
import itertools
import random


def swap_loop(flags):
    a = list(range(len(flags)))
    i, j = 0, len(a)
    while i < j:
        if flags[a[i]]:
            i += 1
        else:
            a[i], a[j - 1] = a[j - 1], a[i]
            j -= 1
    return a, i


def closed_form(flags):
    """Mirrors the kernels: scan = exclusive prefix sum of the flags; two tables, then one destination per element."""
    n = len(flags)
    scan = [0]
    for f in flags:
        scan.append(scan[-1] + (1 if f else 0))
    nl = scan[n]
    hole_pos, tail_pos = [None] * n, [None] * n
    for q in range(n):  # ref_tables
        lb = scan[q]
        if flags[q]:
            if q >= nl:
                tail_pos[(nl - lb) - 1] = q
        elif q <= nl:
            hole_pos[q - lb] = q
    out = [None] * n
    for q in range(n):  # ref_scatter
        lb = scan[q]
        if flags[q]:
            dst = q if q < nl else hole_pos[(nl - lb) - 1]
        elif q <= nl:
            k = q - lb
            dst = (n if k == 0 else tail_pos[k - 1]) - 1
        else:
            dst = q - 1
        assert out[dst] is None
        out[dst] = q
    return out, nl


def test_closed_form_equals_swap_loop_exhaustively():
    for n in range(0, 13):
        for flags in itertools.product((False, True), repeat=n):
            assert closed_form(flags) == swap_loop(flags), flags


def test_closed_form_equals_swap_loop_on_random_ranges():
    rng = random.Random(7)
    for _ in range(300):
        n = rng.randint(13, 400)
        p = rng.choice((0.02, 0.3, 0.5, 0.7, 0.98))
        flags = [rng.random() < p for _ in range(n)]
        assert closed_form(flags) == swap_loop(flags)
