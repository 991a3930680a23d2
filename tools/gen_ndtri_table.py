#!/usr/bin/env python
"""Generate ``csrc/ndtri_table.inc``: piecewise-polynomial inverse normal CDF on the 30-bit
Sobol grid, laid out for conflict-free 128-bit shared-memory reads.

Why a table: inside a warp the 32 consecutive Sobol points of any dimension are stratified
(exactly one point per 1/32 interval), so EVERY warp holds central and tail arguments at every
step and a branchy rational approximation (Cephes / AS241) would execute both branches always.
One uniform instruction stream instead: index a row, run one Horner chain.

A row is 32 bytes — c0, c1 in fp64 and c2..c5 in fp32 — i.e. TWO 128-bit shared-memory reads per
point (the path kernels are limited by instruction issue and by shared-memory wavefronts, not by
arithmetic).  Degree 5 on 16 sub-intervals per octave is accurate to 5e-12 relative (printed below,
recorded in the generated file): five orders of magnitude inside the 1e-6 the path values have to
meet, and far below the Monte-Carlo error of any estimator built on them.

Argument decomposition (the device code mirrors this, see ``table_eval()`` in ``csrc/orp_device.cuh``):

    x in [0, 2^30)          integer Sobol point, u = x / 2^30
    v = min(x, 2^30 - x)    fold to the nearer end (side 1 = upper half: sign baked into the row)
    d = (double) v          ONE conversion does the normalisation: exponent e = floor(log2 v),
                            mantissa m = v / 2^e in [1, 2)
    slab = max(hi32(d) >> 16, BASE) - BASE      = 1 + 16 e + j,  j = top four mantissa bits;
                                                  v = 0 has hi32 = 0 and clamps to slab 0: the -inf row
    s = m - (1 + (j + 1/2) / 16)  in [-1/32, 1/32)      (exact in fp64)
    z = c0 + s (c1 + s (c2 + s (c3 + s (c4 + s c5))))    c2.. and their Horner steps in fp32

Placement: 128-byte slabs of eight 16-byte units.  A quarter-warp (8 lanes = 8 consecutive aligned
points) has pairwise distinct g = x >> 27, and the row of group g keeps (c0, c1) in unit g and
(c2..c5) in unit g ^ 2 of its slab: each of the two LDS.128 of a quarter-warp touches eight different
bank groups.  Groups that meet in a slab — (0,7), (1,6,7), (2,5,6), (3,4,5): the fold makes the upper half
straddle one row — never differ by exactly 2 (asserted below), so their units are disjoint and no per-group
offsets are needed: byte address = slab * 128 + g * 16, second read at address ^ 32.

Accuracy is verified against mpmath (40 digits) and printed; the maximum error is recorded in
the generated file.
"""
from __future__ import annotations

import os
import sys

import mpmath as mp
import numpy as np

mp.mp.dps = 45
BITS = 30
J = 16                      # sub-intervals per octave
DEG = 5                     # polynomial degree: 2 fp64 + 4 fp32 coefficients = 32 bytes per row
N_LOW = 2                   # coefficients kept in fp64
N_SLABS = 1 + 16 * 30       # slab 0 = the v = 0 row; e = 0 .. 29
HALF = mp.mpf(1) / 32       # half-width of a sub-interval in units of the mantissa


def ndtri_mp(u):
    return -mp.sqrt(2) * mp.erfinv(1 - 2 * mp.mpf(u))


def row_fn(e, j):
    """s -> ndtri(v / 2^30) with v = 2^e (1 + (j + 1/2)/16 + s), side 0."""
    centre = 1 + (mp.mpf(j) + mp.mpf(1) / 2) / J
    sc = mp.mpf(2) ** (e - BITS)
    return lambda s: ndtri_mp(sc * (centre + s))


def fit_row(e, j):
    f = row_fn(e, j)
    n = DEG + 1
    nodes = [HALF * mp.cos(mp.pi * (2 * i + 1) / (2 * n)) for i in range(n)]
    A = mp.matrix(n, n)
    for i, x in enumerate(nodes):
        for k in range(n):
            A[i, k] = x ** k
    c = mp.lu_solve(A, mp.matrix([f(x) for x in nodes]))
    c = [c[k] for k in range(n)]
    # the high coefficients are stored in fp32: round them first, then refit the fp64 ones to the remainder
    hi = [mp.mpf(float(np.float32(float(x)))) for x in c[N_LOW:]]
    nodes2 = [HALF * mp.cos(mp.pi * (2 * i + 1) / (2 * N_LOW)) for i in range(N_LOW)]
    rem = [f(x) - sum(h * x ** (N_LOW + k) for k, h in enumerate(hi)) for x in nodes2]
    B = mp.matrix(N_LOW, N_LOW)
    for i, x in enumerate(nodes2):
        for k in range(N_LOW):
            B[i, k] = x ** k
    low = mp.lu_solve(B, mp.matrix(rem))
    return [float(low[k]) for k in range(N_LOW)] + [float(h) for h in hi]


def device_eval(cd, s):
    """Bit-faithful emulation of table_eval(): fp32 Horner for c2..c5, then two fp64 FMAs."""
    f = np.float32
    s32 = f(s)
    hi = f(f(cd[5]) * s32 + f(cd[4]))
    hi = f(hi * s32 + f(cd[3]))
    hi = f(hi * s32 + f(cd[2]))
    acc = float(hi) * s + cd[1]
    return acc * s + cd[0]


def check_row(e, j, cd, ntest=9):
    f = row_fn(e, j)
    worst = 0.0
    for s in np.linspace(-1 / 32, 1 / 32, ntest, endpoint=False):
        s = float(s)
        ex = f(mp.mpf(s))
        err = abs(mp.mpf(device_eval(cd, s)) - ex)
        worst = max(worst, float(err / max(abs(ex), mp.mpf(1))))      # relative in the tails, absolute near 0
    return worst


def slab_of(v):
    if v == 0:
        return 0
    e = v.bit_length() - 1
    j = (v >> (e - 4)) & 15 if e >= 4 else (v << (4 - e)) & 15
    return 1 + 16 * e + j


def groups_of(e, j):
    """The groups g = x >> 27 that can index row (e, j): integers v in [2^e (1 + j/16), 2^e (1 + (j+1)/16))."""
    lo = -((-(2 ** e) * (J + j)) // J)                 # ceil
    hi = -((-(2 ** e) * (J + j + 1)) // J) - 1         # last integer below the upper end
    gs = set()
    if lo > hi:
        return gs
    if hi > 2 ** 29:
        hi = 2 ** 29
    if lo > hi:
        return gs
    for v in (lo, hi):
        assert slab_of(v) == 1 + 16 * e + j, (e, j, v)
    # side 0: x = v (< 2^29); side 1: x = 2^30 - v (>= 2^29); g is monotone in v on each side
    if lo < 2 ** 29:
        gs.update(range(lo >> 27, (min(hi, 2 ** 29 - 1) >> 27) + 1))
    x_hi, x_lo = 2 ** BITS - lo, 2 ** BITS - hi
    if x_hi >= 2 ** 29 and x_hi < 2 ** BITS:
        gs.update(range(max(x_lo, 2 ** 29) >> 27, (x_hi >> 27) + 1))
    return gs


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(
        os.path.dirname(os.path.abspath(__file__)), "..",
        "option-replicating-portfolio-with-neural-networks_b200", "csrc", "ndtri_table.inc")
    table = np.zeros((N_SLABS, 8, 2), dtype=np.float64)        # raw 16-byte units, viewed as 2 doubles
    t32 = table.view(np.float32).reshape(N_SLABS, 8, 4)
    used = np.zeros((N_SLABS, 8), dtype=bool)
    owner = np.zeros(N_SLABS, dtype=np.uint8)                  # bit g set: group g owns a row in this slab

    def place(slab, g, cd):
        sign = 1.0 if g < 4 else -1.0
        for unit in (g, g ^ 2):
            assert not used[slab, unit], (slab, g, unit)
            used[slab, unit] = True
        owner[slab] |= 1 << g
        table[slab, g, 0] = sign * cd[0]
        table[slab, g, 1] = sign * cd[1]
        t32[slab, g ^ 2, :] = [np.float32(sign * c) for c in cd[2:6]]

    worst = 0.0
    place(0, 0, [float("-inf"), 0.0, 0.0, 0.0, 0.0, 0.0])      # x = 0: ndtri(0) = -inf (scipy.special.ndtri)
    for e in range(30):
        for j in range(J):
            gs = groups_of(e, j)
            if not gs:
                continue
            if e == 29:
                cd = [0.0] * 6                                  # only v = 2^29 (u = 1/2) lands here: exactly 0
            else:
                cd = fit_row(e, j)
                worst = max(worst, check_row(e, j, cd))
            for g in sorted(gs):
                place(1 + 16 * e + j, g, cd)
    flat = table.reshape(-1)
    with open(out_path, "w") as f:
        f.write("// GENERATED by tools/gen_ndtri_table.py — do not edit.\n")
        f.write(f"// piecewise degree-{DEG} inverse normal CDF, {J} sub-intervals per octave, 30-bit grid, 32-byte rows\n")
        f.write(f"// max error vs mpmath (relative for |z|>1, absolute otherwise): {worst:.3e}\n")
        f.write(f"#define ORP_NDTRI_SLABS {N_SLABS}\n")
        f.write(f"#define ORP_NDTRI_MAX_ERR {worst:.3e}\n")
        # which groups own a row in each slab: needed to find the rows when the table is transformed
        f.write("ORP_NDTRI_OWNER_DECL = {\n")
        for i in range(0, owner.size, 32):
            f.write("  " + ", ".join(str(int(v)) for v in owner[i:i + 32]) + ",\n")
        f.write("};\n")
        # emitted as raw 64-bit words: unit g ^ 2 of a row holds four fp32 values
        words = flat.view(np.uint64)
        f.write("ORP_NDTRI_TABLE_DECL = {\n")
        for i in range(0, words.size, 4):
            f.write("  " + ", ".join(f"0x{int(v):016x}ull" for v in words[i:i + 4]) + ",\n")
        f.write("};\n")
    print(f"wrote {out_path}: {N_SLABS} slabs, {flat.size * 8 / 1024:.1f} KiB, max err {worst:.3e}, "
          f"cells used {used.sum()} / {used.size}")


if __name__ == "__main__":
    main()
