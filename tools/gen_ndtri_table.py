#!/usr/bin/env python
"""Generate ``csrc/ndtri_table.inc``: piecewise-polynomial inverse normal CDF on the 30-bit
Sobol grid, laid out for conflict-free 128-bit shared-memory reads.

Why a table: inside a warp the 32 consecutive Sobol points of any dimension are stratified
(exactly one point per 1/32 interval), so EVERY warp holds central and tail arguments at every
step and a branchy rational approximation (Cephes / AS241) would execute both branches always.
One uniform instruction stream — index a row, run one Horner chain.  A row is 48 bytes: c0..c3 in fp64 and
c4..c7 in fp32 (the top four terms contribute < 1e-7 of the value, so fp32 there costs < 1e-14): three
128-bit shared-memory reads, 3 FFMA + 1 F2F + 1 DADD + 4 DFMA per point.

Argument decomposition (device code mirrors this, see ``ndtri_bits()`` in ``csrc/orp_device.cuh``):

    x in [0, 2^30)          integer Sobol point, u = x / 2^30
    side = x >> 29          v = side ? 2^30 - x : x          (fold to the nearest end)
    lz = clz(v); oct = lz - 3; mant = (v << lz) << 1         (leading one dropped; v = 0: lz = 32, mant = 0)
    j = mant >> 28 (16 sub-intervals per octave);  f = (mant << 4) / 2^32;  t = f - 1/2
    rho = (oct + 1) * 16 + j                                  (rho = 0: the exact-zero row, u = 1/2)
    z = sum_k c[g][rho][k] t^k,  g = x >> 27                  (sign baked into the row per group)
    unit 0 = (c0, c1) fp64, unit 1 = (c2, c3) fp64, unit 2 = (c4, c5, c6, c7) fp32, unit 3 unused

Placement: 128-byte slabs of eight 16-byte units.  A quarter-warp (8 lanes = 8 consecutive
aligned points) has pairwise distinct g (top three digits), and row (g, rho) keeps coefficient
unit k at position ``(g & 4) | ((g & 3) ^ k)`` of slab ``rho + OFFS[g]``: the k-th LDS.128 of the
quarter-warp therefore touches eight different bank groups.  Groups g and g^4 share slabs.

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
DEG = 7                     # polynomial degree  (8 coefficients = 64 bytes per row)
N_OCT = 29                  # oct = -1 .. 28  ->  rho in [0, 480)
N_RHO = (N_OCT + 1) * J
RHO_ZERO_ARG = N_RHO              # x = 0: clz = 32 lands one octave further; that row is the constant -inf
OFFS = [-48, 848, 880, 432, 0, 864, 880, 400]     # slab = rho + OFFS[g]; multiples of 16
N_SLABS = 928


def ndtri_mp(u):
    return -mp.sqrt(2) * mp.erfinv(1 - 2 * mp.mpf(u))      # valid for u in (0, 1/2]


def row_interval(rho):
    """[lo, hi) in units of u for row rho >= 16."""
    octv, j = rho // J - 1, rho % J
    sc = mp.mpf(2) ** (-(octv + 2))
    return sc * (1 + mp.mpf(j) / J), sc * (1 + mp.mpf(j + 1) / J)


def fit_row(rho):
    """Monomial coefficients in t in [-1/2, 1/2) of ndtri on the row's interval (side 0 sign)."""
    lo, hi_ = row_interval(rho)
    n = DEG + 1
    nodes = [mp.cos(mp.pi * (2 * i + 1) / (2 * n)) / 2 for i in range(n)]
    vals = [ndtri_mp(lo + (hi_ - lo) * (x + mp.mpf(1) / 2)) for x in nodes]
    A = mp.matrix(n, n)
    for i, x in enumerate(nodes):
        for k in range(n):
            A[i, k] = x ** k
    c = mp.lu_solve(A, mp.matrix(vals))
    c = [c[k] for k in range(n)]
    # c4..c7 are stored in fp32: round them first, then refit c0..c3 to the remainder so that the cubic part
    # absorbs what a cubic can of the rounding (the best cubic leaves 1/8 of a t^4 perturbation on [-1/2, 1/2])
    hi = [mp.mpf(float(np.float32(float(x)))) for x in c[4:]]
    nodes4 = [mp.cos(mp.pi * (2 * i + 1) / 8) / 2 for i in range(4)]
    f = lambda x: ndtri_mp(lo + (hi_ - lo) * (x + mp.mpf(1) / 2))
    rem = [f(x) - sum(h * x ** (4 + k) for k, h in enumerate(hi)) for x in nodes4]
    B = mp.matrix(4, 4)
    for i, x in enumerate(nodes4):
        for k in range(4):
            B[i, k] = x ** k
    low = mp.lu_solve(B, mp.matrix(rem))
    return [low[k] for k in range(4)] + hi


def device_eval(cd, tt):
    """Bit-faithful emulation of ndtri_bits(): fp32 Horner for c4..c7, then fp64."""
    f = np.float32
    t32 = f(tt)
    hi = f(f(cd[7]) * t32 + f(cd[6]))
    hi = f(hi * t32 + f(cd[5]))
    hi = f(hi * t32 + f(cd[4]))
    acc = float(hi) * tt + cd[3]
    acc = acc * tt + cd[2]
    acc = acc * tt + cd[1]
    return acc * tt + cd[0]


def check_row(rho, cd, ntest=12):
    lo, hi = row_interval(rho)
    worst = 0.0
    for tt in np.linspace(-0.5, 0.5, ntest, endpoint=False):
        tt = float(tt)
        acc = device_eval(cd, tt)
        ex = ndtri_mp(lo + (hi - lo) * (mp.mpf(tt) + mp.mpf(1) / 2))
        err = abs(mp.mpf(acc) - ex)
        # relative in the tails, absolute near the centre where z -> 0
        worst = max(worst, float(err / max(abs(ex), mp.mpf(1))))
    return worst


def reachable_rows(g):
    """Rows rho the device can index from group g = x >> 27.  The fold v = 2^30 - x shifts the
    side-1 cell boundaries by one grid unit, so side-1 groups straddle one extra row."""
    from fractions import Fraction
    rows = set()
    if g == 4:
        rows.add(0)                                   # x = 2^29 (u = 1/2): the exact-zero row
    for rho in range(16, N_RHO):
        octv, j = rho // J - 1, rho % J
        lo = Fraction((J + j) * 2 ** 28, J * 2 ** octv)
        hi = Fraction((J + j + 1) * 2 ** 28, J * 2 ** octv)
        v_lo = -((-lo.numerator) // lo.denominator)               # ceil
        v_hi = -((-hi.numerator) // hi.denominator) - 1            # last integer < hi
        if v_lo > v_hi:
            continue
        for v in (v_lo, v_hi):
            assert rho_of(v) == rho
            if (v >> 27) == g and v < (1 << 29):
                rows.add(rho)                          # side 0: x = v
            x = (1 << BITS) - v
            if x >= (1 << 29) and (x >> 27) == g:
                assert rho_of(x) == rho
                rows.add(rho)                          # side 1
    return sorted(rows)


def rho_of(x):
    side = x >> 29
    v = ((1 << BITS) - x) if side else x
    lz = 32 - v.bit_length()
    mant = ((v << lz) << 1) & 0xFFFFFFFF
    return (lz - 3 + 1) * J + (mant >> 28)


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(
        os.path.dirname(os.path.abspath(__file__)), "..",
        "option-replicating-portfolio-with-neural-networks_b200", "csrc", "ndtri_table.inc")
    coef = {}
    worst = 0.0
    for rho in range(16, N_RHO):
        c = fit_row(rho)
        cd = [float(x) for x in c]
        worst = max(worst, check_row(rho, cd))
        coef[rho] = cd
    coef[0] = [0.0] * (DEG + 1)
    coef[RHO_ZERO_ARG] = [float('-inf')] + [0.0] * DEG       # ndtri(0) = -inf (scipy.special.ndtri)
    table = np.zeros((N_SLABS, 8, 2), dtype=np.float64)        # raw 16-byte units, viewed as 2 doubles
    t32 = table.view(np.float32).reshape(N_SLABS, 8, 4)
    used = np.zeros((N_SLABS, 8), dtype=bool)
    q_of = np.zeros((N_SLABS, 2), dtype=np.uint8)              # q = g & 3 of the row stored in each half-slab
    for g in range(8):
        sign = 1.0 if g < 4 else -1.0
        for rho in reachable_rows(g) + ([RHO_ZERO_ARG] if g == 0 else []):
            slab = rho + OFFS[g]
            assert 0 <= slab < N_SLABS, (g, rho, slab)
            q_of[slab, g >> 2] = g & 3
            for k in range(3):
                unit = (g & 4) | ((g & 3) ^ k)
                assert not used[slab, unit], (g, rho, k)
                used[slab, unit] = True
                if k < 2:
                    table[slab, unit, 0] = sign * coef[rho][2 * k]
                    table[slab, unit, 1] = sign * coef[rho][2 * k + 1]
                else:
                    t32[slab, unit, :] = [np.float32(sign * c) for c in coef[rho][4:8]]
    flat = table.reshape(-1)
    with open(out_path, "w") as f:
        f.write("// GENERATED by tools/gen_ndtri_table.py — do not edit.\n")
        f.write(f"// piecewise degree-{DEG} inverse normal CDF, {J} sub-intervals per octave, 30-bit grid\n")
        f.write(f"// max error vs mpmath (relative for |z|>1, absolute otherwise): {worst:.3e}\n")
        f.write(f"#define ORP_NDTRI_SLABS {N_SLABS}\n")
        f.write("#define ORP_NDTRI_OFFS_BIAS 48\n")
        packed = [(o + 48) // 16 for o in OFFS]
        lo = sum(p << (8 * i) for i, p in enumerate(packed[:4]))
        hi = sum(p << (8 * i) for i, p in enumerate(packed[4:]))
        f.write(f"#define ORP_NDTRI_OFFS_LO 0x{lo:08x}u\n#define ORP_NDTRI_OFFS_HI 0x{hi:08x}u\n")
        f.write(f"#define ORP_NDTRI_MAX_ERR {worst:.3e}\n")
        # q = g & 3 of the row in each half-slab: needed to un-swizzle rows when the table is transformed
        f.write("ORP_NDTRI_QMAP_DECL = {\n")
        qq = q_of.reshape(-1)
        for i in range(0, qq.size, 32):
            f.write("  " + ", ".join(str(int(v)) for v in qq[i:i + 32]) + ",\n")
        f.write("};\n")
        # emitted as raw 64-bit words: unit 2 of a row holds four fp32 values
        words = flat.view(np.uint64)
        f.write("ORP_NDTRI_TABLE_DECL = {\n")
        for i in range(0, words.size, 4):
            f.write("  " + ", ".join(f"0x{int(v):016x}ull" for v in words[i:i + 4]) + ",\n")
        f.write("};\n")
    print(f"wrote {out_path}: {N_SLABS} slabs, {flat.size * 8 / 1024:.1f} KiB, max err {worst:.3e}, "
          f"cells used {used.sum()} / {used.size}")


if __name__ == "__main__":
    main()
