"""Bit-level IEEE-754 binary32 addition in pure integer arithmetic (numpy uint64), round-to-nearest-even.

TEST INFRASTRUCTURE ONLY (see oracle/gpuprobe_oracle.h).  The reference holds no golden vectors for this
path, so the oracle's arithmetic (`c[i] = a[i] + b[i]` executed by the host FPU) is additionally anchored
here against the standard itself: this module never uses a floating-point add.  It follows the textbook
algorithm of IEEE 754-2019 section 4.3.1 / 6.3 for binary32:

  1. unpack sign, biased exponent, fraction; subnormals have exponent field 0, effective exponent 1 and no
     hidden bit; infinities are handled apart (NaN operands are outside the probe's domain and rejected);
  2. order the operands by magnitude, append three extra bits (guard, round, sticky) to both significands,
     shift the smaller one right by the exponent difference, OR-ing every bit shifted out into sticky;
  3. add or subtract the aligned significands according to the signs;
  4. normalise (one right shift after a carry, left shifts after cancellation, never below exponent 1);
  5. round to nearest, ties to even, on the three extra bits; renormalise if the rounding carried;
  6. overflow gives infinity; an exact zero is +0 unless both operands were -0 (roundTiesToEven rule);
  7. pack: a significand without the hidden bit is stored with exponent field 0 (subnormal or zero).
"""
from __future__ import annotations

import numpy as np

_U = np.uint64


def f32_add_bits(a_bits: np.ndarray, b_bits: np.ndarray) -> np.ndarray:
    """uint32 bit patterns in, uint32 bit pattern of fl(a + b) out.  No NaN operands; inf + (-inf) is rejected."""
    a = np.asarray(a_bits, dtype=np.uint32).astype(_U)
    b = np.asarray(b_bits, dtype=np.uint32).astype(_U)
    sa, ea, fa = a >> _U(31), (a >> _U(23)) & _U(0xFF), a & _U(0x7FFFFF)
    sb, eb, fb = b >> _U(31), (b >> _U(23)) & _U(0xFF), b & _U(0x7FFFFF)
    if np.any((ea == 0xFF) & (fa != 0)) or np.any((eb == 0xFF) & (fb != 0)):
        raise ValueError("NaN operands are outside the probe's domain")
    a_inf, b_inf = ea == 0xFF, eb == 0xFF
    if np.any(a_inf & b_inf & (sa != sb)):
        raise ValueError("inf + (-inf) is outside the probe's domain")

    # 1. effective exponents and 24-bit significands
    xa = np.where(ea == 0, _U(1), ea)
    xb = np.where(eb == 0, _U(1), eb)
    ma = np.where(ea == 0, fa, fa | _U(0x800000))
    mb = np.where(eb == 0, fb, fb | _U(0x800000))

    # 2. order by magnitude: (exponent, significand) lexicographic
    a_big = (xa > xb) | ((xa == xb) & (ma >= mb))
    ex = np.where(a_big, xa, xb)
    ey = np.where(a_big, xb, xa)
    mx = np.where(a_big, ma, mb) << _U(3)
    my = np.where(a_big, mb, ma) << _U(3)
    sx = np.where(a_big, sa, sb)
    sy = np.where(a_big, sb, sa)
    d = np.minimum(ex - ey, _U(40))  # beyond 27 bits everything is sticky anyway
    lost = (my & ((_U(1) << d) - _U(1))) != 0
    my = (my >> d) | lost.astype(_U)

    # 3. add or subtract
    same = sx == sy
    s = np.where(same, mx + my, mx - my)
    e = ex.astype(np.int64)

    # 4. normalise: carry out of bit 26 -> one right shift keeping sticky
    carry = (s >> _U(27)) != 0
    s = np.where(carry, (s >> _U(1)) | (s & _U(1)), s)
    e = np.where(carry, e + 1, e)
    #    cancellation -> left shifts while the hidden-bit position (bit 26) is clear and the exponent can drop
    for _ in range(27):
        need = (s != 0) & ((s >> _U(26)) == 0) & (e > 1)
        if not need.any():
            break
        s = np.where(need, s << _U(1), s)
        e = np.where(need, e - 1, e)

    # 5. round to nearest, ties to even, on the guard/round/sticky bits
    grs = s & _U(7)
    lsb = (s >> _U(3)) & _U(1)
    up = (grs > 4) | ((grs == 4) & (lsb == 1))
    m = (s >> _U(3)) + up.astype(_U)
    ovf = (m >> _U(24)) != 0  # rounding carried out of the 24-bit significand
    m = np.where(ovf, m >> _U(1), m)
    e = np.where(ovf, e + 1, e)

    # 6./7. pack
    hidden = (m >> _U(23)) & _U(1)
    exp_field = np.where(hidden == 1, e, 0).astype(np.int64)
    is_inf = exp_field >= 255
    frac = np.where(is_inf, _U(0), m & _U(0x7FFFFF))
    exp_field = np.where(is_inf, 255, exp_field).astype(_U)
    sign = sx
    zero = m == 0
    sign = np.where(zero, np.where((sa == 1) & (sb == 1), _U(1), _U(0)), sign)  # x + (-x) = +0; -0 + -0 = -0
    out = (sign << _U(31)) | (exp_field << _U(23)) | frac

    # infinities: inf + finite = inf, inf + inf (same sign) = inf
    out = np.where(a_inf, a, out)
    out = np.where(b_inf & ~a_inf, b, out)
    return out.astype(np.uint32)
