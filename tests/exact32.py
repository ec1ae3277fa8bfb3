"""Hand-derivation helper for the known-answer tests: IEEE binary32 arithmetic carried out in exact rationals.

Every operation takes float32 values, computes the EXACT result as a `fractions.Fraction` and rounds it once to the nearest
binary32 (ties to even) — the definition of an IEEE operation, independent of any float unit, compiler flag or library.
The KATs in tests/kats.py evaluate the reference's formulas (cited there line by line) with these operations along ONE
straight-line path each; the oracle, the numpy restatement and the GPU must reproduce the resulting bit patterns.
sqrt is taken to 80 decimal digits before the single rounding; acos and pow go through double like the reference's
System.Math calls (Unity.Mathematics math.acos / math.pow) and are rounded to float.
"""
import math
from decimal import Decimal, getcontext
from fractions import Fraction

import numpy as np

getcontext().prec = 80


def rn(q):
    """round an exact rational to the nearest binary32, ties to even -> np.float32"""
    q = Fraction(q)
    if q == 0:
        return np.float32(0.0)
    sign = -1 if q < 0 else 1
    a = abs(q)
    e = a.numerator.bit_length() - a.denominator.bit_length()        # 2^e <= a < 2^(e+2) roughly
    while Fraction(2) ** e > a:
        e -= 1
    while Fraction(2) ** (e + 1) <= a:
        e += 1
    e = max(e, -126)                                                  # subnormals share the exponent of the smallest normal
    ulp = Fraction(2) ** (e - 23)
    n = a / ulp
    k = n.numerator // n.denominator
    rem = n - k
    if rem > Fraction(1, 2) or (rem == Fraction(1, 2) and (k & 1)):
        k += 1
    v = sign * k * ulp
    if abs(v) >= Fraction(2) ** 128:
        return np.float32(sign * np.inf)
    return np.float32(float(v))                                       # exact: v has <= 24 significant bits


def fr(x):
    return Fraction(float(x))


def add(a, b): return rn(fr(a) + fr(b))
def sub(a, b): return rn(fr(a) - fr(b))
def mul(a, b): return rn(fr(a) * fr(b))
def div(a, b): return rn(fr(a) / fr(b))


def sqrt(a):
    f = fr(a)
    d = (Decimal(f.numerator) / Decimal(f.denominator)).sqrt()
    return rn(Fraction(d))


def dot(a, b):                     # math.dot: x*x' + y*y' + z*z', left to right
    return add(add(mul(a[0], b[0]), mul(a[1], b[1])), mul(a[2], b[2]))


def vsub(a, b): return tuple(sub(x, y) for x, y in zip(a, b))
def vadd(a, b): return tuple(add(x, y) for x, y in zip(a, b))
def vmul(a, b): return tuple(mul(x, y) for x, y in zip(a, b))
def vscale(a, s): return tuple(mul(x, s) for x in a)
def vdiv(a, s): return tuple(div(x, s) for x in a)


def normalize(v):                  # math.normalize: rsqrt(dot(v, v)) * v, rsqrt(x) = 1 / sqrt(x)
    r = div(np.float32(1), sqrt(dot(v, v)))
    return tuple(mul(r, x) for x in v)


def acos_deg(x):                   # degrees(acos(x)) = (float)Math.Acos(x) * 57.29578f
    return mul(np.float32(math.acos(float(x))), np.float32(57.295779513))


def powf(x, y):                    # (float)Math.Pow(x, y)
    return np.float32(math.pow(float(x), float(y)))


def v(*xs):
    return tuple(np.float32(x) for x in xs)
