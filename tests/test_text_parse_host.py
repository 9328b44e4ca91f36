"""CPU: the device's text-field parser (csrc/text_parse.inl), compiled for the host by
tests/native/text_parse_host.cu, against the C library's sscanf("%f") as the oracle calls it
(HadoopUtils::toFloat, c++/utils/impl/StringUtils.cc:49-56).  Bit-exact, including which
strings are rejected."""
import ctypes as C
import os
import random
import shutil
import struct
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "native", "text_parse_host.cu")
INL = os.path.join(HERE, "..", "hadoop-1.0.3-gpu_b200", "csrc", "text_parse.inl")
SO = os.path.join(HERE, "native", "_build", "libtext_parse_host.so")


@pytest.fixture(scope="module")
def host_parser():
    stale = not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(SRC), os.path.getmtime(INL))
    if stale:
        nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
        if not os.path.exists(nvcc):
            pytest.skip("nvcc is needed to compile the parser for the host")
        os.makedirs(os.path.dirname(SO), exist_ok=True)
        subprocess.run([nvcc, "-O2", "-shared", "-Xcompiler", "-fPIC", "-Wno-deprecated-gpu-targets", "-o", SO, SRC],
                       check=True, capture_output=True)
    lib = C.CDLL(SO)

    def parse_many(fields):
        buf = b"".join(fields)
        off = np.zeros(len(fields) + 1, np.int64)
        np.cumsum([len(f) for f in fields], out=off[1:])
        bits = np.zeros(len(fields), np.uint32)
        rc = np.zeros(len(fields), np.int32)
        lib.text_parse_host_many(buf, off.ctypes.data_as(C.c_void_p), C.c_longlong(len(fields)),
                                 bits.ctypes.data_as(C.c_void_p), rc.ctypes.data_as(C.c_void_p))
        return bits, rc

    return parse_many


def compare(host_parser, oracle, fields):
    fields = [f if isinstance(f, bytes) else f.encode() for f in fields]
    bits, rc = host_parser(fields)
    obits, orc = oracle.to_float_many(fields)
    assert not (rc == 1).any(), "bignum ran out of room for %r" % fields[int(np.argmax(rc == 1))]
    bad = np.nonzero((rc != orc) | ((rc == 0) & (bits != obits)))[0]
    assert len(bad) == 0, "%d differ, first: %r ours rc=%d %08x, sscanf rc=%d %08x" % (
        len(bad), fields[bad[0]], rc[bad[0]], bits[bad[0]], orc[bad[0]], obits[bad[0]])


GRAMMAR = ["1e", "1e+", "1e-", ".5", "5.", ".", "+", "-", "+.e1", "1.5f", "0x", "0x1p", "0x1p3", "0x.8", "0x1.8", "0X1P-2",
           "0x1.ffffffp127", "0x1.fffffep127", "0x1.ffffff0p127", "0x1.fffffefp127", "0x1p-150", "0x1.000001p-149",
           "0x1p-149", "0x3p-150", "0x1.8p-149", "0x0.0000000000000000001p0", "0x123456789abcdef0123p-40",
           "0x1.00000100000000000001p0", "0x1.000001p0", "0x1.000003p0", "1_0", "1e5000", "1e-5000", "0e999999999999",
           "infinit", "infx", "nanx", "nan(abc)", "nan()", "NaN", "-nan", "+inf", "INFINITY", "-Infinity", "inf", "iNf",
           "1e+05", "1E5", "1.e5", ".e5", "+.5", "-.5e-3", "--1", "+-1", "\v1.5", "\f2", "\v\f-3", "1\v", "00012", "0.000",
           "-0", "-0.0e10", "1e40", "3.4028236e38", "3.4028235e38", "3.40282356779733661637539395458142568448e38",
           "3.40282356779733661637539395458142568447e38", "3.40282356779733661637539395458142568449e38",
           "340282356779733661637539395458142568448", "1e-46", "7.006492321624085e-46", "7.006492321624086e-46",
           "7.00649232162408535461864791644958065640130970938257885878534141944895541342930300743319094181060791015625e-46",
           "7.00649232162408535461864791644958065640130970938257885878534141944895541342930300743319094181060791015626e-46",
           "1.4e-45", "16777217", "16777219", "16777217.0000000000000000000000000000000000001",
           "1.00000017881393432617187500", "1.00000017881393432617187501", "1.000000178813934326171875",
           "1.00000005960464477539062500", "1.00000005960464477539062501", "1.00000005960464477539062499",
           "9007199791611905.0", "1e23", "123456789012345678901234567890", "0.1", "0xg", "0x.", "-0x.", "0x.p1",
           "1e99999999999999999999", "1e-99999999999999999999", "nan(", "in", "i", "n", "na", "+nan", "1..2", "1.2.3",
           "1e1e1", "1d5", "0x1p+", "0x1.8p", "0x1.8p1.5", "0x1p1p1", "infinityx", "0x1e1", "0x1e+1", "0x1.p", "0x.0p0",
           "0x0p9999999999", "e5", "E", "0e", "0e+", "0x1P", "1.17549435e-38", "1.17549428e-38", "1.1754943e-38",
           "5.8774717e-39", "0." + "0" * 100 + "1", "1" + "0" * 38, "1" + "0" * 39, "9" * 39, "0" * 120 + "1.5",
           "1.175494350822287507968736537222245677818665556772087521508751706278417259454727172851560500000000000000000e-38",
           "\xc2\xa01", "١", "1\x00", ""]


def test_grammar_and_boundary_cases(host_parser, oracle):
    compare(host_parser, oracle, [g for g in GRAMMAR if g != "" and "\x00" not in g])


def rand_decimal(rng):
    style = rng.randrange(8)
    sign = rng.choice(["", "", "-", "+"])
    if style == 0:   # what "%f" / "%g" print
        return sign + "%.*f" % (rng.randrange(0, 9), rng.uniform(-1, 1) * 10 ** rng.randrange(-3, 6))
    if style == 1:
        return "%.*e" % (rng.randrange(0, 12), rng.uniform(-1, 1) * 10.0 ** rng.randrange(-50, 42))
    if style == 2:   # digit soup
        a = "".join(rng.choice("0123456789") for _ in range(rng.randrange(0, 30)))
        b = "".join(rng.choice("0123456789") for _ in range(rng.randrange(0, 30)))
        e = rng.choice(["", "", "e%d" % rng.randrange(-60, 50), "E+%d" % rng.randrange(0, 45)])
        return sign + a + rng.choice([".", ""]) + b + e
    if style == 3:   # an exact float, printed with too few / enough / too many digits
        f = struct.unpack("<f", struct.pack("<I", rng.getrandbits(32) & 0x7F7FFFFF))[0]
        return sign + "%.*e" % (rng.choice([6, 7, 8, 9, 17, 40]), f)
    if style == 4:   # exactly half way between two neighbouring floats, and one digit either side
        u = rng.getrandbits(31) % 0x7F7FFFFF
        lo = struct.unpack("<f", struct.pack("<I", u))[0]
        hi = struct.unpack("<f", struct.pack("<I", u + 1))[0]
        from fractions import Fraction
        mid = (Fraction(lo) + Fraction(hi)) / 2
        if mid == 0:
            return "0"
        # exact decimal expansion of mid (finite: denominator is a power of two)
        import math
        e10 = 0
        while mid.denominator != 1:
            mid *= 10
            e10 -= 1
        digits = str(mid.numerator)
        if len(digits) > 110:
            return "%.*e" % (20, float(lo))
        tweak = rng.choice([0, 0, 1, -1])
        digits = str(mid.numerator * 10 + tweak) if tweak else digits
        e10 -= 1 if tweak else 0
        return sign + digits + "e%d" % e10
    if style == 5:   # hexadecimal
        a = "".join(rng.choice("0123456789abcdefABCDEF") for _ in range(rng.randrange(0, 20)))
        b = "".join(rng.choice("0123456789abcdef") for _ in range(rng.randrange(0, 20)))
        p = rng.choice(["", "p%d" % rng.randrange(-200, 140), "P+%d" % rng.randrange(0, 130), "p"])
        return sign + rng.choice(["0x", "0X"]) + a + rng.choice([".", ""]) + b + p
    if style == 6:   # near the ends of the range
        return sign + "%d.%de%d" % (rng.randrange(1, 10), rng.randrange(0, 10 ** 9), rng.choice([-46, -45, -44, -39, -38, -37, 37, 38, 39]))
    # mutations of valid text
    base = "%.6f" % rng.uniform(-100, 100)
    pos = rng.randrange(len(base) + 1)
    return base[:pos] + rng.choice("eE.+-xXpPnNiI 0123456789a\v\f_") .strip(" ") + base[pos:]


def test_random_literals_round_like_sscanf(host_parser, oracle):
    rng = random.Random(20261106)
    fields = []
    while len(fields) < 300000:
        f = rand_decimal(rng)
        if 0 < len(f.encode()) <= 127 and not any(c in f for c in " ,\t\r\n"):
            fields.append(f)
    compare(host_parser, oracle, fields)


def test_every_float_survives_nine_digits(host_parser, oracle):
    """%.9g round-trips binary32: a sample of all exponents, every mantissa pattern class."""
    rng = random.Random(7)
    bits = [rng.getrandbits(32) for _ in range(100000)] + list(range(0, 64)) + [0x7F7FFFFF, 0x00800000, 0x007FFFFF, 1]
    bits = [b for b in bits if (b & 0x7F800000) != 0x7F800000]
    fields = [("%.9g" % struct.unpack("<f", struct.pack("<I", b))[0]).encode() for b in bits]
    got, rc = host_parser(fields)
    assert (rc == 0).all()
    np.testing.assert_array_equal(got, np.array(bits, np.uint32))


def test_the_fuzz_reaches_every_branch(host_parser, oracle):
    """Guards the tests above against a generator that only makes easy cases."""
    rng = random.Random(99)
    fields = [rand_decimal(rng).encode() for _ in range(40000)]
    fields = [f for f in fields if 0 < len(f) <= 127 and not any(c in f for c in b" ,\t\r\n")]
    bits, rc = oracle.to_float_many(fields)
    ok = rc == 0
    assert 0.05 < (~ok).mean() < 0.5                       # rejected strings
    mag = bits[ok] & 0x7FFFFFFF
    assert (mag == 0x7F800000).sum() > 50                  # overflow to inf
    assert ((mag > 0) & (mag < 0x00800000)).sum() > 50     # denormals
    assert (mag == 0).sum() > 50                           # zeros and underflow
    assert sum(1 for f in fields if len(f) > 40) > 1000    # long digit strings (exact ties and their neighbours)
    assert sum(1 for f in fields if b"x" in f.lower()) > 1000
