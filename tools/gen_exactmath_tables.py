"""Constants of include/b2r_exactmath.h as double-double pairs (mpmath, 200 bits): atan(k/16), 1/(2k+1), 1/n!, pi pieces.
python tools/gen_exactmath_tables.py > /tmp/tables.h   (the output is pasted into the header; rerun to verify)"""
import mpmath as mp

mp.mp.prec = 300


def dd(x):
    hi = float(x)
    lo = float(x - mp.mpf(hi))
    return hi, lo


def td(x):
    a = float(x); r = x - mp.mpf(a)
    b = float(r); r = r - mp.mpf(b)
    c = float(r)
    return a, b, c


def emit(name, vals):
    print(f"B2R_XM_CONST double {name}[{len(vals)}][2] = {{")
    for hi, lo in vals:
        print(f"    {{{hi!r}, {lo!r}}},")
    print("};")


emit("kAtanTab", [dd(mp.atan(mp.mpf(k) / 16)) for k in range(17)])
emit("kOddInv", [dd(mp.mpf(1) / (2 * k + 1)) for k in range(14)])
emit("kInvFact", [dd(mp.mpf(1) / mp.factorial(n)) for n in range(30)])
for nm, v in (("kPi", mp.pi), ("kPiHalf", mp.pi / 2), ("kPiQuarter", mp.pi / 4)):
    hi, lo = dd(v)
    print(f"B2R_XM_CONST double {nm}[2] = {{{hi!r}, {lo!r}}};")
a, b, c = td(mp.pi / 2)
print(f"B2R_XM_CONST double kPiHalf3[3] = {{{a!r}, {b!r}, {c!r}}};")
print(f"// 2/pi = {float(2 / mp.pi)!r}")
