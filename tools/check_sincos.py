"""Accuracy of the reduced-range double sincos of csrc/arm_kernel.cuh (sincos_r<double>) restated in Python floats
(the device version uses fma for the reduction and the polynomials, so it is at least as accurate)."""
import math, random
S = [-1.66666666666666324348e-01, 8.33333333332248946124e-03, -1.98412698298579493134e-04, 2.75573137070700676789e-06, -2.50507602534068634195e-08, 1.58969099521155010221e-10]
C = [4.16666666666666019037e-02, -1.38888888888741095749e-03, 2.48015872894767294178e-05, -2.75573143513906633035e-07, 2.08757232129817482790e-09, -1.13596475577881948265e-11]
def sincos(x):
    n = round(x * 0.63661977236758134308)
    r = x - n * 1.5707963267948966; r -= n * 6.123233995736766e-17
    z = r * r
    ps = S[5]; pc = C[5]
    for k in (4, 3, 2, 1, 0): ps = ps * z + S[k]; pc = pc * z + C[k]
    s, c = r + r * z * ps, 1.0 + z * (z * pc - 0.5)
    return [(s, c), (c, -s), (-s, -c), (-c, s)][int(n) & 3]
random.seed(1); worst = 0.0
for _ in range(400000):
    x = random.uniform(-7, 7); s, c = sincos(x)
    worst = max(worst, abs(s - math.sin(x)), abs(c - math.cos(x)))
print("max abs error over [-7, 7]:", worst)
assert worst < 3e-16
