"""`div_z` (lajolla_public_b200/csrc/lj_math.cuh): the zero-numerator shortcut the grid-medium tracker divides through must be
IEEE division bit for bit -- signs of zero, infinities, NaNs, denormals, zero and infinite denominators included.  The header
is compiled with nvcc as plain host code and run here, no GPU (-fmad=false like the library; x86 has no contraction either)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NVCC = "/usr/local/cuda/bin/nvcc"

SRC = r'''
#include <cstdio>
#include <cstring>
#include <cstdint>
#include <cmath>
#include <limits>
#include "lj_math.cuh"
static uint64_t bits(double v) { uint64_t b; memcpy(&b, &v, 8); return b; }
static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static uint64_t next() { rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17; return rng_state; }
int main() {
    const double inf = std::numeric_limits<double>::infinity(), nan = std::numeric_limits<double>::quiet_NaN();
    const double dmin = std::numeric_limits<double>::denorm_min(), nmin = std::numeric_limits<double>::min(), big = std::numeric_limits<double>::max();
    const double special[] = {0.0, -0.0, 1.0, -1.0, dmin, -dmin, nmin, -nmin, big, -big, inf, -inf, nan, 3.0, 1e-300, -1e300, 0.1, -7.25};
    long long n = 0;
    auto check = [&](double x, double d) {
        volatile double vx = x, vd = d;             // keep the compiler from folding the reference quotient
        const double ref = vx / vd, got = div_z(x, d);
        const bool both_nan = ref != ref && got != got;
        if (!both_nan && bits(ref) != bits(got)) { printf("FAIL %a / %a: %a (%016llx) vs div_z %a (%016llx)\n", x, d, ref, (unsigned long long)bits(ref), got, (unsigned long long)bits(got)); return false; }
        n++;
        return true;
    };
    for (double x : special) for (double d : special) if (!check(x, d)) return 1;
    for (int i = 0; i < 2000000; i++) {
        double x, d; uint64_t a = next(), b = next();
        memcpy(&x, &a, 8); memcpy(&d, &b, 8);       // every bit pattern: NaN payloads, denormals, both signs
        if (!check(x, d)) return 1;
        if (!check(0.0, d) || !check(-0.0, d)) return 1;    // the shortcut's own case against every kind of denominator
        if (!check(x, 0.0) || !check(x, -0.0)) return 1;
    }
    // d3 / d3 goes through div_z per component
    const d3 q = mk3(0.0, -0.0, 5.0) / mk3(-2.0, 4.0, 0.0);
    if (bits(q.x) != bits(-0.0) || bits(q.y) != bits(-0.0) || !(q.z == inf)) { printf("FAIL d3 / d3\n"); return 1; }
    printf("OK %lld\n", n);
    return 0;
}
'''


@pytest.fixture(scope="module")
def math_binary(tmp_path_factory):
    if not os.path.exists(NVCC):
        pytest.skip("nvcc not present")
    d = tmp_path_factory.mktemp("ljmath")
    src = d / "math_host.cu"
    src.write_text(SRC)
    exe = d / "math_host"
    r = subprocess.run([NVCC, "-std=c++17", "-O2", "-fmad=false", "-gencode", "arch=compute_100a,code=sm_100a",
                        "-I", os.path.join(ROOT, "lajolla_public_b200", "csrc"), "-o", str(exe), str(src)],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout[-3000:]
    return str(exe)


def test_div_z_is_ieee_division(math_binary):
    r = subprocess.run([math_binary], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    print(r.stdout[-500:])
    assert r.returncode == 0 and r.stdout.strip().startswith("OK"), r.stdout[-2000:]
