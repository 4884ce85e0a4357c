// TEST INFRASTRUCTURE: accuracy of mpde_sincosf (meta_pde_b200/csrc/compat.cuh, the branch-free sin/cos of every SIREN
// activation) against double-precision sin/cos.  Prints one line per decade: max abs error of sin and cos; then the
// special values.  Built and run by tests/test_sincos_accuracy.py.
#define MPDE_EMU 1
#include "compat.cuh"
#include <cmath>
#include <cstdint>
#include <cstdio>

int main() {
  uint64_t st = 0x9E3779B97F4A7C15ull;
  auto next = [&]() { st ^= st << 13; st ^= st >> 7; st ^= st << 17; return (double)(st >> 11) / 9007199254740992.0; };
  const double ranges[] = {1.0, 10.0, 100.0, 1e3, 1e4, 1e5, 1e6};
  for (double R : ranges) {
    double es = 0, ec = 0;
    for (int i = 0; i < 2000000; ++i) {
      const float x = (float)((2.0 * next() - 1.0) * R);
      float s, c;
      mpde_sincosf(x, &s, &c);
      const double ds = std::fabs((double)s - std::sin((double)x)), dc = std::fabs((double)c - std::cos((double)x));
      if (ds > es) es = ds;
      if (dc > ec) ec = dc;
    }
    std::printf("range %g sin %.4e cos %.4e\n", R, es, ec);
  }
  float s, c;
  mpde_sincosf(0.0f, &s, &c);      std::printf("zero %g %g\n", s, c);
  mpde_sincosf(3e9f, &s, &c);      std::printf("huge %d\n", (std::fabs(s) <= 1.01f && std::fabs(c) <= 1.01f) ? 1 : 0);
  mpde_sincosf(INFINITY, &s, &c);  std::printf("inf %d\n", (std::isnan(s) && std::isnan(c)) ? 1 : 0);
  mpde_sincosf(NAN, &s, &c);       std::printf("nan %d\n", (std::isnan(s) && std::isnan(c)) ? 1 : 0);
  return 0;
}
