/* Array<T>. Same ground as the reference's test/test_cpu_array.cpp (ctor, resize, index). */
#include "harness.hpp"
#define STENCIL_USE_CUDA 0
#include "stencil/array.hpp"

TEST_CASE("array cpu", "[array]") {
  Array<int> a;
  REQUIRE(a.size() == 0);
  REQUIRE(a.data() == nullptr);
  a.resize(10);
  REQUIRE(a.size() == 10);
  REQUIRE(a.data() != nullptr);
  for (size_t i = 0; i < a.size(); ++i) a[i] = int(i) * 3;
  a.resize(20);
  REQUIRE(a[9] == 27);
  Array<int> b(a);
  REQUIRE(b.size() == 20);
  REQUIRE(b[5] == 15);
  Array<int> c(4, 9);
  REQUIRE(c[3] == 9);
  c.swap(b);
  REQUIRE(c.size() == 20);
  REQUIRE(b.size() == 4);
}
