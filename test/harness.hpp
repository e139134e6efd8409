#pragma once
/* Tiny self-registering test harness (the reference vendors Catch2 v2; this library has no third-party code).
   TEST_CASE("name", "[tags]") { REQUIRE(cond); }   run:  ./test_cpu [substring-of-name-or-tag] */
#include <cstdio>
#include <cstring>
#include <functional>
#include <string>
#include <vector>

namespace harness {
struct Case {
  std::string name, tags;
  std::function<void()> fn;
};
inline std::vector<Case> &cases() {
  static std::vector<Case> c;
  return c;
}
struct Failure {};
struct State {
  int checks = 0, failed = 0;
  const char *current = "";
};
inline State &state() {
  static State s;
  return s;
}
struct Reg {
  Reg(const char *n, const char *t, std::function<void()> f) { cases().push_back({n, t, std::move(f)}); }
};
inline int run(int argc, char **argv) {
  const char *filter = argc > 1 ? argv[1] : nullptr;
  int ran = 0, bad = 0;
  for (Case &c : cases()) {
    if (filter && c.name.find(filter) == std::string::npos && c.tags.find(filter) == std::string::npos) continue;
    ++ran;
    state().current = c.name.c_str();
    const int before = state().failed;
    try {
      c.fn();
    } catch (const Failure &) {
    }
    const bool ok = state().failed == before;
    if (!ok) ++bad;
    std::printf("%s %s\n", ok ? "[ ok ]" : "[FAIL]", c.name.c_str());
    std::fflush(stdout);
  }
  std::printf("%d test cases, %d failed, %d assertions\n", ran, bad, state().checks);
  return bad ? 1 : 0;
}
} // namespace harness

#define HARNESS_CAT2(a, b) a##b
#define HARNESS_CAT(a, b) HARNESS_CAT2(a, b)
#define TEST_CASE(...) HARNESS_TEST_CASE(HARNESS_CAT(harness_case_, __COUNTER__), __VA_ARGS__, "")
#define HARNESS_TEST_CASE(id, name, tags, ...)                                                                         \
  static void id();                                                                                                    \
  static harness::Reg HARNESS_CAT(id, _reg)(name, tags, id);                                                           \
  static void id()
#define REQUIRE(cond)                                                                                                  \
  do {                                                                                                                 \
    ++harness::state().checks;                                                                                         \
    if (!(cond)) {                                                                                                     \
      ++harness::state().failed;                                                                                       \
      std::printf("  %s:%d: REQUIRE(%s) failed in '%s'\n", __FILE__, __LINE__, #cond, harness::state().current);       \
      throw harness::Failure();                                                                                        \
    }                                                                                                                  \
  } while (0)
#define CHECK(cond)                                                                                                    \
  do {                                                                                                                 \
    ++harness::state().checks;                                                                                         \
    if (!(cond)) {                                                                                                     \
      ++harness::state().failed;                                                                                       \
      std::printf("  %s:%d: CHECK(%s) failed in '%s'\n", __FILE__, __LINE__, #cond, harness::state().current);         \
    }                                                                                                                  \
  } while (0)
