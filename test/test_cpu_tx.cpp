/* Message ordering and tag construction (reference: test/test_cpu_tx.cpp checks tag uniqueness). */
#include "harness.hpp"
#include <algorithm>
#include <set>
#include "stencil/tx_common.hpp"

TEST_CASE("tags are unique and non-negative", "[tx]") {
  std::set<int> seen;
  for (int payload = 0; payload < 64; ++payload)
    for (int z = -1; z <= 1; ++z)
      for (int y = -1; y <= 1; ++y)
        for (int x = -1; x <= 1; ++x) {
          const int t[4] = {make_tag<MsgKind::ColocatedEvt>(payload, Dim3(x, y, z)), make_tag<MsgKind::ColocatedMem>(payload, Dim3(x, y, z)),
                            make_tag<MsgKind::ColocatedDev>(payload, Dim3(x, y, z)),
                            make_tag<MsgKind::ColocatedNotify>(payload, Dim3(x, y, z))};
          for (int v : t) {
            REQUIRE(v >= 0);
            REQUIRE(v < (1 << 24));
            REQUIRE(seen.insert(v).second);
          }
        }
  std::set<int> legacy;
  for (int gpu = 0; gpu < 4; ++gpu)
    for (int idx = 0; idx < 4; ++idx)
      for (int z = -1; z <= 1; ++z)
        for (int y = -1; y <= 1; ++y)
          for (int x = -1; x <= 1; ++x) {
            const int t = make_tag(gpu, idx, Dim3(x, y, z));
            REQUIRE(t >= 0);
            REQUIRE(legacy.insert(t).second);
          }
}

TEST_CASE("messages sort by direction only", "[tx]") {
  std::vector<Message> m = {Message(Dim3(1, 0, 0), 3, 1), Message(Dim3(-1, 1, 0), 0, 2), Message(Dim3(0, -1, 1), 5, 5), Message(Dim3(-1, 0, 0), 9, 9)};
  std::sort(m.begin(), m.end());
  REQUIRE(m[0].dir_ == Dim3(-1, 0, 0));
  REQUIRE(m[1].dir_ == Dim3(-1, 1, 0));
  REQUIRE(m[2].dir_ == Dim3(0, -1, 1));
  REQUIRE(m[3].dir_ == Dim3(1, 0, 0));
  REQUIRE(Message(Dim3(1, 0, 0), 1, 2) == Message(Dim3(1, 0, 0), 1, 2));
  REQUIRE(!(Message(Dim3(1, 0, 0), 1, 2) == Message(Dim3(1, 0, 0), 2, 1)));
}
