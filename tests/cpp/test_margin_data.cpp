// CollisionMarginData semantics with the numbers of the reference's unit test
// (tesseract_collision/test/collision_margin_data_unit.cpp:13-245): construction, default / pair setters and the
// maximum they imply, incrementMargins / scaleMargins, and the five CollisionMarginOverrideType modes of apply().
// Runs on the CPU tier against the signature shim the plugin is built with (plugin/tesseract_shim.h); with the real
// Tesseract headers (T4_USE_TESSERACT_HEADERS) the same file exercises the reference class.
#include <cmath>
#include <cstdio>
#include <functional>
#include <limits>
#include <vector>

#include "../../tesseract_b200/csrc/plugin/b200_contact_managers.h"

using tesseract_collision::CollisionMarginData;
using tesseract_collision::CollisionMarginOverrideType;

struct Expect
{
  double dflt, max, pair12;
  size_t n_pairs;
  double pair13;  // < 0: not checked
};
struct Case
{
  const char* name;
  std::function<CollisionMarginData()> make;
  Expect want;
};

int main()
{
  const double tol = std::numeric_limits<double>::epsilon();
  const double d = 0.0254, p = 0.5;
  auto with_pair = [&](double dflt, double pair) {
    CollisionMarginData m(dflt);
    m.setPairCollisionMargin("link_1", "link_2", pair);
    return m;
  };
  auto applied = [&](CollisionMarginData base, CollisionMarginData over, CollisionMarginOverrideType t) {
    base.apply(over, t);
    return base;
  };
  std::vector<Case> cases = {
    { "default constructor", [&] { return CollisionMarginData(); }, { 0, 0, 0, 0, -1 } },
    { "non-zero default", [&] { return CollisionMarginData(d); }, { d, d, d, 0, -1 } },
    { "changing the default", [&] { CollisionMarginData m; m.setDefaultCollisionMargin(d); return m; }, { d, d, d, 0, -1 } },
    { "pair above the default", [&] { return with_pair(d, p); }, { d, p, p, 1, -1 } },
    { "pair below the default", [&] { return with_pair(d, 0.01); }, { d, d, 0.01, 1, -1 } },
    { "default raised above the maximum", [&] { auto m = with_pair(d, p); m.setDefaultCollisionMargin(2 * p); return m; },
      { 2 * p, 2 * p, p, 1, -1 } },
    { "pair lowered again: maximum falls back to the default",
      [&] { auto m = with_pair(d, p); m.setPairCollisionMargin("link_1", "link_2", d); return m; }, { d, d, d, 1, -1 } },
    { "default lowered below the pair", [&] { auto m = with_pair(0.05, d); m.setDefaultCollisionMargin(0.0); return m; },
      { 0.0, d, d, 1, -1 } },
    { "increment > 0", [&] { auto m = with_pair(d, p); m.incrementMargins(0.01); return m; }, { d + 0.01, p + 0.01, p + 0.01, 1, -1 } },
    { "increment < 0", [&] { auto m = with_pair(d, p); m.incrementMargins(-0.01); return m; }, { d - 0.01, p - 0.01, p - 0.01, 1, -1 } },
    { "scale > 1", [&] { auto m = with_pair(d, p); m.scaleMargins(1.5); return m; }, { d * 1.5, p * 1.5, p * 1.5, 1, -1 } },
    { "scale < 1", [&] { auto m = with_pair(d, p); m.scaleMargins(0.5); return m; }, { d * 0.5, p * 0.5, p * 0.5, 1, -1 } },
    { "apply OVERRIDE_DEFAULT_MARGIN",
      [&] { return applied(with_pair(d, p), CollisionMarginData(3 * d), CollisionMarginOverrideType::OVERRIDE_DEFAULT_MARGIN); },
      { 3 * d, p, p, 1, -1 } },
    { "apply OVERRIDE_PAIR_MARGIN",
      [&] { return applied(with_pair(d, p), with_pair(3 * d, 3 * p), CollisionMarginOverrideType::OVERRIDE_PAIR_MARGIN); },
      { d, 3 * p, 3 * p, 1, -1 } },
    { "apply REPLACE",
      [&] { return applied(with_pair(d, p), with_pair(3 * d, 3 * p), CollisionMarginOverrideType::REPLACE); },
      { 3 * d, 3 * p, 3 * p, 1, -1 } },
    { "apply NONE", [&] { return applied(with_pair(d, p), with_pair(3 * d, 3 * p), CollisionMarginOverrideType::NONE); },
      { d, p, p, 1, -1 } },
    { "apply MODIFY_PAIR_MARGIN, new pair",
      [&] {
        CollisionMarginData over(3 * d);
        over.setPairCollisionMargin("link_1", "link_3", 3 * p);
        return applied(with_pair(d, p), over, CollisionMarginOverrideType::MODIFY_PAIR_MARGIN);
      },
      { d, 3 * p, p, 2, 3 * p } },
    { "apply MODIFY_PAIR_MARGIN, existing pair",
      [&] {
        CollisionMarginData over(3 * d);
        over.setPairCollisionMargin("link_1", "link_2", 3 * p);
        over.setPairCollisionMargin("link_1", "link_3", 9 * p);
        return applied(with_pair(d, p), over, CollisionMarginOverrideType::MODIFY_PAIR_MARGIN);
      },
      { d, 9 * p, 3 * p, 2, 9 * p } },
  };
  int fail = 0, checks = 0;
  auto near = [&](const char* name, const char* what, double got, double want) {
    ++checks;
    if (!(std::fabs(got - want) <= tol))
    {
      ++fail;
      std::printf("FAIL %s: %s = %.17g, expected %.17g\n", name, what, got, want);
    }
  };
  for (const Case& c : cases)
  {
    const CollisionMarginData m = c.make();
    near(c.name, "default", m.getDefaultCollisionMargin(), c.want.dflt);
    near(c.name, "max", m.getMaxCollisionMargin(), c.want.max);
    near(c.name, "pair(link_1, link_2)", m.getPairCollisionMargin("link_1", "link_2"), c.want.pair12);
    near(c.name, "pair(link_2, link_1)", m.getPairCollisionMargin("link_2", "link_1"), c.want.pair12);
    if (c.want.pair13 >= 0)
      near(c.name, "pair(link_1, link_3)", m.getPairCollisionMargin("link_1", "link_3"), c.want.pair13);
    ++checks;
    if (m.getPairCollisionMargins().size() != c.want.n_pairs)
    {
      ++fail;
      std::printf("FAIL %s: %zu pair margins, expected %zu\n", c.name, m.getPairCollisionMargins().size(), c.want.n_pairs);
    }
  }
  std::printf("%d checks, %d failures\n", checks, fail);
  return fail ? 1 : 0;
}
