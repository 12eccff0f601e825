/*
 * test_managers.cpp -- the reference's own manager tests, run against B200DiscreteManager / B200CastManager through
 * the C++ plugin interface.  Scenarios and expected numbers follow
 *   tesseract_collision/include/tesseract_collision/test_suite/collision_box_box_unit.hpp        (runTest, :14-270)
 *   tesseract_collision/include/tesseract_collision/test_suite/collision_box_box_cast_unit.hpp   (runTest, :14-176)
 *   tesseract_collision/include/tesseract_collision/test_suite/collision_clone_unit.hpp          (clone contract)
 * with the reference's tolerances.  Built by tests/test_cpp_managers.py twice: against the host simulator of the
 * kernels (CPU tier) and against libt4c.so (GPU tier).  Exit code 0 = all expectations hold.
 */
#include <cmath>
#include <cstdio>
#include <stdexcept>

#include "../../tesseract_b200/csrc/plugin/b200_contact_managers.h"

using namespace tesseract_collision;
using tesseract_b200::B200CastManager;
using tesseract_b200::B200DiscreteManager;

static int g_fail = 0, g_checks = 0;
#define EXPECT_TRUE(x)                                                                                                 \
  do                                                                                                                   \
  {                                                                                                                    \
    ++g_checks;                                                                                                        \
    if (!(x))                                                                                                          \
    {                                                                                                                  \
      ++g_fail;                                                                                                        \
      std::printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #x);                                                         \
    }                                                                                                                  \
  } while (0)
#define EXPECT_FALSE(x) EXPECT_TRUE(!(x))
#define EXPECT_NEAR(a, b, tol)                                                                                         \
  do                                                                                                                   \
  {                                                                                                                    \
    ++g_checks;                                                                                                        \
    const double _a = (a), _b = (b);                                                                                   \
    if (!(std::fabs(_a - _b) <= (tol)))                                                                                \
    {                                                                                                                  \
      ++g_fail;                                                                                                        \
      std::printf("FAIL %s:%d: %s = %.9g, expected %.9g +- %g\n", __FILE__, __LINE__, #a, _a, _b, (double)(tol));      \
    }                                                                                                                  \
  } while (0)

static Eigen::Isometry3d translation(double x, double y, double z)
{
  Eigen::Isometry3d T = Eigen::Isometry3d::Identity();
  T(0, 3) = x, T(1, 3) = y, T(2, 3) = z;
  return T;
}

static bool approxIdentity(const Eigen::Isometry3d& T, const Eigen::Isometry3d& R = Eigen::Isometry3d::Identity())
{
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 4; ++c)
      if (std::fabs(T(r, c) - R(r, c)) > 1e-5)
        return false;
  return true;
}

static void flatten(const ContactResultMap& m, ContactResultVector& v)
{
  v.clear();
  for (const auto& kv : m)
    v.insert(v.end(), kv.second.begin(), kv.second.end());
}

static void addOne(DiscreteContactManager& c, const char* name, tesseract_geometry::Geometry::ConstPtr g, bool enabled = true)
{
  c.addCollisionObject(name, 0, { g }, { Eigen::Isometry3d::Identity() }, enabled);
}

// collision_box_box_unit.hpp:14-118
static void boxBoxAdd(DiscreteContactManager& checker)
{
  addOne(checker, "box_link", std::make_shared<tesseract_geometry::Box>(1, 1, 1), false);
  checker.enableCollisionObject("box_link");
  addOne(checker, "thin_box_link", std::make_shared<tesseract_geometry::Box>(0.1, 1, 1));
  checker.disableCollisionObject("thin_box_link");
  addOne(checker, "second_box_link", std::make_shared<tesseract_geometry::Box>(2, 2, 2));
  addOne(checker, "remove_box_link", std::make_shared<tesseract_geometry::Box>(0.1, 1, 1));
  EXPECT_TRUE(checker.getCollisionObjects().size() == 4);
  EXPECT_TRUE(checker.hasCollisionObject("remove_box_link"));
  checker.removeCollisionObject("remove_box_link");
  EXPECT_FALSE(checker.hasCollisionObject("remove_box_link"));
  EXPECT_FALSE(checker.removeCollisionObject("link_does_not_exist"));
  EXPECT_FALSE(checker.enableCollisionObject("link_does_not_exist"));
  EXPECT_FALSE(checker.disableCollisionObject("link_does_not_exist"));
  EXPECT_FALSE(checker.addCollisionObject("empty_link", 0, CollisionShapesConst(), tesseract_common::VectorIsometry3d()));
  EXPECT_TRUE(checker.getCollisionObjects().size() == 3);
  for (const auto& co : checker.getCollisionObjects())
  {
    EXPECT_TRUE(checker.getCollisionObjectGeometries(co).size() == 1);
    EXPECT_TRUE(checker.getCollisionObjectGeometriesTransforms(co).size() == 1);
    for (const auto& cgt : checker.getCollisionObjectGeometriesTransforms(co))
      EXPECT_TRUE(approxIdentity(cgt));
  }
  EXPECT_TRUE(checker.getCollisionObjectGeometries("link_does_not_exist").empty());
}

// collision_box_box_unit.hpp:120-262
static void boxBoxTyped(DiscreteContactManager& checker, ContactTestType test_type)
{
  checker.setActiveCollisionObjects({ "box_link", "second_box_link" });
  checker.setCollisionMarginData(CollisionMarginData(0.1));
  EXPECT_NEAR(checker.getCollisionMarginData().getPairCollisionMargin("box_link", "second_box_link"), 0.1, 1e-5);
  tesseract_common::TransformMap location;
  location["box_link"] = translation(0.2, 0.1, 0);
  location["second_box_link"] = Eigen::Isometry3d::Identity();
  checker.setCollisionObjectsTransform(location);

  ContactResultMap result;
  checker.contactTest(result, ContactRequest(test_type));
  ContactResultVector rv;
  flatten(result, rv);
  EXPECT_TRUE(!rv.empty());
  if (rv.empty())
    return;
  EXPECT_NEAR(rv[0].distance, -1.30, 0.001);
  EXPECT_NEAR(rv[0].nearest_points[0][1], rv[0].nearest_points[1][1], 0.001);
  EXPECT_NEAR(rv[0].nearest_points[0][2], rv[0].nearest_points[1][2], 0.001);
  int idx[3] = { 0, 1, 1 };
  if (rv[0].link_names[0] != "box_link")
    idx[0] = 1, idx[1] = 0, idx[2] = -1;
  EXPECT_NEAR(rv[0].nearest_points[(size_t)idx[0]][0], -0.3, 0.001);
  EXPECT_NEAR(rv[0].nearest_points[(size_t)idx[1]][0], 1.0, 0.001);
  EXPECT_NEAR(rv[0].normal[0], idx[2] * -1.0, 0.001);
  EXPECT_NEAR(rv[0].normal[1], 0.0, 0.001);
  EXPECT_NEAR(rv[0].normal[2], 0.0, 0.001);
  EXPECT_TRUE(approxIdentity(rv[0].transform[(size_t)idx[0]], location["box_link"]));

  // outside the contact distance
  location["box_link"] = translation(1.60, 0, 0);
  result.clear();
  checker.setCollisionObjectsTransform(std::vector<std::string>{ "box_link" },
                                       tesseract_common::VectorIsometry3d{ location["box_link"] });
  checker.contactTest(result, ContactRequest(test_type));
  flatten(result, rv);
  EXPECT_TRUE(rv.empty());

  // a larger margin on an unrelated pair raises the max margin only
  {
    CollisionMarginData data = checker.getCollisionMarginData();
    data.setPairCollisionMargin("not_box_link", "also_not_box_link", 1.7);
    checker.setCollisionMarginData(data);
    EXPECT_NEAR(checker.getCollisionMarginData().getMaxCollisionMargin(), 1.7, 0);
    EXPECT_NEAR(checker.getCollisionMarginData().getPairCollisionMargin("box_link", "second_box_link"), 0.1, 1e-5);
    result.clear();
    checker.contactTest(result, ContactRequest(test_type));
    flatten(result, rv);
    EXPECT_TRUE(rv.empty());
  }
  // inside the contact distance for this pair only
  {
    result.clear();
    CollisionMarginData data(0.1);
    data.setPairCollisionMargin("box_link", "second_box_link", 0.25);
    checker.setCollisionMarginData(data);
    EXPECT_NEAR(checker.getCollisionMarginData().getMaxCollisionMargin(), 0.25, 1e-5);
    checker.contactTest(result, ContactRequest(test_type));
    flatten(result, rv);
    EXPECT_TRUE(!rv.empty());
    if (rv.empty())
      return;
    EXPECT_NEAR(rv[0].distance, 0.1, 0.001);
    idx[0] = 0, idx[1] = 1, idx[2] = 1;
    if (rv[0].link_names[0] != "box_link")
      idx[0] = 1, idx[1] = 0, idx[2] = -1;
    EXPECT_NEAR(rv[0].nearest_points[(size_t)idx[0]][0], 1.1, 0.001);
    EXPECT_NEAR(rv[0].nearest_points[(size_t)idx[1]][0], 1.0, 0.001);
    EXPECT_NEAR(rv[0].normal[0], idx[2] * -1.0, 0.001);
    // the caller's map is not cleared by contactTest (bullet_discrete_bvh_manager.cpp:269)
    checker.contactTest(result, ContactRequest(ContactTestType::ALL));
    EXPECT_TRUE(result.begin()->second.size() == 2);
  }
  // IsContactAllowedFn and is_valid
  {
    checker.setIsContactAllowedFn([](const std::string& a, const std::string& b) { return a != b; });
    result.clear();
    checker.contactTest(result, ContactRequest(test_type));
    EXPECT_TRUE(result.empty());
    checker.setIsContactAllowedFn(nullptr);
    ContactRequest rq(test_type);
    rq.is_valid = [](const ContactResult&) { return false; };
    checker.contactTest(result, rq);
    EXPECT_TRUE(result.empty());
    rq.is_valid = [](const ContactResult& c) { return c.distance < 0.2; };
    checker.contactTest(result, rq);
    EXPECT_TRUE(result.size() == 1);
  }
}

// collision_clone_unit.hpp: a clone answers exactly like its source and is independent afterwards
static void cloneContract(DiscreteContactManager& checker)
{
  checker.setCollisionMarginData(CollisionMarginData(0.1));
  checker.setCollisionObjectsTransform("box_link", translation(0.2, 0.1, 0));
  auto clone = checker.clone();
  ContactResultMap a, b;
  checker.contactTest(a, ContactRequest(ContactTestType::ALL));
  clone->contactTest(b, ContactRequest(ContactTestType::ALL));
  ContactResultVector va, vb;
  flatten(a, va);
  flatten(b, vb);
  EXPECT_TRUE(va.size() == vb.size() && !va.empty());
  for (size_t i = 0; i < va.size() && i < vb.size(); ++i)
    EXPECT_NEAR(va[i].distance, vb[i].distance, 1e-9);
  EXPECT_TRUE(clone->getActiveCollisionObjects() == checker.getActiveCollisionObjects());
  EXPECT_TRUE(clone->getCollisionObjects() == checker.getCollisionObjects());
  clone->setCollisionObjectsTransform("box_link", translation(5, 0, 0));
  b.clear();
  clone->contactTest(b, ContactRequest(ContactTestType::ALL));
  EXPECT_TRUE(b.empty());
  a.clear();
  checker.contactTest(a, ContactRequest(ContactTestType::ALL));
  EXPECT_TRUE(!a.empty());
}

// collision_box_box_cast_unit.hpp:14-176
static void boxBoxCast(ContinuousContactManager& checker)
{
  const auto I = Eigen::Isometry3d::Identity();
  checker.addCollisionObject("static_box_link", 0, { std::make_shared<tesseract_geometry::Box>(1, 1, 1) }, { I }, false);
  checker.enableCollisionObject("static_box_link");
  checker.addCollisionObject("thin_box_link", 0, { std::make_shared<tesseract_geometry::Box>(0.1, 1, 1) }, { I });
  checker.disableCollisionObject("thin_box_link");
  const Eigen::Isometry3d moving_box_pose = translation(0.5, -0.5, 0);
  checker.addCollisionObject("moving_box_link", 0, { std::make_shared<tesseract_geometry::Box>(0.25, 0.25, 0.25) },
                             { moving_box_pose });
  checker.addCollisionObject("remove_box_link", 0, { std::make_shared<tesseract_geometry::Box>(0.1, 1, 1) }, { I });
  EXPECT_TRUE(checker.getCollisionObjects().size() == 4);
  checker.removeCollisionObject("remove_box_link");
  checker.removeCollisionObject("remove_box_link");
  EXPECT_FALSE(checker.hasCollisionObject("remove_box_link"));
  EXPECT_TRUE(checker.getCollisionObjects().size() == 3);
  EXPECT_TRUE(approxIdentity(checker.getCollisionObjectGeometriesTransforms("moving_box_link")[0], moving_box_pose));

  checker.setActiveCollisionObjects({ "moving_box_link" });
  checker.setCollisionMarginData(CollisionMarginData(0.1));
  EXPECT_NEAR(checker.getCollisionMarginData().getMaxCollisionMargin(), 0.1, 1e-5);
  checker.setCollisionObjectsTransform(std::vector<std::string>{ "static_box_link" },
                                       tesseract_common::VectorIsometry3d{ I });
  const Eigen::Isometry3d start_pos = translation(-1.9, 0.0, 0), end_pos = translation(1.9, 3.8, 0);
  checker.setCollisionObjectsTransform("moving_box_link", start_pos, end_pos);
  for (ContactTestType t : { ContactTestType::ALL, ContactTestType::CLOSEST, ContactTestType::FIRST })
  {
    ContactResultMap result;
    checker.contactTest(result, ContactRequest(t));
    ContactResultVector rv;
    flatten(result, rv);
    EXPECT_TRUE(!rv.empty());
    if (rv.empty())
      continue;
    EXPECT_NEAR(rv[0].distance, -0.2475, 0.001);
    EXPECT_NEAR(rv[0].cc_time[0], -1.0, 0.001);
    EXPECT_NEAR(rv[0].cc_time[1], 0.25, 0.001);
    EXPECT_TRUE(rv[0].cc_type[0] == ContinuousCollisionType::CCType_None);
    EXPECT_TRUE(rv[0].cc_type[1] == ContinuousCollisionType::CCType_Between);
    EXPECT_NEAR(rv[0].nearest_points[0][0], -0.5, 0.001);
    EXPECT_NEAR(rv[0].nearest_points[0][1], 0.5, 0.001);
    EXPECT_NEAR(rv[0].nearest_points[0][2], 0.0, 0.001);
    EXPECT_NEAR(rv[0].nearest_points[1][0], -0.325, 0.001);
    EXPECT_NEAR(rv[0].nearest_points[1][1], 0.325, 0.001);
    EXPECT_NEAR(rv[0].nearest_points[1][2], 0.0, 0.001);
    const Eigen::Vector3d p0 = rv[0].transform[1] * rv[0].nearest_points_local[1];
    EXPECT_NEAR(p0[0], -1.275, 0.001);
    EXPECT_NEAR(p0[1], -0.625, 0.001);
    EXPECT_NEAR(p0[2], 0.0, 0.001);
    const Eigen::Vector3d p1 = rv[0].cc_transform[1] * rv[0].nearest_points_local[1];
    EXPECT_NEAR(p1[0], 2.525, 0.001);
    EXPECT_NEAR(p1[1], 3.175, 0.001);
    EXPECT_NEAR(p1[2], 0.0, 0.001);
  }
}

int main()
{
  try
  {
    for (ContactTestType t : { ContactTestType::FIRST, ContactTestType::CLOSEST, ContactTestType::ALL })
    {
      B200DiscreteManager checker;
      boxBoxAdd(checker);
      boxBoxTyped(checker, t);
    }
    {
      auto checker = B200DiscreteManager::create();
      EXPECT_TRUE(B200DiscreteManager::name() == "B200DiscreteManager");
      boxBoxAdd(*checker);
      checker->setActiveCollisionObjects({ "box_link", "second_box_link" });
      cloneContract(*checker);
    }
    {
      std::unique_ptr<ContinuousContactManager> checker(tesseract_b200_create_cast_manager());
      boxBoxCast(*checker);
      auto clone = checker->clone();
      ContactResultMap r;
      clone->contactTest(r, ContactRequest(ContactTestType::CLOSEST));
      EXPECT_TRUE(r.size() == 1);
    }
  }
  catch (const std::exception& e)
  {
    std::printf("EXCEPTION: %s\n", e.what());
    return 2;
  }
  std::printf("%d checks, %d failures\n", g_checks, g_fail);
  return g_fail ? 1 : 0;
}
