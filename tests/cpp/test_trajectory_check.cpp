/*
 * test_trajectory_check.cpp -- the batched checkTrajectory / checkTrajectories / isValid of
 * tesseract_b200/csrc/plugin/b200_trajectory_check.h against the reference's per-state loop.
 *
 * referenceCheckTrajectory() below is the control flow of tesseract_environment/core/utils.h:172-428 written against
 * the manager interface only (getState -> setCollisionObjectsTransform per active link -> contactTest -> push_back of
 * non-empty maps, FIRST stops at the first colliding state / sub-segment), with the state solver replaced by one-state
 * t4c_fk calls.  The reference snapshot has no test that calls checkTrajectory, so the expected values are produced
 * by that loop through the classic one-state contactTest() of the same managers.
 */
#include <cmath>
#include <cstdio>
#include <stdexcept>
#include <type_traits>

#include "../../tesseract_b200/csrc/plugin/b200_trajectory_check.h"

using namespace tesseract_collision;
using namespace tesseract_b200;

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
#define EXPECT_THROW(stmt)                                                                                             \
  do                                                                                                                   \
  {                                                                                                                    \
    ++g_checks;                                                                                                        \
    bool thrown = false;                                                                                               \
    try                                                                                                                \
    {                                                                                                                  \
      stmt;                                                                                                            \
    }                                                                                                                  \
    catch (const std::runtime_error&)                                                                                  \
    {                                                                                                                  \
      thrown = true;                                                                                                   \
    }                                                                                                                  \
    if (!thrown)                                                                                                       \
    {                                                                                                                  \
      ++g_fail;                                                                                                        \
      std::printf("FAIL %s:%d: no exception from %s\n", __FILE__, __LINE__, #stmt);                                    \
    }                                                                                                                  \
  } while (0)

static Eigen::Isometry3d translation(double x, double y, double z)
{
  Eigen::Isometry3d T = Eigen::Isometry3d::Identity();
  T(0, 3) = x, T(1, 3) = y, T(2, 3) = z;
  return T;
}

// a 4-joint arm (revolute z, revolute y, prismatic x, revolute y) on a pedestal; "elbow" has no collision geometry
static std::vector<JointSpec> armJoints()
{
  std::vector<JointSpec> J(6);
  J[0].name = "world_to_base", J[0].type = JointSpec::FIXED, J[0].parent_link = "world", J[0].child_link = "base";
  J[0].origin = translation(0, 0, 0.1);
  J[1].name = "j1", J[1].type = JointSpec::REVOLUTE, J[1].parent_link = "base", J[1].child_link = "shoulder";
  J[1].origin = translation(0, 0, 0.2), J[1].axis = Eigen::Vector3d(0, 0, 1);
  J[2].name = "j2", J[2].type = JointSpec::REVOLUTE, J[2].parent_link = "shoulder", J[2].child_link = "elbow";
  J[2].origin = translation(0, 0, 0.3), J[2].axis = Eigen::Vector3d(0, 1, 0);
  J[3].name = "j3", J[3].type = JointSpec::PRISMATIC, J[3].parent_link = "elbow", J[3].child_link = "forearm";
  J[3].origin = translation(0.25, 0, 0), J[3].axis = Eigen::Vector3d(1, 0, 0);
  J[4].name = "j4", J[4].type = JointSpec::REVOLUTE, J[4].parent_link = "forearm", J[4].child_link = "hand";
  J[4].origin = translation(0.3, 0, 0), J[4].axis = Eigen::Vector3d(0, 1, 0);
  J[5].name = "hand_to_tool", J[5].type = JointSpec::FIXED, J[5].parent_link = "hand", J[5].child_link = "tool";
  J[5].origin = translation(0.12, 0, 0);
  // given out of order on purpose: the solver sorts parent-before-child
  std::swap(J[1], J[4]);
  return J;
}

template <class Manager>
static void addArm(Manager& m)
{
  using namespace tesseract_geometry;
  const Eigen::Isometry3d I = Eigen::Isometry3d::Identity();
  m.addCollisionObject("base", 0, { std::make_shared<Box>(0.3, 0.3, 0.2) }, { I });
  m.addCollisionObject("shoulder", 0, { std::make_shared<Box>(0.12, 0.12, 0.3) }, { translation(0, 0, 0.15) });
  m.addCollisionObject("forearm", 0, { std::make_shared<Box>(0.3, 0.08, 0.08) }, { translation(0.15, 0, 0) });
  m.addCollisionObject("hand", 0, { std::make_shared<Box>(0.1, 0.1, 0.06) }, { translation(0.05, 0, 0) });
  m.addCollisionObject("tool", 0, { std::make_shared<Sphere>(0.04) }, { I });
  // static obstacles (not in the kinematic tree)
  m.addCollisionObject("post", 0, { std::make_shared<Box>(0.1, 0.1, 1.2) }, { I });
  m.setCollisionObjectsTransform("post", translation(0.55, 0.35, 0.6));
  m.addCollisionObject("shelf", 0, { std::make_shared<Box>(0.5, 0.6, 0.04) }, { I });
  m.setCollisionObjectsTransform("shelf", translation(0.2, -0.55, 0.75));
  m.addCollisionObject("ball", 0, { std::make_shared<Sphere>(0.12) }, { I });
  m.setCollisionObjectsTransform("ball", translation(-0.5, 0.2, 0.8));
  m.setCollisionObjectsTransform("base", translation(0, 0, 0.1));
  m.setActiveCollisionObjects({ "shoulder", "forearm", "hand", "tool" });
  m.setCollisionMarginData(CollisionMarginData(0.02));
  m.setIsContactAllowedFn([](const std::string& a, const std::string& b) {
    auto adj = [&](const char* x, const char* y) { return (a == x && b == y) || (a == y && b == x); };
    return adj("base", "shoulder") || adj("shoulder", "forearm") || adj("forearm", "hand") || adj("hand", "tool") ||
           adj("forearm", "tool");
  });
}

static std::vector<double> linspaced(long cnt, double lo, double hi)
{
  // Eigen::VectorXd::LinSpaced
  std::vector<double> v((size_t)cnt);
  const double step = cnt == 1 ? 0.0 : (hi - lo) / double(cnt - 1);
  const bool flip = std::fabs(hi) < std::fabs(lo);
  for (long i = 0; i < cnt; ++i)
  {
    if (cnt == 1)
      v[(size_t)i] = hi;
    else if (flip)
      v[(size_t)i] = (i == 0) ? lo : hi - double(cnt - 1 - i) * step;
    else
      v[(size_t)i] = (i == cnt - 1) ? hi : lo + double(i) * step;
  }
  return v;
}

// getState + setCollisionObjectsTransform for the active links (core/utils.h:92-94,136-137)
template <class Manager>
static void setState(Manager& m, const t4c_fk_model* model, const std::vector<double>& q0, const std::vector<double>* q1)
{
  const std::vector<std::string>& names = m.getCollisionObjects();
  std::vector<double> T0(names.size() * 12), T1(names.size() * 12);
  if (t4c_fk(model, m.core().scene(), q0.data(), 1, T0.data()) != T4C_OK)
    throw std::runtime_error(t4c_last_error());
  if (q1 && t4c_fk(model, m.core().scene(), q1->data(), 1, T1.data()) != T4C_OK)
    throw std::runtime_error(t4c_last_error());
  for (const std::string& link : m.getActiveCollisionObjects())
  {
    const int i = m.core().objectIndex(link);
    if constexpr (std::is_same<Manager, B200CastManager>::value)
      m.setCollisionObjectsTransform(link, fromPose12(&T0[(size_t)i * 12]), fromPose12(&T1[(size_t)i * 12]));
    else
      m.setCollisionObjectsTransform(link, fromPose12(&T0[(size_t)i * 12]));
  }
}

template <class Manager>
static bool referenceCheckTrajectory(std::vector<ContactResultMap>& contacts, Manager& m, const t4c_fk_model* model,
                                     const tesseract_common::TrajArray& traj, const CollisionCheckConfig& config)
{
  constexpr bool cast = std::is_same<Manager, B200CastManager>::value;
  const bool lvs = config.type == CollisionEvaluatorType::LVS_DISCRETE || config.type == CollisionEvaluatorType::LVS_CONTINUOUS;
  const bool first = config.contact_request.type == ContactTestType::FIRST;
  auto row = [&](long r) {
    std::vector<double> q((size_t)traj.cols());
    for (long c = 0; c < traj.cols(); ++c)
      q[(size_t)c] = traj(r, c);
    return q;
  };
  auto test = [&](const std::vector<double>& a, const std::vector<double>& b) {
    setState(m, model, a, cast ? &b : nullptr);
    ContactResultMap collisions;
    m.contactTest(collisions, config.contact_request);
    if (collisions.empty())
      return false;
    contacts.push_back(collisions);
    return true;
  };
  bool found = false;
  const long last = cast ? traj.rows() - 1 : traj.rows();
  for (long s = 0; s < last; ++s)
  {
    double dist = -1;
    if (s < traj.rows() - 1)
    {
      double acc = 0;
      for (long c = 0; c < traj.cols(); ++c)
        acc += (traj(s + 1, c) - traj(s, c)) * (traj(s + 1, c) - traj(s, c));
      dist = std::sqrt(acc);
    }
    if (lvs && dist > 0 && dist > config.longest_valid_segment_length)
    {
      const long cnt = (long)std::ceil(dist / config.longest_valid_segment_length) + 1;
      std::vector<std::vector<double>> cols;
      for (long c = 0; c < traj.cols(); ++c)
        cols.push_back(linspaced(cnt, traj(s, c), traj(s + 1, c)));
      for (long i = 0; i < cnt - 1; ++i)
      {
        std::vector<double> a((size_t)traj.cols()), b((size_t)traj.cols());
        for (long c = 0; c < traj.cols(); ++c)
          a[(size_t)c] = cols[(size_t)c][(size_t)i], b[(size_t)c] = cols[(size_t)c][(size_t)i + 1];
        if (test(a, b))
          found = true;
        if (found && first)
          break;
      }
    }
    else if (test(row(s), s + 1 < traj.rows() ? row(s + 1) : row(s)))
      found = true;
    if (found && first)
      break;
  }
  return found;
}

static void expectSame(const std::vector<ContactResultMap>& got, const std::vector<ContactResultMap>& want, bool first)
{
  EXPECT_TRUE(got.size() == want.size());
  for (size_t i = 0; i < got.size() && i < want.size(); ++i)
  {
    if (first)
    {
      // FIRST reports whichever contact the traversal meets first: only "some contact of a colliding pair" is defined
      EXPECT_TRUE(got[i].size() == 1 && want[i].size() == 1);
      continue;
    }
    EXPECT_TRUE(got[i].size() == want[i].size());
    for (const auto& kv : want[i])
    {
      const auto it = got[i].find(kv.first);
      EXPECT_TRUE(it != got[i].end());
      if (it == got[i].end())
        continue;
      EXPECT_TRUE(it->second.size() == kv.second.size());
      if (it->second.size() != kv.second.size())
        continue;
      for (size_t k = 0; k < kv.second.size(); ++k)
      {
        const ContactResult& a = it->second[k];
        const ContactResult& b = kv.second[k];
        EXPECT_NEAR(a.distance, b.distance, 1e-6);
        EXPECT_TRUE(a.link_names[0] == b.link_names[0] && a.link_names[1] == b.link_names[1]);
        for (int j = 0; j < 2; ++j)
        {
          EXPECT_NEAR((a.nearest_points[j] - b.nearest_points[j]).norm(), 0.0, 1e-6);
          EXPECT_NEAR((a.nearest_points_local[j] - b.nearest_points_local[j]).norm(), 0.0, 2e-6);
          EXPECT_NEAR((a.transform[j].translation() - b.transform[j].translation()).norm(), 0.0, 1e-9);
          EXPECT_NEAR(a.cc_time[j], b.cc_time[j], 1e-6);
          EXPECT_TRUE(a.cc_type[j] == b.cc_type[j]);
          EXPECT_NEAR((a.cc_transform[j].translation() - b.cc_transform[j].translation()).norm(), 0.0, 1e-9);
        }
        EXPECT_NEAR((a.normal - b.normal).norm(), 0.0, 1e-6);
      }
    }
  }
}

static tesseract_common::TrajArray sweep(int rows, int variant)
{
  // swings the arm through the post, under the shelf and past the ball
  tesseract_common::TrajArray t(rows, 4);
  for (int r = 0; r < rows; ++r)
  {
    const double u = rows == 1 ? 0.3 : double(r) / double(rows - 1);
    t(r, 0) = -1.9 + 3.9 * u + 0.2 * variant;
    t(r, 1) = -0.5 + 0.35 * std::sin(5.0 * u + variant);
    t(r, 2) = 0.05 + 0.1 * u;
    t(r, 3) = 0.4 * std::cos(3.0 * u);
  }
  return t;
}

template <class Manager>
static void runCases(bool cast)
{
  const std::vector<std::string> names{ "j1", "j2", "j3", "j4" };
  B200StateSolver solver(armJoints(), "world");
  EXPECT_TRUE(solver.getActiveJointNames() == names);
  Manager batched, classic;
  addArm(batched);
  addArm(classic);
  const t4c_fk_model* model = solver.model(classic.core(), names);
  EXPECT_TRUE(model == solver.model(classic.core(), names));  // cached
  const CollisionEvaluatorType plain = cast ? CollisionEvaluatorType::CONTINUOUS : CollisionEvaluatorType::DISCRETE;
  const CollisionEvaluatorType lvs = cast ? CollisionEvaluatorType::LVS_CONTINUOUS : CollisionEvaluatorType::LVS_DISCRETE;
  size_t colliding = 0;
  for (ContactTestType tt : { ContactTestType::FIRST, ContactTestType::CLOSEST, ContactTestType::ALL })
    for (CollisionEvaluatorType et : { plain, lvs })
    {
      CollisionCheckConfig config(0.02, ContactRequest(tt), et, 0.05);
      const tesseract_common::TrajArray traj = sweep(12, 0);
      std::vector<ContactResultMap> got, want;
      const bool fg = checkTrajectory(got, batched, solver, names, traj, config);
      const bool fw = referenceCheckTrajectory(want, classic, model, traj, config);
      EXPECT_TRUE(fg == fw);
      EXPECT_TRUE(fw);
      expectSame(got, want, tt == ContactTestType::FIRST);
      colliding += want.size();
    }
  EXPECT_TRUE(colliding > 20);
  {
    // the caller's vector is appended to, never cleared (contacts.push_back, core/utils.h:114,155)
    CollisionCheckConfig config(0.02, ContactRequest(ContactTestType::CLOSEST), lvs, 0.05);
    std::vector<ContactResultMap> got(3);
    checkTrajectory(got, batched, solver, names, sweep(6, 1), config);
    EXPECT_TRUE(got.size() > 3 && got[0].empty() && got[2].empty());
  }
  {
    // a contact-free motion: nothing pushed, false returned
    CollisionCheckConfig config(0.02, ContactRequest(ContactTestType::ALL), lvs, 0.05);
    tesseract_common::TrajArray traj(3, 4);
    for (int r = 0; r < 3; ++r)
      traj(r, 0) = 2.2 + 0.1 * r, traj(r, 1) = -1.2, traj(r, 2) = 0.0, traj(r, 3) = 0.0;
    std::vector<ContactResultMap> got, want;
    EXPECT_TRUE(!checkTrajectory(got, batched, solver, names, traj, config));
    EXPECT_TRUE(!referenceCheckTrajectory(want, classic, model, traj, config));
    EXPECT_TRUE(got.empty() && want.empty());
  }
  {
    // user validity callback: sees every candidate before the FIRST / CLOSEST reduction (core/common.h:117-119)
    ContactRequest rq(ContactTestType::CLOSEST);
    rq.is_valid = [](const ContactResult& c) { return c.link_names[0] != "post" && c.link_names[1] != "post"; };
    CollisionCheckConfig config(0.02, rq, lvs, 0.05);
    std::vector<ContactResultMap> got, want;
    const bool fg = checkTrajectory(got, batched, solver, names, sweep(12, 0), config);
    const bool fw = referenceCheckTrajectory(want, classic, model, sweep(12, 0), config);
    EXPECT_TRUE(fg == fw);
    expectSame(got, want, false);
    for (const auto& m : got)
      for (const auto& kv : m)
        EXPECT_TRUE(kv.first.first != "post" && kv.first.second != "post");
  }
  {
    // a set of trajectories in one batch == one call per trajectory
    CollisionCheckConfig config(0.02, ContactRequest(ContactTestType::CLOSEST), lvs, 0.05);
    std::vector<tesseract_common::TrajArray> set{ sweep(8, 0), sweep(8, 1), sweep(8, 2), sweep(8, 3) };
    // ... one of them collision free
    for (int r = 0; r < 8; ++r)
      set[2](r, 0) = 2.2 + 0.01 * r, set[2](r, 1) = -1.2, set[2](r, 2) = 0.0, set[2](r, 3) = 0.0;
    std::vector<std::vector<ContactResultMap>> per;
    const std::vector<uint8_t> found = checkTrajectories(per, batched, solver, names, set, config);
    EXPECT_TRUE(found.size() == 4 && per.size() == 4);
    for (size_t t = 0; t < set.size(); ++t)
    {
      std::vector<ContactResultMap> want;
      const bool fw = referenceCheckTrajectory(want, classic, model, set[t], config);
      EXPECT_TRUE((found[t] != 0) == fw);
      expectSame(per[t], want, false);
    }
    EXPECT_TRUE(found[2] == 0 && found[0] == 1);
  }
  // argument errors (core/utils.h:179-186, 316-319)
  {
    CollisionCheckConfig wrong(0.02, ContactRequest(ContactTestType::FIRST),
                               cast ? CollisionEvaluatorType::LVS_DISCRETE : CollisionEvaluatorType::LVS_CONTINUOUS, 0.05);
    std::vector<ContactResultMap> got;
    EXPECT_THROW(checkTrajectory(got, batched, solver, names, sweep(4, 0), wrong));
    CollisionCheckConfig ok(0.02, ContactRequest(ContactTestType::FIRST), lvs, 0.05);
    EXPECT_THROW(checkTrajectory(got, batched, solver, { "j1", "j2", "j3" }, sweep(4, 0), ok));
    EXPECT_THROW(checkTrajectory(got, batched, solver, { "j1", "j2", "j3", "nope" }, sweep(4, 0), ok));
    if (cast)
      EXPECT_THROW(checkTrajectory(got, batched, solver, names, sweep(1, 0), ok));
  }
}

static void discreteOnly()
{
  const std::vector<std::string> names{ "j1", "j2", "j3", "j4" };
  B200StateSolver solver(armJoints(), "world");
  B200DiscreteManager batched, classic;
  addArm(batched);
  addArm(classic);
  const t4c_fk_model* model = solver.model(classic.core(), names);
  {
    // one-row trajectory = checkTrajectoryState of that state (core/utils.h:321-325)
    CollisionCheckConfig config(0.02, ContactRequest(ContactTestType::ALL), CollisionEvaluatorType::DISCRETE, 0.05);
    tesseract_common::TrajArray one(1, 4);
    one(0, 0) = 0.58, one(0, 1) = -0.35, one(0, 2) = 0.1, one(0, 3) = 0.2;
    std::vector<ContactResultMap> got, want;
    const bool fg = checkTrajectory(got, batched, solver, names, one, config);
    const bool fw = referenceCheckTrajectory(want, classic, model, one, config);
    EXPECT_TRUE(fg == fw);
    expectSame(got, want, false);
  }
  {
    // OMPL-style validity batch == "FIRST contactTest is empty" per state
    const tesseract_common::TrajArray states = sweep(64, 1);
    std::vector<uint8_t> valid;
    isValid(valid, batched, solver, names, states);
    EXPECT_TRUE(valid.size() == 64);
    int n_invalid = 0;
    for (int r = 0; r < 64; ++r)
    {
      std::vector<double> q{ states(r, 0), states(r, 1), states(r, 2), states(r, 3) };
      setState(classic, model, q, nullptr);
      ContactResultMap m;
      classic.contactTest(m, ContactRequest(ContactTestType::FIRST));
      EXPECT_TRUE((valid[(size_t)r] != 0) == m.empty());
      n_invalid += m.empty() ? 0 : 1;
    }
    EXPECT_TRUE(n_invalid > 3 && n_invalid < 64);
    // the same batch over every GPU of the box (contiguous shards, one NCCL gather to the first device): identical
    // flags, and the gathered records carry global state indices.  Needs two devices: skipped elsewhere.
    if (t4c_device_count() >= 2)
    {
      std::vector<int> devs;
      for (int d = 0; d < std::min(t4c_device_count(), 4); ++d)
        devs.push_back(d);
      B200MultiDeviceChecker multi(batched, devs);
      const tesseract_common::TrajArray many = sweep(1001, 1);
      std::vector<uint8_t> v1, vm;
      isValid(v1, batched, solver, names, many);
      multi.isValid(vm, solver, names, many);
      EXPECT_TRUE(v1 == vm);
      std::vector<uint8_t> hit;
      std::vector<t4c_contact> recs;
      multi.checkStates(solver, names, many, ContactRequest(ContactTestType::FIRST), hit, recs);
      size_t colliding = 0;
      for (size_t r = 0; r < hit.size(); ++r)
      {
        EXPECT_TRUE((hit[r] != 0) == (v1[r] == 0));
        colliding += hit[r] ? 1 : 0;
      }
      EXPECT_TRUE(recs.size() == colliding);
      for (const t4c_contact& c : recs)
        EXPECT_TRUE(c.state >= 0 && c.state < 1001 && hit[(size_t)c.state]);
      EXPECT_TRUE(multi.lastStats().states == 1001);
      std::printf("multi-device validator: %d devices, %zu of 1001 states colliding\n", multi.devices(), colliding);
    }
  }
  {
    // a commanded subset: the other joints stay at zero (StateSolver::getState(joint_names, values))
    CollisionCheckConfig config(0.02, ContactRequest(ContactTestType::CLOSEST), CollisionEvaluatorType::LVS_DISCRETE, 0.05);
    tesseract_common::TrajArray sub(5, 2), full(5, 4);
    for (int r = 0; r < 5; ++r)
    {
      sub(r, 0) = full(r, 0) = -0.2 + 0.3 * r;
      sub(r, 1) = full(r, 2) = 0.02 * r;
    }
    std::vector<ContactResultMap> a, b;
    const bool fa = checkTrajectory(a, batched, solver, { "j1", "j3" }, sub, config);
    const bool fb = checkTrajectory(b, batched, solver, names, full, config);
    EXPECT_TRUE(fa == fb);
    expectSame(a, b, false);
  }
  // kinematic tree errors
  {
    std::vector<JointSpec> J = armJoints();
    J.push_back(J[0]);
    J.back().name = "dup";
    EXPECT_THROW(B200StateSolver(J, "world"));
    std::vector<JointSpec> K = armJoints();
    K[0].parent_link = "nowhere";
    EXPECT_THROW(B200StateSolver(K, "world"));
  }
}

int main()
{
  try
  {
    runCases<B200DiscreteManager>(false);
    runCases<B200CastManager>(true);
    discreteOnly();
  }
  catch (const std::exception& e)
  {
    std::printf("EXCEPTION: %s\n", e.what());
    return 2;
  }
  std::printf("%d checks, %d failures\n", g_checks, g_fail);
  return g_fail ? 1 : 0;
}
