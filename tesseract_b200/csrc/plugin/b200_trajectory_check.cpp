/*
 * b200_trajectory_check.cpp -- see b200_trajectory_check.h.  Host code over the C ABI (include/t4c.h); the contact
 * work itself is ManagerCore::checkTrajectories -> t4c_check_trajectories_* (CUDA, no CPU path).
 */
#include "b200_trajectory_check.h"

#include <algorithm>
#include <cstring>
#include <limits>
#include <stdexcept>

namespace tesseract_b200
{
struct B200StateSolver::Cached
{
  const ManagerCore* core = nullptr;
  std::vector<std::string> objects, joint_names;
  t4c_fk_model* model = nullptr;
  ~Cached()
  {
    if (model)
      t4c_fk_model_destroy(model);
  }
};

B200StateSolver::B200StateSolver(std::vector<JointSpec> joints, std::string root_link) : root_(std::move(root_link))
{
  // parent-before-child order (t4c_fk_model_desc), starting at the root; also rejects loops and orphans
  std::vector<char> used(joints.size(), 0);
  std::vector<std::string> frontier{ root_ };
  while (!frontier.empty())
  {
    const std::string link = frontier.back();
    frontier.pop_back();
    for (size_t j = 0; j < joints.size(); ++j)
      if (!used[j] && joints[j].parent_link == link)
      {
        for (const JointSpec& k : joints_)
          if (k.child_link == joints[j].child_link)
            throw std::runtime_error("B200StateSolver: link '" + k.child_link + "' has two parent joints");
        used[j] = 1;
        joints_.push_back(joints[j]);
        frontier.push_back(joints[j].child_link);
      }
  }
  for (size_t j = 0; j < joints.size(); ++j)
    if (!used[j])
      throw std::runtime_error("B200StateSolver: joint '" + joints[j].name + "' is not connected to root link '" +
                               root_ + "'");
}

B200StateSolver::~B200StateSolver() = default;

std::vector<std::string> B200StateSolver::getActiveJointNames() const
{
  std::vector<std::string> names;
  for (const JointSpec& j : joints_)
    if (j.type != JointSpec::FIXED)
      names.push_back(j.name);
  return names;
}

const t4c_fk_model* B200StateSolver::model(ManagerCore& core, const std::vector<std::string>& joint_names) const
{
  const std::vector<std::string>& objects = core.getCollisionObjects();
  for (const auto& c : cache_)
    if (c->core == &core && c->objects == objects && c->joint_names == joint_names)
      return c->model;
  for (const std::string& n : joint_names)
  {
    const auto it = std::find_if(joints_.begin(), joints_.end(), [&](const JointSpec& j) { return j.name == n; });
    if (it == joints_.end() || it->type == JointSpec::FIXED)
      throw std::runtime_error("B200StateSolver: '" + n + "' is not a movable joint of the kinematic tree");
  }
  // link indices: collision objects keep the manager's index; links without collision geometry follow them
  std::unordered_map<std::string, int> index;
  for (size_t i = 0; i < objects.size(); ++i)
    index[objects[i]] = (int)i;
  int n_links = (int)objects.size();
  auto linkIndex = [&](const std::string& name) {
    auto it = index.find(name);
    if (it != index.end())
      return it->second;
    index[name] = n_links;
    return n_links++;
  };
  std::vector<t4c_joint> dj(joints_.size());
  const int root = linkIndex(root_);
  for (size_t j = 0; j < joints_.size(); ++j)
  {
    const JointSpec& J = joints_[j];
    t4c_joint& D = dj[j];
    std::memset(&D, 0, sizeof(D));
    D.type = J.type == JointSpec::FIXED ? T4C_JOINT_FIXED :
                                          (J.type == JointSpec::PRISMATIC ? T4C_JOINT_PRISMATIC : T4C_JOINT_REVOLUTE);
    D.parent_link = linkIndex(J.parent_link);
    D.child_link = linkIndex(J.child_link);
    D.q_index = -1;
    const auto it = std::find(joint_names.begin(), joint_names.end(), J.name);
    if (it != joint_names.end())
      D.q_index = (int32_t)(it - joint_names.begin());
    else if (J.type != JointSpec::FIXED)
      D.type = T4C_JOINT_FIXED;  // not commanded: stays at zero (getState(joint_names, values))
    toPose12(J.origin, D.origin);
    D.axis[0] = J.axis.x(), D.axis[1] = J.axis.y(), D.axis[2] = J.axis.z();
  }
  t4c_fk_model_desc desc;
  std::memset(&desc, 0, sizeof(desc));
  desc.n_links = n_links;
  desc.n_joints = (int32_t)dj.size();
  desc.joints = dj.data();
  desc.n_q = (int32_t)joint_names.size();
  desc.root_link = root;
  std::unique_ptr<Cached> c(new Cached());
  if (t4c_fk_model_create(&desc, core.scene(), &c->model) != T4C_OK)
    throw std::runtime_error(std::string("B200StateSolver: t4c_fk_model_create: ") + t4c_last_error());
  c->core = &core;
  c->objects = objects;
  c->joint_names = joint_names;
  cache_.push_back(std::move(c));  // never freed before the solver: the scene caches per-model data by address
  return cache_.back()->model;
}

namespace
{
double lvsOf(const CollisionCheckConfig& config, bool cast)
{
  const bool plain = config.type == (cast ? CollisionEvaluatorType::CONTINUOUS : CollisionEvaluatorType::DISCRETE);
  const bool lvs = config.type == (cast ? CollisionEvaluatorType::LVS_CONTINUOUS : CollisionEvaluatorType::LVS_DISCRETE);
  if (!plain && !lvs)
    throw std::runtime_error("checkTrajectory was given an CollisionEvaluatorType that is inconsistent with the "
                             "ContactManager type");  // core/utils.h:179-182, 316-319
  if (plain)
    return std::numeric_limits<double>::infinity();
  if (!(config.longest_valid_segment_length > 0))
    throw std::runtime_error("checkTrajectory: longest_valid_segment_length must be positive");
  return config.longest_valid_segment_length;
}

std::vector<uint8_t> runSet(std::vector<std::vector<tesseract_collision::ContactResultMap>>& contacts,
                            ManagerCore& core, const B200StateSolver& solver,
                            const std::vector<std::string>& joint_names, const std::vector<TrajArray>& trajs,
                            const CollisionCheckConfig& config)
{
  const double lvs = lvsOf(config, core.isCast());
  std::vector<uint8_t> found;
  contacts.clear();
  if (trajs.empty())
    return found;
  const long rows = trajs[0].rows(), cols = trajs[0].cols();
  if (cols != (long)joint_names.size())
    throw std::runtime_error("checkTrajectory: trajectory has " + std::to_string(cols) + " columns for " +
                             std::to_string(joint_names.size()) + " joint names");
  std::vector<double> flat;
  const double* data = trajs[0].data();
  if (trajs.size() > 1)
  {
    flat.reserve(trajs.size() * (size_t)(rows * cols));
    for (const TrajArray& t : trajs)
    {
      if (t.rows() != rows || t.cols() != cols)
        throw std::runtime_error("checkTrajectories: all trajectories of a set must have the same shape");
      flat.insert(flat.end(), t.data(), t.data() + rows * cols);
    }
    data = flat.data();
  }
  core.checkTrajectories(solver.model(core, joint_names), data, (int64_t)trajs.size(), (int32_t)rows, (int32_t)cols,
                         lvs, config.contact_request, contacts, found);
  return found;
}

bool runOne(std::vector<tesseract_collision::ContactResultMap>& contacts, ManagerCore& core,
            const B200StateSolver& solver, const std::vector<std::string>& joint_names, const TrajArray& traj,
            const CollisionCheckConfig& config)
{
  lvsOf(config, core.isCast());  // argument errors first, like the reference
  if (core.isCast() && traj.rows() == 1)
    throw std::runtime_error("checkTrajectory was given continuous contact manager with a trajectory that only has "
                             "one state.");
  std::vector<std::vector<tesseract_collision::ContactResultMap>> per;
  const std::vector<uint8_t> found = runSet(per, core, solver, joint_names, std::vector<TrajArray>{ traj }, config);
  // the reference appends to the caller's vector (contacts.push_back, core/utils.h:114,155)
  for (auto& m : per[0])
    contacts.push_back(std::move(m));
  return found[0] != 0;
}
}  // namespace

bool checkTrajectory(std::vector<tesseract_collision::ContactResultMap>& contacts, B200DiscreteManager& manager,
                     const B200StateSolver& state_solver, const std::vector<std::string>& joint_names,
                     const TrajArray& traj, const CollisionCheckConfig& config)
{
  return runOne(contacts, manager.core(), state_solver, joint_names, traj, config);
}

bool checkTrajectory(std::vector<tesseract_collision::ContactResultMap>& contacts, B200CastManager& manager,
                     const B200StateSolver& state_solver, const std::vector<std::string>& joint_names,
                     const TrajArray& traj, const CollisionCheckConfig& config)
{
  return runOne(contacts, manager.core(), state_solver, joint_names, traj, config);
}

std::vector<uint8_t> checkTrajectories(std::vector<std::vector<tesseract_collision::ContactResultMap>>& contacts,
                                       B200DiscreteManager& manager, const B200StateSolver& state_solver,
                                       const std::vector<std::string>& joint_names, const std::vector<TrajArray>& trajs,
                                       const CollisionCheckConfig& config)
{
  return runSet(contacts, manager.core(), state_solver, joint_names, trajs, config);
}

std::vector<uint8_t> checkTrajectories(std::vector<std::vector<tesseract_collision::ContactResultMap>>& contacts,
                                       B200CastManager& manager, const B200StateSolver& state_solver,
                                       const std::vector<std::string>& joint_names, const std::vector<TrajArray>& trajs,
                                       const CollisionCheckConfig& config)
{
  return runSet(contacts, manager.core(), state_solver, joint_names, trajs, config);
}

void isValid(std::vector<uint8_t>& valid, B200DiscreteManager& manager, const B200StateSolver& state_solver,
             const std::vector<std::string>& joint_names, const TrajArray& states)
{
  ManagerCore& core = manager.core();
  if (states.cols() != (long)joint_names.size())
    throw std::runtime_error("isValid: states have " + std::to_string(states.cols()) + " columns for " +
                             std::to_string(joint_names.size()) + " joint names");
  const int64_t n = states.rows();
  valid.assign((size_t)n, 1);
  if (n == 0)
    return;
  const t4c_fk_model* model = state_solver.model(core, joint_names);
  t4c_request rq;
  rq.type = T4C_TEST_FIRST;
  rq.calculate_penetration = 1;
  rq.calculate_distance = 1;
  std::vector<uint8_t> hit((size_t)n, 0);
  t4c_result r;
  std::memset(&r, 0, sizeof(r));
  r.hit = hit.data();
  if (t4c_check_discrete(core.scene(), model, states.data(), n, &rq, &r) != T4C_OK)
    throw std::runtime_error(std::string("tesseract_b200: t4c_check_discrete: ") + t4c_last_error());
  for (int64_t i = 0; i < n; ++i)
    valid[(size_t)i] = hit[(size_t)i] ? 0 : 1;
}

// ---------------------------------------------------------------------------------------------------------------
// several GPUs
// ---------------------------------------------------------------------------------------------------------------
B200MultiDeviceChecker::B200MultiDeviceChecker(B200DiscreteManager& manager, const std::vector<int>& devices)
{
  if (devices.empty())
    throw std::runtime_error("B200MultiDeviceChecker: empty device list");
  for (int d : devices)
    cores_.push_back(manager.core().clone(d));
  std::vector<int32_t> devs(devices.begin(), devices.end());
  comms_.assign(devices.size(), nullptr);
  if (t4c_comm_create_local(devs.data(), (int32_t)devs.size(), comms_.data()) != T4C_OK)
    throw std::runtime_error(std::string("tesseract_b200: t4c_comm_create_local: ") + t4c_last_error());
}

B200MultiDeviceChecker::~B200MultiDeviceChecker()
{
  for (t4c_comm* c : comms_)
    t4c_comm_destroy(c);
}

void B200MultiDeviceChecker::checkStates(const B200StateSolver& state_solver, const std::vector<std::string>& joint_names,
                                         const TrajArray& states, const tesseract_collision::ContactRequest& request,
                                         std::vector<uint8_t>& hit, std::vector<t4c_contact>& contacts)
{
  if (states.cols() != (long)joint_names.size())
    throw std::runtime_error("checkStates: states have " + std::to_string(states.cols()) + " columns for " +
                             std::to_string(joint_names.size()) + " joint names");
  const int64_t n = states.rows();
  hit.assign((size_t)n, 0);
  contacts.clear();
  if (n == 0)
    return;
  // the joint table does not depend on the device: one FK model, built against the first replica
  const t4c_fk_model* model = state_solver.model(*cores_[0], joint_names);
  std::vector<t4c_scene*> scenes;
  for (auto& c : cores_)
    scenes.push_back(c->scene());
  t4c_request rq;
  rq.type = request.type == tesseract_collision::ContactTestType::FIRST ?
                T4C_TEST_FIRST :
                (request.type == tesseract_collision::ContactTestType::CLOSEST ? T4C_TEST_CLOSEST : T4C_TEST_ALL);
  rq.calculate_penetration = request.calculate_penetration ? 1 : 0;
  rq.calculate_distance = request.calculate_distance ? 1 : 0;
  int64_t cap = std::max<int64_t>(256, 4 * n);
  for (;;)
  {
    contacts.resize((size_t)cap);
    t4c_result r;
    std::memset(&r, 0, sizeof(r));
    r.hit = hit.data();
    r.contacts = contacts.data();
    r.capacity = cap;
    const t4c_status st = t4c_check_discrete_multi(scenes.data(), model, comms_.data(), (int32_t)scenes.size(),
                                                   states.data(), n, &rq, &r);
    stats_ = r.stats;
    if (st == T4C_ERR_CAPACITY)
    {
      cap = r.n_contacts;
      continue;
    }
    if (st != T4C_OK)
      throw std::runtime_error(std::string("tesseract_b200: t4c_check_discrete_multi: ") + t4c_last_error());
    contacts.resize((size_t)r.n_contacts);
    return;
  }
}

void B200MultiDeviceChecker::isValid(std::vector<uint8_t>& valid, const B200StateSolver& state_solver,
                                     const std::vector<std::string>& joint_names, const TrajArray& states)
{
  const int64_t n = states.rows();
  valid.assign((size_t)n, 1);
  if (n == 0)
    return;
  const t4c_fk_model* model = state_solver.model(*cores_[0], joint_names);
  std::vector<t4c_scene*> scenes;
  for (auto& c : cores_)
    scenes.push_back(c->scene());
  t4c_request rq;
  rq.type = T4C_TEST_FIRST;
  rq.calculate_penetration = 1;
  rq.calculate_distance = 1;
  std::vector<uint8_t> hit((size_t)n, 0);
  t4c_result r;
  std::memset(&r, 0, sizeof(r));
  r.hit = hit.data();
  if (t4c_check_discrete_multi(scenes.data(), model, comms_.data(), (int32_t)scenes.size(), states.data(), n, &rq, &r) !=
      T4C_OK)
    throw std::runtime_error(std::string("tesseract_b200: t4c_check_discrete_multi: ") + t4c_last_error());
  stats_ = r.stats;
  for (int64_t i = 0; i < n; ++i)
    valid[(size_t)i] = hit[(size_t)i] ? 0 : 1;
}
}  // namespace tesseract_b200
