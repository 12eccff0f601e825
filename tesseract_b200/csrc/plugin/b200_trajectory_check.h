/*
 * b200_trajectory_check.h -- planner-facing validators on top of the B200 contact managers.
 *
 * Twins of tesseract_environment::checkTrajectory (tesseract_environment/include/tesseract_environment/core/utils.h:
 * 172-297 continuous, :309-428 discrete): same arguments, same exceptions, same contents of `contacts` and the same
 * return value -- but the LVS subdivision, the forward kinematics and every contact test of the trajectory run as ONE
 * device batch instead of one getState() + contactTest() per sub-state.  checkTrajectories() takes many trajectories at
 * once (a sampling planner's motion validator, a TrajOpt seed set) and isValid() is the OMPL-style boolean state
 * validity batch (SURVEY 8(d) mode A: per-state flag only).
 *
 * B200StateSolver stands where the reference takes a StateSolver (tesseract_environment/core/state_solver.h): it holds
 * the joint list of the scene graph (what KDLStateSolver builds its tree from, kdl_state_solver.cpp:144-161) and
 * compiles it into a device FK model for a given manager and joint-name order.
 */
#ifndef B200_TRAJECTORY_CHECK_H_
#define B200_TRAJECTORY_CHECK_H_

#include "b200_contact_managers.h"

namespace tesseract_b200
{
using tesseract_collision::CollisionCheckConfig;
using tesseract_collision::CollisionEvaluatorType;
using tesseract_common::TrajArray;

// tesseract_scene_graph::Joint, reduced to what forward kinematics reads (joint.h: type, parent / child link names,
// parent_to_joint_origin_transform, axis)
struct JointSpec
{
  enum Type
  {
    FIXED = 0,
    REVOLUTE = 1,  // also CONTINUOUS
    PRISMATIC = 2
  };
  std::string name;
  Type type = FIXED;
  std::string parent_link, child_link;
  Eigen::Isometry3d origin = Eigen::Isometry3d::Identity();
  Eigen::Vector3d axis = Eigen::Vector3d(0, 0, 1);
};

class B200StateSolver
{
public:
  // joints in any order; root_link is the link no joint has as its child (SceneGraph::getRoot())
  B200StateSolver(std::vector<JointSpec> joints, std::string root_link);
  ~B200StateSolver();
  B200StateSolver(const B200StateSolver&) = delete;
  B200StateSolver& operator=(const B200StateSolver&) = delete;

  const std::vector<JointSpec>& getJoints() const { return joints_; }
  const std::string& getRootLinkName() const { return root_; }
  // joint names with a degree of freedom, tree order
  std::vector<std::string> getActiveJointNames() const;

  // device FK model for `core`'s collision objects with the columns of a trajectory in joint_names order; joints not
  // named stay at zero (StateSolver::getState(joint_names, values) semantics).  Cached; owned by the solver.
  // Throws std::runtime_error for unknown joint names (kdl_state_solver.cpp:281-290 logs an error and skips; here
  // a misnamed column would silently check the wrong motion).
  const t4c_fk_model* model(ManagerCore& core, const std::vector<std::string>& joint_names) const;

private:
  std::vector<JointSpec> joints_;  // parent-before-child
  std::string root_;
  struct Cached;
  mutable std::vector<std::unique_ptr<Cached>> cache_;
};

// checkTrajectory (core/utils.h:309-428): config.type must be DISCRETE or LVS_DISCRETE
bool checkTrajectory(std::vector<tesseract_collision::ContactResultMap>& contacts, B200DiscreteManager& manager,
                     const B200StateSolver& state_solver, const std::vector<std::string>& joint_names,
                     const TrajArray& traj, const CollisionCheckConfig& config);
// checkTrajectory (core/utils.h:172-297): config.type must be CONTINUOUS or LVS_CONTINUOUS; a one-row trajectory throws
bool checkTrajectory(std::vector<tesseract_collision::ContactResultMap>& contacts, B200CastManager& manager,
                     const B200StateSolver& state_solver, const std::vector<std::string>& joint_names,
                     const TrajArray& traj, const CollisionCheckConfig& config);

// the same for a set of trajectories with equal shape (rows x cols), one device batch for all of them;
// contacts[t] / the returned flags[t] are what checkTrajectory gives for trajs[t]
std::vector<uint8_t> checkTrajectories(std::vector<std::vector<tesseract_collision::ContactResultMap>>& contacts,
                                       B200DiscreteManager& manager, const B200StateSolver& state_solver,
                                       const std::vector<std::string>& joint_names, const std::vector<TrajArray>& trajs,
                                       const CollisionCheckConfig& config);
std::vector<uint8_t> checkTrajectories(std::vector<std::vector<tesseract_collision::ContactResultMap>>& contacts,
                                       B200CastManager& manager, const B200StateSolver& state_solver,
                                       const std::vector<std::string>& joint_names, const std::vector<TrajArray>& trajs,
                                       const CollisionCheckConfig& config);

// OMPL-style state validity: valid[i] = 1 when row i of `states` is collision free (no contact within the margins of
// the manager); flags only, no contact records leave the device
void isValid(std::vector<uint8_t>& valid, B200DiscreteManager& manager, const B200StateSolver& state_solver,
             const std::vector<std::string>& joint_names, const TrajArray& states);

// The same validators over SEVERAL GPUs of one box: the manager is replicated on every device (the
// reference's "one clone per thread", discrete_contact_manager.h:59-65, with a device per clone), a batch is cut into
// contiguous shards, and flags + contact records + counters come back to the first device in one grouped NCCL exchange
// (t4c_check_discrete_multi, include/t4c.h).
class B200MultiDeviceChecker
{
public:
  // devices: CUDA device indices, devices[0] receives the gathered results.  The manager's current objects, transforms
  // of static links, active set, margins and ACM function are copied; later changes of `manager` are not followed.
  B200MultiDeviceChecker(B200DiscreteManager& manager, const std::vector<int>& devices);
  ~B200MultiDeviceChecker();
  B200MultiDeviceChecker(const B200MultiDeviceChecker&) = delete;
  B200MultiDeviceChecker& operator=(const B200MultiDeviceChecker&) = delete;

  // isValid() over all devices
  void isValid(std::vector<uint8_t>& valid, const B200StateSolver& state_solver,
               const std::vector<std::string>& joint_names, const TrajArray& states);
  // per-state flags and raw records (t4c_contact::state = row of `states`) of a discrete check over all devices
  void checkStates(const B200StateSolver& state_solver, const std::vector<std::string>& joint_names,
                   const TrajArray& states, const tesseract_collision::ContactRequest& request,
                   std::vector<uint8_t>& hit, std::vector<t4c_contact>& contacts);
  const t4c_stats& lastStats() const { return stats_; }
  int devices() const { return (int)cores_.size(); }

private:
  std::vector<std::unique_ptr<ManagerCore>> cores_;
  std::vector<t4c_comm*> comms_;
  t4c_stats stats_{};
};
}  // namespace tesseract_b200
#endif
