/*
 * b200_contact_managers.h -- the contact-manager plugin: tesseract_collision::DiscreteContactManager and
 * ContinuousContactManager implemented on the C ABI of include/t4c.h (CUDA, sm_100a; no CPU path).
 *
 * Drop-in twins of BulletDiscreteBVHManager / BulletCastBVHManager
 * (tesseract_collision/include/tesseract_collision/bullet/bullet_discrete_bvh_manager.h:52-138,
 *  bullet_cast_bvh_manager.h:52-160): same virtuals, same return conventions (false on empty / mismatched shapes,
 * silent no-op for transforms of unknown links, re-adding a name replaces the object), same clone() contract, the
 * conventional statics name() / create() used with Environment::registerDiscreteContactManager
 * (tesseract_environment/src/core/environment.cpp:518-532).
 *
 * One manager instance per thread (discrete_contact_manager.h:59-65).  The classic contactTest() is the latency path
 * (batch of one state); planners that check many states use the batched entries (checkStates / checkSegments below,
 * or t4c_check_discrete / t4c_check_cast / t4c_check_trajectories_* directly).
 */
#ifndef B200_CONTACT_MANAGERS_H_
#define B200_CONTACT_MANAGERS_H_

#include "../../../include/t4c.h"
#include "tesseract_shim.h"

namespace tesseract_b200
{
using tesseract_collision::CollisionMarginData;
using tesseract_collision::CollisionMarginOverrideType;
using tesseract_collision::CollisionShapesConst;
using tesseract_collision::ContactRequest;
using tesseract_collision::ContactResultMap;
using tesseract_collision::IsContactAllowedFn;

// state + logic shared by the two managers
class ManagerCore
{
public:
  ManagerCore(bool cast, int device);
  ~ManagerCore();
  ManagerCore(const ManagerCore&) = delete;
  ManagerCore& operator=(const ManagerCore&) = delete;
  // device < 0: same device, sharing the compiled scene image; otherwise a replica of this manager on that device
  std::unique_ptr<ManagerCore> clone(int device = -1) const;
  int device() const { return device_; }

  bool addCollisionObject(const std::string& name, int mask_id, const CollisionShapesConst& shapes,
                          const tesseract_common::VectorIsometry3d& shape_poses, bool enabled);
  const CollisionShapesConst& getGeometries(const std::string& name) const;
  const tesseract_common::VectorIsometry3d& getGeometriesTransforms(const std::string& name) const;
  bool hasCollisionObject(const std::string& name) const;
  bool removeCollisionObject(const std::string& name);
  bool setEnabled(const std::string& name, bool enabled);
  void setTransform(const std::string& name, const Eigen::Isometry3d& pose);
  void setTransform(const std::string& name, const Eigen::Isometry3d& pose1, const Eigen::Isometry3d& pose2);
  const std::vector<std::string>& getCollisionObjects() const { return names_; }
  void setActive(const std::vector<std::string>& names);
  const std::vector<std::string>& getActive() const { return active_; }
  void setMarginData(CollisionMarginData data, CollisionMarginOverrideType type);
  void setDefaultMargin(double m);
  void setPairMargin(const std::string& a, const std::string& b, double m);
  const CollisionMarginData& getMarginData() const { return margins_; }
  void setIsContactAllowedFn(IsContactAllowedFn fn) { fn_ = std::move(fn); }
  IsContactAllowedFn getIsContactAllowedFn() const { return fn_; }
  void contactTest(ContactResultMap& collisions, const ContactRequest& request);

  // batched FK-bypass entry: poses = n x n_active x 12 doubles (active links in getCollisionObjects() order);
  // returns per-state flags and raw records (t4c_contact.state = state index)
  void checkLinkPoses(const double* poses0, const double* poses1, int64_t n, const ContactRequest& request,
                      std::vector<uint8_t>& hit, std::vector<t4c_contact>& contacts);
  // planner-facing batched validator: every LVS work item of n_traj trajectories (each n_wp x n_q row-major) in one
  // device batch; contacts[t] receives what checkTrajectory (tesseract_environment/core/utils.h:172-428) pushes for
  // trajectory t (one ContactResultMap per colliding state / sub-segment, in visiting order; FIRST stops after the
  // first), found[t] its return value.  lvs = +infinity: no subdivision (DISCRETE / CONTINUOUS).
  void checkTrajectories(const t4c_fk_model* model, const double* traj, int64_t n_traj, int32_t n_wp, int32_t n_q,
                         double lvs, const ContactRequest& request,
                         std::vector<std::vector<ContactResultMap>>& contacts, std::vector<uint8_t>& found);
  int objectIndex(const std::string& name) const { return find(name); }
  bool isCast() const { return cast_; }
  const t4c_stats& lastStats() const { return stats_; }
  t4c_scene* scene();  // synchronised device scene (for the joint-space batched entries of t4c.h)

private:
  struct Object
  {
    std::string name;
    int mask_id = 0;
    CollisionShapesConst shapes;
    tesseract_common::VectorIsometry3d shape_poses;
    bool enabled = true;
    Eigen::Isometry3d pose0 = Eigen::Isometry3d::Identity();
    Eigen::Isometry3d pose1 = Eigen::Isometry3d::Identity();
  };
  int find(const std::string& name) const;
  bool isActive(int i) const;
  void sync();  // rebuild / update the device scene to match the host state
  void run(const double* p0, const double* p1, int64_t n, const ContactRequest& request, int device_type,
           std::vector<uint8_t>* hit, std::vector<t4c_contact>& contacts, std::vector<float>& local);
  // device records of ONE state -> ContactResult -> processResult; poses0/1 = that state's link poses (null: the
  // objects' own); returns true when FIRST is satisfied
  bool processRecords(ContactResultMap& collisions, const ContactRequest& request, const t4c_contact* const* recs,
                      const float* const* locals, size_t n, const double* poses0, const double* poses1) const;

  bool cast_;
  int device_;
  std::vector<Object> objects_;
  std::vector<std::string> names_;
  std::unordered_map<std::string, int> index_;
  std::vector<std::string> active_;
  CollisionMarginData margins_;
  IsContactAllowedFn fn_;
  t4c_scene* scene_ = nullptr;
  bool dirty_structure_ = true, dirty_soft_ = true;
  std::vector<int32_t> acm_uploaded_;
  bool acm_valid_ = false;
  t4c_stats stats_{};
};

class B200DiscreteManager : public tesseract_collision::DiscreteContactManager
{
public:
  using Ptr = std::shared_ptr<B200DiscreteManager>;
  explicit B200DiscreteManager(int device = 0);
  static std::string name() { return "B200DiscreteManager"; }
  static tesseract_collision::DiscreteContactManager::Ptr create() { return std::make_shared<B200DiscreteManager>(); }

  std::shared_ptr<tesseract_collision::DiscreteContactManager> clone() const override;
  bool addCollisionObject(const std::string& name, const int& mask_id, const CollisionShapesConst& shapes,
                          const tesseract_common::VectorIsometry3d& shape_poses, bool enabled = true) override;
  const CollisionShapesConst& getCollisionObjectGeometries(const std::string& name) const override;
  const tesseract_common::VectorIsometry3d& getCollisionObjectGeometriesTransforms(const std::string& name) const override;
  bool hasCollisionObject(const std::string& name) const override;
  bool removeCollisionObject(const std::string& name) override;
  bool enableCollisionObject(const std::string& name) override;
  bool disableCollisionObject(const std::string& name) override;
  void setCollisionObjectsTransform(const std::string& name, const Eigen::Isometry3d& pose) override;
  void setCollisionObjectsTransform(const std::vector<std::string>& names,
                                    const tesseract_common::VectorIsometry3d& poses) override;
  void setCollisionObjectsTransform(const tesseract_common::TransformMap& transforms) override;
  const std::vector<std::string>& getCollisionObjects() const override;
  void setActiveCollisionObjects(const std::vector<std::string>& names) override;
  const std::vector<std::string>& getActiveCollisionObjects() const override;
  void setCollisionMarginData(CollisionMarginData collision_margin_data,
                              CollisionMarginOverrideType override_type = CollisionMarginOverrideType::REPLACE) override;
  void setDefaultCollisionMarginData(double default_collision_margin) override;
  void setPairCollisionMarginData(const std::string& name1, const std::string& name2, double collision_margin) override;
  const CollisionMarginData& getCollisionMarginData() const override;
  void setIsContactAllowedFn(IsContactAllowedFn fn) override;
  IsContactAllowedFn getIsContactAllowedFn() const override;
  void contactTest(ContactResultMap& collisions, const ContactRequest& request) override;

  ManagerCore& core() { return *core_; }

private:
  explicit B200DiscreteManager(std::unique_ptr<ManagerCore> core);
  std::unique_ptr<ManagerCore> core_;
};

class B200CastManager : public tesseract_collision::ContinuousContactManager
{
public:
  using Ptr = std::shared_ptr<B200CastManager>;
  explicit B200CastManager(int device = 0);
  static std::string name() { return "B200CastManager"; }
  static tesseract_collision::ContinuousContactManager::Ptr create() { return std::make_shared<B200CastManager>(); }

  std::shared_ptr<tesseract_collision::ContinuousContactManager> clone() const override;
  bool addCollisionObject(const std::string& name, const int& mask_id, const CollisionShapesConst& shapes,
                          const tesseract_common::VectorIsometry3d& shape_poses, bool enabled = true) override;
  const CollisionShapesConst& getCollisionObjectGeometries(const std::string& name) const override;
  const tesseract_common::VectorIsometry3d& getCollisionObjectGeometriesTransforms(const std::string& name) const override;
  bool hasCollisionObject(const std::string& name) const override;
  bool removeCollisionObject(const std::string& name) override;
  bool enableCollisionObject(const std::string& name) override;
  bool disableCollisionObject(const std::string& name) override;
  void setCollisionObjectsTransform(const std::string& name, const Eigen::Isometry3d& pose) override;
  void setCollisionObjectsTransform(const std::vector<std::string>& names,
                                    const tesseract_common::VectorIsometry3d& poses) override;
  void setCollisionObjectsTransform(const tesseract_common::TransformMap& transforms) override;
  void setCollisionObjectsTransform(const std::string& name, const Eigen::Isometry3d& pose1,
                                    const Eigen::Isometry3d& pose2) override;
  void setCollisionObjectsTransform(const std::vector<std::string>& names, const tesseract_common::VectorIsometry3d& pose1,
                                    const tesseract_common::VectorIsometry3d& pose2) override;
  void setCollisionObjectsTransform(const tesseract_common::TransformMap& pose1,
                                    const tesseract_common::TransformMap& pose2) override;
  const std::vector<std::string>& getCollisionObjects() const override;
  void setActiveCollisionObjects(const std::vector<std::string>& names) override;
  const std::vector<std::string>& getActiveCollisionObjects() const override;
  void setCollisionMarginData(CollisionMarginData collision_margin_data,
                              CollisionMarginOverrideType override_type = CollisionMarginOverrideType::REPLACE) override;
  void setDefaultCollisionMarginData(double default_collision_margin) override;
  void setPairCollisionMarginData(const std::string& name1, const std::string& name2, double collision_margin) override;
  const CollisionMarginData& getCollisionMarginData() const override;
  void setIsContactAllowedFn(IsContactAllowedFn fn) override;
  IsContactAllowedFn getIsContactAllowedFn() const override;
  void contactTest(ContactResultMap& collisions, const ContactRequest& request) override;

  ManagerCore& core() { return *core_; }

private:
  explicit B200CastManager(std::unique_ptr<ManagerCore> core);
  std::unique_ptr<ManagerCore> core_;
};
}  // namespace tesseract_b200

// plain-C factory symbols for dlopen()-style loaders that do not use TESSERACT_ADD_PLUGIN (class_loader.h:93-95)
extern "C" {
tesseract_collision::DiscreteContactManager* tesseract_b200_create_discrete_manager(void);
tesseract_collision::ContinuousContactManager* tesseract_b200_create_cast_manager(void);
}
#endif
