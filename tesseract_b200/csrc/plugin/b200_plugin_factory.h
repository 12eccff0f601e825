/*
 * b200_plugin_factory.h -- adapter for the plugin API of Tesseract releases AFTER the 0.5.0 snapshot this repository
 * tracks.
 *
 * 0.5.0 (the snapshot):  Environment::registerDiscreteContactManager(name, &Class::create)
 *                        (tesseract_environment/src/core/environment.cpp:518-532) and, for shared libraries,
 *                        TESSERACT_ADD_PLUGIN(Class, alias) (tesseract_common/class_loader.h:93-95) -- see the end of
 *                        b200_contact_managers.cpp and INTEGRATION.md.
 * later releases:        contact managers are created by factories that a ContactManagersPluginFactory loads by name
 *                        from the YAML "contact_manager_plugins" section:
 *                            class DiscreteContactManagerFactory   { virtual DiscreteContactManager::UPtr
 *                                                                     create(const std::string& name, const YAML::Node& config) const = 0; };
 *                            class ContinuousContactManagerFactory { ... create(...) const = 0; };
 *                        exported with TESSERACT_ADD_DISCRETE_MANAGER_PLUGIN / TESSERACT_ADD_CONTINUOUS_MANAGER_PLUGIN
 *                        (tesseract_collision/core/contact_managers_plugin_factory.h).  The manager interface itself is
 *                        the same set of virtuals this plugin implements.
 *
 * The later headers are not in this container, so the adapter is written against their PUBLISHED signatures and compiled
 * only when T4_HAVE_CONTACT_MANAGERS_PLUGIN_FACTORY is defined by a build that has them; nothing else in the repository
 * depends on it.  YAML config keys understood:  device: <CUDA device index> (default 0).
 */
#ifndef B200_PLUGIN_FACTORY_H_
#define B200_PLUGIN_FACTORY_H_

#ifdef T4_HAVE_CONTACT_MANAGERS_PLUGIN_FACTORY
#include <tesseract_collision/core/contact_managers_plugin_factory.h>

#include "b200_contact_managers.h"

namespace tesseract_b200
{
inline int deviceFromConfig(const YAML::Node& config)
{
  if (config && config["device"])
    return config["device"].as<int>();
  return 0;
}

class B200DiscreteManagerFactory : public tesseract_collision::DiscreteContactManagerFactory
{
public:
  std::unique_ptr<tesseract_collision::DiscreteContactManager> create(const std::string& /*name*/,
                                                                      const YAML::Node& config) const override
  {
    return std::make_unique<B200DiscreteManager>(deviceFromConfig(config));
  }
};

class B200CastManagerFactory : public tesseract_collision::ContinuousContactManagerFactory
{
public:
  std::unique_ptr<tesseract_collision::ContinuousContactManager> create(const std::string& /*name*/,
                                                                        const YAML::Node& config) const override
  {
    return std::make_unique<B200CastManager>(deviceFromConfig(config));
  }
};
}  // namespace tesseract_b200

// NOLINTNEXTLINE
TESSERACT_ADD_DISCRETE_MANAGER_PLUGIN(tesseract_b200::B200DiscreteManagerFactory, B200DiscreteManagerFactory)
// NOLINTNEXTLINE
TESSERACT_ADD_CONTINUOUS_MANAGER_PLUGIN(tesseract_b200::B200CastManagerFactory, B200CastManagerFactory)
#endif  // T4_HAVE_CONTACT_MANAGERS_PLUGIN_FACTORY

#endif
