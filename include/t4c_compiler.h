/*
 * t4c_compiler.h -- C ABI of the scene compiler (libt4c_compiler.so, host code only): Tesseract's input files -> the
 * PODs of t4c.h.  SURVEY 8(f) rank 2.
 *
 * Reference interfaces this replaces (paths relative to the reference checkout):
 *   t4c_compile_urdf          tesseract_urdf::parseURDFFile + tesseract_scene_graph::SRDFModel::initFile
 *                             (tesseract_urdf/src/urdf_parser.cpp, tesseract_scene_graph/src/srdf_model.cpp), the
 *                             Environment's population of a manager (tesseract_environment/src/core/environment.cpp:613-640)
 *                             and, for meshes_as_convex_hulls, makeConvexMesh (tesseract_collision/core/common.h:181-264,
 *                             tesseract_urdf/src/convex_mesh.cpp:80-100)
 *   t4c_compiled_add_octomap  createShapePrimitive(Octree) (tesseract_collision/src/bullet/bullet_utils.cpp:160-243)
 *   t4c_compiled_add_object   DiscreteContactManager::addCollisionObject (discrete_contact_manager.h:75-79)
 */
#ifndef T4C_COMPILER_H_
#define T4C_COMPILER_H_

#include "t4c.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct t4c_compiled t4c_compiled;

const char* t4c_compiler_last_error(void);
/* package_root: directory that "package://" URIs are resolved against (package://tesseract_support/x -> root/tesseract_support/x).
 * meshes_as_convex_hulls != 0: <mesh> collision geometry is loaded as a ConvexMesh of its hull vertices (BASELINE configs
 * 1, 2, 4), else as one triangle shape per face.  Links below a moving joint are active; link poses are the q = 0 poses. */
t4c_status t4c_compile_urdf(const char* urdf_path, const char* srdf_path, const char* package_root,
                            int32_t meshes_as_convex_hulls, double default_margin, t4c_compiled** out);
/* octomap binary tree file -> one box (sub_type 0), inscribed sphere (1) or circumscribing-in-plane sphere (2) per occupied
 * leaf, as a new geometry of `link_name` (created when absent, static); pose12 = world pose of the link or NULL */
t4c_status t4c_compiled_add_octomap(t4c_compiled* c, const char* link_name, const char* bt_path, int32_t sub_type,
                                    const double* pose12);
/* ready-made shape records (link field ignored, vertex offsets relative to `vertices`) as a new object */
t4c_status t4c_compiled_add_object(t4c_compiled* c, const char* link_name, int32_t n_shapes, const t4c_shape* shapes,
                                   int32_t n_vertices, const double* vertices, const double* pose12);
/* the descriptions t4c_scene_create / t4c_fk_model_create take; valid until the next t4c_compiled_* mutation */
const t4c_scene_desc* t4c_compiled_scene(const t4c_compiled* c);
const t4c_fk_model_desc* t4c_compiled_fk(const t4c_compiled* c);
int32_t t4c_compiled_link_index(const t4c_compiled* c, const char* name);
const char* t4c_compiled_joint_name(const t4c_compiled* c, int32_t joint); /* in t4c_fk_model_desc::joints order */
t4c_status t4c_compiled_limits(const t4c_compiled* c, double* lower, double* upper); /* n_q entries each */
void t4c_compiled_destroy(t4c_compiled* c);
/* building blocks, exposed for tests */
t4c_status t4c_convex_hull_vertices(const double* xyz, int32_t n, int32_t* indices /* n entries */, int32_t* n_out);
t4c_status t4c_read_octomap_bt(const char* bt_path, int64_t capacity, double* leaves /* x y z edge */, int64_t* n_leaves);

#ifdef __cplusplus
}
#endif
#endif
