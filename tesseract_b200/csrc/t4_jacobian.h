/*
 * t4_jacobian.h -- geometric Jacobians and contact-distance gradients on the device.
 *
 * What the planners do with a contact on the host today (TrajOpt's collision terms): take the contact's link, point
 * and normal, ask the kinematics for the Jacobian of that link at that point
 *     ForwardKinematics::calcJacobian(q, link_name [, link_point])   tesseract_kinematics/src/kdl/kdl_fwd_kin_chain.cpp:127-178
 * (KDL JntToJac: 6 x n_q, rows 0-2 linear / 3-5 angular velocity of the link frame per unit joint velocity, in the base
 * frame; the reference point is moved to the contact with jacobianChangeRefPoint), and form
 *     d(distance)/dq = n . (J_lin,1(p1) - J_lin,0(p0))              n = contact normal, from link 0 towards link 1.
 * Here both live next to the batched check: the contact records never have to leave the device for an optimiser.
 *
 * Column of a joint on the chain from the root to the link, joint axis a (world) through the point o (world):
 *     revolute    linear = a x (p - o), angular = a        prismatic    linear = a, angular = 0
 * and zero for every joint that is not an ancestor of the link.  Joint frames come from the same FK as the link poses
 * (fkOne, t4_kernels.h): the joint frame Tj[j] = T_parent * T_origin * motion(q_j), its axis is constant in it, so
 * a = R_j * axis and o = the frame's origin.
 */
#ifndef T4_JACOBIAN_H_
#define T4_JACOBIAN_H_

#include "t4_kernels.h"

namespace t4
{
struct JacParams
{
  KParams P;             // FK model + q (q0, q_layout, q_stride_n, n_q, joints)
  int32_t link_joint;    // t4c_jacobian: joint that drives the queried link (-1: fixed to the world)
  double point[3];       // ... point in the link frame
  int32_t has_point;
  const double* link_pose_static;  // pose of the link when link_joint < 0 (unused: the Jacobian is zero)
  double* J;             // [n][6][n_q]
  // contact gradients
  const t4c_contact* contacts;
  int64_t n_contacts;
  const int32_t* link_to_joint;  // [n_links] joint driving each link or -1
  double* grad;                  // [n_contacts][n_q]
};

// accumulate the columns of the chain that ends in joint `j` for the world point p; sign * (n . linear) into grad, or
// the full 6-vector into J (row-major 6 x n_q)
T4_KD void chainColumns(const KParams& P, const Xf<double>* Tj, int j, const V3<double>& p, double* J, const V3<double>* n,
                        double sign, double* grad)
{
  while (j >= 0)
  {
    const DJoint& D = P.joints[j];
    if (D.type != T4C_JOINT_FIXED)
    {
      const Xf<double>& X = Tj[j];
      const V3<double> a = X.rot(V3<double>(D.axis[0], D.axis[1], D.axis[2]));
      V3<double> lin, ang(0, 0, 0);
      if (D.type == T4C_JOINT_REVOLUTE)
      {
        lin = a.cross(p - X.p);
        ang = a;
      }
      else
        lin = a;
      const int c = D.q_index;
      if (J)
      {
        J[0 * P.n_q + c] += lin.x, J[1 * P.n_q + c] += lin.y, J[2 * P.n_q + c] += lin.z;
        J[3 * P.n_q + c] += ang.x, J[4 * P.n_q + c] += ang.y, J[5 * P.n_q + c] += ang.z;
      }
      if (grad)
        grad[c] += sign * n->dot(lin);
    }
    j = D.parent_joint;
  }
}

// item = state: 6 x n_q Jacobian of one link (at a point given in the link frame) for every state
template <int MAXJ>
T4_KD void jacobianBody(int64_t s, const JacParams& Q)
{
  const KParams& P = Q.P;
  double* J = Q.J + (size_t)s * 6 * P.n_q;
  for (int k = 0; k < 6 * P.n_q; ++k)
    J[k] = 0.0;
  if (Q.link_joint < 0)
    return;
  Xf<double> Tj[MAXJ];
  fkOne(P, P.q0, s, Tj);
  const Xf<double>& L = Tj[Q.link_joint];
  const V3<double> p = Q.has_point ? L(V3<double>(Q.point[0], Q.point[1], Q.point[2])) : L.p;
  chainColumns(P, Tj, Q.link_joint, p, J, nullptr, 0.0, nullptr);
}

// item = contact record: d(distance)/dq of that contact (world points of the record, normal from link 0 to link 1)
template <int MAXJ>
T4_KD void contactGradBody(int64_t i, const JacParams& Q)
{
  const KParams& P = Q.P;
  const t4c_contact& c = Q.contacts[i];
  double* g = Q.grad + (size_t)i * P.n_q;
  for (int k = 0; k < P.n_q; ++k)
    g[k] = 0.0;
  Xf<double> Tj[MAXJ];
  fkOne(P, P.q0, (int64_t)c.state, Tj);
  const V3<double> n((double)c.normal[0], (double)c.normal[1], (double)c.normal[2]);
  for (int side = 0; side < 2; ++side)
  {
    const int j = Q.link_to_joint[c.link[side]];
    if (j < 0)
      continue;
    const V3<double> p((double)c.point[side][0], (double)c.point[side][1], (double)c.point[side][2]);
    chainColumns(P, Tj, j, p, nullptr, &n, side == 0 ? -1.0 : 1.0, g);
  }
}

}  // namespace t4
#endif
