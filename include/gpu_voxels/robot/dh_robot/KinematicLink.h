// gpu_voxels/robot/dh_robot/KinematicLink.h -- DH parameters of one link (GV/robot/dh_robot/KinematicLink.h:34-118).
// The DH matrix itself (DHParameters::convertDHtoM, KinematicLink.cu:30-84) is evaluated in libgvx.so for whole batches
// of configurations (gvx_dh_chain_transforms), bit-identical to the reference's host arithmetic.
#ifndef GPU_VOXELS_B200_ROBOT_DH_ROBOT_KINEMATICLINK_H
#define GPU_VOXELS_B200_ROBOT_DH_ROBOT_KINEMATICLINK_H

#include <cstring>
#include <memory>

#include <gpu_voxels/helpers/GvxCall.h>
#include <gpu_voxels/helpers/cuda_datatypes.h>

namespace gpu_voxels {
namespace robot {

typedef enum
{
  REVOLUTE = 0,
  PRISMATIC
} DHJointType;

class DHParameters
{
public:
  float d;
  float theta;  // initial rotation
  float a;      // initial translation
  float alpha;
  float value;  // joint value (rotation for revolute joints, translation for prismatic joints)
  DHJointType joint_type;

  DHParameters() : d(0.0), theta(0.0), a(0.0), alpha(0.0), value(0.0), joint_type(REVOLUTE) {}
  DHParameters(float _d, float _theta, float _a, float _alpha, float _value, DHJointType _joint_type)
    : d(_d), theta(_theta), a(_a), alpha(_alpha), value(_value), joint_type(_joint_type) {}

  // GV/robot/dh_robot/KinematicLink.cu:30-84, through the batched kernel with one configuration and one link
  void convertDHtoM(Matrix4f& m) const
  {
    gvxEnsureInit();
    // a two-link chain: the matrix handed to link 1's cloud is link 0's DH matrix
    const float dh4[8] = {d, theta, a, alpha, 0.f, 0.f, 0.f, 0.f};
    const int32_t jt[2] = {int32_t(joint_type), 0};
    const float jv[2] = {value, 0.f};
    float out[32];
    if (gvxCheck(gvx_dh_chain_transforms(dh4, jt, 2, jv, 1, out, NULL), "convertDHtoM")) std::memcpy(m.data(), out + 16, 64);
  }
};

class KinematicLink
{
public:
  KinematicLink(const DHParameters& dh_parameters) : m_dh_parameters(dh_parameters) {}
  KinematicLink(float d, float theta, float a, float alpha, float joint_value, DHJointType joint_type)
    : m_dh_parameters(d, theta, a, alpha, joint_value, joint_type) {}

  void setJointValue(float value) { m_dh_parameters.value = value; }
  DHJointType getJointType() { return m_dh_parameters.joint_type; }
  // GV/robot/dh_robot/KinematicLink.h:82-93 (PRISMATIC reports a + value there, kept)
  float getJointValue()
  {
    if (m_dh_parameters.joint_type == PRISMATIC) return m_dh_parameters.a + m_dh_parameters.value;
    if (m_dh_parameters.joint_type == REVOLUTE) return m_dh_parameters.theta + m_dh_parameters.value;
    return 0;
  }
  DHParameters getDHParam() { return m_dh_parameters; }
  void setDHParam(float d, float theta, float a, float alpha, float joint_value, DHJointType joint_type)
  {
    m_dh_parameters = DHParameters(d, theta, a, alpha, joint_value, joint_type);
  }
  void setDHParam(DHParameters dh_parameters) { m_dh_parameters = dh_parameters; }
  void getMatrixRepresentation(Matrix4f& m) { m_dh_parameters.convertDHtoM(m); }

private:
  DHParameters m_dh_parameters;
};
typedef std::shared_ptr<KinematicLink> KinematicLinkSharedPtr;

}  // namespace robot
}  // namespace gpu_voxels
#endif
