// gpu_voxels/robot/robot_interface.h -- RobotInterface of the reference (GV/robot/robot_interface.h:30-95): the part of
// it the DH-chain robot implements. URDF robots need KDL/urdfdom and are out of scope.
#ifndef GPU_VOXELS_B200_ROBOT_ROBOT_INTERFACE_H
#define GPU_VOXELS_B200_ROBOT_ROBOT_INTERFACE_H

#include <iostream>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include <gpu_voxels/helpers/MetaPointCloud.h>

namespace gpu_voxels {
namespace robot {

// joint name -> joint value (radians for revolute, metres for prismatic joints)
using JointValueMap = std::map<std::string, float>;
using JointValuePair = std::pair<std::string, float>;
using JointValueMapIterator = JointValueMap::iterator;
using JointValueMapConstIterator = JointValueMap::const_iterator;

// What GpuVoxels needs from a robot: take a configuration, hand out the link clouds transformed to it.
class RobotInterface
{
public:
  virtual ~RobotInterface() {}

  // configuration in, transformed link clouds out
  virtual void setConfiguration(const JointValueMap& jointmap) = 0;
  virtual const MetaPointCloud* getTransformedClouds() = 0;

  // introspection
  virtual void getConfiguration(JointValueMap& jointmap) = 0;
  virtual void getJointNames(std::vector<std::string>& jointnames) = 0;
  // DH chains carry no joint limits in the reference either: KinematicChain logs an error and leaves the map alone
  virtual void getLowerJointLimits(JointValueMap&) { std::cerr << "<Error Robot> getLowerJointLimits not implemented for DH-Robot!" << std::endl; }
  virtual void getUpperJointLimits(JointValueMap&) { std::cerr << "<Error Robot> getUpperJointLimits not implemented for DH-Robot!" << std::endl; }

  // replace the cloud of one link (e.g. a gripper that changed shape)
  virtual void updatePointcloud(const std::string& link_name, const std::vector<Vector3f>& cloud) = 0;
};

}  // namespace robot

typedef std::shared_ptr<robot::RobotInterface> RobotInterfaceSharedPtr;

}  // namespace gpu_voxels
#endif
