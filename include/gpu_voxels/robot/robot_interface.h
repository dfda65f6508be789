// gpu_voxels/robot/robot_interface.h -- RobotInterface of the reference (GV/robot/robot_interface.h:30-95): the part of
// it the DH-chain robot implements. URDF robots need KDL/urdfdom and are out of scope.
#ifndef GPU_VOXELS_B200_ROBOT_ROBOT_INTERFACE_H
#define GPU_VOXELS_B200_ROBOT_ROBOT_INTERFACE_H

#include <map>
#include <memory>
#include <string>
#include <vector>

#include <gpu_voxels/helpers/MetaPointCloud.h>

namespace gpu_voxels {
namespace robot {

typedef std::pair<std::string, float> JointValuePair;
typedef std::map<std::string, float> JointValueMap;
typedef JointValueMap::iterator JointValueMapIterator;
typedef JointValueMap::const_iterator JointValueMapConstIterator;

class RobotInterface
{
public:
  virtual ~RobotInterface() {}
  virtual void getJointNames(std::vector<std::string>& jointnames) = 0;
  virtual void setConfiguration(const JointValueMap& jointmap) = 0;
  virtual void getConfiguration(JointValueMap& jointmap) = 0;
  virtual const MetaPointCloud* getTransformedClouds() = 0;
  virtual void updatePointcloud(const std::string& link_name, const std::vector<Vector3f>& cloud) = 0;
};

}  // namespace robot

typedef std::shared_ptr<robot::RobotInterface> RobotInterfaceSharedPtr;

}  // namespace gpu_voxels
#endif
