// gpu_voxels/robot/dh_robot/KinematicChain.h -- the DH-chain robot of the reference (GV/robot/dh_robot/KinematicChain.h:
// 45-176, KinematicChain.cu:31-215) over the C ABI: setConfiguration evaluates the whole chain in ONE kernel
// (gvx_dh_chain_transforms, bit-identical to the reference's host loop) and transforms each link's sub-cloud on the
// device. For planners that test many configurations, GpuVoxels::collideRobotConfigurations / gvx_batch_robot_collide_dh
// skip the transformed cloud altogether.
#ifndef GPU_VOXELS_B200_ROBOT_DH_ROBOT_KINEMATICCHAIN_H
#define GPU_VOXELS_B200_ROBOT_DH_ROBOT_KINEMATICCHAIN_H

#include <map>
#include <string>
#include <vector>

#include <gpu_voxels/logging/logging_gpu_voxels.h>
#include <gpu_voxels/robot/dh_robot/KinematicLink.h>
#include <gpu_voxels/robot/robot_interface.h>

namespace gpu_voxels {
namespace robot {

class KinematicChain : public RobotInterface
{
public:
  // GV/robot/dh_robot/KinematicChain.cu:46-83. (The constructor that loads point-cloud files needs the file loaders,
  // which are out of scope: build the MetaPointCloud and use this one.)
  KinematicChain(const std::vector<std::string>& linknames, const std::vector<robot::DHParameters>& dh_params,
                 const MetaPointCloud& pointclouds)
    : m_linknames(linknames), m_links_meta_cloud(new MetaPointCloud(pointclouds)),
      m_transformed_links_meta_cloud(new MetaPointCloud(pointclouds))
  {
    if (linknames.size() != dh_params.size())
      LOGGING_ERROR_C(RobotLog, KinematicChain, "Number of linknames does not fit number of DH parameters. EXITING!" << endl);
    for (size_t i = 0; i < dh_params.size() && i < linknames.size(); i++)
      m_links[linknames[i]] = KinematicLinkSharedPtr(new KinematicLink(dh_params[i]));
    LOGGING_INFO_C(RobotLog, KinematicChain, "now handling " << m_links.size() << " links." << endl);
  }
  // link clouds from .xyz / .pcd / .binvox files, each cloud NAMED BY ITS PATH: a link moves the cloud whose path equals
  // the link's name (GV/robot/dh_robot/KinematicChain.cu:35-43)
  KinematicChain(const std::vector<std::string>& linknames, const std::vector<robot::DHParameters>& dh_params,
                 const std::vector<std::string>& paths_to_pointclouds, const bool use_model_path)
    : m_linknames(linknames),
      m_links_meta_cloud(new MetaPointCloud(paths_to_pointclouds, paths_to_pointclouds, use_model_path)),
      m_transformed_links_meta_cloud(new MetaPointCloud(*m_links_meta_cloud))
  {
    if (linknames.size() != dh_params.size())
      LOGGING_ERROR_C(RobotLog, KinematicChain, "Number of linknames does not fit number of DH parameters. EXITING!" << endl);
    for (size_t i = 0; i < dh_params.size() && i < linknames.size(); i++)
      m_links[linknames[i]] = KinematicLinkSharedPtr(new KinematicLink(dh_params[i]));
    LOGGING_INFO_C(RobotLog, KinematicChain, "now handling " << m_links.size() << " links." << endl);
  }
  virtual ~KinematicChain()
  {
    delete m_links_meta_cloud;
    delete m_transformed_links_meta_cloud;
  }

  // GV/robot/dh_robot/KinematicChain.cu:93-126
  virtual void setConfiguration(const JointValueMap& jointmap)
  {
    for (std::map<std::string, KinematicLinkSharedPtr>::const_iterator it = m_links.begin(); it != m_links.end(); ++it)
      if (jointmap.count(it->first) != 0) it->second->setJointValue(jointmap.at(it->first));

    const uint32_t n = uint32_t(m_linknames.size());
    if (!n) return;
    std::vector<float> dh4(4 * n), values(n), out(16 * size_t(n));
    std::vector<int32_t> types(n);
    for (uint32_t i = 0; i < n; i++)
    {
      const DHParameters p = m_links[m_linknames[i]]->getDHParam();
      dh4[4 * i] = p.d; dh4[4 * i + 1] = p.theta; dh4[4 * i + 2] = p.a; dh4[4 * i + 3] = p.alpha;
      values[i] = p.value;
      types[i] = int32_t(p.joint_type);
    }
    if (!gvxCheck(gvx_dh_chain_transforms(&dh4[0], &types[0], n, &values[0], 1, &out[0], NULL), "setConfiguration")) return;
    // link i's cloud moves with the product of the DH matrices BEFORE it; links without a cloud only pass it on
    for (uint32_t i = 0; i < n; i++)
    {
      const int16_t pc_num = m_links_meta_cloud->getCloudNumber(m_linknames[i]);
      if (pc_num == -1) continue;
      Matrix4f transformation;
      std::memcpy(transformation.data(), &out[16 * size_t(i)], 64);
      m_links_meta_cloud->transformSubCloud(uint8_t(pc_num), &transformation, m_transformed_links_meta_cloud);
    }
    gvx_sync();
  }

  virtual void getJointNames(std::vector<std::string>& jointnames)
  {
    jointnames.clear();
    for (std::map<std::string, KinematicLinkSharedPtr>::const_iterator it = m_links.begin(); it != m_links.end(); ++it)
      jointnames.push_back(it->first);
  }
  virtual void getConfiguration(JointValueMap& jointmap)
  {
    for (std::map<std::string, KinematicLinkSharedPtr>::const_iterator it = m_links.begin(); it != m_links.end(); ++it)
      jointmap[it->first] = it->second->getJointValue();
  }
  // the DH chain carries no joint limits (GV/robot/dh_robot/KinematicChain.cu:150-163)
  virtual void getLowerJointLimits(JointValueMap&) { LOGGING_ERROR_C(RobotLog, KinematicChain, "getLowerJointLimits not implemented for DH-Robot!" << endl); }
  virtual void getUpperJointLimits(JointValueMap&) { LOGGING_ERROR_C(RobotLog, KinematicChain, "getUpperJointLimits not implemented for DH-Robot!" << endl); }

  virtual const MetaPointCloud* getTransformedClouds() { return m_transformed_links_meta_cloud; }
  const MetaPointCloud* getLinkClouds() const { return m_links_meta_cloud; }
  const std::vector<std::string>& getLinkNames() const { return m_linknames; }
  KinematicLinkSharedPtr getLink(const std::string& name) { return m_links[name]; }

  // GV/robot/dh_robot/KinematicChain.cu:165-176: the geometry of one link changes (tool change, grasped object)
  virtual void updatePointcloud(const std::string& link_name, const std::vector<Vector3f>& cloud)
  {
    m_links_meta_cloud->updatePointCloud(link_name, cloud, true);
    delete m_transformed_links_meta_cloud;
    m_transformed_links_meta_cloud = new MetaPointCloud(*m_links_meta_cloud);
  }

private:
  std::vector<std::string> m_linknames;
  std::map<std::string, KinematicLinkSharedPtr> m_links;
  MetaPointCloud* m_links_meta_cloud;
  MetaPointCloud* m_transformed_links_meta_cloud;
};

}  // namespace robot
}  // namespace gpu_voxels
#endif
