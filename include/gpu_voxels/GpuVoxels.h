// gpu_voxels/GpuVoxels.h -- the GpuVoxels singleton of the reference's API for the hot path
// (GV/GpuVoxels.h:91-415, GV/GpuVoxels.cpp:33-78,164-313,443-559,594-609), header-only over the C ABI (include/gvx.h).
// Source-compatible for: getInstance, initialize, addMap, delMap, getMap, clearMap x2, insertPointCloudIntoMap x2,
// insertMetaPointCloudIntoMap x2, insertPointCloudFromFile, insertBoxIntoMap, getDimensions x2, getVoxelSideLength; the returned maps provide
// collideWith / collideWithTypes / collideWithTypeMask / collideWithBitcheck, the swept-volume shift and the file
// formats; DH-chain robots (addRobot with a MetaPointCloud, setRobotConfiguration, insertRobotIntoMap). Visualisation,
// URDF robots, octrees, distance maps and primitive arrays are outside this build
// (SURVEY.md section 8).
#ifndef GPU_VOXELS_B200_GPUVOXELS_H
#define GPU_VOXELS_B200_GPUVOXELS_H

#include <map>
#include <memory>
#include <string>
#include <vector>

#include <gpu_voxels/helpers/MathHelpers.h>
#include <gpu_voxels/helpers/cuda_handling.h>
#include <gpu_voxels/GpuVoxelsMap.h>
#include <gpu_voxels/helpers/GeometryGeneration.h>
#include <gpu_voxels/robot/dh_robot/KinematicChain.h>
#include <gpu_voxels/voxellist/VoxelList.h>
#include <gpu_voxels/voxelmap/VoxelMap.h>

namespace gpu_voxels {

class GpuVoxels;
typedef std::shared_ptr<GpuVoxels> GpuVoxelsSharedPtr;

struct ManagedMap
{
  GpuVoxelsMapSharedPtr map_shared_ptr;
};
typedef std::map<std::string, ManagedMap> ManagedMaps;
typedef ManagedMaps::iterator ManagedMapsIterator;

class GpuVoxels
{
public:
  ~GpuVoxels() { m_managed_robots.clear(); m_managed_maps.clear(); }

  // one shared instance as long as somebody holds it (weak_ptr singleton, GV/GpuVoxels.cpp:67-78)
  static GpuVoxelsSharedPtr getInstance()
  {
    static std::weak_ptr<GpuVoxels> master;
    static std::mutex mu;
    std::lock_guard<std::mutex> lk(mu);
    GpuVoxelsSharedPtr p = master.lock();
    if (!p)
    {
      p.reset(new GpuVoxels());
      master = p;
    }
    return p;
  }

  void initialize(const uint32_t dim_x, const uint32_t dim_y, const uint32_t dim_z, const float voxel_side_length)
  {
    if (m_dim.x == 0 || m_dim.y == 0 || m_dim.z == 0 || m_voxel_side_length == 0)
    {
      m_dim = Vector3ui(dim_x, dim_y, dim_z);
      m_voxel_side_length = voxel_side_length;
    }
    else  // GV/GpuVoxels.cpp:52-65
    {
      LOGGING_WARNING(Gpu_voxels, "Do not try to initialize GpuVoxels multiple times. Parameters remain unchanged." << endl);
    }
  }

  GpuVoxelsMapSharedPtr addMap(const MapType map_type, const std::string& map_name)
  {
    if (m_managed_maps.find(map_name) != m_managed_maps.end())  // GV/GpuVoxels.cpp:169-176
    {
      LOGGING_ERROR_C(Gpu_voxels, GpuVoxels, "Map with name '" << map_name << "' already exists." << endl);
      return GpuVoxelsMapSharedPtr();
    }
    GpuVoxelsMapSharedPtr map_shared_ptr;
    switch (map_type)
    {
      case MT_BITVECTOR_VOXELMAP:
        map_shared_ptr.reset(new voxelmap::BitVectorVoxelMap(m_dim, m_voxel_side_length, MT_BITVECTOR_VOXELMAP));
        break;
      case MT_PROBAB_VOXELMAP:
        map_shared_ptr.reset(new voxelmap::ProbVoxelMap(m_dim, m_voxel_side_length, MT_PROBAB_VOXELMAP));
        break;
      case MT_BITVECTOR_VOXELLIST:
        map_shared_ptr.reset(new voxellist::BitVectorVoxelList(m_dim, m_voxel_side_length, MT_BITVECTOR_VOXELLIST));
        break;
      default:  // the reference throws a std::string for unimplemented types (GV/GpuVoxels.cpp:210-211)
        throw GPU_VOXELS_MAP_TYPE_NOT_IMPLEMENTED;
    }
    ManagedMap m;
    m.map_shared_ptr = map_shared_ptr;
    m_managed_maps.insert(std::make_pair(map_name, m));
    return map_shared_ptr;
  }

  bool delMap(const std::string& map_name)
  {
    ManagedMapsIterator it = findMap(map_name);
    if (it == m_managed_maps.end()) return false;
    m_managed_maps.erase(it);
    return true;
  }

  GpuVoxelsMapSharedPtr getMap(const std::string& map_name)
  {
    ManagedMapsIterator it = findMap(map_name);
    return it == m_managed_maps.end() ? GpuVoxelsMapSharedPtr() : it->second.map_shared_ptr;
  }

  // GV/GpuVoxels.cpp:561-570. The visualizer (a separate OpenGL process fed through shared memory) is out of scope
  // and unavailable on a headless box: the call keeps the reference's contract for the map name (false + error for an
  // unknown map) and is otherwise a successful no-op, so sensor / planner loops written for the reference
  // compile and run unchanged (there, too, the call succeeds whether or not a visualizer process is listening).
  bool visualizeMap(const std::string& map_name, const bool force_repaint = true)
  {
    (void)force_repaint;
    return findMap(map_name) != m_managed_maps.end();
  }

  bool clearMap(const std::string& map_name)
  {
    ManagedMapsIterator it = findMap(map_name);
    if (it == m_managed_maps.end()) return false;
    it->second.map_shared_ptr->clearMap();
    return true;
  }
  bool clearMap(const std::string& map_name, BitVoxelMeaning voxel_meaning)
  {
    ManagedMapsIterator it = findMap(map_name);
    if (it == m_managed_maps.end()) return false;
    it->second.map_shared_ptr->clearBitVoxelMeaning(voxel_meaning);
    return true;
  }

  bool insertPointCloudIntoMap(const PointCloud& cloud, std::string map_name, const BitVoxelMeaning voxel_meaning)
  {
    ManagedMapsIterator it = findMap(map_name);
    if (it == m_managed_maps.end()) return false;
    it->second.map_shared_ptr->insertPointCloud(cloud, voxel_meaning);
    return true;
  }
  bool insertPointCloudIntoMap(const std::vector<Vector3f>& cloud, std::string map_name, const BitVoxelMeaning voxel_meaning)
  {
    ManagedMapsIterator it = findMap(map_name);
    if (it == m_managed_maps.end()) return false;
    it->second.map_shared_ptr->insertPointCloud(cloud, voxel_meaning);
    return true;
  }
  // GV/GpuVoxels.cpp:427-440
  bool insertPointCloudFromFile(const std::string map_name, const std::string path, const bool use_model_path,
                                const BitVoxelMeaning voxel_meaning, const bool shift_to_zero = false,
                                const Vector3f& offset_XYZ = Vector3f(), const float scaling = 1.0)
  {
    ManagedMapsIterator it = findMap(map_name);
    if (it == m_managed_maps.end()) return false;
    return it->second.map_shared_ptr->insertPointCloudFromFile(path, use_model_path, voxel_meaning, shift_to_zero,
                                                               offset_XYZ, scaling);
  }
  bool insertMetaPointCloudIntoMap(const MetaPointCloud& meta_point_cloud, std::string map_name,
                                   const std::vector<BitVoxelMeaning>& voxel_meanings)
  {
    ManagedMapsIterator it = findMap(map_name);
    if (it == m_managed_maps.end()) return false;
    it->second.map_shared_ptr->insertMetaPointCloud(meta_point_cloud, voxel_meanings);
    return true;
  }
  bool insertMetaPointCloudIntoMap(const MetaPointCloud& meta_point_cloud, std::string map_name,
                                   const BitVoxelMeaning voxel_meaning)
  {
    ManagedMapsIterator it = findMap(map_name);
    if (it == m_managed_maps.end()) return false;
    it->second.map_shared_ptr->insertMetaPointCloud(meta_point_cloud, voxel_meaning);
    return true;
  }

  // GV/GpuVoxels.cpp:519-535: a box sampled with delta = side / points_per_voxel
  bool insertBoxIntoMap(const Vector3f& corner_min, const Vector3f& corner_max, std::string map_name,
                        const BitVoxelMeaning voxel_meaning, uint16_t points_per_voxel = 1)
  {
    ManagedMapsIterator it = findMap(map_name);
    if (it == m_managed_maps.end()) return false;
    const float delta = m_voxel_side_length / points_per_voxel;
    it->second.map_shared_ptr->insertPointCloud(geometry_generation::createBoxOfPoints(corner_min, corner_max, delta), voxel_meaning);
    return true;
  }

  // Extension for planners (BASELINE.json configs[2]): the planner's per-state sequence clearMap(robot) /
  // transformSubCloud per link / insertMetaPointCloud / collideWith(env) (gvl_ompl_planner_helper.cpp:169-202)
  // for MANY states in one call. transforms[c * n_links + l] is link l's pose in state c.
  bool collideRobotConfigurations(const MetaPointCloud& robot_links, const std::vector<BitVoxelMeaning>& link_meanings,
                                  const std::vector<Matrix4f>& transforms, const std::string& env_map_name,
                                  std::vector<size_t>& collisions_per_configuration, float coll_threshold = 1.0,
                                  const BitVectorVoxel* types_to_check = NULL)
  {
    ManagedMapsIterator it = findMap(env_map_name);
    if (it == m_managed_maps.end()) return false;
    const uint32_t n_links = robot_links.getNumberOfPointclouds();
    if (!n_links || link_meanings.size() != n_links || transforms.size() % n_links)
    {
      LOGGING_ERROR_C(Gpu_voxels, GpuVoxels, "Number of link meanings / transforms does not fit the robot" << endl);
      return false;
    }
    const uint32_t n_cfg = uint32_t(transforms.size() / n_links);
    const std::vector<int32_t> mg = detail::meaningsToInts(link_meanings);
    std::vector<uint64_t> counts(n_cfg);
    lock_guard guard(it->second.map_shared_ptr->m_mutex);
    const int rc = gvx_batch_robot_collide(it->second.map_shared_ptr->getHandle(),
                                           reinterpret_cast<const float*>(robot_links.getDeviceBasePointer()), GVX_MEM_DEVICE,
                                           &robot_links.getPointcloudSizes()[0], n_links, &mg[0],
                                           n_cfg ? transforms[0].data() : NULL, n_cfg,
                                           types_to_check ? types_to_check->bitVector().bytes() : NULL, coll_threshold,
                                           n_cfg ? &counts[0] : NULL);
    if (!gvxCheck(rc, "collideRobotConfigurations")) return false;
    collisions_per_configuration.assign(counts.begin(), counts.end());
    return true;
  }

  // ---- robots (DH chains; GV/GpuVoxels.cpp addRobot / getRobot / setRobotConfiguration / insertRobotIntoMap) ----
  bool addRobot(const std::string& robot_name, const std::vector<std::string>& link_names,
                const std::vector<robot::DHParameters>& dh_params, const MetaPointCloud& pointclouds)
  {
    if (m_managed_robots.find(robot_name) != m_managed_robots.end())
    {
      LOGGING_ERROR_C(Gpu_voxels, GpuVoxels, "Robot with name '" << robot_name << "' already exists." << endl);
      return false;
    }
    m_managed_robots[robot_name] = RobotInterfaceSharedPtr(new robot::KinematicChain(link_names, dh_params, pointclouds));
    return true;
  }
  // link clouds loaded from files (GV/GpuVoxels.cpp:316-336)
  bool addRobot(const std::string& robot_name, const std::vector<std::string>& link_names,
                const std::vector<robot::DHParameters>& dh_params, const std::vector<std::string>& paths_to_pointclouds,
                const bool use_model_path)
  {
    if (m_managed_robots.find(robot_name) != m_managed_robots.end())
    {
      LOGGING_ERROR_C(Gpu_voxels, GpuVoxels, "Robot with name '" << robot_name << "' already exists." << endl);
      return false;
    }
    m_managed_robots[robot_name] =
      RobotInterfaceSharedPtr(new robot::KinematicChain(link_names, dh_params, paths_to_pointclouds, use_model_path));
    return true;
  }
  RobotInterfaceSharedPtr getRobot(const std::string& rob_name)
  {
    ManagedRobots::iterator it = m_managed_robots.find(rob_name);
    if (it == m_managed_robots.end())
    {
      LOGGING_ERROR_C(Gpu_voxels, GpuVoxels, "Could not find robot '" << rob_name << "'" << endl);
      return RobotInterfaceSharedPtr();
    }
    return it->second;
  }
  bool setRobotConfiguration(std::string robot_name, const robot::JointValueMap& jointmap)
  {
    RobotInterfaceSharedPtr r = getRobot(robot_name);
    if (!r) return false;
    r->setConfiguration(jointmap);
    return true;
  }
  bool getRobotConfiguration(const std::string& robot_name, robot::JointValueMap& jointmap)
  {
    RobotInterfaceSharedPtr r = getRobot(robot_name);
    if (!r) return false;
    r->getConfiguration(jointmap);
    return true;
  }
  bool updateRobotPart(std::string robot_name, const std::string& link_name, const std::vector<Vector3f> pointcloud)
  {
    RobotInterfaceSharedPtr r = getRobot(robot_name);
    if (!r) return false;
    r->updatePointcloud(link_name, pointcloud);
    return true;
  }
  bool insertRobotIntoMap(std::string robot_name, std::string map_name, const BitVoxelMeaning voxel_meaning)
  {
    RobotInterfaceSharedPtr r = getRobot(robot_name);
    if (!r) return false;
    ManagedMapsIterator it = findMap(map_name);
    if (it == m_managed_maps.end()) return false;
    it->second.map_shared_ptr->insertMetaPointCloud(*r->getTransformedClouds(), voxel_meaning);
    return true;
  }
  // Extension: collideRobotConfigurations with the forward kinematics of a managed DH robot done on the device.
  // joint_values[c * n_links + l] is the joint value of link l (in link_names order) in state c.
  bool collideRobotConfigurations(const std::string& robot_name, const std::vector<BitVoxelMeaning>& link_meanings,
                                  const std::vector<float>& joint_values, const std::string& env_map_name,
                                  std::vector<size_t>& collisions_per_configuration, float coll_threshold = 1.0,
                                  const BitVectorVoxel* types_to_check = NULL)
  {
    RobotInterfaceSharedPtr r = getRobot(robot_name);
    robot::KinematicChain* chain = r ? dynamic_cast<robot::KinematicChain*>(r.get()) : NULL;
    ManagedMapsIterator it = findMap(env_map_name);
    if (!chain || it == m_managed_maps.end()) return false;
    const MetaPointCloud& clouds = *chain->getLinkClouds();
    const std::vector<std::string>& names = chain->getLinkNames();
    const uint32_t n_links = uint32_t(names.size()), n_clouds = clouds.getNumberOfPointclouds();
    if (!n_links || !n_clouds || link_meanings.size() != n_clouds || joint_values.size() % n_links)
    {
      LOGGING_ERROR_C(Gpu_voxels, GpuVoxels, "Number of link meanings / joint values does not fit the robot" << endl);
      return false;
    }
    std::vector<float> dh4(4 * n_links);
    std::vector<int32_t> types(n_links), link_cloud(n_links);
    for (uint32_t i = 0; i < n_links; i++)
    {
      const robot::DHParameters p = chain->getLink(names[i])->getDHParam();
      dh4[4 * i] = p.d; dh4[4 * i + 1] = p.theta; dh4[4 * i + 2] = p.a; dh4[4 * i + 3] = p.alpha;
      types[i] = int32_t(p.joint_type);
      link_cloud[i] = clouds.getCloudNumber(names[i]);
    }
    const uint32_t n_cfg = uint32_t(joint_values.size() / n_links);
    const std::vector<int32_t> mg = detail::meaningsToInts(link_meanings);
    std::vector<uint64_t> counts(n_cfg);
    lock_guard guard(it->second.map_shared_ptr->m_mutex);
    const int rc = gvx_batch_robot_collide_dh(it->second.map_shared_ptr->getHandle(),
                                              reinterpret_cast<const float*>(clouds.getDeviceBasePointer()), GVX_MEM_DEVICE,
                                              &clouds.getPointcloudSizes()[0], n_clouds, &mg[0], &dh4[0], &types[0], n_links,
                                              &link_cloud[0], n_cfg ? &joint_values[0] : NULL, n_cfg,
                                              types_to_check ? types_to_check->bitVector().bytes() : NULL, coll_threshold,
                                              n_cfg ? &counts[0] : NULL);
    if (!gvxCheck(rc, "collideRobotConfigurations")) return false;
    collisions_per_configuration.assign(counts.begin(), counts.end());
    return true;
  }

  void getDimensions(uint32_t& dim_x, uint32_t& dim_y, uint32_t& dim_z) { dim_x = m_dim.x; dim_y = m_dim.y; dim_z = m_dim.z; }
  void getDimensions(Vector3ui& dim) { dim = m_dim; }
  void getVoxelSideLength(float& voxel_side_length) { voxel_side_length = m_voxel_side_length; }

protected:
  GpuVoxels() : m_dim(0, 0, 0), m_voxel_side_length(0) { gvxEnsureInit(); }

private:
  ManagedMapsIterator findMap(const std::string& map_name)
  {
    ManagedMapsIterator it = m_managed_maps.find(map_name);
    if (it == m_managed_maps.end())
      LOGGING_ERROR_C(Gpu_voxels, GpuVoxels, "Could not find map '" << map_name << "'" << endl);
    return it;
  }

  typedef std::map<std::string, RobotInterfaceSharedPtr> ManagedRobots;
  ManagedMaps m_managed_maps;
  ManagedRobots m_managed_robots;
  Vector3ui m_dim;
  float m_voxel_side_length;
};

}  // namespace gpu_voxels
#endif
