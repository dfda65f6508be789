// tests/cpp/test_facade.cpp -- the reference's own known-answer tests for the hot path, written against the
// header-only C++ facade (include/gpu_voxels/) exactly as a user of the reference would write them:
//   voxelmap suite   <- GV/test/testing_voxelmap.cu:95-167, 312-398
//   voxellists suite <- GV/test/testing_voxellist.cu:109-338, 739-864
//   facade suite     <- example_how_to_link/gvl_linkage_test_lib.cpp:20-39, GV/GpuVoxels.cpp error conventions
// plus the planner-batch extension checked against the per-configuration call sequence.
// Build: g++ -std=c++14 -Iinclude tests/cpp/test_facade.cpp -Lgpu_voxels_b200 -lgvx  (see tests/test_cpp_facade.py)
#include <cmath>
#include <sstream>

#include <gpu_voxels/GpuVoxels.h>

#include "mini_boost_test.h"

using namespace gpu_voxels;
using namespace gpu_voxels::voxelmap;
using namespace gpu_voxels::voxellist;
using namespace gpu_voxels::geometry_generation;

// GV/test/testing_fixtures.hpp:131-166
static const uint32_t dimX = 89, dimY = 123, dimZ = 74;
static const int numberOfPoints = 10000;

BOOST_AUTO_TEST_SUITE(voxelmap_tests)

BOOST_AUTO_TEST_CASE(collision)
{
  const float side_length = 1.f;
  ProbVoxelMap map_1(Vector3ui(dimX, dimY, dimZ), side_length, MT_PROBAB_VOXELMAP);
  ProbVoxelMap map_2(Vector3ui(dimX, dimY, dimZ), side_length, MT_PROBAB_VOXELMAP);
  std::vector<Vector3f> pts;
  createEquidistantPointsInBox(numberOfPoints, Vector3ui(dimX, dimY, dimZ), side_length, pts);
  map_1.insertPointCloud(pts, eBVM_OCCUPIED);
  map_2.insertPointCloud(pts, eBVM_OCCUPIED);
  BOOST_CHECK_MESSAGE(map_1.collideWith(&map_2, 0.1) == (size_t)numberOfPoints, "All collisions detected.");
  // the same through the collider-object entry points (GV/voxelmap/TemplateVoxelMap.hpp:477-519)
  BOOST_CHECK(map_1.collisionCheckWithCounter(&map_2, DefaultCollider(0.1f)) == (uint32_t)numberOfPoints);
  BOOST_CHECK(map_1.collisionCheckWithCounterRelativeTransform(&map_2, DefaultCollider(Probability(-101), Probability(-101))) == (uint32_t)numberOfPoints);
  // DefaultCollider(threshold1, threshold2) with two DIFFERENT thresholds (GV/voxel/DefaultCollider.hpp:45-49): every
  // inserted voxel holds 127, so it collides unless a threshold is out of reach on its side
  BOOST_CHECK(map_1.collisionCheckWithCounter(&map_2, DefaultCollider(Probability(1), Probability(2))) == (uint32_t)numberOfPoints);
  BOOST_CHECK(map_1.collisionCheckWithCounter(&map_2, DefaultCollider(Probability(127), Probability(-127))) == (uint32_t)numberOfPoints);
  ProbabilisticVoxel* raw = map_1.getDeviceDataPtr();
  BOOST_CHECK(raw != NULL && (const void*)raw == map_1.getConstVoidDeviceDataPtr());
}

BOOST_AUTO_TEST_CASE(collision_with_offset)
{
  ProbVoxelMap map_1(Vector3ui(dimX, dimY, dimZ), 1.f, MT_PROBAB_VOXELMAP);
  ProbVoxelMap map_2(Vector3ui(dimX, dimY, dimZ), 1.f, MT_PROBAB_VOXELMAP);
  map_1.insertPointCloud(createBoxOfPoints(Vector3f(2.1, 2.1, 2.1), Vector3f(4.1, 4.1, 4.1), 0.5), eBVM_OCCUPIED);
  map_2.insertPointCloud(createBoxOfPoints(Vector3f(3.1, 3.1, 3.1), Vector3f(5.1, 5.1, 5.1), 0.5), eBVM_OCCUPIED);
  BOOST_CHECK_MESSAGE(map_1.collideWith(&map_2, 0.1) == 8, "All collisions without offset detected.");
  BOOST_CHECK_MESSAGE(map_1.collideWith(&map_2, 0.1, Vector3i(-1, 0, -1)) == 18, "negative offset");
  BOOST_CHECK_MESSAGE(map_2.collideWith(&map_1, 0.1, Vector3i(1, 0, 1)) == 18, "positive offset");
}

BOOST_AUTO_TEST_CASE(no_collision)
{
  std::vector<Vector3f> black, white;
  ProbVoxelMap map_1(Vector3ui(dimX, dimY, dimZ), 1.f, MT_PROBAB_VOXELMAP);
  ProbVoxelMap map_2(Vector3ui(dimX, dimY, dimZ), 1.f, MT_PROBAB_VOXELMAP);
  createNonOverlapping3dCheckerboard(numberOfPoints, Vector3ui(dimX, dimY, dimZ), 1.f, black, white);
  map_1.insertPointCloud(black, eBVM_OCCUPIED);
  map_2.insertPointCloud(white, eBVM_OCCUPIED);
  BOOST_CHECK_MESSAGE(map_1.collideWith(&map_2, 0.1) == 0, "No collisions detected.");
  BOOST_CHECK(!map_1.collisionCheck(&map_2, 0.1));
}

BOOST_AUTO_TEST_CASE(collision_with_types_probab)
{
  BitVectorVoxelMap map_1(Vector3ui(dimX, dimY, dimZ), 1.f, MT_BITVECTOR_VOXELMAP);
  ProbVoxelMap map_2(Vector3ui(dimX, dimY, dimZ), 1.f, MT_PROBAB_VOXELMAP);
  std::vector<Vector3f> points;
  for (int x = 2; x <= 3; ++x)
    for (int y = 2; y <= 3; ++y)
      for (int z = 2; z <= 3; ++z) points.push_back(Vector3f(x, y, z));
  map_1.insertPointCloud(points, BitVoxelMeaning(eBVM_SWEPT_VOLUME_START + 10));
  map_1.insertPointCloud(points, BitVoxelMeaning(eBVM_SWEPT_VOLUME_START + 42));
  map_1.insertPointCloud(points, BitVoxelMeaning(eBVM_SWEPT_VOLUME_START + 99));
  map_2.insertPointCloud(points, eBVM_OCCUPIED);
  BitVectorVoxel coll_types;
  const size_t num_collisions = map_1.collideWithTypes(&map_2, coll_types);
  BOOST_CHECK(num_collisions == points.size());
  BOOST_CHECK(coll_types.bitVector().getBit(eBVM_SWEPT_VOLUME_START + 10));
  BOOST_CHECK(coll_types.bitVector().getBit(eBVM_SWEPT_VOLUME_START + 42));
  BOOST_CHECK(coll_types.bitVector().getBit(eBVM_SWEPT_VOLUME_START + 99));
  coll_types.bitVector().clearBit(eBVM_SWEPT_VOLUME_START + 10);
  coll_types.bitVector().clearBit(eBVM_SWEPT_VOLUME_START + 42);
  coll_types.bitVector().clearBit(eBVM_SWEPT_VOLUME_START + 99);
  BOOST_CHECK(coll_types.bitVector().noneButEmpty());
}

BOOST_AUTO_TEST_CASE(iostream_bitvoxel)
{
  BitVectorVoxel my_voxel;
  my_voxel.insert(eBVM_FREE);
  my_voxel.insert(eBVM_COLLISION);
  my_voxel.insert(eBVM_UNKNOWN);
  my_voxel.insert(eBVM_SWEPT_VOLUME_START);
  my_voxel.insert(BitVoxelMeaning(111));
  my_voxel.insert(eBVM_SWEPT_VOLUME_END);
  my_voxel.insert(eBVM_UNDEFINED);
  const std::string ground_truth = "10111" + std::string(106, '0') + "1" + std::string(142, '0') + "11";
  std::stringstream ss;
  ss << my_voxel;
  BOOST_CHECK_MESSAGE(ss.str() == ground_truth, "Outstream");
  BitVectorVoxel my_voxel2;
  ss.str(ground_truth);
  ss >> my_voxel2;
  std::stringstream s2;
  s2 << my_voxel2;
  BOOST_CHECK_MESSAGE(s2.str() == ground_truth, "Instream");
}

BOOST_AUTO_TEST_CASE(prob_map_rejects_other_meanings_and_marks_collisions)
{
  ProbVoxelMap pm(Vector3ui(32, 32, 32), 1.f);
  BitVectorVoxelMap bm(Vector3ui(32, 32, 32), 1.f);
  std::vector<Vector3f> p(1, Vector3f(3.5f, 4.5f, 5.5f));
  pm.insertPointCloud(p, eBVM_SWEPT_VOLUME_START);  // logged no-op (GV/voxelmap/ProbVoxelMap.hpp:118-140)
  bm.insertPointCloud(p, BitVoxelMeaning(20));
  BOOST_CHECK(bm.collideWith(&pm, 0.1) == 0);
  pm.insertPointCloud(p, eBVM_OCCUPIED);
  BOOST_CHECK(bm.collideWith(&pm, 0.1) == 1);
  // collideWith marked the colliding voxel of `this` with eBVM_COLLISION: clearing meaning 20 leaves it occupied
  bm.clearBitVoxelMeaning(BitVoxelMeaning(20));
  BOOST_CHECK(bm.collideWith(&pm, 0.1) == 1);
  bm.clearBitVoxelMeaning(eBVM_COLLISION);
  BOOST_CHECK(bm.collideWith(&pm, 0.1) == 0);
}

BOOST_AUTO_TEST_SUITE_END()

BOOST_AUTO_TEST_SUITE(voxellists)

static MetaPointCloud twoBoxes()
{
  std::vector<std::vector<Vector3f> > box_clouds;
  box_clouds.push_back(createBoxOfPoints(Vector3f(1.1, 1.1, 1.1), Vector3f(3.9, 3.9, 3.9), 0.1));
  box_clouds.push_back(createBoxOfPoints(Vector3f(2.1, 2.1, 2.1), Vector3f(4.9, 4.9, 4.9), 0.1));
  MetaPointCloud boxes(box_clouds);
  boxes.syncToDevice();
  return boxes;
}
static std::vector<BitVoxelMeaning> meanings1112()
{
  std::vector<BitVoxelMeaning> m;
  m.push_back(BitVoxelMeaning(11));
  m.push_back(BitVoxelMeaning(12));
  return m;
}

BOOST_AUTO_TEST_CASE(collide_bitvoxellist_with_prob_voxelmap)
{
  BitVectorVoxelList* list = new BitVectorVoxelList(Vector3ui(dimX, dimY, dimZ), 1.f, MT_BITVECTOR_VOXELLIST);
  MetaPointCloud boxes = twoBoxes();
  list->insertMetaPointCloud(boxes, meanings1112());
  GpuVoxelsMapSharedPtr map_2(new ProbVoxelMap(Vector3ui(dimX, dimY, dimZ), 1.f, MT_PROBAB_VOXELMAP));
  map_2->insertMetaPointCloud(boxes, eBVM_OCCUPIED);
  BOOST_CHECK_MESSAGE(list->collideWith(map_2->as<ProbVoxelMap>(), 1.0) == 46, "Number of Collisions == 46");
  delete list;
}

BOOST_AUTO_TEST_CASE(bitvoxellist_collide_with_types_both_map_types)
{
  BitVectorVoxelList list(Vector3ui(dimX, dimY, dimZ), 1.f, MT_BITVECTOR_VOXELLIST);
  MetaPointCloud boxes = twoBoxes();
  list.insertMetaPointCloud(boxes, meanings1112());
  ProbVoxelMap pmap(Vector3ui(dimX, dimY, dimZ), 1.f);
  BitVectorVoxelMap bmap(Vector3ui(dimX, dimY, dimZ), 1.f);
  pmap.insertMetaPointCloud(boxes, eBVM_OCCUPIED);
  bmap.insertMetaPointCloud(boxes, eBVM_OCCUPIED);
  BitVectorVoxel types;
  BOOST_CHECK(list.collideWithTypes(&pmap, types, 1.0) == 46);
  BOOST_CHECK(types.bitVector().getBit(11) && types.bitVector().getBit(12));
  BitVectorVoxel types_b;
  BOOST_CHECK(list.collideWithTypes(&bmap, types_b, 1.0) == 46);
  BOOST_CHECK(types_b.bitVector().getBit(11) && types_b.bitVector().getBit(12));
  std::vector<size_t> per_meaning;
  const size_t total = list.collideCountingPerMeaning(&bmap, per_meaning);
  BOOST_CHECK(per_meaning.size() == 256 && per_meaning[11] == 27 && per_meaning[12] == 27);
  BOOST_CHECK(total == 27 + 27 + per_meaning[eBVM_COLLISION]);
}

BOOST_AUTO_TEST_CASE(collide_bitvoxellist_with_prob_voxelmap_shifting)
{
  BitVectorVoxelList list(Vector3ui(dimX, dimY, dimZ), 1.f, MT_BITVECTOR_VOXELLIST);
  std::vector<std::vector<Vector3f> > box_cloud;
  box_cloud.push_back(createBoxOfPoints(Vector3f(1.1, 1.1, 1.1), Vector3f(3.9, 3.9, 3.9), 0.1));
  MetaPointCloud box(box_cloud);
  box.syncToDevice();
  list.insertMetaPointCloud(box, eBVM_OCCUPIED);
  GpuVoxelsMapSharedPtr map_2(new ProbVoxelMap(Vector3ui(dimX, dimY, dimZ), 1.f, MT_PROBAB_VOXELMAP));
  map_2->insertMetaPointCloud(box, eBVM_OCCUPIED);
  for (float shift = 0.0; shift < 4.0; shift += 0.5)
  {
    const size_t num_colls = list.collideWith(map_2->as<ProbVoxelMap>(), 1.0, Vector3i(shift, 0, 0));
    const size_t want = shift < 1.0 ? 27 : shift < 2.0 ? 18 : shift < 3.0 ? 9 : 0;
    BOOST_CHECK_MESSAGE(num_colls == want, "shift " << shift << ": " << num_colls << " != " << want);
  }
}

BOOST_AUTO_TEST_CASE(bitvoxellist_insert_metapointcloud)
{
  BitVectorVoxelList list(Vector3ui(dimX, dimY, dimZ), 1, MT_BITVECTOR_VOXELLIST);
  MetaPointCloud boxes = twoBoxes();
  list.insertMetaPointCloud(boxes, meanings1112());
  std::vector<uint32_t> ids;
  std::vector<Vector3ui> coords;
  std::vector<BitVectorVoxel> voxels;
  list.copyToHost(ids, coords, voxels);
  BOOST_CHECK_MESSAGE(ids.size() == 46, "Number of reduced cubes == 46");
  BOOST_CHECK(list.getDimensions().x == 46);
  BitVector<BIT_VECTOR_LENGTH> ref_bv;
  ref_bv.setBit(11);
  ref_bv.setBit(12);
  size_t both = 0;
  for (size_t i = 0; i < ids.size(); i++)
    if (voxels[i].bitVector() == ref_bv)
    {
      ++both;
      const Vector3ui& c = coords[i];
      BOOST_CHECK_MESSAGE(c.x >= 2 && c.x <= 3 && c.y >= 2 && c.y <= 3 && c.z >= 2 && c.z <= 3, "merged voxel in the centre");
    }
  BOOST_CHECK_MESSAGE(both == 8, "Detect 8 overlapping voxels.");
}

BOOST_AUTO_TEST_CASE(bitmasked_collision)
{
  // GV/test/testing_voxellist.cu:739-864: three boxes with meanings 34 / 63 / 102 against one obstacle, side 0.025
  const float side = 0.025f;
  BitVectorVoxelList list(Vector3ui(dimX, dimY, dimZ), side, MT_BITVECTOR_VOXELLIST);
  ProbVoxelMap map(Vector3ui(dimX, dimY, dimZ), side, MT_PROBAB_VOXELMAP);
  map.insertPointCloud(createBoxOfPoints(Vector3f(0.111, 0.111, 0.111), Vector3f(0.799, 0.399, 0.399), 0.005), eBVM_OCCUPIED);
  list.insertPointCloud(createBoxOfPoints(Vector3f(0.111, 0.311, 0.111), Vector3f(0.399, 0.599, 0.399), 0.005), BitVoxelMeaning(34));
  list.insertPointCloud(createBoxOfPoints(Vector3f(0.311, 0.311, 0.111), Vector3f(0.599, 0.599, 0.399), 0.005), BitVoxelMeaning(63));
  list.insertPointCloud(createBoxOfPoints(Vector3f(0.511, 0.311, 0.111), Vector3f(0.799, 0.599, 0.399), 0.005), BitVoxelMeaning(102));
  const Vector3i offs[3] = {Vector3i(0, 0, 0), Vector3i(0, -4, 0), Vector3i(0, -4, -4)};
  const size_t expect[3][5] = {{0, 0, 9, 15, 21}, {0, 0, 18, 30, 42}, {0, 0, 12, 20, 28}};
  const int bits[4] = {33, 34, 63, 102};
  for (int o = 0; o < 3; ++o)
  {
    BitVectorVoxel mask;
    BOOST_CHECK(list.collideWithTypeMask(&map, mask, 1.0, offs[o]) == expect[o][0] * 64);
    for (int b = 0; b < 4; ++b)
    {
      mask.bitVector().setBit(bits[b]);
      const size_t got = list.collideWithTypeMask(&map, mask, 1.0, offs[o]);
      BOOST_CHECK_MESSAGE(got == expect[o][b + 1] * 64, "offset " << o << " bit " << bits[b] << ": " << got);
    }
  }
}

BOOST_AUTO_TEST_SUITE_END()

BOOST_AUTO_TEST_SUITE(facade)

BOOST_AUTO_TEST_CASE(linkage_example)
{
  // example_how_to_link/gvl_linkage_test_lib.cpp: two overlapping boxes in two ProbVoxelMaps
  GpuVoxelsSharedPtr gvl = GpuVoxels::getInstance();
  gvl->initialize(100, 100, 100, 0.01);
  gvl->initialize(10, 10, 10, 1.0);  // warns, parameters remain unchanged
  uint32_t dx, dy, dz;
  gvl->getDimensions(dx, dy, dz);
  BOOST_CHECK(dx == 100 && dy == 100 && dz == 100);
  BOOST_CHECK(gvl->addMap(MT_PROBAB_VOXELMAP, "myFirstMap") != NULL);
  BOOST_CHECK(gvl->addMap(MT_PROBAB_VOXELMAP, "mySecondMap") != NULL);
  BOOST_CHECK(gvl->addMap(MT_PROBAB_VOXELMAP, "mySecondMap") == NULL);  // duplicate name
  const Vector3f center(0.5, 0.5, 0.5), box_size(0.4, 0.4, 0.4), overlap(0.1, 0.1, 0.1);
  const Vector3f c1 = center - overlap, c2 = c1 + box_size, c3 = center + overlap, c4 = c3 - box_size;
  BOOST_CHECK(gvl->insertBoxIntoMap(c1, c2, "myFirstMap", eBVM_OCCUPIED, 1));
  BOOST_CHECK(gvl->insertBoxIntoMap(c4, c3, "mySecondMap", eBVM_OCCUPIED, 1));
  BOOST_CHECK(!gvl->insertBoxIntoMap(c4, c3, "noSuchMap", eBVM_OCCUPIED, 1));
  const size_t n = gvl->getMap("myFirstMap")->as<ProbVoxelMap>()->collideWith(gvl->getMap("mySecondMap")->as<ProbVoxelMap>());
  // overlap = [0.4, 0.6]^3 sampled at 1 cm: count the voxels both clouds touch with the same float accumulation
  std::vector<Vector3f> a = createBoxOfPoints(c1, c2, 0.01f), b = createBoxOfPoints(c4, c3, 0.01f);
  std::vector<char> va(100 * 100 * 100, 0), vb(100 * 100 * 100, 0);
  for (size_t i = 0; i < a.size(); ++i) va[size_t(floorf(a[i].z / 0.01f)) * 10000 + size_t(floorf(a[i].y / 0.01f)) * 100 + size_t(floorf(a[i].x / 0.01f))] = 1;
  for (size_t i = 0; i < b.size(); ++i) vb[size_t(floorf(b[i].z / 0.01f)) * 10000 + size_t(floorf(b[i].y / 0.01f)) * 100 + size_t(floorf(b[i].x / 0.01f))] = 1;
  size_t want = 0;
  for (size_t i = 0; i < va.size(); ++i) want += va[i] && vb[i];
  BOOST_CHECK_MESSAGE(n == want && n > 0, n << " collisions, expected " << want);
  BOOST_CHECK(gvl->clearMap("myFirstMap"));
  BOOST_CHECK(gvl->getMap("myFirstMap")->as<ProbVoxelMap>()->collideWith(gvl->getMap("mySecondMap")->as<ProbVoxelMap>()) == 0);
  BOOST_CHECK(!gvl->clearMap("nope"));
  bool thrown = false;
  try { gvl->addMap(MT_BITVECTOR_OCTREE, "tree"); } catch (const std::string&) { thrown = true; }
  BOOST_CHECK(thrown);
  BOOST_CHECK(gvl->delMap("myFirstMap") && !gvl->delMap("myFirstMap"));
  BOOST_CHECK(gvl->getMap("mySecondMap")->is<ProbVoxelMap>() && !gvl->getMap("mySecondMap")->is<BitVectorVoxelMap>());
}

BOOST_AUTO_TEST_CASE(batch_equals_planner_sequence)
{
  GpuVoxelsSharedPtr gvl = GpuVoxels::getInstance();
  gvl->initialize(100, 100, 100, 0.01);
  gvl->addMap(MT_BITVECTOR_VOXELMAP, "env");
  gvl->addMap(MT_BITVECTOR_VOXELMAP, "robot");
  gvl->insertBoxIntoMap(Vector3f(0.3, 0.3, 0.3), Vector3f(0.6, 0.6, 0.6), "env", eBVM_OCCUPIED, 1);
  std::vector<std::vector<Vector3f> > links;
  links.push_back(createBoxOfPoints(Vector3f(-0.05, -0.05, 0), Vector3f(0.05, 0.05, 0.3), 0.004));
  links.push_back(createBoxOfPoints(Vector3f(-0.03, -0.03, 0), Vector3f(0.03, 0.03, 0.2), 0.004));
  MetaPointCloud robot(links);
  robot.syncToDevice();
  std::vector<BitVoxelMeaning> mg;
  mg.push_back(BitVoxelMeaning(10));
  mg.push_back(BitVoxelMeaning(11));
  const int n_cfg = 9;
  std::vector<Matrix4f> T;
  for (int c = 0; c < n_cfg; ++c)
  {
    const Matrix4f base = Matrix4f::createFromRotationAndTranslation(Matrix3f::createFromRPY(0.1f * c, 0.2f * c, 0.3f * c),
                                                                    Vector3f(0.1f + 0.08f * c, 0.45f, 0.2f));
    const Matrix4f joint = Matrix4f::createFromRotationAndTranslation(Matrix3f::createFromRPY(0.5f, 0, 0), Vector3f(0, 0, 0.3f));
    T.push_back(base);
    T.push_back(base * joint);
  }
  std::vector<size_t> counts;
  BOOST_CHECK(gvl->collideRobotConfigurations(robot, mg, T, "env", counts));
  BOOST_CHECK(counts.size() == size_t(n_cfg));
  BitVectorVoxelMap* rob = gvl->getMap("robot")->as<BitVectorVoxelMap>();
  BitVectorVoxelMap* env = gvl->getMap("env")->as<BitVectorVoxelMap>();
  size_t sum = 0;
  for (int c = 0; c < n_cfg; ++c)
  {
    // the reference's sequence: clear, transform each link, insert, collide
    gvl->clearMap("robot");
    MetaPointCloud moved(robot);
    moved.transformSubCloud(0, &T[2 * c], &moved);
    moved.transformSubCloud(1, &T[2 * c + 1], &moved);
    gvl->insertMetaPointCloudIntoMap(moved, "robot", mg);
    const size_t want = rob->collideWith(env, 1.0);
    BOOST_CHECK_MESSAGE(counts[c] == want, "configuration " << c << ": " << counts[c] << " != " << want);
    sum += want;
  }
  BOOST_CHECK(sum > 0);
  gvl->delMap("env");
  gvl->delMap("robot");
}

BOOST_AUTO_TEST_SUITE_END()

// ---- swept volumes, list x list, files, robots ------------------------------------------------------
BOOST_AUTO_TEST_SUITE(next_rows)

// GV/test/testing_voxelmap.cu:173-229
BOOST_AUTO_TEST_CASE(collision_bitshift)
{
  float side_length = 1.f;
  BitVectorVoxelMap map_1(Vector3ui(dimX, dimY, dimZ), side_length, MT_BITVECTOR_VOXELMAP);
  BitVectorVoxelMap map_2(Vector3ui(dimX, dimY, dimZ), side_length, MT_BITVECTOR_VOXELMAP);
  std::vector<Vector3f> points;
  for (int x = 2; x <= 3; ++x) for (int y = 2; y <= 3; ++y) for (int z = 2; z <= 3; ++z) points.push_back(Vector3f(x, y, z));
  const BitVoxelMeaning type_1 = BitVoxelMeaning(eBVM_SWEPT_VOLUME_START + 40);
  map_1.insertPointCloud(points, type_1);
  size_t window_size = 20;
  SVCollider collider(1.0, window_size);
  BitVectorVoxel types_voxel;
  for (size_t i = type_1 - window_size - 9; i <= type_1 + window_size + 9; ++i)
  {
    map_2.clearMap();
    map_2.insertPointCloud(points, BitVoxelMeaning(i));
    size_t num_collisions = map_1.collisionCheckBitvector(&map_2, collider, types_voxel.bitVector());
    if (num_collisions > 0)
    {
      types_voxel.bitVector().clearBit(type_1);
      BOOST_CHECK(types_voxel.bitVector().isZero());
    }
    if (i < type_1 - window_size || i > type_1 + window_size) BOOST_CHECK(num_collisions == 0);
    else BOOST_CHECK(num_collisions == points.size());
  }
}

// GV/test/testing_voxelmap.cu:238-304
BOOST_AUTO_TEST_CASE(bitvoxelmap_bitshift)
{
  BitVectorVoxelMap map_1(Vector3ui(dimX, dimY, dimZ), 1.f, MT_BITVECTOR_VOXELMAP);
  BitVectorVoxelMap map_2(Vector3ui(dimX, dimY, dimZ), 1.f, MT_BITVECTOR_VOXELMAP);
  std::vector<Vector3f> points;
  points.push_back(Vector3f(0.3, 0.3, 0.3)); points.push_back(Vector3f(1.3, 1.3, 1.3));
  points.push_back(Vector3f(2.3, 2.3, 2.3)); points.push_back(Vector3f(3.3, 3.3, 3.3));
  uint32_t shift_size = 0;
  const uint32_t shift_start = 50;
  const uint32_t type_int = eBVM_SWEPT_VOLUME_START + shift_start;
  while (shift_size < shift_start + eBVM_SWEPT_VOLUME_START - 5)
  {
    map_1.clearMap(); map_2.clearMap();
    map_2.insertPointCloud(points, BitVoxelMeaning(type_int));
    map_1.insertPointCloud(points, BitVoxelMeaning(type_int - shift_size));
    map_2.shiftLeftSweptVolumeIDs(shift_size);
    BitVectorVoxel types_voxel;
    const size_t window_size = 1;
    SVCollider collider(1.0, window_size);  // as GV/test/testing_voxelmap.cu:197 writes it
    size_t num_collisions = map_1.collisionCheckBitvector(&map_2, collider, types_voxel.bitVector());
    if (shift_size <= shift_start) BOOST_CHECK(num_collisions == points.size());
    else BOOST_CHECK(num_collisions == 0);
    shift_size++;
  }
}

// GV/test/testing_voxellist.cu:341-410
BOOST_AUTO_TEST_CASE(bitvoxellist_findMatchingVoxels)
{
  float side_length = 0.01;
  GpuVoxelsMapSharedPtr voxellist1_shrd_ptr(new BitVectorVoxelList(Vector3ui(dimX, dimY, dimZ), side_length, MT_BITVECTOR_VOXELLIST));
  GpuVoxelsMapSharedPtr voxellist2_shrd_ptr(new BitVectorVoxelList(Vector3ui(dimX, dimY, dimZ), side_length, MT_BITVECTOR_VOXELLIST));
  std::vector<Vector3f> box1 = createBoxOfPoints(Vector3f(2.1, 0.1, 5.1), Vector3f(4.9, 4.9, 9.9), 0.02f);
  std::vector<Vector3f> box2 = createBoxOfPoints(Vector3f(5.1, 5.1, 10.1), Vector3f(9.9, 9.9, 14.9), 0.3f);
  std::vector<Vector3f> cp1, cp2, cp3;
  for (int j = 0; j < 40; j++) cp1.push_back(Vector3f(j / 40.0 * 5.0, 1.0, 1.0));
  for (int j = 0; j < 30; j++) cp2.push_back(Vector3f(j / 30.0 * 5.0, 2.0, 1.0));
  for (int j = 0; j < 20; j++) cp3.push_back(Vector3f(j / 20.0 * 5.0, 3.0, 1.0));
  voxellist1_shrd_ptr->insertPointCloud(box1, BitVoxelMeaning(10));
  voxellist1_shrd_ptr->insertPointCloud(cp1, BitVoxelMeaning(11));
  voxellist1_shrd_ptr->insertPointCloud(cp2, BitVoxelMeaning(12));
  voxellist1_shrd_ptr->insertPointCloud(cp3, BitVoxelMeaning(13));
  voxellist2_shrd_ptr->insertPointCloud(box2, BitVoxelMeaning(20));
  voxellist2_shrd_ptr->insertPointCloud(cp1, BitVoxelMeaning(21));
  voxellist2_shrd_ptr->insertPointCloud(cp2, BitVoxelMeaning(22));
  voxellist2_shrd_ptr->insertPointCloud(cp3, BitVoxelMeaning(23));
  std::vector<size_t> collisions_per_meaning(BIT_VECTOR_LENGTH, 0);
  size_t num_coll = voxellist1_shrd_ptr->as<BitVectorVoxelList>()->collideCountingPerMeaning(voxellist2_shrd_ptr, collisions_per_meaning);
  BOOST_CHECK(num_coll == 90);
  BOOST_CHECK(collisions_per_meaning[11] == 40 && collisions_per_meaning[12] == 30 && collisions_per_meaning[13] == 20);
  BOOST_CHECK(collisions_per_meaning[21] == 0 && collisions_per_meaning[22] == 0 && collisions_per_meaning[23] == 0);
  BOOST_CHECK(collisions_per_meaning[10] == 0 && collisions_per_meaning[20] == 0);
  collisions_per_meaning.assign(BIT_VECTOR_LENGTH, 0);
  num_coll = voxellist2_shrd_ptr->as<BitVectorVoxelList>()->collideCountingPerMeaning(voxellist1_shrd_ptr, collisions_per_meaning);
  BOOST_CHECK(num_coll == 90);
  BOOST_CHECK(collisions_per_meaning[21] == 40 && collisions_per_meaning[22] == 30 && collisions_per_meaning[23] == 20);
  BOOST_CHECK(collisions_per_meaning[11] == 0 && collisions_per_meaning[12] == 0 && collisions_per_meaning[13] == 0);
}

// GV/test/testing_voxellist.cu:412-460
BOOST_AUTO_TEST_CASE(bitvoxellist_bitshift_collision)
{
  BitVectorVoxelList map_1(Vector3ui(dimX, dimY, dimZ), 1.f, MT_BITVECTOR_VOXELLIST);
  GpuVoxelsMapSharedPtr map_2(new BitVectorVoxelList(Vector3ui(dimX, dimY, dimZ), 1.f, MT_BITVECTOR_VOXELLIST));
  BitVectorVoxelList* map2_base_ptr = map_2->as<BitVectorVoxelList>();
  std::vector<Vector3f> points;
  points.push_back(Vector3f(0.3, 0.3, 0.3));
  uint32_t shift_size = 0;
  const uint32_t shift_start = 50;
  const uint32_t type_int = eBVM_SWEPT_VOLUME_START + shift_start;
  while (shift_size < shift_start + eBVM_SWEPT_VOLUME_START - 5)
  {
    map_1.clearMap(); map_2->clearMap();
    map_2->insertPointCloud(points, BitVoxelMeaning(type_int));
    map_1.insertPointCloud(points, BitVoxelMeaning(type_int - shift_size));
    map2_base_ptr->shiftLeftSweptVolumeIDs(shift_size);
    size_t num_collisions = map_1.collideWithBitcheck(map_2->as<voxellist::BitVectorVoxelList>(), 1);
    if (shift_size <= shift_start) BOOST_CHECK(num_collisions == 1);
    else BOOST_CHECK(num_collisions == 0);
    shift_size++;
  }
}

// GV/test/testing_voxellist.cu:462-572: equals, disk io, subtract
BOOST_AUTO_TEST_CASE(voxellist_equals_disk_io_subtract)
{
  BitVectorVoxelList list1(Vector3ui(dimX, dimY, dimZ), 1, MT_BITVECTOR_VOXELLIST);
  BitVectorVoxelList list2(Vector3ui(dimX, dimY, dimZ), 1, MT_BITVECTOR_VOXELLIST);
  std::vector<BitVoxelMeaning> voxel_meanings;
  voxel_meanings.push_back(BitVoxelMeaning(11));
  voxel_meanings.push_back(BitVoxelMeaning(12));
  std::vector<std::vector<Vector3f> > box_clouds;
  box_clouds.push_back(createBoxOfPoints(Vector3f(1.1, 1.1, 1.1), Vector3f(3.9, 3.9, 3.9), 0.1));
  box_clouds.push_back(createBoxOfPoints(Vector3f(2.1, 2.1, 2.1), Vector3f(4.9, 4.9, 4.9), 0.1));
  MetaPointCloud boxes(box_clouds);
  boxes.syncToDevice();
  list1.insertMetaPointCloud(boxes, voxel_meanings);
  list2.insertMetaPointCloud(boxes, voxel_meanings);
  BOOST_CHECK_MESSAGE(list1.equals(list2), "Lists are equal.");
  BOOST_CHECK(list1.writeToDisk("/tmp/gvx_temp_list.lst"));
  BitVectorVoxelList list3(Vector3ui(dimX, dimY, dimZ), 1, MT_BITVECTOR_VOXELLIST);
  BOOST_CHECK(list3.readFromDisk("/tmp/gvx_temp_list.lst"));
  BOOST_CHECK_MESSAGE(list1.equals(list3), "List from Disk equals original list.");
  std::vector<Vector3f> outliers;
  outliers.push_back(Vector3f(9.22, 9.22, 9.22));
  list1.insertPointCloud(outliers, BitVoxelMeaning(11));
  list2.insertPointCloud(outliers, BitVoxelMeaning(12));
  BOOST_CHECK_MESSAGE(!list1.equals(list2), "Lists differ.");

  BitVectorVoxelList a(Vector3ui(dimX, dimY, dimZ), 1, MT_BITVECTOR_VOXELLIST);
  BitVectorVoxelList b(Vector3ui(dimX, dimY, dimZ), 1, MT_BITVECTOR_VOXELLIST);
  a.insertPointCloud(box_clouds[0], BitVoxelMeaning(11));
  b.insertPointCloud(box_clouds[1], BitVoxelMeaning(12));
  a.subtract(&b, Vector3f());
  BOOST_CHECK_MESSAGE(a.size() == 19, "Number of cubes after subtract == 19");
}

// TemplateVoxelList<BitVoxel>::merge (GV/voxellist/TemplateVoxelList.hpp:548-608): the two 27-voxel boxes of
// bitvoxellist_insert_metapointcloud as two lists -> 46 entries; with an offset and a new meaning; maps are refused
BOOST_AUTO_TEST_CASE(voxellist_merge)
{
  std::shared_ptr<BitVectorVoxelList> a(new BitVectorVoxelList(Vector3ui(dimX, dimY, dimZ), 1, MT_BITVECTOR_VOXELLIST));
  std::shared_ptr<BitVectorVoxelList> b(new BitVectorVoxelList(Vector3ui(dimX, dimY, dimZ), 1, MT_BITVECTOR_VOXELLIST));
  a->insertPointCloud(createBoxOfPoints(Vector3f(1.1, 1.1, 1.1), Vector3f(3.9, 3.9, 3.9), 0.1), BitVoxelMeaning(11));
  b->insertPointCloud(createBoxOfPoints(Vector3f(2.1, 2.1, 2.1), Vector3f(4.9, 4.9, 4.9), 0.1), BitVoxelMeaning(12));
  BOOST_CHECK(a->merge(b));
  BOOST_CHECK_MESSAGE(a->size() == 46, "27 + 27 - 8 shared voxels");
  BOOST_CHECK(b->size() == 27);
  BitVectorVoxel types;
  BOOST_CHECK(a->collideWithTypes(b.get(), types, 1.0) == 27);
  BOOST_CHECK(types.bitVector().getBit(11) && types.bitVector().getBit(12));
  // shifted by (10, 0, 0) with a new meaning: 27 new, disjoint voxels carrying only bit 40
  const BitVoxelMeaning forty = BitVoxelMeaning(40);
  BOOST_CHECK(a->merge(b, Vector3i(10, 0, 0), &forty));
  BOOST_CHECK(a->size() == 46 + 27);
  BOOST_CHECK(a->merge(b, Vector3f(10.5f, 0.f, 0.f), &forty));  // metric: floor(10.5 / 1) = 10 -> nothing new
  BOOST_CHECK(a->size() == 46 + 27);
  GpuVoxelsMapSharedPtr map(new ProbVoxelMap(Vector3ui(dimX, dimY, dimZ), 1.f, MT_PROBAB_VOXELMAP));
  BOOST_CHECK(!a->merge(map));
}

// TemplateVoxelMap::writeToDisk / readFromDisk (GV/voxelmap/TemplateVoxelMap.hpp:666-799)// TemplateVoxelMap::writeToDisk / readFromDisk (GV/voxelmap/TemplateVoxelMap.hpp:666-799)
BOOST_AUTO_TEST_CASE(voxelmap_disk_io)
{
  ProbVoxelMap m1(Vector3ui(40, 30, 20), 0.5f, MT_PROBAB_VOXELMAP), m2(Vector3ui(40, 30, 20), 0.5f, MT_PROBAB_VOXELMAP);
  BitVectorVoxelMap wrong(Vector3ui(40, 30, 20), 0.5f, MT_BITVECTOR_VOXELMAP);
  m1.insertPointCloud(createBoxOfPoints(Vector3f(2.1, 2.1, 2.1), Vector3f(6.1, 6.1, 6.1), 0.25), eBVM_OCCUPIED);
  BOOST_CHECK(m1.writeToDisk("/tmp/gvx_temp_map.bin"));
  BOOST_CHECK(m2.readFromDisk("/tmp/gvx_temp_map.bin"));
  BOOST_CHECK(!wrong.readFromDisk("/tmp/gvx_temp_map.bin"));
  BOOST_CHECK(m1.collideWith(&m2, 0.5) == 9 * 9 * 9);
}

// PointcloudFileHandler / insertPointCloudFromFile / PointCloud(file) / MetaPointCloud(files)
// (GV/helpers/PointcloudFileHandler.cpp:64-137, GV/GpuVoxelsMap.cpp:36-49, GV/helpers/PointCloud.cu:84-105,
// GV/helpers/MetaPointCloud.cu:115-151); files of tests/golden/files, point counts as the reference's readers give them
BOOST_AUTO_TEST_CASE(pointcloud_files)
{
  const char* dir = std::getenv("GVX_TEST_FILES");
  BOOST_CHECK_MESSAGE(dir != NULL, "GVX_TEST_FILES names tests/golden/files");
  if (!dir) return;
  setenv("GPU_VOXELS_MODEL_PATH", dir, 1);
  std::vector<Vector3f> pts;
  file_handling::PointcloudFileHandler* fh = file_handling::PointcloudFileHandler::Instance();
  BOOST_CHECK(fh->loadPointCloud("cube.binvox", true, pts));
  BOOST_CHECK_MESSAGE(pts.size() == 52, "cube.binvox holds 52 voxels");
  BOOST_CHECK(fh->loadPointCloud("messy.xyz", true, pts, true, Vector3f(1, 1, 1), 2.0f));  // appends, then post-processes all
  BOOST_CHECK_MESSAGE(pts.size() == 52 + 11, "messy.xyz adds 11 points");
  float lo = pts[0].x;
  for (size_t i = 0; i < pts.size(); ++i) lo = std::min(lo, pts[i].x);
  BOOST_CHECK_MESSAGE(lo == 1.0f, "shifted to zero, then scaled and offset by 1");
  BOOST_CHECK(!fh->loadPointCloud("bad_magic.binvox", true, pts));
  BOOST_CHECK(!fh->loadPointCloud("cloud.txt", true, pts));
  BOOST_CHECK(pts.size() == 63);

  PointCloud pc("grid32.binvox", true);
  BOOST_CHECK(pc.getPointCloudSize() == 1697);
  std::vector<std::string> files;
  files.push_back("cube.binvox");
  files.push_back("simple.xyz");
  MetaPointCloud mpc(files, true);
  BOOST_CHECK(mpc.getNumberOfPointclouds() == 2 && mpc.getPointcloudSize(0) == 52 && mpc.getPointcloudSize(1) == 200);
  files.push_back("missing.xyz");
  MetaPointCloud none(files, true);  // one unreadable file: nothing is added
  BOOST_CHECK(none.getNumberOfPointclouds() == 0);

  // grid32.binvox (0.77 / 32 per cell) scaled to one 1 cm voxel per cell: every one of its 1697 voxels is distinct
  GpuVoxelsSharedPtr gvl = GpuVoxels::getInstance();
  gvl->initialize(160, 160, 160, 0.01f);
  gvl->addMap(MT_BITVECTOR_VOXELMAP, "from_file");
  gvl->addMap(MT_BITVECTOR_VOXELLIST, "list_from_file");
  BOOST_CHECK(gvl->insertPointCloudFromFile("from_file", "grid32.binvox", true, eBVM_OCCUPIED, true, Vector3f(0.005f, 0.005f, 0.005f), 0.415584f));
  BOOST_CHECK(gvl->insertPointCloudFromFile("list_from_file", "grid32.binvox", true, eBVM_OCCUPIED, true, Vector3f(0.005f, 0.005f, 0.005f), 0.415584f));
  BOOST_CHECK(!gvl->insertPointCloudFromFile("no_such_map", "grid32.binvox", true, eBVM_OCCUPIED));
  BOOST_CHECK(!gvl->insertPointCloudFromFile("from_file", "no_dim.binvox", true, eBVM_OCCUPIED));
  voxellist::BitVectorVoxelList* list = gvl->getMap("list_from_file")->as<voxellist::BitVectorVoxelList>();
  BOOST_CHECK_MESSAGE(list->size() == 1697, "one list entry per binvox voxel");
  BOOST_CHECK(list->collideWith(gvl->getMap("from_file")->as<voxelmap::BitVectorVoxelMap>(), 1.0f) == 1697);

  // a DH robot whose link clouds come from files (GV/GpuVoxels.cpp:316-336): the clouds are named by their paths and
  // a link moves the cloud of its own name -- it must equal the same robot built from the loaded clouds
  std::vector<std::string> paths;
  paths.push_back("simple.xyz");
  paths.push_back("cube.binvox");
  std::vector<robot::DHParameters> dh;
  dh.push_back(robot::DHParameters(0.3f, 0.f, 0.1f, 1.5707963f, 0.f, robot::REVOLUTE));
  dh.push_back(robot::DHParameters(0.0f, 0.f, 0.3f, 0.f, 0.f, robot::REVOLUTE));
  BOOST_CHECK(gvl->addRobot("file_arm", paths, dh, paths, true));
  BOOST_CHECK(!gvl->addRobot("file_arm", paths, dh, paths, true));
  MetaPointCloud loaded(paths, paths, true);
  BOOST_CHECK(gvl->addRobot("cloud_arm", paths, dh, loaded));
  gvl->addMap(MT_BITVECTOR_VOXELMAP, "rob_a");
  gvl->addMap(MT_BITVECTOR_VOXELMAP, "rob_b");
  robot::JointValueMap jm;
  jm["simple.xyz"] = 0.4f;
  jm["cube.binvox"] = -0.7f;
  BOOST_CHECK(gvl->setRobotConfiguration("file_arm", jm) && gvl->setRobotConfiguration("cloud_arm", jm));
  BOOST_CHECK(gvl->insertRobotIntoMap("file_arm", "rob_a", eBVM_OCCUPIED));
  BOOST_CHECK(gvl->insertRobotIntoMap("cloud_arm", "rob_b", eBVM_OCCUPIED));
  voxelmap::BitVectorVoxelMap* ra = gvl->getMap("rob_a")->as<voxelmap::BitVectorVoxelMap>();
  voxelmap::BitVectorVoxelMap* rb = gvl->getMap("rob_b")->as<voxelmap::BitVectorVoxelMap>();
  const size_t occupied = ra->collideWith(ra, 1.0f);
  BOOST_CHECK_MESSAGE(occupied > 20, "the robot occupies voxels: " << occupied);
  BOOST_CHECK(ra->collideWith(rb, 1.0f) == occupied && rb->collideWith(rb, 1.0f) == occupied);
  gvl->delMap("rob_a");
  gvl->delMap("rob_b");
  gvl->delMap("from_file");
  gvl->delMap("list_from_file");
}

// ray casting through the reference's call: the cells between the sensor and a wall become free, the wall occupied
BOOST_AUTO_TEST_CASE(insert_sensor_data_with_raycasting_and_robot_cut)
{
  const Vector3ui dim(64, 64, 32);
  ProbVoxelMap env(dim, 0.1f, MT_PROBAB_VOXELMAP), probe(dim, 0.1f, MT_PROBAB_VOXELMAP);
  BitVectorVoxelMap robot(dim, 0.1f, MT_BITVECTOR_VOXELMAP);
  std::vector<Vector3f> wall;
  for (int y = -10; y <= 10; ++y) for (int z = -5; z <= 5; ++z) wall.push_back(Vector3f(0.1f * y, 0.1f * z, 3.05f));  // sensor frame
  // sensor at (0.55, 3.25, 1.65) looking along +x: sensor z -> world x, sensor x -> world y, sensor y -> world z
  Sensor sensor(Vector3f(0.55f, 3.25f, 1.65f), Matrix3f(0, 0, 1, 1, 0, 0, 0, 1, 0), uint32_t(wall.size()), 1);
  env.initSensorSettings(sensor);
  env.insertSensorData(&wall[0], true, false, eBVM_OCCUPIED);
  // the wall plane x = 36 is occupied, the cells in front of it on the central ray are free
  std::vector<Vector3f> hit(1, Vector3f(3.65f, 3.25f, 1.65f)), mid(1, Vector3f(2.05f, 3.25f, 1.65f));
  probe.insertPointCloud(hit, eBVM_OCCUPIED);
  BOOST_CHECK(env.collideWith(&probe, 0.5) == 0);      // -127 -10 .. +72 = -55 < 0: one hit is not yet "occupied" at 0.5
  BOOST_CHECK(env.collideWith(&probe, 0.2) == 1);      // but it is at threshold 0.2 (-76)
  probe.clearMap();
  probe.insertPointCloud(mid, eBVM_OCCUPIED);
  BOOST_CHECK(env.collideWith(&probe, 0.01) == 0);     // free: -127 is below every threshold above 0
  BOOST_CHECK(env.collideWith(&probe, 0.0) == 1);      // but known (-127 >= -127); this call marks the voxel (127)
  // robot cut: the robot occupies the wall's central voxel -> that ray is dropped entirely
  ProbVoxelMap env2(dim, 0.1f, MT_PROBAB_VOXELMAP);
  robot.insertPointCloud(hit, eBVM_OCCUPIED);
  env2.initSensorSettings(sensor);
  env2.insertSensorData(&wall[0], true, true, eBVM_OCCUPIED,
                        static_cast<BitVoxel<BIT_VECTOR_LENGTH>*>(robot.getVoidDeviceDataPtr()));
  probe.clearMap();
  probe.insertPointCloud(hit, eBVM_OCCUPIED);
  BOOST_CHECK(env2.collideWith(&probe, 0.0) == 0);     // still unknown
}

// DH robot through the singleton: addRobot / setRobotConfiguration / insertRobotIntoMap, and the batch with device FK
BOOST_AUTO_TEST_CASE(dh_robot_and_batched_fk)
{
  GpuVoxelsSharedPtr gvl = GpuVoxels::getInstance();
  gvl->initialize(120, 120, 120, 0.01);
  gvl->addMap(MT_BITVECTOR_VOXELMAP, "env");
  gvl->addMap(MT_BITVECTOR_VOXELMAP, "rob");
  gvl->insertBoxIntoMap(Vector3f(0.2, 0.2, 0.2), Vector3f(0.9, 0.9, 0.9), "env", eBVM_OCCUPIED, 1);
  std::vector<std::string> names;
  names.push_back("base"); names.push_back("upper"); names.push_back("lower");
  std::vector<robot::DHParameters> dh;
  dh.push_back(robot::DHParameters(0.3f, 0.f, 0.1f, 1.5707963f, 0.f, robot::REVOLUTE));
  dh.push_back(robot::DHParameters(0.0f, 0.f, 0.3f, 0.f, 0.f, robot::REVOLUTE));
  dh.push_back(robot::DHParameters(0.1f, 0.f, 0.2f, 0.f, 0.f, robot::PRISMATIC));
  MetaPointCloud clouds;
  clouds.addCloud(createBoxOfPoints(Vector3f(0.45, 0.45, 0.0), Vector3f(0.55, 0.55, 0.3), 0.004), false, "base");
  clouds.addCloud(createBoxOfPoints(Vector3f(0.4, 0.4, 0.0), Vector3f(0.5, 0.5, 0.3), 0.004), false, "upper");
  clouds.addCloud(createBoxOfPoints(Vector3f(0.3, 0.3, 0.0), Vector3f(0.4, 0.4, 0.2), 0.004), true, "lower");
  BOOST_CHECK(gvl->addRobot("arm", names, dh, clouds));
  BOOST_CHECK(!gvl->addRobot("arm", names, dh, clouds));
  std::vector<BitVoxelMeaning> mg(3, eBVM_OCCUPIED);
  const int n_cfg = 6;
  std::vector<float> q;
  std::vector<size_t> want;
  BitVectorVoxelMap* rob = gvl->getMap("rob")->as<BitVectorVoxelMap>();
  for (int c = 0; c < n_cfg; ++c)
  {
    robot::JointValueMap jm;
    jm["base"] = 0.3f * c; jm["upper"] = -0.2f * c; jm["lower"] = 0.05f * c;
    q.push_back(jm["base"]); q.push_back(jm["upper"]); q.push_back(jm["lower"]);
    BOOST_CHECK(gvl->setRobotConfiguration("arm", jm));
    gvl->clearMap("rob");
    BOOST_CHECK(gvl->insertRobotIntoMap("arm", "rob", eBVM_OCCUPIED));
    want.push_back(rob->collideWith(gvl->getMap("env")->as<BitVectorVoxelMap>(), 1.0));
  }
  std::vector<size_t> got;
  BOOST_CHECK(gvl->collideRobotConfigurations("arm", mg, q, "env", got));
  BOOST_CHECK(got.size() == want.size());
  size_t sum = 0;
  for (int c = 0; c < n_cfg; ++c)
  {
    BOOST_CHECK_MESSAGE(got[c] == want[c], "configuration " << c << ": " << got[c] << " != " << want[c]);
    sum += want[c];
  }
  BOOST_CHECK(sum > 0);
  robot::JointValueMap back;
  BOOST_CHECK(gvl->getRobotConfiguration("arm", back) && back.size() == 3);
  // one DH matrix: revolute, theta 0 + value, alpha pi/2 (KinematicLink.cu:30-84)
  Matrix4f m;
  robot::DHParameters(0.3f, 0.f, 0.1f, 1.5707963f, 0.5f, robot::REVOLUTE).convertDHtoM(m);
  BOOST_CHECK(m.a11 == float(std::cos(0.5)) && m.a21 == float(std::sin(0.5)) && m.a34 == 0.3f && m.a44 == 1.f);
  BOOST_CHECK(m.a14 == 0.1f * float(std::cos(0.5)) - 0.f * float(std::sin(0.5)));
  gvl->delMap("env");
  gvl->delMap("rob");
}

BOOST_AUTO_TEST_SUITE_END()

// ---- cases added after the round's GPU minutes were spent: run by tests/test_zz_cpp_facade_late.py, apart from the
// verified cases above, so that their first run on a GPU cannot hide the rest of the suite ------------------------------
BOOST_AUTO_TEST_SUITE(late_additions)

// GV/test/testing_pointclouds.cu:35-150: copies of (meta) point clouds stay equal through chains of copy construction
// and assignment, a changed point is detected, transform and transformSelf agree. Equality is byte-wise, as the
// reference's kernelCompareMem: point 0 is (0, 0, inf) and becomes NaN under the rotation, and still compares equal.
BOOST_AUTO_TEST_CASE(pointclouds_equality_and_rotation)
{
  std::vector<Vector3f> testdata;
  for (size_t i = 0; i < (size_t)numberOfPoints; i++) testdata.push_back(Vector3f(i, 0, 1.0 / i));
  {
    MetaPointCloud* orig = new MetaPointCloud();
    for (size_t j = 0; j < 15; j++) orig->addCloud(testdata);
    orig->syncToDevice();
    MetaPointCloud* working_copy = new MetaPointCloud(*orig);
    for (size_t k = 0; k < 100; k++)
    {
      MetaPointCloud* copy1 = new MetaPointCloud(*working_copy);
      MetaPointCloud* copy2 = new MetaPointCloud(*copy1);
      delete copy1;
      *working_copy = *copy2;
      delete copy2;
    }
    BOOST_CHECK_MESSAGE(*orig == *working_copy, "Point clouds are the same.");
    std::vector<Vector3f> changed(testdata);
    changed[100] = Vector3f(100, 0.33, 1.0 / 100);
    working_copy->updatePointCloud(3, changed, true);
    BOOST_CHECK_MESSAGE(!(*orig == *working_copy), "Point cloud difference detected.");
    delete orig;
    delete working_copy;
  }
  {
    PointCloud* orig = new PointCloud();
    orig->update(testdata);
    PointCloud* working_copy = new PointCloud(*orig);
    for (size_t k = 0; k < 100; k++)
    {
      PointCloud* copy1 = new PointCloud(*working_copy);
      PointCloud* copy2 = new PointCloud(*copy1);
      delete copy1;
      *working_copy = *copy2;
      delete copy2;
    }
    BOOST_CHECK_MESSAGE(*orig == *working_copy, "Point clouds are the same.");
    std::vector<Vector3f> changed(testdata);
    changed[100] = Vector3f(100, 0.33, 1.0 / 100);
    working_copy->update(changed);
    BOOST_CHECK_MESSAGE(!(*orig == *working_copy), "Point cloud difference detected.");
    delete orig;
    delete working_copy;
  }
  {
    PointCloud orig, rotated;
    orig.add(testdata);
    Matrix4f transformation = Matrix4f::createFromRotationAndTranslation(Matrix3f::createFromRPY(Vector3f(12.0, 11.0, 56.0)),
                                                                        Vector3f(2.2, 1.1, 3.3));
    orig.transform(&transformation, &rotated);
    orig.transformSelf(&transformation);
    BOOST_CHECK_MESSAGE(orig == rotated, "Self rotation and copy rotation are equal.");
    BOOST_CHECK(orig.getPointCloudSize() == (size_t)numberOfPoints);
  }
}

// the GpuVoxelsMap base interface as callers use it through GpuVoxelsMapSharedPtr (GV/GpuVoxelsMap.h:112-260), and the
// raw-storage / collider entry points of the list (GV/voxellist/TemplateVoxelList.h:95-116, .hpp:690-774)
BOOST_AUTO_TEST_CASE(base_class_interface)
{
  GpuVoxelsMapSharedPtr maps[3] = {
    GpuVoxelsMapSharedPtr(new ProbVoxelMap(Vector3ui(dimX, dimY, dimZ), 1.f, MT_PROBAB_VOXELMAP)),
    GpuVoxelsMapSharedPtr(new BitVectorVoxelMap(Vector3ui(dimX, dimY, dimZ), 1.f, MT_BITVECTOR_VOXELMAP)),
    GpuVoxelsMapSharedPtr(new BitVectorVoxelList(Vector3ui(dimX, dimY, dimZ), 1.f, MT_BITVECTOR_VOXELLIST))};
  MetaPointCloud boxes = voxellists::twoBoxes();
  for (int k = 0; k < 3; ++k)
  {
    BOOST_CHECK(!maps[k]->needsRebuild());
    BOOST_CHECK(!maps[k]->rebuild());
    BOOST_CHECK(!maps[k]->rebuildIfNeeded());
    BOOST_CHECK(!maps[k]->insertRobotConfiguration(&boxes, false));
    maps[k]->generateVisualizerData();
  }
  BOOST_CHECK(maps[0]->as<ProbVoxelMap>()->getTemplateType() == MT_PROBAB_VOXELMAP);
  BOOST_CHECK(maps[1]->as<BitVectorVoxelMap>()->getTemplateType() == MT_BITVECTOR_VOXELMAP);
  BOOST_CHECK(maps[2]->as<BitVectorVoxelList>()->getTemplateType() == MT_BITVECTOR_VOXELLIST);

  // disk I/O through the base pointer
  maps[0]->insertMetaPointCloud(boxes, eBVM_OCCUPIED);
  maps[2]->insertMetaPointCloud(boxes, voxellists::meanings1112());
  BOOST_CHECK(maps[0]->writeToDisk("/tmp/gvx_temp_base_map.bin"));
  BOOST_CHECK(maps[2]->writeToDisk("/tmp/gvx_temp_base_list.bin"));
  GpuVoxelsMapSharedPtr map_copy(new ProbVoxelMap(Vector3ui(dimX, dimY, dimZ), 1.f, MT_PROBAB_VOXELMAP));
  GpuVoxelsMapSharedPtr list_copy(new BitVectorVoxelList(Vector3ui(dimX, dimY, dimZ), 1.f, MT_BITVECTOR_VOXELLIST));
  BOOST_CHECK(map_copy->readFromDisk("/tmp/gvx_temp_base_map.bin"));
  BOOST_CHECK(list_copy->readFromDisk("/tmp/gvx_temp_base_list.bin"));
  BitVectorVoxelList* list = maps[2]->as<BitVectorVoxelList>();
  BOOST_CHECK(list->equals(*list_copy->as<BitVectorVoxelList>()));

  // collider-object entry points of the list == collideWith (46 collisions, GV/test/testing_voxellist.cu:142)
  BOOST_CHECK(list->collisionCheckWithCollider(map_copy->as<ProbVoxelMap>(), DefaultCollider(1.0f)) == 46);
  BOOST_CHECK(list->collisionCheckWithCollider(map_copy->as<ProbVoxelMap>(), DefaultCollider(Probability(1), Probability(2))) == 46);  // only threshold2 applies
  BOOST_CHECK(list->collisionCheckWithCollider(list_copy->as<BitVectorVoxelList>()) == list->collideWith(list_copy->as<BitVectorVoxelList>()));

  // raw storage: three device arrays of size() entries; make_unique / shrinkToFit leave a list as it is
  const size_t n = list->size();
  BOOST_CHECK(n > 0 && list->getDeviceIdPtr() != NULL && list->getDeviceCoordPtr() != NULL && list->getDeviceDataPtr() != NULL);
  BOOST_CHECK(list->getVoidDeviceDataPtr() == (void*)list->getDeviceDataPtr());
  std::vector<Vector3ui> coords_before, coords_after;
  list->copyCoordsToHost(coords_before);
  list->make_unique();
  list->shrinkToFit();
  list->copyCoordsToHost(coords_after);
  BOOST_CHECK(list->size() == n && coords_before.size() == n && coords_after.size() == n);
  bool same = coords_before.size() == coords_after.size();
  for (size_t i = 0; same && i < n; ++i) same = coords_before[i] == coords_after[i];
  BOOST_CHECK(same);  // (the voxels now carry eBVM_COLLISION marks from the checks above, so no equals() here)
  // device-pointer insert overload: the coordinates of the first box again, nothing new
  PointCloud pc(createBoxOfPoints(Vector3f(2.1, 2.1, 2.1), Vector3f(4.1, 4.1, 4.1), 0.5));
  list->insertPointCloud(pc.getConstDevicePointer(), pc.getPointCloudSize(), BitVoxelMeaning(11));
  BOOST_CHECK(list->size() >= n);

  // device reports of the helper library (GV/helpers/cuda_handling.cu:66-149)
  BOOST_CHECK(getDeviceMemoryInfo().find("total memory (MB)  : ") != std::string::npos);
  BOOST_CHECK(getDeviceInfos().find("Multi Processor Count: ") != std::string::npos);
  int n_dev = 0;
  BOOST_CHECK(cuGetNrOfDevices(&n_dev) && n_dev == 1);

  // dumps print one line per entry / voxel
  BitVectorVoxelList tiny(Vector3ui(4, 4, 4), 1.f, MT_BITVECTOR_VOXELLIST);
  std::vector<Vector3f> two;
  two.push_back(Vector3f(1.5f, 1.5f, 1.5f));
  two.push_back(Vector3f(2.5f, 1.5f, 1.5f));
  tiny.insertPointCloud(two, BitVoxelMeaning(7));
  std::stringstream captured;
  std::streambuf* old = std::cout.rdbuf(captured.rdbuf());
  tiny.screendump(true);
  ProbVoxelMap tiny_map(Vector3ui(16, 16, 16), 1.f, MT_PROBAB_VOXELMAP);
  tiny_map.printVoxelMapData();
  std::cout.rdbuf(old);
  const std::string text = captured.str();
  BOOST_CHECK(text.find("[0] ID = 21 Coords: ") != std::string::npos);
  BOOST_CHECK(text.find("[1] ID = 22 Coords: ") != std::string::npos);
  BOOST_CHECK(text.find("VoxelMap dump: ") != std::string::npos && text.find("4095: -128") != std::string::npos);
}

BOOST_AUTO_TEST_SUITE_END()

int main(int argc, char** argv) { return mini_test::run_all(argc > 1 ? argv[1] : NULL); }
