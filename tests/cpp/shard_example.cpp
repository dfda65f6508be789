// tests/cpp/shard_example.cpp -- caller code written against the reference's public API (the shape of
// example_how_to_link/gvl_linkage_test_lib.cpp:20-39 plus the map / list calls of GV/test/testing_voxelmap.cu and
// testing_voxellist.cu), with NOTHING about GPUs, ranks or slabs in it. Run once it prints the single-GPU answers; run N
// times with GVX_SHARD=z GVX_WORLD=N GVX_RANK=r in the environment every copy holds one z-slab of every map and prints
// the SAME lines (tests/test_gpu_shard_api.py compares them; tests/multi_gpu_check job runs it on 2 / 8 GPUs).
#include <cstdio>

#include <gpu_voxels/GpuVoxels.h>

using namespace gpu_voxels;

int main()
{
  GpuVoxelsSharedPtr gvl = GpuVoxels::getInstance();
  gvl->initialize(150, 150, 150, 0.01f);
  gvl->addMap(MT_PROBAB_VOXELMAP, "myFirstMap");
  gvl->addMap(MT_PROBAB_VOXELMAP, "mySecondMap");
  gvl->addMap(MT_BITVECTOR_VOXELMAP, "myBitMap");
  gvl->addMap(MT_BITVECTOR_VOXELLIST, "myList");

  // example_how_to_link: two overlapping boxes ("about 8000 collisions", README.txt:21)
  const Vector3f center(0.5f, 0.5f, 0.5f), box_size(0.4f, 0.4f, 0.4f), overlap(0.1f, 0.1f, 0.1f);
  const Vector3f c1 = center - overlap, c2 = c1 + box_size, c3 = center + overlap, c4 = c3 - box_size;
  gvl->insertBoxIntoMap(c1, c2, "myFirstMap", eBVM_OCCUPIED, 1);
  gvl->insertBoxIntoMap(c4, c3, "mySecondMap", eBVM_OCCUPIED, 1);
  voxelmap::ProbVoxelMap* first = gvl->getMap("myFirstMap")->as<voxelmap::ProbVoxelMap>();
  voxelmap::ProbVoxelMap* second = gvl->getMap("mySecondMap")->as<voxelmap::ProbVoxelMap>();
  std::printf("linkage %zu\n", first->collideWith(second));
  std::printf("linkage_again %zu\n", first->collideWith(second));  // marks of the first call change nothing
  std::printf("check %d\n", int(first->collisionCheck(second)));

  // per-meaning clouds in a BitVectorVoxelMap, collideWithTypes against the ProbVoxelMap
  voxelmap::BitVectorVoxelMap* bits = gvl->getMap("myBitMap")->as<voxelmap::BitVectorVoxelMap>();
  bits->insertPointCloud(geometry_generation::createBoxOfPoints(Vector3f(0.2f, 0.2f, 0.1f), Vector3f(0.6f, 0.6f, 0.9f), 0.01f),
                         BitVoxelMeaning(14));
  bits->insertPointCloud(geometry_generation::createBoxOfPoints(Vector3f(0.5f, 0.5f, 0.4f), Vector3f(0.9f, 0.9f, 1.4f), 0.01f),
                         BitVoxelMeaning(46));
  BitVectorVoxel types;
  std::printf("types %zu", bits->collideWithTypes(second, types, 1.0f));
  for (uint32_t k = 0; k < BIT_VECTOR_LENGTH; ++k) if (types.bitVector().getBit(k)) std::printf(" %u", k);
  std::printf("\n");
  std::printf("bit_x_prob %zu\n", bits->collideWith(first, 1.0f));
  std::printf("prob_x_bit %zu\n", first->collideWith(bits, 1.0f));

  gvl->addMap(MT_BITVECTOR_VOXELMAP, "myBitMap2");
  voxelmap::BitVectorVoxelMap* bits2 = gvl->getMap("myBitMap2")->as<voxelmap::BitVectorVoxelMap>();
  bits2->insertPointCloud(geometry_generation::createBoxOfPoints(Vector3f(0.1f, 0.1f, 0.55f), Vector3f(0.8f, 0.8f, 1.45f), 0.01f),
                          BitVoxelMeaning(33));

  // a voxel list (replicated on every rank) against the maps
  voxellist::BitVectorVoxelList* list = gvl->getMap("myList")->as<voxellist::BitVectorVoxelList>();
  list->insertPointCloud(geometry_generation::createBoxOfPoints(Vector3f(0.3f, 0.3f, 0.05f), Vector3f(0.7f, 0.7f, 1.45f), 0.02f),
                         BitVoxelMeaning(11));
  list->insertPointCloud(geometry_generation::createBoxOfPoints(Vector3f(0.6f, 0.3f, 0.7f), Vector3f(0.8f, 0.7f, 1.2f), 0.02f),
                         BitVoxelMeaning(12));
  std::printf("list_x_prob %zu\n", list->collideWith(second, 1.0f));
  std::printf("list_x_bit %zu\n", list->collideWith(bits, 1.0f));
  BitVectorVoxel ltypes;
  std::printf("list_types %zu", list->collideWithTypes(bits, ltypes, 1.0f));
  for (uint32_t k = 0; k < BIT_VECTOR_LENGTH; ++k) if (ltypes.bitVector().getBit(k)) std::printf(" %u", k);
  std::printf("\n");
  BitVectorVoxel mask;
  mask.bitVector().setBit(12);
  std::printf("list_mask %zu\n", list->collideWithTypeMask(bits, mask, 1.0f));

  // collideWith with an offset: voxel i + offset of this map meets voxel i of the other. Sharded, the partners of a slab's
  // voxels live in other ranks' slabs (z offsets: all of them; x / y offsets: the slab's edge layers) and are read from
  // the peers. 150 layers / 3 ranks = slabs of 50: an offset of 55 layers reaches past the neighbouring slab.
  std::printf("off_z %zu\n", first->collideWith(second, 1.0f, Vector3i(0, 0, 7)));
  std::printf("off_xyz %zu\n", first->collideWith(second, 1.0f, Vector3i(3, -2, -11)));
  std::printf("off_far %zu\n", second->collideWith(first, 1.0f, Vector3i(1, 0, -55)));
  std::printf("off_xy %zu\n", first->collideWith(second, 1.0f, Vector3i(-9, 5, 0)));
  std::printf("off_bit_prob %zu\n", bits->collideWith(first, 1.0f, Vector3i(-5, 4, 23)));
  std::printf("off_prob_bit %zu\n", first->collideWith(bits, 1.0f, Vector3i(2, 1, -37)));
  std::printf("off_bit_bit %zu\n", bits->collideWith(bits2, 1.0f, Vector3i(0, -3, 52)));
  std::printf("off_thresholds %u\n", first->collisionCheckWithCounterRelativeTransform(second, DefaultCollider(Probability(40), Probability(100)), Vector3i(4, 4, -19)));
  std::printf("off_none %zu\n", first->collideWith(second, 1.0f, Vector3i(0, 0, 149)));

  gvl->clearMap("mySecondMap");
  std::printf("after_clear %zu\n", first->collideWith(second));
  int rank = 0, world = 1, fused = 0;
  gvx_shard_info(&rank, &world, &fused);
  std::fprintf(stderr, "rank %d of %d, in-kernel all-reduce %d\n", rank, world, fused);
  return 0;
}
