// tests/cpp/lanes_example.cpp -- planner threads checking collisions at the same time.
// Part 1, C ABI: T threads, each with a collision lane of its own, check T different robot maps against ONE environment
// (read-only, so they may share it); every result must equal the serial answer, and the wall time per check is printed
// beside the same number of checks issued serially from one thread.
// Part 2, facade: worker threads call map->collisionCheck(other) on disjoint map pairs -- the case the reference's
// per-map m_mutex lets run concurrently (GV/voxelmap/TemplateVoxelMap.hpp:329-414) -- and get lanes behind the scenes.
// tests/test_gpu_lanes.py runs it.
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>

#include <gpu_voxels/GpuVoxels.h>

using namespace gpu_voxels;

static void need(int rc, const char* what)
{
  if (rc != GVX_OK) { std::fprintf(stderr, "%s: %s\n", what, gvx_last_error()); std::exit(1); }
}

static std::vector<Vector3f> cloud(uint32_t seed, size_t n, float lo, float hi)
{
  std::vector<Vector3f> v(n);
  uint32_t s = seed * 2654435761u + 12345u;
  for (size_t i = 0; i < n; ++i)
  {
    float c[3];
    for (int k = 0; k < 3; ++k) { s = s * 1664525u + 1013904223u; c[k] = lo + (hi - lo) * float(s >> 8) / 16777216.0f; }
    v[i] = Vector3f(c[0], c[1], c[2]);
  }
  return v;
}

int main()
{
  const int T = 4, K = 1500;
  const uint32_t D = 256;
  const float side = 0.01f;
  GpuVoxelsSharedPtr gvl = GpuVoxels::getInstance();
  gvl->initialize(D, D, D, side);

  // ---- part 1: C ABI
  gvl->addMap(MT_PROBAB_VOXELMAP, "env");
  voxelmap::ProbVoxelMap* env = gvl->getMap("env")->as<voxelmap::ProbVoxelMap>();
  env->insertPointCloud(cloud(1, 300000, 0.0f, 2.56f), eBVM_OCCUPIED);
  std::vector<voxelmap::BitVectorVoxelMap*> robots;
  std::vector<uint64_t> want(T);
  for (int t = 0; t < T; ++t)
  {
    const std::string name = "robot" + std::to_string(t);
    gvl->addMap(MT_BITVECTOR_VOXELMAP, name);
    robots.push_back(gvl->getMap(name)->as<voxelmap::BitVectorVoxelMap>());
    robots[t]->insertPointCloud(cloud(10 + t, 40000 + 7000 * t, 0.3f * t, 1.2f + 0.3f * t), BitVoxelMeaning(20 + t));
    need(gvx_collide_map_map(robots[t]->getHandle(), env->getHandle(), 1.0f, NULL, 0, &want[t]), "serial collide");
    if (want[t] == 0) { std::fprintf(stderr, "robot %d does not collide: bad test data\n", t); return 1; }
  }
  // serial: T*K checks from this thread (after a warm-up that also builds the coarse masks)
  uint64_t n = 0;
  for (int k = 0; k < 50; ++k) for (int t = 0; t < T; ++t) gvx_collide_map_map(robots[t]->getHandle(), env->getHandle(), 1.0f, NULL, 0, &n);
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  long mismatches = 0;
  for (int k = 0; k < K; ++k)
    for (int t = 0; t < T; ++t)
    {
      need(gvx_collide_map_map(robots[t]->getHandle(), env->getHandle(), 1.0f, NULL, 0, &n), "serial collide");
      mismatches += n != want[t];
    }
  const double serial_us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count() / (T * K);
  // lanes: T threads x K checks
  std::vector<gvx_handle_t> lanes(T);
  for (int t = 0; t < T; ++t) need(gvx_lane_create(&lanes[t]), "gvx_lane_create");
  std::atomic<long> bad(0);
  std::atomic<int> ready(0);
  std::atomic<bool> go(false);
  std::vector<std::thread> th;
  for (int t = 0; t < T; ++t)
    th.emplace_back([&, t]() {
      uint64_t c = 0;
      for (int k = 0; k < 50; ++k) gvx_lane_collide_map_map(lanes[t], robots[t]->getHandle(), env->getHandle(), 1.0f, NULL, &c);
      ++ready;
      while (!go.load()) std::this_thread::yield();
      for (int k = 0; k < K; ++k)
      {
        if (gvx_lane_collide_map_map(lanes[t], robots[t]->getHandle(), env->getHandle(), 1.0f, NULL, &c) != GVX_OK || c != want[t]) ++bad;
        if ((k & 255) == 255)
        {
          int flag = 0;
          if (gvx_lane_collide_map_map_any(lanes[t], robots[t]->getHandle(), env->getHandle(), 1.0f, &flag) != GVX_OK || !flag) ++bad;
        }
      }
    });
  while (ready.load() < T) std::this_thread::yield();
  t0 = std::chrono::steady_clock::now();
  go.store(true);
  for (size_t i = 0; i < th.size(); ++i) th[i].join();
  const double lanes_us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count() / (T * (K + K / 256));
  for (int t = 0; t < T; ++t) need(gvx_lane_destroy(lanes[t]), "gvx_lane_destroy");
  std::printf("threads %d\nchecks_per_thread %d\nmismatches %ld\nserial_us_per_check %.3f\nlanes_us_per_check %.3f\n", T, K,
              mismatches + bad.load(), serial_us, lanes_us);

  // ---- part 2: the facade from worker threads, disjoint map pairs
  std::vector<voxelmap::ProbVoxelMap*> envs;
  std::vector<int> expect(T);
  for (int t = 0; t < T; ++t)
  {
    const std::string name = "env" + std::to_string(t);
    gvl->addMap(MT_PROBAB_VOXELMAP, name);
    envs.push_back(gvl->getMap(name)->as<voxelmap::ProbVoxelMap>());
    if (t & 1) envs[t]->insertPointCloud(cloud(10 + t, 40000 + 7000 * t, 0.3f * t, 1.2f + 0.3f * t), eBVM_OCCUPIED);  // same points as the robot
    else envs[t]->insertPointCloud(cloud(77 + t, 1000, 2.4f, 2.5f), eBVM_OCCUPIED);                                  // far from every robot
    expect[t] = robots[t]->collisionCheck(envs[t]) ? 1 : 0;   // first thread: the serial path
    if (expect[t] != (t & 1)) { std::fprintf(stderr, "facade pair %d: unexpected serial answer\n", t); return 1; }
  }
  std::atomic<long> fbad(0);
  std::atomic<int> max_lanes(0);
  std::vector<std::thread> fth;
  for (int t = 0; t < T; ++t)
    fth.emplace_back([&, t]() {
      for (int k = 0; k < 300; ++k)
      {
        if (int(robots[t]->collisionCheck(envs[t])) != expect[t]) ++fbad;
        if (k == 150)
        {
          uint32_t nl = 0;
          gvx_lane_count(&nl);
          int cur = max_lanes.load();
          while (int(nl) > cur && !max_lanes.compare_exchange_weak(cur, int(nl))) {}
        }
      }
    });
  for (size_t i = 0; i < fth.size(); ++i) fth[i].join();
  uint32_t left = 99;
  gvx_lane_count(&left);
  std::printf("facade_mismatches %ld\nfacade_lanes %d\nlanes_after_threads %u\n", fbad.load(), max_lanes.load(), left);
  return (mismatches + bad.load() + fbad.load()) == 0 && left == 0 ? 0 : 1;
}
