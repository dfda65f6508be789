// tests/cpp/host_helpers.cpp -- host-only part of the C++ facade (no GPU, no gvx_init): prints what
// gpu_voxels::computeLinearLoad / interpolateLinear (include/gpu_voxels/helpers/MathHelpers.h) return for the cases on
// its command line, one per line, floats as raw bit patterns; tests/test_host_helpers.py compares the lines with
// tests/golden/math_helpers.npz (produced by the reference's MathHelpers.cpp compiled for the host).
//   host_helpers load <n>...            -> "<blocks> <threads>"
//   host_helpers lerpf <a> <b> <r> ...  (uint32 bit patterns)  -> "<bits>"
//   host_helpers lerpd <a> <b> <r> ...  (uint64 bit patterns)  -> "<bits>"
//   host_helpers geom <kind> <bits>...  -> binary: rotation matrices / generated clouds (tests/golden/geometry.npz)
//   host_helpers prob                   -> 256 lines of updateOccupancy results, one line of insert results, "ok"
#include <cinttypes>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <sstream>
#include <vector>

#include <icl_core_config/Config.h>
#include <gpu_voxels/helpers/BoostCompat.h>
#include <gpu_voxels/helpers/GeometryGeneration.h>
#include <gpu_voxels/helpers/MathHelpers.h>
#include <gpu_voxels/voxel/BitVoxel.h>
#include <gpu_voxels/voxel/DefaultCollider.h>
#include <gpu_voxels/voxel/ProbabilisticVoxel.h>
#include <gpu_voxels/voxel/SVCollider.h>

template <class T, class U> static T bits_to(U u) { T t; static_assert(sizeof(T) == sizeof(U), "size"); std::memcpy(&t, &u, sizeof t); return t; }

int main(int argc, char** argv)
{
  if (argc < 2) return 2;
  const std::string mode = argv[1];
  if (mode == "load")
    for (int i = 2; i < argc; ++i)
    {
      uint32_t blocks = 0, threads = 0;
      gpu_voxels::computeLinearLoad(uint32_t(std::strtoul(argv[i], NULL, 10)), &blocks, &threads);
      std::printf("%u %u\n", blocks, threads);
    }
  else if (mode == "lerpf")
    for (int i = 2; i + 2 < argc; i += 3)
    {
      const float a = bits_to<float>(uint32_t(std::strtoul(argv[i], NULL, 10)));
      const float b = bits_to<float>(uint32_t(std::strtoul(argv[i + 1], NULL, 10)));
      const float r = bits_to<float>(uint32_t(std::strtoul(argv[i + 2], NULL, 10)));
      std::printf("%u\n", bits_to<uint32_t>(gpu_voxels::interpolateLinear(a, b, r)));
    }
  else if (mode == "lerpd")
    for (int i = 2; i + 2 < argc; i += 3)
    {
      const double a = bits_to<double>(uint64_t(std::strtoull(argv[i], NULL, 10)));
      const double b = bits_to<double>(uint64_t(std::strtoull(argv[i + 1], NULL, 10)));
      const double r = bits_to<double>(uint64_t(std::strtoull(argv[i + 2], NULL, 10)));
      std::printf("%" PRIu64 "\n", bits_to<uint64_t>(gpu_voxels::interpolateLinear(a, b, r)));
    }
  else if (mode == "prob")
  {  // ProbabilisticVoxel: updateOccupancy for every (occupancy, delta) pair, then insert for every meaning
    for (int occ = -128; occ <= 127; ++occ)
    {
      for (int d = -128; d <= 127; ++d)
      {
        gpu_voxels::ProbabilisticVoxel v;
        v.occupancy() = gpu_voxels::Probability(occ);
        std::printf("%d ", int(v.updateOccupancy(gpu_voxels::Probability(d))));
      }
      std::printf("\n");
    }
    for (int m = 0; m < 256; ++m)
    {
      gpu_voxels::ProbabilisticVoxel v;
      v.occupancy() = 5;
      v.insert(gpu_voxels::BitVoxelMeaning(m));
      std::printf("%d ", int(v.getOccupancy()));
    }
    gpu_voxels::ProbabilisticVoxel u, w;
    w.occupancy() = 100;
    const bool ok = u.getOccupancy() == -128 && !u.isOccupied(-127.5f) && u.isOccupied(-128.f) && w.isOccupied(99.5f) &&
                    !w.isOccupied(100.5f) && gpu_voxels::ProbabilisticVoxel::reduce(w, w).getOccupancy() == 127;
    std::printf("\n%s\n", ok ? "ok" : "MISMATCH");
  }
  else if (mode == "sv")
  {  // lines of hex on stdin: "m <v1> <v2> <coll_in> <margin> <sv_offset>" -> "<result> <coll_out>",
     //                        "s <v> <shift>" -> "<shifted>"
    char kind;
    char h1[80], h2[80], h3[80];
    unsigned p1, p2;
    auto parse = [](const char* hex, gpu_voxels::BitVector<256>& v) {
      for (int k = 0; k < 32; ++k) { unsigned byte; std::sscanf(hex + 2 * k, "%2x", &byte); v.bytes()[k] = uint8_t(byte); }
    };
    auto print = [](const gpu_voxels::BitVector<256>& v) { for (int k = 0; k < 32; ++k) std::printf("%02x", v.bytes()[k]); };
    while (std::scanf(" %c", &kind) == 1)
    {
      if (kind == 'm')
      {
        if (std::scanf("%79s %79s %79s %u %u", h1, h2, h3, &p1, &p2) != 5) return 3;
        gpu_voxels::BitVector<256> v1, v2, coll;
        parse(h1, v1); parse(h2, v2); parse(h3, coll);
        const bool r = gpu_voxels::bitMarginCollisionCheck(v1, v2, &coll, uint8_t(p1), p2);
        std::printf("%d ", r ? 1 : 0); print(coll); std::printf("\n");
      }
      else
      {
        if (std::scanf("%79s %u", h1, &p1) != 2) return 3;
        gpu_voxels::BitVector<256> v;
        parse(h1, v);
        gpu_voxels::performLeftShift(v, uint8_t(p1));
        print(v); std::printf("\n");
      }
    }
  }
  else if (mode == "f2p")
  {  // DefaultCollider::floatToProbability of float bit patterns, and the float the facade hands to the C ABI for a
     // collider built from that Probability: "<probability> <round trip ok>"
    for (int i = 2; i < argc; ++i)
    {
      const float thr = bits_to<float>(uint32_t(std::strtoul(argv[i], NULL, 10)));
      const gpu_voxels::Probability p = gpu_voxels::DefaultCollider::floatToProbability(thr);
      float back = 0.f;
      const bool ok = gpu_voxels::DefaultCollider(p, p).asFloatThreshold(&back) &&
                      gpu_voxels::DefaultCollider::floatToProbability(back) == p &&
                      gpu_voxels::DefaultCollider(thr).getThreshold1() == p;
      std::printf("%d %d\n", int(p), ok ? 1 : 0);
    }
    // predicates on value types
    gpu_voxels::ProbabilisticVoxel lo, hi;
    lo.occupancy() = -5; hi.occupancy() = 5;
    gpu_voxels::BitVectorVoxel freeOnly, occ;
    freeOnly.insert(gpu_voxels::eBVM_FREE); occ.insert(gpu_voxels::BitVoxelMeaning(77));
    const gpu_voxels::DefaultCollider half(0.5f);  // threshold 0
    const bool ok = !half.collide(lo, hi) && half.collide(hi, hi) && half.collide(hi) && !half.collide(lo) &&
                    half.collide(hi, occ) && !half.collide(hi, freeOnly) && !half.collide(lo, occ) &&
                    half.collide(occ, hi) && !half.collide(freeOnly, hi) && half.collide(occ, occ) &&
                    !half.collide(occ, freeOnly) && !gpu_voxels::DefaultCollider(gpu_voxels::Probability(1), gpu_voxels::Probability(9)).asFloatThreshold(NULL);
    // SVCollider (GV/voxel/SVCollider.hpp:54-116): the second Prob voxel is compared STRICTLY; the types variants
    // OR the colliding voxel's meanings into the output vector; Bit x Bit is the windowed ID test
    const gpu_voxels::SVCollider sv(0.5f, 1);  // thresholds 0, window 1
    gpu_voxels::ProbabilisticVoxel zero;
    zero.occupancy() = 0;
    gpu_voxels::BitVector<256> seen;
    gpu_voxels::BitVectorVoxel id20, id21, id23;
    id20.insert(gpu_voxels::BitVoxelMeaning(20)); id21.insert(gpu_voxels::BitVoxelMeaning(21)); id23.insert(gpu_voxels::BitVoxelMeaning(23));
    bool ok2 = sv.collide(zero, hi) && !sv.collide(hi, zero) && sv.collide(zero) && sv.collide(hi, occ) && sv.collide(occ, hi) &&
               !sv.collide(freeOnly, hi) && sv.collide(occ, hi, &seen) && seen.getBit(77) && !sv.collide(occ, lo, &seen) &&
               sv.collide(hi, id20, &seen) && seen.getBit(20);
    gpu_voxels::BitVector<256> win;
    ok2 = ok2 && sv.collide(id20, id21, &win) && !sv.collide(id20, id23, &win) && win.isZero() &&
          gpu_voxels::SVCollider(gpu_voxels::Probability(3), gpu_voxels::Probability(3), std::size_t(2)).getWindowSize() == 2;
    std::printf("%s\n", ok && ok2 ? "ok" : "MISMATCH");
  }
  else if (mode == "vec")
  {  // the vector and map overloads against the scalar one
    std::vector<float> a(3), b(3);
    a[0] = 0.1f; a[1] = -2.f; a[2] = 3.5f; b[0] = 1.1f; b[1] = 4.f; b[2] = -3.5f;
    std::vector<float> v = gpu_voxels::interpolateLinear(a, b, 0.3f);
    std::map<std::string, float> ma, mb;
    ma["j0"] = a[0]; ma["j1"] = a[1]; mb["j0"] = b[0]; mb["j1"] = b[1];
    std::map<std::string, float> m = gpu_voxels::interpolateLinear(ma, mb, 0.3f);
    bool ok = v.size() == 3 && m.size() == 2;
    for (int k = 0; k < 3 && ok; ++k) ok = v[k] == gpu_voxels::interpolateLinear(a[k], b[k], 0.3f);
    ok = ok && m["j0"] == v[0] && m["j1"] == v[1];
    std::vector<double> da(1, 0.25), db(1, 0.75);
    ok = ok && gpu_voxels::interpolateLinear(da, db, 0.5)[0] == 0.5;
    std::printf("%s\n", ok ? "ok" : "MISMATCH");
  }
  else if (mode == "geom")
  {  // geom <kind> <float bit patterns...>: binary to stdout (uint64 counts, raw float32), see tests/test_host_helpers.py
    using namespace gpu_voxels;
    using namespace gpu_voxels::geometry_generation;
    if (argc < 3) return 2;
    const std::string kind = argv[2];
    std::vector<float> f;
    for (int i = 3; i < argc; ++i) f.push_back(bits_to<float>(uint32_t(std::strtoul(argv[i], NULL, 10))));
    struct Out
    {
      static void cloud(const std::vector<Vector3f>& c)
      {
        const uint64_t n = c.size();
        std::fwrite(&n, 8, 1, stdout);
        if (n) std::fwrite(&c[0], 12, n, stdout);
      }
    };
    if (kind == "rpy")
      for (size_t i = 0; i + 2 < f.size(); i += 3)
      {
        const Matrix3f m = Matrix3f::createFromRPY(Vector3f(f[i], f[i + 1], f[i + 2]));
        std::fwrite(&m, 36, 1, stdout);
      }
    else if (kind == "box" && f.size() == 7)
      Out::cloud(createBoxOfPoints(Vector3f(f[0], f[1], f[2]), Vector3f(f[3], f[4], f[5]), f[6]));
    else if (kind == "sphere" && f.size() == 5)
      Out::cloud(createSphereOfPoints(Vector3f(f[0], f[1], f[2]), f[3], f[4]));
    else if (kind == "cylinder" && f.size() == 6)
      Out::cloud(createCylinderOfPoints(Vector3f(f[0], f[1], f[2]), f[3], f[4], f[5]));
    else if (kind == "obox" && f.size() == 11)
    {
      OrientedBoxParams p;
      p.dim = Vector3f(f[0], f[1], f[2]); p.center = Vector3f(f[3], f[4], f[5]); p.rot = Vector3f(f[6], f[7], f[8]);
      const Matrix4f m = geometry_generation::detail::orientedBoxTransform(p);
      std::fwrite(&m, 64, 1, stdout);
      Out::cloud(f[10] != 0.f ? geometry_generation::detail::orientedBoxEdgesCloud(p, f[9]) : geometry_generation::detail::orientedBoxCloud(p, f[9]));
    }
    else if (kind == "eq" && argc == 8)
    {  // geom eq <max points> <dx> <dy> <dz> <side bits>: equidistant points, then the two checkerboard clouds
      const size_t max_n = size_t(std::strtoull(argv[3], NULL, 10));
      const Vector3ui dims(uint32_t(std::strtoul(argv[4], NULL, 10)), uint32_t(std::strtoul(argv[5], NULL, 10)),
                           uint32_t(std::strtoul(argv[6], NULL, 10)));
      const float side = bits_to<float>(uint32_t(std::strtoul(argv[7], NULL, 10)));
      std::vector<Vector3f> pts, black, white;
      createEquidistantPointsInBox(max_n, dims, side, pts);
      createNonOverlapping3dCheckerboard(max_n, dims, side, black, white);
      Out::cloud(pts); Out::cloud(black); Out::cloud(white);
    }
    else return 2;
  }
  else if (mode == "bitvector")
  {  // the reference's host-side BitVector tests (GV/test/testing_bitvector.cu:36-194) against the facade's BitVector
    using namespace gpu_voxels;
    typedef BitVector<BIT_VECTOR_LENGTH> BV;
    int failed = 0;
#define GVX_EXPECT(cond) do { if (!(cond)) { ++failed; std::printf("FAILED: %s\n", #cond); } } while (0)
    std::string a_bits(256, '0'), not_a_bits(256, '1');
    const int set_bits[] = {0, 2, 3, 4, 111, 254, 255};  // FREE, COLLISION, UNKNOWN, SV_START, 111, SV_END, UNDEFINED
    for (int k = 0; k < 7; ++k) { a_bits[set_bits[k]] = '1'; not_a_bits[set_bits[k]] = '0'; }
    BV a, same_a, not_a, b, c, tmp;
    { std::stringstream in(a_bits); in >> a; }
    { std::stringstream in(a_bits); in >> same_a; }
    { std::stringstream in(not_a_bits); in >> not_a; }
    { std::stringstream in(a_bits); in >> b; }
    { std::stringstream out; out << a; GVX_EXPECT(out.str() == a_bits); }
    GVX_EXPECT(tmp.isZero()); GVX_EXPECT(!a.isZero()); GVX_EXPECT(a == same_a); GVX_EXPECT(!(not_a == a));
    tmp = a; GVX_EXPECT(a == tmp);
    tmp.clear(); GVX_EXPECT(tmp.isZero());
    tmp = BV(a); GVX_EXPECT(a == tmp);
    c = ~a; GVX_EXPECT(c == not_a);
    GVX_EXPECT((a & not_a).isZero()); GVX_EXPECT((~(a | not_a)).isZero());
    tmp.clear();
    tmp.setBit(eBVM_FREE); tmp.setBit(eBVM_COLLISION); tmp.setBit(eBVM_UNKNOWN); tmp.setBit(eBVM_SWEPT_VOLUME_START);
    tmp.setBit(BitVoxelMeaning(111)); tmp.setBit(eBVM_SWEPT_VOLUME_END); tmp.setBit(eBVM_UNDEFINED);
    GVX_EXPECT(tmp == b);
    for (uint32_t i = 0; i < BIT_VECTOR_LENGTH; ++i)
    {
      bool want = false;
      for (int k = 0; k < 7; ++k) want = want || uint32_t(set_bits[k]) == i;
      GVX_EXPECT(b.getBit(i) == want);
    }
    tmp.clear();
    tmp.setByte(0 * 8, uint8_t(29)); tmp.setByte(13 * 8, uint8_t(128)); tmp.setByte(31 * 8, uint8_t(192));
    GVX_EXPECT(tmp == b);
    GVX_EXPECT(tmp.getByte(13 * 8) == 128);
    b.clearBit(eBVM_COLLISION); GVX_EXPECT(!b.getBit(eBVM_COLLISION));
    a |= not_a; GVX_EXPECT((~a).isZero());
    tmp.clear(); GVX_EXPECT(tmp.noneButEmpty());
    tmp.setBit(eBVM_FREE); GVX_EXPECT(tmp.noneButEmpty());
    tmp.setBit(123); GVX_EXPECT(!tmp.noneButEmpty());
    for (uint32_t shift = 3; shift < 20; shift += 10)  // bitvector_bitshift, :139-194
    {
      BV v;
      const uint32_t bits[] = {0, 1, 2, 3, 4, 14, 24, 254, 255};
      for (int k = 0; k < 9; ++k) v.setBit(bits[k]);
      performLeftShift(v, uint8_t(shift));
      for (int k = 0; k < 9; ++k) GVX_EXPECT(!v.getBit(bits[k]));
      GVX_EXPECT(v.getBit(24 - shift) == (24 - shift > 3));
      GVX_EXPECT(v.getBit(254 - shift)); GVX_EXPECT(v.getBit(255 - shift));
    }
#undef GVX_EXPECT
    std::printf("%s\n", failed ? "MISMATCH" : "ok");
  }
  else if (mode == "compat")
  {  // the icl_core option stand-in and the boost aliases: host_helpers compat [--device-identifier=X | -id X] [--flag]
    icl_core::config::GetoptParameter ident("device-identifier:", "id", "Identifer of the kinect device");
    icl_core::config::addParameter(ident);
    icl_core::config::addParameter(icl_core::config::GetoptParameter("flag", "f", "a switch"));
    icl_core::logging::initialize(argc, argv);
    const std::string id = icl_core::config::Getopt::instance().paramOpt("device-identifier");
    boost::shared_ptr<std::string> p = boost::make_shared<std::string>(id);
    std::printf("%s|%d|%d|%d\n", p->c_str(), int(icl_core::config::Getopt::instance().paramOptPresent("device-identifier")),
                int(icl_core::config::Getopt::instance().paramOptPresent("flag")),
                int(icl_core::config::Getopt::instance().paramOptPresent("unknown")));
  }
  else if (mode == "cudamath")
  {  // the reference's host-side matrix tests (GV/test/testing_cudaMath.cu:33-195) against the facade's types
    using namespace gpu_voxels;
    int failed = 0;
#define GVX_EXPECT(cond) do { if (!(cond)) { ++failed; std::printf("FAILED: %s\n", #cond); } } while (0)
    const Matrix4f matrix(0.9751700, -0.218711, -0.0347626, 10, 0.1976770, 0.930432, -0.3085770, 11,
                          0.0998334, 0.294044, 0.9505640, 12, 0, 0, 0, 1);
    GVX_EXPECT(matrix == matrix);
    { Matrix4f a = matrix, b = matrix; b.a13 += 0.0001; GVX_EXPECT(!(a == b)); }
    { Matrix4f a = matrix, b = matrix; b.a22 -= 0.00001; GVX_EXPECT(a.apprx_equal(b, 0.000011)); }
    { Matrix4f a = matrix, b = matrix; b.a22 += 0.00001; GVX_EXPECT(a.apprx_equal(b, 0.000011)); }
    { Matrix4f a = matrix, b = matrix; b.a22 += 0.000013; GVX_EXPECT(!a.apprx_equal(b, 0.000011)); }
    GVX_EXPECT(matrix.transpose() == Matrix4f(0.9751700, 0.1976770, 0.0998334, 0, -0.218711, 0.930432, 0.294044, 0,
                                              -0.0347626, -0.3085770, 0.9505640, 0, 10, 11, 12, 1));
    {
      const Matrix4f a(1, 2, 3, 4, 4, 3, 2, 1, 2, 1, 3, 4, 2, 3, 1, 4), b(7, 4, 5, 6, 6, 5, 7, 4, 4, 6, 5, 7, 4, 5, 6, 7);
      GVX_EXPECT(a * b == Matrix4f(47, 52, 58, 63, 58, 48, 57, 57, 48, 51, 56, 65, 52, 49, 60, 59));
    }
    {
      Matrix4f inverse;
      GVX_EXPECT(matrix.invertMatrix(inverse));
      GVX_EXPECT(Matrix4f::createIdentity().apprx_equal(matrix * inverse, 0.00000011));
    }
    {  // matrix_rpy: both decompositions of 1000 random rotations rebuild the rotation (difference < 0.005)
      uint64_t state = 88172645463325252ull;  // xorshift64: any fixed sequence of angles in [-2 pi, 2 pi) will do
      for (int n = 0; n < 1000; ++n)
      {
        float ang[3];
        for (int k = 0; k < 3; ++k)
        {
          state ^= state << 13; state ^= state >> 7; state ^= state << 17;
          ang[k] = float((double(state >> 11) / 9007199254740992.0 * 4.0 - 2.0) * M_PI);
        }
        const Matrix3f a = Matrix3f::createFromRPY(Vector3f(ang[0], ang[1], ang[2]));
        const Matrix3f b = Matrix3f::createFromRPY(a.toRPY(0)), c = Matrix3f::createFromRPY(a.toRPY(1));
        GVX_EXPECT(a.orientationMatrixDiff(b).apprx_equal(Vector3f(0.0f), 0.005));
        GVX_EXPECT(a.orientationMatrixDiff(c).apprx_equal(Vector3f(0.0f), 0.005));
      }
    }
    { Matrix4f z, mz; mz.a12 = -0.0f; GVX_EXPECT(z == mz); z.a33 = mz.a33 = std::nanf(""); GVX_EXPECT(!(z == mz)); }
#undef GVX_EXPECT
    std::printf("%s\n", failed ? "MISMATCH" : "ok");
  }
  else if (mode == "mathfile")
  {  // mathfile <kind> <path>: raw float32 records in, raw float32 results out (value API of Vector3f / Matrix3f /
     // Matrix4f against tests/golden/geometry.npz; record layouts in tests/test_host_helpers.py)
    using namespace gpu_voxels;
    if (argc < 4) return 2;
    const std::string kind = argv[2];
    std::vector<float> in;
    {
      FILE* fp = std::fopen(argv[3], "rb");
      if (!fp) return 3;
      float buf[4096];
      size_t n;
      while ((n = std::fread(buf, 4, 4096, fp)) > 0) in.insert(in.end(), buf, buf + n);
      std::fclose(fp);
    }
    struct IO
    {
      static Matrix3f m3(const float* p) { Matrix3f m; std::memcpy(static_cast<void*>(&m), p, 36); return m; }
      static Matrix4f m4(const float* p) { Matrix4f m; std::memcpy(static_cast<void*>(&m), p, 64); return m; }
      static void put(const Vector3f& v) { std::fwrite(&v, 12, 1, stdout); }
      static void put(float v) { std::fwrite(&v, 4, 1, stdout); }
    };
    if (kind == "torpy")
      for (size_t i = 0; i + 9 <= in.size(); i += 9)
        for (unsigned sol = 0; sol < 3; ++sol) IO::put(IO::m3(&in[i]).toRPY(sol));
    else if (kind == "diff")
    {
      const size_t n = in.size() / 9;
      for (size_t i = 0; i < n; ++i) IO::put(IO::m3(&in[9 * i]).orientationMatrixDiff(IO::m3(&in[9 * ((i + 1) % n)])));
    }
    else if (kind == "inv3")
      for (size_t i = 0; i + 9 <= in.size(); i += 9)
      {
        Matrix3f inv;
        const bool ok = IO::m3(&in[i]).invertMatrix(inv);
        if (!ok) inv = Matrix3f();
        IO::put(ok ? 1.f : 0.f);
        std::fwrite(&inv, 36, 1, stdout);
      }
    else if (kind == "inv4")
      for (size_t i = 0; i + 16 <= in.size(); i += 16)
      {
        Matrix4f inv;
        const bool ok = IO::m4(&in[i]).invertMatrix(inv);
        if (!ok) inv = Matrix4f();
        IO::put(ok ? 1.f : 0.f);
        std::fwrite(&inv, 64, 1, stdout);
      }
    else if (kind == "vec")
      for (size_t i = 0; i + 22 <= in.size(); i += 22)
      {
        const Vector3f a(in[i], in[i + 1], in[i + 2]), b(in[i + 3], in[i + 4], in[i + 5]);
        IO::put(a.length()); IO::put(a.dot(b)); IO::put(a.angleBetween(b));
        IO::put(a.cross(b)); IO::put(a.normalized()); IO::put(a.abs());
        Vector3f n = a;
        n.normalize();
        IO::put(n);
        IO::put(IO::m4(&in[i + 6]) * a);
        IO::put(Vector3f::compVec(a, b) ? 1.f : 0.f);
        IO::put(a.apprx_equal(b, 1e-3) ? 1.f : 0.f);
      }
    else if (kind == "mat")
      for (size_t i = 0; i + 35 <= in.size(); i += 35)
      {
        const Matrix4f a = IO::m4(&in[i]), b = IO::m4(&in[i + 16]);
        const Matrix4f t = a.transpose(), d = a - b, pr = a * b;
        Matrix4f lm = a;
        lm.leftMultiply(b);
        std::fwrite(&t, 64, 1, stdout); std::fwrite(&d, 64, 1, stdout); std::fwrite(&pr, 64, 1, stdout);
        IO::put(a.getRotation() * Vector3f(in[i + 32], in[i + 33], in[i + 34]));
        IO::put(a.apprx_equal(b, 1e-3) ? 1.f : 0.f);
        Matrix4f back = Matrix4f::createIdentity();
        back.setRotation(a.getRotation());
        back.setTranslation(a.getTranslation());
        const bool affine = a.a41 == 0.f && a.a42 == 0.f && a.a43 == 0.f && a.a44 == 1.f;
        IO::put(lm == pr && (!affine || back == a) ? 1.f : 0.f);
      }
    else if (kind == "fmt")
      for (size_t i = 0; i + 17 <= in.size(); i += 17)
      {
        std::stringstream ss;
        if (in[i] == 0.f) ss << Vector3f(in[i + 1], in[i + 2], in[i + 3]);
        else if (in[i] == 1.f) ss << IO::m3(&in[i + 1]);
        else ss << IO::m4(&in[i + 1]);
        const std::string text = ss.str();
        std::fwrite(text.c_str(), 1, text.size() + 1, stdout);
      }
    else return 2;
  }
  else return 2;
  return 0;
}
