// gvx_kinematics.cu -- batched forward kinematics of a DH chain: per-link Matrix4f for every configuration in one launch.
//
// Replaces the host loop of KinematicChain::setConfiguration (GV/robot/dh_robot/KinematicChain.cu:93-126), which per
// link calls DHParameters::convertDHtoM (GV/robot/dh_robot/KinematicLink.cu:30-84) and one host
// Matrix4f * Matrix4f (GV/helpers/cuda_matrices.h:772-803), then copies the matrix to the device and launches a
// transform kernel -- per link and per configuration.
//
// Exactness. The reference evaluates sin/cos in DOUBLE (libm, < 1 ulp) on the float angle and rounds to float; every
// later operation is a float multiply or add with its own rounding (host g++, no FMA). Here the products use
// __fmul_rn / __fadd_rn / __fsub_rn in the reference's association order, so they are bit-identical by construction.
// CUDA's double sin/cos are accurate to 2 ulp, i.e. they can differ from libm's by up to 3 ulp of the DOUBLE; after
// rounding to float that is visible only if a float rounding boundary lies between the two (probability ~1e-8 per
// value). The kernel brackets each value by +-|v| * 2^-check_bits (default 2^-50 >= 4 ulp): if both ends round to the
// same float the result is proven equal to the reference's; otherwise the configuration is reported in `ambiguous` and
// the host recomputes that configuration with libm exactly as the reference does (dh_chain_host below).
#include <cmath>
#include "gvx_internal.h"

namespace gvx {

__device__ __forceinline__ float round_checked(double v, double rel, bool& amb) {
  const float r = __double2float_rn(v);
  if (v == v) {
    const double e = fabs(v) * rel;
    amb |= __double2float_rn(v - e) != __double2float_rn(v + e);
  }
  return r;
}

// x * y with the reference's left-to-right sums, one rounding per operation
__device__ __forceinline__ void mat4_mul_rn(const float* x, const float* y, float* r) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float s = __fmul_rn(x[4 * i], y[j]);
      s = __fadd_rn(s, __fmul_rn(x[4 * i + 1], y[4 + j]));
      s = __fadd_rn(s, __fmul_rn(x[4 * i + 2], y[8 + j]));
      r[4 * i + j] = __fadd_rn(s, __fmul_rn(x[4 * i + 3], y[12 + j]));
    }
}

// one thread per configuration; out: n_cfg x out_stride x 16, link l writes slot links[l].out_slot (< 0: the link has
// no cloud, its DH matrix still enters the chain -- KinematicChain.cu:107-112)
__global__ void __launch_bounds__(128)
k_dh_chain(const DhLinkDev* __restrict__ links, uint32_t n_links, const float* __restrict__ joint_values, uint32_t n_cfg,
           float* __restrict__ out, uint32_t out_stride, double rel, uint32_t* __restrict__ ambiguous, uint32_t amb_cap) {
  const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_cfg) return;
  float acc[16] = {1.f, 0.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 0.f, 1.f};
  bool amb = false;
  for (uint32_t l = 0; l < n_links; ++l) {
    const DhLinkDev k = links[l];
    if (k.out_slot >= 0) {
      float4* o = reinterpret_cast<float4*>(out + ((size_t)c * out_stride + k.out_slot) * 16);
      o[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
      o[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
      o[2] = make_float4(acc[8], acc[9], acc[10], acc[11]);
      o[3] = make_float4(acc[12], acc[13], acc[14], acc[15]);
    }
    const float value = joint_values[(size_t)c * n_links + l];
    float d_cur = k.d, theta_cur = k.theta;
    if (k.joint_type == 1) d_cur = __fadd_rn(k.d, value);
    else if (k.joint_type == 0) theta_cur = __fadd_rn(k.theta, value);
    double sd, cd;
    sincos((double)theta_cur, &sd, &cd);
    const float ct = round_checked(cd, rel, amb), st = round_checked(sd, rel, amb);
    const float b = 0.f;  // "currently only b = 0", KinematicLink.cu:35 -- kept so that -0 and NaN behave as written there
    float m[16];
    m[0] = ct;  m[1] = __fmul_rn(-st, k.ca); m[2] = __fmul_rn(st, k.sa);
    m[3] = __fsub_rn(__fmul_rn(k.a, ct), __fmul_rn(b, st));
    m[4] = st;  m[5] = __fmul_rn(ct, k.ca);  m[6] = __fmul_rn(-ct, k.sa);
    m[7] = __fadd_rn(__fmul_rn(k.a, st), __fmul_rn(b, ct));
    m[8] = 0.f; m[9] = k.sa; m[10] = k.ca; m[11] = d_cur;
    m[12] = 0.f; m[13] = 0.f; m[14] = 0.f; m[15] = 1.f;
    float t[16];
    mat4_mul_rn(acc, m, t);
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = t[i];
  }
  if (amb) {
    const uint32_t slot = atomicAdd(ambiguous, 1u);
    if (slot < amb_cap) ambiguous[1 + slot] = c;
  }
}

cudaError_t launch_dh_chain(const LaunchCtx& c, const DhLinkDev* links_d, uint32_t n_links, const float* joint_values_d,
                            uint32_t n_cfg, float* out_d, uint32_t out_stride, int check_bits, uint32_t* ambiguous_d,
                            uint32_t amb_cap) {
  if (n_cfg == 0 || n_links == 0) return cudaSuccess;
  ++*c.launches;
  k_dh_chain<<<(n_cfg + 127) / 128, 128, 0, c.stream>>>(links_d, n_links, joint_values_d, n_cfg, out_d, out_stride,
                                                        ldexp(1.0, -check_bits), ambiguous_d, amb_cap);
  return cudaGetLastError();
}

// ---- host: the reference's own arithmetic (libm double sin/cos, float products without contraction) ------------
#if defined(__GNUC__) && !defined(__clang__)
#define GVX_NO_CONTRACT __attribute__((optimize("fp-contract=off")))
#else
#define GVX_NO_CONTRACT
#endif

DhLinkDev dh_link_prepare(float d, float theta, float a, float alpha, int joint_type, int out_slot) {
  DhLinkDev k;
  k.out_slot = out_slot;
  k.d = d; k.theta = theta; k.a = a;
  k.ca = (float)std::cos((double)alpha);  // per link, not per configuration: libm on the host, exactly the reference's value
  k.sa = (float)std::sin((double)alpha);
  k.joint_type = joint_type;
  return k;
}

GVX_NO_CONTRACT void dh_chain_host(const DhLinkDev* links, uint32_t n_links, const float* joint_values, float* out) {
  float acc[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
  for (uint32_t l = 0; l < n_links; ++l) {
    const DhLinkDev& k = links[l];
    if (k.out_slot >= 0)
      for (int i = 0; i < 16; ++i) out[(size_t)k.out_slot * 16 + i] = acc[i];
    volatile float d_cur = k.d, theta_cur = k.theta;
    if (k.joint_type == 1) d_cur = k.d + joint_values[l];
    else if (k.joint_type == 0) theta_cur = k.theta + joint_values[l];
    const float ct = (float)std::cos((double)theta_cur), st = (float)std::sin((double)theta_cur), b = 0.f;
    volatile float m[16];
    m[0] = ct;  m[1] = -st * k.ca; m[2] = st * k.sa;   m[3] = k.a * ct - b * st;
    m[4] = st;  m[5] = ct * k.ca;  m[6] = -ct * k.sa;  m[7] = k.a * st + b * ct;
    m[8] = 0;   m[9] = k.sa;       m[10] = k.ca;       m[11] = d_cur;
    m[12] = 0;  m[13] = 0;         m[14] = 0;          m[15] = 1;
    float t[16];
    for (int i = 0; i < 4; ++i)
      for (int j = 0; j < 4; ++j) {
        volatile float s = acc[4 * i] * m[j];
        s = s + acc[4 * i + 1] * m[4 + j];
        s = s + acc[4 * i + 2] * m[8 + j];
        s = s + acc[4 * i + 3] * m[12 + j];
        t[4 * i + j] = s;
      }
    for (int i = 0; i < 16; ++i) acc[i] = t[i];
  }
}

}  // namespace gvx
