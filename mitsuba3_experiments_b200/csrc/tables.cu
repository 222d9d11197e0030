// tables.cu -- scene-load precomputation of the rough-plastic tables on the device.
//
// Replaces RoughPlastic::parameters_changed (mitsuba3/src/bsdfs/roughplastic.cpp:232-262): external transmittance at
// 64 incident cosines (eval_transmittance, mitsuba/render/microfacet.h:521-570) and the internal diffuse reflectance
// (mean of eval_reflectance * cos at 1 / eta, :463-519, times 2), both Gauss-Legendre quadratures (mitsuba/core/quad.h:27-86)
// over the unit square of the visible-normal sampling routine.  One block per (table, cosine): 128 blocks.
#include "b2_build.h"
#include "b2_shade.cuh"
#include <cmath>
#include <vector>

namespace b2 {

namespace {
void legendre_pd(int l, double x, double &lc, double &dc) {           // math.h:93-120
    lc = 0; dc = 0;
    if (l > 1) {
        double l_p_pred = 1, l_pred = x, d_p_pred = 0, d_pred = 1, k0 = 3, k1 = 2, k2 = 1;
        for (int ki = 2; ki <= l; ++ki) {
            lc = (k0 * x * l_pred - k2 * l_p_pred) / k1; dc = d_p_pred + k0 * l_pred;
            l_p_pred = l_pred; l_pred = lc; d_p_pred = d_pred; d_pred = dc; k2 = k1; k0 += 2; k1 += 1;
        }
    } else if (l == 0) { lc = 1; dc = 0; } else { lc = x; dc = 1; }
}
void gauss_legendre(int n, std::vector<float> &nodes, std::vector<float> &weights) {   // quad.h:27-86
    nodes.assign(n, 0.f); weights.assign(n, 0.f);
    n--;
    if (n == 0) { weights[0] = 2.f; return; }
    const int m = (n + 1) / 2;
    for (int i = 0; i < m; ++i) {
        double x = -std::cos((double) (2 * i + 1) / (double) (2 * n + 2) * 3.14159265358979323846), lc, dc;
        for (int it = 0; it < 20; ++it) {
            legendre_pd(n + 1, x, lc, dc);
            const double step = lc / dc; x -= step;
            if (std::fabs(step) <= 4 * std::fabs(x) * 1.1102230246251565e-16) break;
        }
        legendre_pd(n + 1, x, lc, dc);
        weights[i] = weights[n - i] = (float) (2 / ((1 - x * x) * (dc * dc)));
        nodes[i] = (float) x; nodes[n - i] = (float) -x;
    }
    if ((n % 2) == 0) { double lc, dc; legendre_pd(n + 1, 0.0, lc, dc); weights[n / 2] = (float) (2.0 / (dc * dc)); nodes[n / 2] = 0.f; }
}
}  // namespace

constexpr int kRoughRes = 64;       // MI_ROUGH_TRANSMITTANCE_RES

// quad[0..31] / quad[32..63]: nodes / weights of the 32-point rule; quad[64..191] / quad[192..319]: the 128-point rule
__global__ void __launch_bounds__(256) k_rough_tables(int type, float alpha, float eta, const float *__restrict__ quad, float *__restrict__ out) {
    __shared__ float s_sum[256];
    const int i = blockIdx.x & (kRoughRes - 1);
    const bool transmit = blockIdx.x < kRoughRes;
    const float e = transmit ? eta : 1.f / eta;
    const int res = e > 1.f ? 32 : 128;
    const float *nodes = res == 32 ? quad : quad + 64, *weights = nodes + res;
    const float mu = fmaxf(1e-6f, (float) i * (1.f / (float) (kRoughRes - 1)));
    const f3 wi(sqrtf(1.f - mu * mu), 0.f, mu);
    const Microfacet distr(type, alpha, alpha, true);
    float sum = 0.f;
    for (int j = threadIdx.x; j < res * res; j += blockDim.x) {
        const int jx = j % res, jy = j / res;
        const f2 node{ fma_(nodes[jx], .5f, .5f), fma_(nodes[jy], .5f, .5f) };
        float pdf; const f3 m = distr.sample(wi, node, pdf);
        float f, cos_t, eta_it, eta_ti; fresnel(dot(wi, m), e, f, cos_t, eta_it, eta_ti);
        float smith;
        if (transmit) {
            const f3 wo = fmadd(m, fma_(dot(wi, m), eta_ti, cos_t), -(wi * eta_ti));     // refract(wi, m, cos_t, eta_ti)
            smith = distr.smith_g1(wo, m) * (1.f - f);
            if (wo.z * wi.z >= 0.f) smith = 0.f;
        } else {
            const f3 wo = fmadd(m, 2.f * dot(wi, m), -wi);                                // reflect(wi, m)
            smith = distr.smith_g1(wo, m) * f;
            if (wo.z <= 0.f || wi.z <= 0.f) smith = 0.f;
        }
        sum += smith * (weights[jx] * weights[jy]) * 0.25f;
    }
    s_sum[threadIdx.x] = sum;
    __syncthreads();
    for (int o = 128; o; o >>= 1) { if ((int) threadIdx.x < o) s_sum[threadIdx.x] += s_sum[threadIdx.x + o]; __syncthreads(); }
    if (threadIdx.x == 0) out[blockIdx.x] = transmit ? s_sum[0] : s_sum[0] * wi.z;
}

bool rough_tables_build(int type, float alpha, float eta, float *trans64, float *internal_reflectance, std::string &err) {
    std::vector<float> n32, w32, n128, w128, quad;
    gauss_legendre(32, n32, w32); gauss_legendre(128, n128, w128);
    quad.insert(quad.end(), n32.begin(), n32.end()); quad.insert(quad.end(), w32.begin(), w32.end());
    quad.insert(quad.end(), n128.begin(), n128.end()); quad.insert(quad.end(), w128.begin(), w128.end());
    float *d_quad = nullptr, *d_out = nullptr, h_out[2 * kRoughRes];
    cudaError_t e = cudaMalloc((void **) &d_quad, quad.size() * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc((void **) &d_out, sizeof(h_out));
    if (e == cudaSuccess) e = cudaMemcpy(d_quad, quad.data(), quad.size() * sizeof(float), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) { k_rough_tables<<<2 * kRoughRes, 256>>>(type, alpha, eta, d_quad, d_out); e = cudaGetLastError(); }
    if (e == cudaSuccess) e = cudaMemcpy(h_out, d_out, sizeof(h_out), cudaMemcpyDeviceToHost);
    if (d_quad) cudaFree(d_quad);
    if (d_out) cudaFree(d_out);
    if (e != cudaSuccess) { err = std::string("rough_tables_build: ") + cudaGetErrorString(e); return false; }
    float acc = 0.f;
    for (int i = 0; i < kRoughRes; ++i) { trans64[i] = h_out[i]; acc += h_out[kRoughRes + i]; }
    *internal_reflectance = acc / (float) kRoughRes * 2.f;
    return true;
}

}  // namespace b2
