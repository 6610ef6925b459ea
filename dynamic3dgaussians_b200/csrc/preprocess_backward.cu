// (d) per-Gaussian backward: from the per-(view, Gaussian) gradient records of the blend backward to the gradients
// of means3D / scales / rotations / opacity / colours (/ cov3D / SH), summed over the V views of the batch.
// Replaces computeCov2DCUDA + BACKWARD::preprocessCUDA + computeCov3D/computeColorFromSH backward
// (R/cuda_rasterizer/backward.cu:20-139,144-291,295-429) -- two kernels per view in the reference, one kernel here.
//
// One thread per Gaussian loops over the views: the Gaussian's parameters and 3D covariance are loaded/derived once,
// the per-view records are read coalesced ([V][P][16] floats), and the sum over views is a deterministic register
// accumulation (no atomics).  Gradient conventions follow autograd through the forward as written:
// the clamp of t.x/t.z passes gradient only when not clamped, det is inverted without the reference's +1e-7.
#include "geom_math.cuh"

namespace gsr {

namespace {

struct SHGrad {
    float dmean[3];
};

// Backward of computeColorFromSH (forward.cu:20-71): writes dL/dsh for this view (accumulating over views is done by
// the caller through `accumulate`), returns the gradient w.r.t. the mean through the view direction.
__device__ void sh_backward(int deg, int M, const float* __restrict__ sh_g, float3 pos, const float* campos,
                            const uint8_t* __restrict__ clamped, const float* dL_dcolor, float* __restrict__ dL_dsh_g,
                            bool first, float* dmean) {
    const float dx = pos.x - campos[0], dy = pos.y - campos[1], dz = pos.z - campos[2];
    const float inv_len = rsqrtf(dx * dx + dy * dy + dz * dz);
    const float x = dx * inv_len, y = dy * inv_len, z = dz * inv_len;
    float dRGB[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) dRGB[c] = clamped[c] ? 0.0f : dL_dcolor[c];
    float dRGBdx[3] = {0, 0, 0}, dRGBdy[3] = {0, 0, 0}, dRGBdz[3] = {0, 0, 0};
    auto SH = [&](int k, int c) { return sh_g[k * 3 + c]; };
    auto put = [&](int k, float coef) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float g = coef * dRGB[c];
            if (first) dL_dsh_g[k * 3 + c] = g; else dL_dsh_g[k * 3 + c] += g;
        }
    };
    put(0, GSR_SH_C0);
    if (deg > 0) {
        put(1, -GSR_SH_C1 * y);
        put(2, GSR_SH_C1 * z);
        put(3, -GSR_SH_C1 * x);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            dRGBdx[c] = -GSR_SH_C1 * SH(3, c);
            dRGBdy[c] = -GSR_SH_C1 * SH(1, c);
            dRGBdz[c] = GSR_SH_C1 * SH(2, c);
        }
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            put(4, GSR_SH_C2_0 * xy);
            put(5, GSR_SH_C2_1 * yz);
            put(6, GSR_SH_C2_2 * (2.f * zz - xx - yy));
            put(7, GSR_SH_C2_3 * xz);
            put(8, GSR_SH_C2_4 * (xx - yy));
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                dRGBdx[c] += GSR_SH_C2_0 * y * SH(4, c) + GSR_SH_C2_2 * 2.f * -x * SH(6, c) + GSR_SH_C2_3 * z * SH(7, c) +
                             GSR_SH_C2_4 * 2.f * x * SH(8, c);
                dRGBdy[c] += GSR_SH_C2_0 * x * SH(4, c) + GSR_SH_C2_1 * z * SH(5, c) + GSR_SH_C2_2 * 2.f * -y * SH(6, c) +
                             GSR_SH_C2_4 * 2.f * -y * SH(8, c);
                dRGBdz[c] += GSR_SH_C2_1 * y * SH(5, c) + GSR_SH_C2_2 * 4.f * z * SH(6, c) + GSR_SH_C2_3 * x * SH(7, c);
            }
            if (deg > 2) {
                put(9, GSR_SH_C3_0 * y * (3.f * xx - yy));
                put(10, GSR_SH_C3_1 * xy * z);
                put(11, GSR_SH_C3_2 * y * (4.f * zz - xx - yy));
                put(12, GSR_SH_C3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy));
                put(13, GSR_SH_C3_4 * x * (4.f * zz - xx - yy));
                put(14, GSR_SH_C3_5 * z * (xx - yy));
                put(15, GSR_SH_C3_6 * x * (xx - 3.f * yy));
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    dRGBdx[c] += GSR_SH_C3_0 * SH(9, c) * 6.f * xy + GSR_SH_C3_1 * SH(10, c) * yz +
                                 GSR_SH_C3_2 * SH(11, c) * -2.f * xy + GSR_SH_C3_3 * SH(12, c) * -6.f * xz +
                                 GSR_SH_C3_4 * SH(13, c) * (-3.f * xx + 4.f * zz - yy) +
                                 GSR_SH_C3_5 * SH(14, c) * 2.f * xz + GSR_SH_C3_6 * SH(15, c) * 3.f * (xx - yy);
                    dRGBdy[c] += GSR_SH_C3_0 * SH(9, c) * 3.f * (xx - yy) + GSR_SH_C3_1 * SH(10, c) * xz +
                                 GSR_SH_C3_2 * SH(11, c) * (-3.f * yy + 4.f * zz - xx) +
                                 GSR_SH_C3_3 * SH(12, c) * -6.f * yz + GSR_SH_C3_4 * SH(13, c) * -2.f * xy +
                                 GSR_SH_C3_5 * SH(14, c) * -2.f * yz + GSR_SH_C3_6 * SH(15, c) * -6.f * xy;
                    dRGBdz[c] += GSR_SH_C3_1 * SH(10, c) * xy + GSR_SH_C3_2 * SH(11, c) * 8.f * yz +
                                 GSR_SH_C3_3 * SH(12, c) * 3.f * (2.f * zz - xx - yy) +
                                 GSR_SH_C3_4 * SH(13, c) * 8.f * xz + GSR_SH_C3_5 * SH(14, c) * (xx - yy);
                }
            }
        }
    }
    // coefficients above the active degree take no part in the colour (forward.cu:20-71): their gradient is zero (the
    // reference's buffer is zero-filled, rasterize_points.cu:176); written once, with the Gaussian's first visible view
    if (first)
        for (int i = (deg + 1) * (deg + 1) * 3; i < M * 3; ++i) dL_dsh_g[i] = 0.0f;
    // d(dir)/d(mean): normalisation Jacobian (auxiliary.h:108-119)
    const float ddx = dRGBdx[0] * dRGB[0] + dRGBdx[1] * dRGB[1] + dRGBdx[2] * dRGB[2];
    const float ddy = dRGBdy[0] * dRGB[0] + dRGBdy[1] * dRGB[1] + dRGBdy[2] * dRGB[2];
    const float ddz = dRGBdz[0] * dRGB[0] + dRGBdz[1] * dRGB[1] + dRGBdz[2] * dRGB[2];
    const float sum2 = dx * dx + dy * dy + dz * dz;
    const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    dmean[0] += ((sum2 - dx * dx) * ddx - dy * dx * ddy - dz * dx * ddz) * invsum32;
    dmean[1] += (-dx * dy * ddx + (sum2 - dy * dy) * ddy - dz * dy * ddz) * invsum32;
    dmean[2] += (-dx * dz * ddx - dy * dz * ddy + (sum2 - dz * dz) * ddz) * invsum32;
}

// TPG threads per Gaussian share the loop over the views (thread `sub` takes views sub, sub + TPG, ...) and add their
// partial sums with shuffles: one thread per Gaussian left the kernel latency-bound (ncu, round 2: 3 CTAs / SM at 80
// registers, 33 % of the warp slots, long-scoreboard stalls 4.7 per issue while it walks 27 dependent record loads).
// The SH path writes dL/dsh straight to global memory, accumulating over the views, and keeps TPG = 1.
#ifndef GSR_PB_MINB
#define GSR_PB_MINB 3
#endif
template <int NCH, int TPG>
__global__ void __launch_bounds__(256, GSR_PB_MINB)
preprocess_backward_kernel(const CamDev* __restrict__ cams, int V, int P, int sh_coeffs,
                           const float* __restrict__ means3D, const float* __restrict__ scales,
                           const float* __restrict__ rotations, const float* __restrict__ cov3D_precomp,
                           const float* __restrict__ sh, const int32_t* __restrict__ radii,
                           const uint8_t* __restrict__ clamped, const float* __restrict__ grad_rec,
                           const float* __restrict__ label, float* __restrict__ dL_dmeans3D,
                           float* __restrict__ dL_dmeans2D, float* __restrict__ dL_dm2c, float* __restrict__ dL_dcolors,
                           float* __restrict__ dL_dextra, float* __restrict__ dL_dopacity,
                           float* __restrict__ dL_dscales, float* __restrict__ dL_drotations,
                           float* __restrict__ dL_dcov3D, float* __restrict__ dL_dsh) {
    pdl_wait();          // (common.cuh) nothing global is touched before the predecessor kernel has completed
    pdl_trigger();
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int sub = t % TPG;
    const bool live = t / TPG < P;                        // (no early return: the whole warp takes part in the shuffles)
    if (TPG == 1 && !live) return;
    const int g = live ? t / TPG : P - 1;
    const float x = means3D[3 * g], y = means3D[3 * g + 1], z = means3D[3 * g + 2];
    float cv[6];
    float3 sc = make_float3(0, 0, 0);
    float4 q = make_float4(1, 0, 0, 0);
    float mod = 1.0f;
    if (cov3D_precomp != nullptr) {
#pragma unroll
        for (int i = 0; i < 6; ++i) cv[i] = cov3D_precomp[6 * (size_t)g + i];
    } else {
        sc = make_float3(scales[3 * g], scales[3 * g + 1], scales[3 * g + 2]);
        q = reinterpret_cast<const float4*>(rotations)[g];
        mod = cams[0].scale_modifier;     // one modifier per batch (checked on the host)
        cov3d_from_scale_rot(sc, mod, q, cv);
    }

    float dmean[3] = {0, 0, 0};
    float dcov[6] = {0, 0, 0, 0, 0, 0};
    float dopac = 0.0f;
    float dcol[NCH];
#pragma unroll
    for (int k = 0; k < NCH; ++k) dcol[k] = 0.0f;
    bool sh_first = true;

    // Software pipeline over the views: the radius of view v + 2 and the gradient record of view v + 1 are in flight while
    // view v's chain rule runs (ncu on the plain loop: 45 % of the stall samples on the radius load and on the first use
    // of the record -- two dependent DRAM round trips per view at 33 % occupancy).
    constexpr int REC = grad_rec_stride(NCH);
    const int v_end = live ? V : 0;
    auto load_radius = [&](int v) -> int { return v < v_end ? radii[(size_t)v * P + g] : 0; };
    auto load_rec = [&](int v, float4* dst) {
        const float4* rp = reinterpret_cast<const float4*>(grad_rec + ((size_t)v * P + g) * REC);
#pragma unroll
        for (int i = 0; i < REC / 4; ++i) dst[i] = rp[i];
    };
    int r_cur = load_radius(sub), r_nxt = load_radius(sub + TPG);
    float4 rec_cur[REC / 4];
    if (r_cur > 0) load_rec(sub, rec_cur);
    for (int v = sub; v < v_end; v += TPG) {
        const size_t idx = (size_t)v * P + g;
        const int r_now = r_cur;
        float4 rec4[REC / 4];
#pragma unroll
        for (int i = 0; i < REC / 4; ++i) rec4[i] = rec_cur[i];
        r_cur = r_nxt;
        if (r_cur > 0) load_rec(v + TPG, rec_cur);
        r_nxt = load_radius(v + 2 * TPG);
        if (!(r_now > 0)) {                                       // backward.cu:158,386
            if (dL_dmeans2D) {
                dL_dmeans2D[3 * idx] = 0.0f;
                dL_dmeans2D[3 * idx + 1] = 0.0f;
                dL_dmeans2D[3 * idx + 2] = 0.0f;
            }
            if (dL_dm2c) {
                dL_dm2c[3 * idx] = 0.0f;
                dL_dm2c[3 * idx + 1] = 0.0f;
                dL_dm2c[3 * idx + 2] = 0.0f;
            }
            continue;
        }
        float rec[REC];
#pragma unroll
        for (int i = 0; i < REC / 4; ++i) {
            const float4 t = rec4[i];
            rec[4 * i] = t.x; rec[4 * i + 1] = t.y; rec[4 * i + 2] = t.z; rec[4 * i + 3] = t.w;
        }
        const CamDev& cam = cams[v];
        Cov2DInter ci;
        cov2d(cam, x, y, z, cv, ci);

        // ---- conic -> 2D covariance (a, b, c) ----
        const float a = ci.a, b = ci.b, c = ci.c;
        const float det = a * c - b * b;
        // the blend backward hands over raw moments (S1 = sum a d, S2 = sum a d d^T, a = opacity G dL/dalpha):
        //   dL/dconic = (-S2xx/2, -S2xy, -S2yy/2)   [the true d/d(conic.y), no 0.5]         backward.cu:644-646
        //   dL/dmean2D = -conic S1, scaled to the reference's NDC units (0.5 W, 0.5 H)         backward.cu:520-521,640-641
        const float dA = -0.5f * rec[GR_CA], dB = -rec[GR_CB], dC = -0.5f * rec[GR_CC];
        const float idet = 1.0f / det;
        const float conA = c * idet, conB = -b * idet, conC = a * idet;
        const float gx = -0.5f * (float)cam.width * (conA * rec[GR_MX] + conB * rec[GR_MY]);
        const float gy = -0.5f * (float)cam.height * (conC * rec[GR_MY] + conB * rec[GR_MX]);
        const float d2 = idet * idet;
        const float da = d2 * (-c * c * dA + b * c * dB - b * b * dC);
        const float dc = d2 * (-a * a * dC + a * b * dB - b * b * dA);
        const float db = d2 * (2.f * b * c * dA - (det + 2.f * b * b) * dB + 2.f * a * b * dC);

        // ---- 2D covariance -> 3D covariance (6 unique entries) ----
        const float T00 = ci.T00, T01 = ci.T01, T02 = ci.T02, T10 = ci.T10, T11 = ci.T11, T12 = ci.T12;
        dcov[0] += T00 * T00 * da + T00 * T10 * db + T10 * T10 * dc;
        dcov[3] += T01 * T01 * da + T01 * T11 * db + T11 * T11 * dc;
        dcov[5] += T02 * T02 * da + T02 * T12 * db + T12 * T12 * dc;
        dcov[1] += 2.f * T00 * T01 * da + (T00 * T11 + T01 * T10) * db + 2.f * T10 * T11 * dc;
        dcov[2] += 2.f * T00 * T02 * da + (T00 * T12 + T02 * T10) * db + 2.f * T10 * T12 * dc;
        dcov[4] += 2.f * T02 * T01 * da + (T01 * T12 + T02 * T11) * db + 2.f * T11 * T12 * dc;

        // ---- 2D covariance -> T -> J -> view-space mean t ----
        const float vt0x = cv[0] * T00 + cv[1] * T01 + cv[2] * T02;     // Vrk * T0
        const float vt0y = cv[1] * T00 + cv[3] * T01 + cv[4] * T02;
        const float vt0z = cv[2] * T00 + cv[4] * T01 + cv[5] * T02;
        const float vt1x = cv[0] * T10 + cv[1] * T11 + cv[2] * T12;     // Vrk * T1
        const float vt1y = cv[1] * T10 + cv[3] * T11 + cv[4] * T12;
        const float vt1z = cv[2] * T10 + cv[4] * T11 + cv[5] * T12;
        const float dT00 = 2.f * vt0x * da + vt1x * db, dT01 = 2.f * vt0y * da + vt1y * db,
                    dT02 = 2.f * vt0z * da + vt1z * db;
        const float dT10 = 2.f * vt1x * dc + vt0x * db, dT11 = 2.f * vt1y * dc + vt0y * db,
                    dT12 = 2.f * vt1z * dc + vt0z * db;
        const float* Vm = cam.view;
        const float dJ00 = Vm[0] * dT00 + Vm[4] * dT01 + Vm[8] * dT02;
        const float dJ02 = Vm[2] * dT00 + Vm[6] * dT01 + Vm[10] * dT02;
        const float dJ11 = Vm[1] * dT10 + Vm[5] * dT11 + Vm[9] * dT12;
        const float dJ12 = Vm[2] * dT10 + Vm[6] * dT11 + Vm[10] * dT12;
        const float tz = ci.tz;
        const float itz = 1.0f / tz, itz2 = itz * itz;
        const bool free_x = ci.txtz > -cam.lim_x_neg && ci.txtz < cam.lim_x_pos;
        const bool free_y = ci.tytz > -cam.lim_y_neg && ci.tytz < cam.lim_y_pos;
        // J02 = -fx * clamp(tx/tz) / tz ,  J12 = -fy * clamp(ty/tz) / tz
        const float dtx = free_x ? -cam.focal_x * itz2 * dJ02 : 0.0f;
        const float dty = free_y ? -cam.focal_y * itz2 * dJ12 : 0.0f;
        float dtz = -cam.focal_x * itz2 * dJ00 - cam.focal_y * itz2 * dJ11 + cam.focal_x * ci.cxl * itz2 * dJ02 +
                    cam.focal_y * ci.cyl * itz2 * dJ12;
        if (free_x) dtz += cam.focal_x * ci.txtz * itz2 * dJ02;
        if (free_y) dtz += cam.focal_y * ci.tytz * itz2 * dJ12;
        // depth output: D = sum depth * alpha * T, depth = t.z
        dtz += rec[GR_DEPTH];
        dmean[0] += Vm[0] * dtx + Vm[1] * dty + Vm[2] * dtz;
        dmean[1] += Vm[4] * dtx + Vm[5] * dty + Vm[6] * dtz;
        dmean[2] += Vm[8] * dtx + Vm[9] * dty + Vm[10] * dtz;

        // ---- screen-space mean (NDC units) -> mean3D through the projection (backward.cu:392-406) ----
        const float* Pm = cam.proj;
        const float hx = Pm[0] * x + Pm[4] * y + Pm[8] * z + Pm[12];
        const float hy = Pm[1] * x + Pm[5] * y + Pm[9] * z + Pm[13];
        const float hw = Pm[3] * x + Pm[7] * y + Pm[11] * z + Pm[15];
        const float mw = 1.0f / (hw + 0.0000001f);
        const float mul1 = hx * mw * mw, mul2 = hy * mw * mw;
        dmean[0] += (Pm[0] * mw - Pm[3] * mul1) * gx + (Pm[1] * mw - Pm[3] * mul2) * gy;
        dmean[1] += (Pm[4] * mw - Pm[7] * mul1) * gx + (Pm[5] * mw - Pm[7] * mul2) * gy;
        dmean[2] += (Pm[8] * mw - Pm[11] * mul1) * gx + (Pm[9] * mw - Pm[11] * mul2) * gy;
        if (dL_dmeans2D) {
            dL_dmeans2D[3 * idx] = gx;
            dL_dmeans2D[3 * idx + 1] = gy;
            dL_dmeans2D[3 * idx + 2] = 0.0f;
        }
        if (dL_dm2c) {
            // the colour render's share (REF/train.py:245): its own moments when extra colours were blended in the same
            // pass (slots GR_COL + NCH, + 1 of the record, blend_backward_kernel<6, true>), the whole gradient otherwise
            float cx = gx, cy = gy;
            if (NCH > 3) {
                const float mxc = rec[GR_COL + NCH], myc = rec[GR_COL + NCH + 1];
                cx = -0.5f * (float)cam.width * (conA * mxc + conB * myc);
                cy = -0.5f * (float)cam.height * (conC * myc + conB * mxc);
            }
            dL_dm2c[3 * idx] = cx;
            dL_dm2c[3 * idx + 1] = cy;
            dL_dm2c[3 * idx + 2] = 0.0f;
        }

        dopac += rec[GR_OP];
        if (sh != nullptr) {
            sh_backward(cam.sh_degree, sh_coeffs, sh + (size_t)g * sh_coeffs * 3, make_float3(x, y, z), cam.campos,
                        clamped + 3 * idx, &rec[GR_COL], dL_dsh + (size_t)g * sh_coeffs * 3, sh_first, dmean);
            sh_first = false;
        } else {
#pragma unroll
            for (int k = 0; k < 3; ++k) dcol[k] += rec[GR_COL + k];
        }
#pragma unroll
        for (int k = 3; k < NCH; ++k) dcol[k] += rec[GR_COL + k];
    }

    if (TPG > 1) {
#pragma unroll
        for (int o = 1; o < TPG; o <<= 1) {
#pragma unroll
            for (int i = 0; i < 3; ++i) dmean[i] += __shfl_xor_sync(0xffffffffu, dmean[i], o);
#pragma unroll
            for (int i = 0; i < 6; ++i) dcov[i] += __shfl_xor_sync(0xffffffffu, dcov[i], o);
            dopac += __shfl_xor_sync(0xffffffffu, dopac, o);
#pragma unroll
            for (int k = 0; k < NCH; ++k) dcol[k] += __shfl_xor_sync(0xffffffffu, dcol[k], o);
        }
        if (sub != 0 || !live) return;
    }
    const float lab = label ? label[g] : 1.0f;
    if (dL_dmeans3D) {
        dL_dmeans3D[3 * g] = dmean[0] * lab;
        dL_dmeans3D[3 * g + 1] = dmean[1] * lab;
        dL_dmeans3D[3 * g + 2] = dmean[2] * lab;
    }
    if (dL_dopacity) dL_dopacity[g] = dopac * lab;
    if (dL_dcolors && sh == nullptr) {
#pragma unroll
        for (int k = 0; k < 3; ++k) dL_dcolors[3 * g + k] = dcol[k] * lab;
    }
    if (NCH > 3 && dL_dextra) {
#pragma unroll
        for (int k = 3; k < NCH; ++k) dL_dextra[3 * g + (k - 3)] = dcol[k] * lab;
    }
    if (sh != nullptr && dL_dsh) {
        float* d = dL_dsh + (size_t)g * sh_coeffs * 3;
        if (sh_first) {
            for (int i = 0; i < sh_coeffs * 3; ++i) d[i] = 0.0f;       // Gaussian visible in no view
        } else if (label) {
            for (int i = 0; i < sh_coeffs * 3; ++i) d[i] *= lab;
        }
    }
    if (cov3D_precomp != nullptr) {
        if (dL_dcov3D) {
#pragma unroll
            for (int i = 0; i < 6; ++i) dL_dcov3D[6 * (size_t)g + i] = dcov[i] * lab;
        }
        return;
    }

    // ---- 3D covariance -> scale / rotation:  Sigma = L L^T with L = Rm * diag(mod * s)  (backward.cu:295-358) ----
    const float r = q.x, qx = q.y, qy = q.z, qz = q.w;
    const float Rm[3][3] = {
        {1.f - 2.f * (qy * qy + qz * qz), 2.f * (qx * qy - r * qz), 2.f * (qx * qz + r * qy)},
        {2.f * (qx * qy + r * qz), 1.f - 2.f * (qx * qx + qz * qz), 2.f * (qy * qz - r * qx)},
        {2.f * (qx * qz - r * qy), 2.f * (qy * qz + r * qx), 1.f - 2.f * (qx * qx + qy * qy)}};
    const float s[3] = {mod * sc.x, mod * sc.y, mod * sc.z};
    // G = dL/dSigma as a full symmetric matrix (off-diagonals halved), dL/dL = 2 G L
    const float Gm[3][3] = {{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]},
                            {0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
                            {0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}};
    float dLm[3][3];   // dL/dL
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int k = 0; k < 3; ++k)
            dLm[i][k] = 2.f * (Gm[i][0] * Rm[0][k] + Gm[i][1] * Rm[1][k] + Gm[i][2] * Rm[2][k]) * s[k];
    float ds[3], dR[3][3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        ds[k] = dLm[0][k] * Rm[0][k] + dLm[1][k] * Rm[1][k] + dLm[2][k] * Rm[2][k];
#pragma unroll
        for (int i = 0; i < 3; ++i) dR[i][k] = dLm[i][k] * s[k];
    }
    if (dL_dscales) {
        dL_dscales[3 * g] = mod * ds[0] * lab;
        dL_dscales[3 * g + 1] = mod * ds[1] * lab;
        dL_dscales[3 * g + 2] = mod * ds[2] * lab;
    }
    if (dL_drotations) {
        const float dr = 2.f * (qz * (dR[1][0] - dR[0][1]) + qy * (dR[0][2] - dR[2][0]) + qx * (dR[2][1] - dR[1][2]));
        const float dqx = 2.f * (qy * (dR[0][1] + dR[1][0]) + qz * (dR[0][2] + dR[2][0]) + r * (dR[2][1] - dR[1][2])) -
                          4.f * qx * (dR[1][1] + dR[2][2]);
        const float dqy = 2.f * (qx * (dR[0][1] + dR[1][0]) + r * (dR[0][2] - dR[2][0]) + qz * (dR[1][2] + dR[2][1])) -
                          4.f * qy * (dR[0][0] + dR[2][2]);
        const float dqz = 2.f * (r * (dR[1][0] - dR[0][1]) + qx * (dR[0][2] + dR[2][0]) + qy * (dR[1][2] + dR[2][1])) -
                          4.f * qz * (dR[0][0] + dR[1][1]);
        reinterpret_cast<float4*>(dL_drotations)[g] = make_float4(dr * lab, dqx * lab, dqy * lab, dqz * lab);
    }
}

}  // namespace

int launch_preprocess_backward(int V, int P, int W, int H, int channels, int sh_coeffs, const float* means3D,
                               const float* opacities, const float* scales, const float* rotations,
                               const float* cov3D_precomp, const float* sh, const int32_t* radii, const void* gws,
                               const Layout& L, const float* label, float* dL_dmeans3D, float* dL_dmeans2D,
                               float* dL_dmeans2D_color_only, float* dL_dcolors, float* dL_dextra_colors,
                               float* dL_dopacity, float* dL_dscales, float* dL_drotations, float* dL_dcov3D,
                               float* dL_dsh, cudaStream_t s) {
    if (P == 0 || V == 0) return GSR_OK;
    const CamDev* cams = at<CamDev>(gws, L.cams);
    const int32_t* rad = radii ? radii : at<int32_t>(gws, L.radii);
    const uint8_t* cl = at<uint8_t>(gws, L.clamped);
    const float* rec = at<float>(gws, L.grad_rec);
    const bool split = sh == nullptr;                    // see the kernel: TPG threads per Gaussian unless SH
#ifndef GSR_PB_TPG
#define GSR_PB_TPG 1      // measured at C2 on B200: 1 thread per Gaussian 0.195 ms, 4 threads 0.215 ms
#endif
    constexpr int TPG = GSR_PB_TPG;
    const int blocks = split ? (int)(((size_t)P * TPG + 255) / 256) : (P + 255) / 256;
#define GSR_PB(NCH_, TPG_)                                                                                            \
    launch_kernel(preprocess_backward_kernel<NCH_, TPG_>, blocks, 256, 0, s, cams, V, P, sh_coeffs, means3D, scales, rotations,  \
                                                                 cov3D_precomp, sh, rad, cl, rec, label, dL_dmeans3D, \
                                                                 dL_dmeans2D, dL_dmeans2D_color_only, dL_dcolors,     \
                                                                 dL_dextra_colors, dL_dopacity, dL_dscales,           \
                                                                 dL_drotations, dL_dcov3D, dL_dsh)
    if (channels == 3) {
        if (split) GSR_PB(3, TPG); else GSR_PB(3, 1);
    } else if (channels == 6) {
        if (split) GSR_PB(6, TPG); else GSR_PB(6, 1);
    } else {
        return GSR_ERR_UNSUPPORTED;
    }
#undef GSR_PB
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

}  // namespace gsr
