// Real spherical-harmonics basis (degree 1..5) in Cartesian form, shared by the stand-alone encoding kernel (sh.cu) and
// the fused-MLP forward, which can write the encoding of the view direction next to the geometry features (mlp_tc.cu).
// tcnn encodings/spherical_harmonics.h:46-152 (Sloan, "Stupid Spherical Harmonics Tricks", appendix A).
#pragma once

namespace {

__device__ __forceinline__ void sh_basis(float x, float y, float z, int degree, float *o) {
    const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
    o[0] = 0.28209479177387814f;
    if (degree <= 1) return;
    o[1] = -0.48860251190291987f * y;
    o[2] = 0.48860251190291987f * z;
    o[3] = -0.48860251190291987f * x;
    if (degree <= 2) return;
    o[4] = 1.0925484305920792f * xy;
    o[5] = -1.0925484305920792f * yz;
    o[6] = 0.94617469575755997f * z2 - 0.31539156525251999f;
    o[7] = -1.0925484305920792f * xz;
    o[8] = 0.54627421529603959f * x2 - 0.54627421529603959f * y2;
    if (degree <= 3) return;
    o[9] = 0.59004358992664352f * y * (-3.0f * x2 + y2);
    o[10] = 2.8906114426405538f * xy * z;
    o[11] = 0.45704579946446572f * y * (1.0f - 5.0f * z2);
    o[12] = 0.3731763325901154f * z * (5.0f * z2 - 3.0f);
    o[13] = 0.45704579946446572f * x * (1.0f - 5.0f * z2);
    o[14] = 1.4453057213202769f * z * (x2 - y2);
    o[15] = 0.59004358992664352f * x * (-x2 + 3.0f * y2);
    if (degree <= 4) return;
    const float x4 = x2 * x2, y4 = y2 * y2, z4 = z2 * z2;
    o[16] = 2.5033429417967046f * xy * (x2 - y2);
    o[17] = 1.7701307697799304f * yz * (-3.0f * x2 + y2);
    o[18] = 0.94617469575756008f * xy * (7.0f * z2 - 1.0f);
    o[19] = 0.66904654355728921f * yz * (3.0f - 7.0f * z2);
    o[20] = -3.1735664074561294f * z2 + 3.7024941420321507f * z4 + 0.31735664074561293f;
    o[21] = 0.66904654355728921f * xz * (3.0f - 7.0f * z2);
    o[22] = 0.47308734787878004f * (x2 - y2) * (7.0f * z2 - 1.0f);
    o[23] = 1.7701307697799304f * xz * (-x2 + 3.0f * y2);
    o[24] = -3.7550144126950569f * x2 * y2 + 0.62583573544917614f * x4 + 0.62583573544917614f * y4;
}


// direction as the reference feeds it: SHEmbedder maps d to (d+1)/2 (remap), the encoding maps back with 2x-1
__device__ __forceinline__ void sh_input(float &x, float &y, float &z, int remap) {
    if (remap) { x = (x + 1.0f) / 2.0f; y = (y + 1.0f) / 2.0f; z = (z + 1.0f) / 2.0f; }
    x = fmaf(x, 2.f, -1.f); y = fmaf(y, 2.f, -1.f); z = fmaf(z, 2.f, -1.f);
}

}  // namespace
