// Degree-3 real spherical harmonics (16 components) in the torch path's convention: evaluated on
// d' = (d+1)/2 directly.  Shared by the head-input assembly kernel (heads.cu) and the tensor-core head MLP, which
// builds its input tile from the same pieces (mlp_tc.cu).
#pragma once

// SH component c of degree<=3 evaluated on d' = (d+1)/2, the torch path's convention
__device__ __forceinline__ float sh_component(int c, float x, float y, float z) {
  float xx = x * x, yy = y * y, zz = z * z;
  switch (c) {
    case 0: return 0.28209479177387814f;
    case 1: return 0.4886025119029199f * y;
    case 2: return 0.4886025119029199f * z;
    case 3: return 0.4886025119029199f * x;
    case 4: return 1.0925484305920792f * x * y;
    case 5: return 1.0925484305920792f * y * z;
    case 6: return 0.9461746957575601f * zz - 0.31539156525251999f;
    case 7: return 1.0925484305920792f * x * z;
    case 8: return 0.5462742152960396f * (xx - yy);
    case 9: return 0.5900435899266435f * y * (3.f * xx - yy);
    case 10: return 2.890611442640554f * x * y * z;
    case 11: return 0.4570457994644658f * y * (5.f * zz - 1.f);
    case 12: return 0.3731763325901154f * z * (5.f * zz - 3.f);
    case 13: return 0.4570457994644658f * x * (5.f * zz - 1.f);
    case 14: return 1.445305721320277f * z * (xx - yy);
    default: return 0.5900435899266435f * x * (xx - 3.f * yy);
  }
}


// all 16 components at once
__device__ __forceinline__ void sh16(float x, float y, float z, float (&o)[16]) {
#pragma unroll
  for (int c = 0; c < 16; ++c) o[c] = sh_component(c, x, y, z);
}
