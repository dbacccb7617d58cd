// rsqrt.cuh -- 1/sqrt for the potential kernels (direct sums in potential.cu, kd-tree walk in ptree.cu)
#pragma once

namespace vrgpu {

// 1/sqrt(r2): hardware seed (rsqrt.approx.f64 = MUFU.RSQ64H on the high word, ~2^-22) + one fp64 Newton step written so
// that only THREE instructions go to the FP64 pipe (t = r2*y0, e = 1 - t*y0, y = y0 + (y0/2)*e with y0/2 formed by an
// integer decrement of the exponent); relative error ~1e-13.  Valid for normal r2 in [2^-963, 2^963): the callers test
// rsqrt_out_of_range() -- an integer compare on the high word, not an FP64-pipe DSETP -- and patch the rare zero /
// denormal / huge / non-finite arguments with the exact 1/sqrt (inf for r2 == 0, like the reference).
__device__ __forceinline__ double rsqrt_fast(double r2)
{
    double y0;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y0) : "d"(r2));
    const double hy0 = __hiloint2double(__double2hiint(y0) - 0x00100000, __double2loint(y0));
    const double t = r2 * y0;
    const double e = fma(-t, y0, 1.0);
    return fma(hy0, e, y0);
}
__device__ __forceinline__ bool rsqrt_out_of_range(double r2)
{
    return (unsigned)(__double2hiint(r2) - 0x03c00000) >= (0x7c200000u - 0x03c00000u);
}
__device__ __forceinline__ double rsqrt_refined(double r2)
{
    return rsqrt_out_of_range(r2) ? 1.0 / sqrt(r2) : rsqrt_fast(r2);
}

} // namespace vrgpu
