// rsqrt.cuh -- 1/sqrt for the potential kernels (direct sums in potential.cu, kd-tree walk in ptree.cu)
#pragma once

namespace vrgpu {

// 1/sqrt(r2): hardware seed (rsqrt.approx.f64 = MUFU.RSQ64H on the high word, ~2^-22) + one fp64 Newton step written so
// that only THREE instructions go to the FP64 pipe: with h = r2/2 (an integer decrement of the exponent),
// t = h*y0, e = 1/2 - t*y0, y = y0 + y0*e; relative error ~1e-13.  Valid for finite r2 >= 2^-1000: the callers test
// rsqrt_seed_bad(y0) -- an integer compare on the seed's high word, not an FP64-pipe DSETP; it catches r2 zero, denormal,
// tiny or NaN -- and patch those rare arguments with the exact 1/sqrt (inf for r2 == 0, like the reference).
__device__ __forceinline__ double rsqrt_seed(double r2)
{
    double y0;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y0) : "d"(r2));
    return y0;
}
__device__ __forceinline__ bool rsqrt_seed_bad(double y0) { return (unsigned)__double2hiint(y0) >= 0x5f300000u; } // y0 > 2^500, inf, nan
__device__ __forceinline__ double rsqrt_newton(double r2, double y0)
{
    const double h = __hiloint2double(__double2hiint(r2) - 0x00100000, __double2loint(r2));
    const double t = h * y0;
    const double e = fma(-t, y0, 0.5);
    return fma(y0, e, y0);
}
__device__ __forceinline__ double rsqrt_refined(double r2)
{
    const double y0 = rsqrt_seed(r2);
    return rsqrt_seed_bad(y0) ? 1.0 / sqrt(r2) : rsqrt_newton(r2, y0);
}

} // namespace vrgpu
