// gs_arith.cuh -- the per-cell Gray-Scott update, in two arithmetics.
//
// ARITH_REFERENCE reproduces, rounding for rounding, the expression the reference CPU solver
// evaluates (src/simulation/Common.jl:13-18 + src/simulation/Simulation_CPU.jl:96-109, with
// Julia's promotion rules): for Float32 fields the six-neighbour sum,
// u*v^2 and (F+K)*v are Float32 and everything else Float64; for Float64 fields the same tree
// in Float64.  Every operation is an explicit round-to-nearest intrinsic, so nvcc can neither
// contract to FMA nor reassociate: results are bit-identical to the CPU solver.
//
// ARITH_NATIVE (Float32 only) evaluates the same formula in Float32 with FMAs; it differs from
// the reference by ~1 ulp per step and is checked against the north-star tolerance (1e-5).
//
// x / 6.0 is computed as q = x*c, r = fma(-6,q,x), q' = fma(r,c,q) with c = RN(1/6): the
// correctly rounded quotient whenever no intermediate underflows (checked against true
// division on 3.2e9 random doubles, build/scratch history in DESIGN.md); tiny and zero x take
// the exact paths below.
#pragma once
#include "gs_common.cuh"

namespace gsb {

__device__ __forceinline__ double div6_rn(double x) {
    const double c = 0x1.5555555555555p-3;  // RN(1/6)
    const double q = __dmul_rn(x, c);
    const double r = __fma_rn(-6.0, q, x);
    double res = __fma_rn(r, c, q);
    // zero and tiny x: exact paths.  (Measured: replacing this test by integer compares on the exponent
    // field makes nvcc evaluate the slow division speculatively for every x = 0 cell -- 2.2x slower.)
    if (fabs(x) < 0x1.0p-800) res = (x == 0.0) ? x : __ddiv_rn(x, 6.0);
    return res;
}

// The same quotient for the Float64 update kernels, with the rare case split off: the fast path (sign of a
// zero transferred with one LOP3) plus a flag "x is tiny but not zero" from integer compares on the exponent
// field; the caller redoes flagged cells with div6_rare, a real call that nvcc cannot evaluate speculatively.
// One branch per cell instead of one compare + branch region per quotient.
static __device__ __noinline__ double div6_rare(double x) { return x == 0.0 ? x : __ddiv_rn(x, 6.0); }

__device__ __forceinline__ double div6_fast(double x, bool& rare) {
    const double c = 0x1.5555555555555p-3;
    const double q = __dmul_rn(x, c);
    const double r = __fma_rn(-6.0, q, x);
    const double res = __fma_rn(r, c, q);
    const int xh = __double2hiint(x), ah = xh & 0x7fffffff;
    rare = rare || (ah < 0x0df00000 && (ah | __double2loint(x)) != 0);      // |x| < 2^-800, x != 0
    int hi;
    asm("lop3.b32 %0, %1, %2, 0x80000000, 0xD8;" : "=r"(hi) : "r"(__double2hiint(res)), "r"(xh));   // +-0 / 6 = +-0
    return __hiloint2double(hi, __double2loint(res));
}

// noise * r with r = ri * 2^-31, ri the integer-valued variate of gs_philox.cuh.  Scaling by a
// power of two is exact, so RN(noise * r) == RN((noise * 2^-31) * ri) unless noise * 2^-31 leaves
// the normal range; the host decides (Params::ref_mode) and both forms give the same bits.
template <bool FAST>
__device__ __forceinline__ double noise_term(const Params& p, double ri) {
    return FAST ? __dmul_rn(p.noise_s, ri) : __dmul_rn(p.noise, __dmul_rn(ri, 0x1.0p-31));
}

// Variants of the reference arithmetic, chosen ONCE per vector by a warp-uniform branch on
// Params::ref_mode.  (Left as per-operation conditions, nvcc turns `if (!dt_is_one) x *= dt` and the
// noise form into compute-both-and-select: 4 DMUL + 5 FSEL per cell on the FP64-bound path, ncu r2.)
enum : int { REF_DT1 = 0,      // dt == 1 exactly (x * 1 == x: multiply skipped) and noise * 2^-31 normal
             REF_DT = 1,       // any dt, noise * 2^-31 normal
             REF_GENERIC = 2 };// any dt, noise so small that noise * 2^-31 would be subnormal

// Same quotient when x is an exact difference of Float32-derived values (the Float32 fields): such an
// x is either zero or at least 2^-149 in magnitude, so no intermediate can underflow and only the sign
// of a zero quotient needs care (+-0 / 6 = +-0).
__device__ __forceinline__ double div6_rn_f32src(double x) {
    const double c = 0x1.5555555555555p-3;
    const double q = __dmul_rn(x, c);
    const double r = __fma_rn(-6.0, q, x);
    const double res = __fma_rn(r, c, q);
    // x = +-0 gives res = +0; the quotient of any other x already carries x's sign.  Copying the sign
    // bit of x is one LOP3 on the high word instead of a compare and two selects.
    int hi;
    asm("lop3.b32 %0, %1, %2, 0x80000000, 0xD8;" : "=r"(hi) : "r"(__double2hiint(res)), "r"(__double2hiint(x)));  // c ? b : a
    return __hiloint2double(hi, __double2loint(res));
}

// su, sv: the left-folded six-neighbour sums ((((W+E)+S)+N)+D)+U computed in T by the caller.
// r: integer-valued noise variate ri in [-2^31, 2^31) (reference) / r32 in [-1,1) (native); ignored
// unless NOISE.

template <bool NOISE, int MODE>
__device__ __forceinline__ void cell_update_ref(float uc, float vc, float su, float sv, double r,
                                                const Params& p, float& un, float& vn) {
    const double u64 = (double)uc, v64 = (double)vc;
    // (Float64(su) - 6.0*Float64(uc)) / 6.0 ; 6.0*uc is exact in Float64, so one FMA == mul, sub
    const double lap_u = div6_rn_f32src(__fma_rn(-6.0, u64, (double)su));
    const double lap_v = div6_rn_f32src(__fma_rn(-6.0, v64, (double)sv));
    const float uvv = __fmul_rn(uc, __fmul_rn(vc, vc));            // u * v^2       (Float32)
    const float fkv = __fmul_rn(p.fFK, vc);                        // (F + K) * v   (Float32)
    double du = __dsub_rn(__dmul_rn(p.Du, lap_u), (double)uvv);
    du = __dadd_rn(du, __dmul_rn(p.F, __dsub_rn(1.0, u64)));
    if (NOISE) du = __dadd_rn(du, noise_term<MODE != REF_GENERIC>(p, r));
    double dv = __dadd_rn(__dmul_rn(p.Dv, lap_v), (double)uvv);
    dv = __dsub_rn(dv, (double)fkv);
    if (MODE != REF_DT1) { du = __dmul_rn(du, p.dt); dv = __dmul_rn(dv, p.dt); }
    un = (float)__dadd_rn(u64, du);                                // one rounding on store
    vn = (float)__dadd_rn(v64, dv);
}

template <bool NOISE, int MODE>
__device__ __forceinline__ void cell_update_ref(double uc, double vc, double su, double sv, double r,
                                                const Params& p, double& un, double& vn) {
    const double xu = __dsub_rn(su, __dmul_rn(6.0, uc)), xv = __dsub_rn(sv, __dmul_rn(6.0, vc));
    bool rare = false;
    double lap_u = div6_fast(xu, rare), lap_v = div6_fast(xv, rare);
    if (rare) { lap_u = div6_rare(xu); lap_v = div6_rare(xv); }   // a thin shell of cells at most
    const double uvv = __dmul_rn(uc, __dmul_rn(vc, vc));
    const double fkv = __dmul_rn(p.FK, vc);
    double du = __dsub_rn(__dmul_rn(p.Du, lap_u), uvv);
    du = __dadd_rn(du, __dmul_rn(p.F, __dsub_rn(1.0, uc)));
    if (NOISE) du = __dadd_rn(du, noise_term<MODE != REF_GENERIC>(p, r));
    double dv = __dadd_rn(__dmul_rn(p.Dv, lap_v), uvv);
    dv = __dsub_rn(dv, fkv);
    if (MODE != REF_DT1) { du = __dmul_rn(du, p.dt); dv = __dmul_rn(dv, p.dt); }
    un = __dadd_rn(uc, du);
    vn = __dadd_rn(vc, dv);
}

// Native Float32: same formula, Float32 FMAs.  r32 = noise_r32(word): the variate times 2^23.
// (Regrouping the update as u' = au * su + (cu * u + dt F) - dt u v^2 with host-side coefficients saves
// three operations per cell but rounds cu * u at the size of u instead of du at the size of du: 3e-5
// off the reference after 1000 steps at L = 64, over the 1e-5 bar.  Measured, not adopted.)
template <bool NOISE>
__device__ __forceinline__ void cell_update_native(float uc, float vc, float su, float sv, float r32,
                                                   const Params& p, float& un, float& vn) {
    // 1/6 of the Laplacian is folded into the diffusion coefficients (fDu6 = Du/6, fDv6 = Dv/6)
    const float lap6_u = fmaf(-6.0f, uc, su);
    const float lap6_v = fmaf(-6.0f, vc, sv);
    const float uvv = uc * vc * vc;
    float du = fmaf(p.fDu6, lap6_u, -uvv);
    du = fmaf(p.fF, 1.0f - uc, du);
    if (NOISE) du = fmaf(p.fnoise_s, r32, du);   // r32: integer-valued, fnoise_s = noise * 2^-23
    const float dv = fmaf(p.fDv6, lap6_v, fmaf(-p.fFK, vc, uvv));
    un = fmaf(du, p.fdt, uc);
    vn = fmaf(dv, p.fdt, vc);
}

// Left-fold neighbour sum in T, the order of Common.jl:14-15: i-1, i+1, j-1, j+1, k-1, k+1.
__device__ __forceinline__ float sum6(float w, float e, float s, float n, float d, float u) {
    return __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(w, e), s), n), d), u);
}
__device__ __forceinline__ double sum6(double w, double e, double s, double n, double d, double u) {
    return __dadd_rn(__dadd_rn(__dadd_rn(__dadd_rn(__dadd_rn(w, e), s), n), d), u);
}

}  // namespace gsb
