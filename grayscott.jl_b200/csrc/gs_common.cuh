// gs_common.cuh -- device-side layout, parameters and small helpers shared by all kernels.
//
// Device layout (ours, not the reference's): one subdomain of sx*sy*sz cells with one ghost
// layer (reference: Array{T,3}(sx+2, sy+2, sz+2), x fastest -- Simulation_CPU.jl:23-28) is
// stored with a padded row pitch so that interior cell x = 0 sits on a 128-byte boundary:
//
//   row:  [ unused ... | W ghost | x=0 ... x=sx-1 | E ghost | pad ... ]
//           0            xoff-1    xoff                       xoff+sx      < px
//
//   xoff = 128 / sizeof(T)            (32 floats / 16 doubles)
//   px   = round_up(xoff + sx + 1, 128 / sizeof(T))   elements per row
//   rows = sy + 2, planes = sz + 2    (ghost row/plane first and last, as in the reference)
//
// Local coordinates used by the kernels are 0-based interior coordinates: x in [-1, sx],
// y in [-1, sy], z in [-1, sz]; -1 and s are the ghost layers (reference indices 1 and s+2).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace gsb {

struct Layout {
    int sx, sy, sz;      // interior sizes (mcd.proc_sizes, communication.jl:73)
    int xoff;            // element index of interior x = 0 within a row
    long long px;        // row pitch in elements
    long long pxy;       // plane pitch in elements = px * (sy + 2)

    __host__ __device__ long long idx(int x, int y, int z) const {
        return (long long)(xoff + x) + px * (long long)(y + 1) + pxy * (long long)(z + 1);
    }
    __host__ __device__ long long elems() const { return pxy * (long long)(sz + 2); }
};

inline Layout make_layout(int sx, int sy, int sz, size_t elem_size) {
    Layout l;
    l.sx = sx; l.sy = sy; l.sz = sz;
    const int line = (int)(128 / elem_size);
    l.xoff = line;
    l.px = ((long long)(l.xoff + sx + 1) + line - 1) / line * line;
    l.pxy = l.px * (long long)(sy + 2);
    return l;
}

// Physics + noise parameters, prepared on the host exactly as calculate! does
// (Simulation_CPU.jl:82-87: every Settings value is first rounded to T).
struct Params {
    // reference arithmetic: T-rounded values widened to double
    double Du, Dv, F, dt, noise;
    double FK;           // (F + K) added in T, then widened   (Simulation_CPU.jl:105)
    double noise_s;      // noise * 2^-31 (exactly), valid when noise_fast
    int noise_fast;
    int ref_mode;        // REF_DT1 / REF_DT / REF_GENERIC (gs_arith.cuh): which variant of the reference arithmetic
    // native Float32 arithmetic
    float fDu, fDv, fF, fFK, fdt, fnoise;
    float fnoise_s;      // noise * 2^-23: multiplies the integer-valued variate of noise_r32
    float fDu6, fDv6;    // Du/6, Dv/6
    int dt_is_one;       // dt == 1 exactly: x * dt == x, multiply skipped (bit-identical)
    // noise: Philox4x32-10 keyed on (seed, step, global cell)
    unsigned seed_lo, seed_hi, step;
    unsigned pk[10][2];  // Philox round keys: seed + r * (0x9E3779B9, 0xBB67AE85)
    int gx0, gy0, gz0;   // global 0-based coordinate of local interior cell (0,0,0) = proc_offsets
};

// A box of interior cells [x0,x1) x [y0,y1) x [z0,z1) to update (interior / shell split).
struct Box {
    int x0, x1, y0, y1, z0, z1;
};

// X faces carried by the update kernel itself (pipelined halo scheme, sx a multiple of the tile width,
// sy a multiple of a 16-byte vector): instead of a separate strided gather/scatter over 2 * sy * sz cells,
// the CTAs of the first / last x tile read their W / E ghosts from the dense receive buffer and write
// their new boundary columns into the dense message buffer (bulk copies issued by the producer warp).
// Layout of a piece [zbase, zbase+nz): [u | v], element y + sy * (z - zbase); the v part starts sy * nz
// elements later.  A side is on when both of its pointers are set.
struct XFaceArgs {
    const void* ghost_lo;   // W ghosts for column x = 0        (null: side off, the array's ghost column is read)
    const void* ghost_hi;   // E ghosts for column x = sx-1
    void* send_lo;          // new values of column x = 0
    void* send_hi;          // new values of column x = sx-1
    int zbase, nz;
};

enum : int { ARITH_REFERENCE = 0, ARITH_NATIVE = 1 };

// 16-byte vector of T: float x4 or double x2.
template <typename T>
struct alignas(16) Vec {
    static constexpr int N = 16 / sizeof(T);
    T a[N];
};

template <typename T>
__device__ __forceinline__ Vec<T> ld_vec(const T* p) {
    return *reinterpret_cast<const Vec<T>*>(p);
}
template <typename T>
__device__ __forceinline__ void st_vec(T* p, const Vec<T>& v) {
    *reinterpret_cast<Vec<T>*>(p) = v;
}
// streaming store (st.global.cs): the results are not read again before the next step, so they should
// not displace the halo rows and planes that neighbouring CTAs are about to re-read from L2
__device__ __forceinline__ void st_vec_stream(float* p, const Vec<float>& v) {
    __stcs(reinterpret_cast<float4*>(p), make_float4(v.a[0], v.a[1], v.a[2], v.a[3]));
}
__device__ __forceinline__ void st_vec_stream(double* p, const Vec<double>& v) {
    __stcs(reinterpret_cast<double2*>(p), make_double2(v.a[0], v.a[1]));
}

}  // namespace gsb
