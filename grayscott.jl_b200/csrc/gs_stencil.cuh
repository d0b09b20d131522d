// gs_stencil.cuh -- fused 7-point-Laplacian + Gray-Scott reaction kernels for sm_100a.
//
// Replaces calculate! (src/simulation/Simulation_CPU.jl:77-113; superseded GPU versions:
// ext/CUDAExt.jl:135-186, src/simulation/Simulation_KA.jl:177-235 -- thread <-> z mapping,
// uncoalesced, 256x over-launched).  Here x (the contiguous dimension) maps to lanes, every
// thread owns one 16-byte vector of x-consecutive cells (4 floats / 2 doubles), marches along z
// and keeps the z-1 / z / z+1 vectors of u and v in registers, so each field is read once and
// written once per step.
//
// Three kernels:
//   stencil_tma_kernel   xy tiles of each z plane staged in shared memory by TMA
//                        (cp.async.bulk.tensor.3d + mbarrier ring, producer warp), the headline
//   stencil_l1_kernel    same marching, y neighbours fetched through L1; no shared memory
//   naive_reflayout_kernel   one thread per cell on arrays in the REFERENCE layout
//                        (unpadded (s+2)^3), for callers that own such device arrays
#pragma once
#include <cuda.h>

#include "gs_arith.cuh"
#include "gs_common.cuh"
#include "gs_philox.cuh"

namespace gsb {

// Update the N cells of one vector.  *m/*0/*p: centre vectors at z-1/z/z+1; *s/*n: rows y-1/y+1
// at z; *w/*e: the scalars left of a[0] and right of a[N-1] at z.
template <typename T, int ARITH, bool NOISE, int MODE>
__device__ __forceinline__ void update_cells(const Vec<T>& um, const Vec<T>& u0, const Vec<T>& up,
                                             const Vec<T>& us, const Vec<T>& un, T uw, T ue,
                                             const Vec<T>& vm, const Vec<T>& v0, const Vec<T>& vp,
                                             const Vec<T>& vs, const Vec<T>& vn, T vw, T ve,
                                             const Params& p, const unsigned (&words)[Vec<T>::N], Vec<T>& uo, Vec<T>& vo) {
    constexpr int N = Vec<T>::N;
#pragma unroll
    for (int i = 0; i < N; ++i) {
        const T uW = i == 0 ? uw : u0.a[i == 0 ? 0 : i - 1];
        const T uE = i == N - 1 ? ue : u0.a[i == N - 1 ? i : i + 1];
        const T vW = i == 0 ? vw : v0.a[i == 0 ? 0 : i - 1];
        const T vE = i == N - 1 ? ve : v0.a[i == N - 1 ? i : i + 1];
        const T su = sum6(uW, uE, us.a[i], un.a[i], um.a[i], up.a[i]);
        const T sv = sum6(vW, vE, vs.a[i], vn.a[i], vm.a[i], vp.a[i]);
        if constexpr (ARITH == ARITH_NATIVE) {
            float r32 = 0.f;
            if (NOISE) r32 = noise_r32(words[i]);
            cell_update_native<NOISE>(u0.a[i], v0.a[i], su, sv, r32, p, uo.a[i], vo.a[i]);
        } else {
            double r = 0.0;
            if (NOISE) r = noise_r64(words[i]);
            cell_update_ref<NOISE, MODE>(u0.a[i], v0.a[i], su, sv, r, p, uo.a[i], vo.a[i]);
        }
    }
}

template <typename T, int ARITH, bool NOISE>
__device__ __forceinline__ void update_vec_w(const Vec<T>& um, const Vec<T>& u0, const Vec<T>& up, const Vec<T>& us,
                                             const Vec<T>& un, T uw, T ue, const Vec<T>& vm, const Vec<T>& v0,
                                             const Vec<T>& vp, const Vec<T>& vs, const Vec<T>& vn, T vw, T ve,
                                             const Params& p, const unsigned (&words)[Vec<T>::N], Vec<T>& uo, Vec<T>& vo);

template <typename T, int ARITH, bool NOISE>
__device__ __forceinline__ void update_vec(const Vec<T>& um, const Vec<T>& u0, const Vec<T>& up,
                                           const Vec<T>& us, const Vec<T>& un, T uw, T ue,
                                           const Vec<T>& vm, const Vec<T>& v0, const Vec<T>& vp,
                                           const Vec<T>& vs, const Vec<T>& vn, T vw, T ve,
                                           const Params& p, int x, int y, int z, Vec<T>& uo, Vec<T>& vo) {
    constexpr int N = Vec<T>::N;
    unsigned words[N];
    if constexpr (NOISE) {
        noise_words<N>(p, (unsigned)(p.gx0 + x), (unsigned)(p.gy0 + y), (unsigned)(p.gz0 + z), words);
    } else {
#pragma unroll
        for (int i = 0; i < N; ++i) words[i] = 0u;
    }
    update_vec_w<T, ARITH, NOISE>(um, u0, up, us, un, uw, ue, vm, v0, vp, vs, vn, vw, ve, p, words, uo, vo);
}

// the same with the noise words of the N cells supplied by the caller
template <typename T, int ARITH, bool NOISE>
__device__ __forceinline__ void update_vec_w(const Vec<T>& um, const Vec<T>& u0, const Vec<T>& up,
                                             const Vec<T>& us, const Vec<T>& un, T uw, T ue,
                                             const Vec<T>& vm, const Vec<T>& v0, const Vec<T>& vp,
                                             const Vec<T>& vs, const Vec<T>& vn, T vw, T ve,
                                             const Params& p, const unsigned (&words)[Vec<T>::N], Vec<T>& uo, Vec<T>& vo) {
    constexpr int N = Vec<T>::N;
    if constexpr (ARITH == ARITH_NATIVE) {
#pragma unroll
        for (int i = 0; i < N; ++i) {
            const T uW = i == 0 ? uw : u0.a[i == 0 ? 0 : i - 1];
            const T uE = i == N - 1 ? ue : u0.a[i == N - 1 ? i : i + 1];
            const T vW = i == 0 ? vw : v0.a[i == 0 ? 0 : i - 1];
            const T vE = i == N - 1 ? ve : v0.a[i == N - 1 ? i : i + 1];
            const T su = sum6(uW, uE, us.a[i], un.a[i], um.a[i], up.a[i]);
            const T sv = sum6(vW, vE, vs.a[i], vn.a[i], vm.a[i], vp.a[i]);
            float r32 = 0.f;
            if (NOISE) r32 = noise_r32(words[i]);
            cell_update_native<NOISE>(u0.a[i], v0.a[i], su, sv, r32, p, uo.a[i], vo.a[i]);
        }
    } else {
        // warp-uniform: one variant of the reference arithmetic for the whole vector
        if (p.ref_mode == REF_DT1)
            update_cells<T, ARITH, NOISE, REF_DT1>(um, u0, up, us, un, uw, ue, vm, v0, vp, vs, vn, vw, ve, p, words, uo, vo);
        else if (p.ref_mode == REF_DT)
            update_cells<T, ARITH, NOISE, REF_DT>(um, u0, up, us, un, uw, ue, vm, v0, vp, vs, vn, vw, ve, p, words, uo, vo);
        else
            update_cells<T, ARITH, NOISE, REF_GENERIC>(um, u0, up, us, un, uw, ue, vm, v0, vp, vs, vn, vw, ve, p, words, uo, vo);
    }
}

// Store the cells of a vector that fall inside [bx0, bx1).  The partial case (a box edge that cuts a
// vector) is a real call: kept inline, its address arithmetic made nvcc carry a second copy of the running
// pointers through the hot loop.
template <typename T>
static __device__ __noinline__ void store_partial(T* pu, T* pv, Vec<T> uo, Vec<T> vo, int x, int bx0, int bx1) {
    constexpr int N = Vec<T>::N;
#pragma unroll
    for (int i = 0; i < N; ++i)
        if (x + i >= bx0 && x + i < bx1) { pu[i] = uo.a[i]; pv[i] = vo.a[i]; }
}

template <typename T>
__device__ __forceinline__ void store_masked(T* pu, T* pv, const Vec<T>& uo, const Vec<T>& vo, int x, int bx0,
                                             int bx1) {
    constexpr int N = Vec<T>::N;
    if (x >= bx0 && x + N <= bx1) {
        st_vec_stream(pu, uo);
        st_vec_stream(pv, vo);
    } else {
        store_partial<T>(pu, pv, uo, vo, x, bx0, bx1);
    }
}

// ------------------------------------------------------------------------------------------
// L1 variant.  block = (32, TY); warp <-> one row; grid = (x tiles, y tiles, z chunks).
// ------------------------------------------------------------------------------------------
template <typename T, int ARITH, bool NOISE, int TY>
__global__ void __launch_bounds__(32 * TY)
stencil_l1_kernel(const T* __restrict__ u, const T* __restrict__ v, T* __restrict__ uo, T* __restrict__ vo,
                  const Layout L, const Params p, const Box box, const int zchunk) {
    constexpr int N = Vec<T>::N;
    const int lane = threadIdx.x;
    const int xt0 = (box.x0 / N) * N;  // vectors are aligned to the absolute grid
    const int x = xt0 + (blockIdx.x * 32 + lane) * N;
    const int y = box.y0 + blockIdx.y * TY + threadIdx.y;
    const int zs = box.z0 + blockIdx.z * zchunk;
    const int ze = min(zs + zchunk, box.z1);
    if (y >= box.y1 || zs >= ze) return;                 // whole warp leaves together
    const bool loadable = (long long)(L.xoff + x + N) <= L.px;
    const bool lane_w = lane == 0, lane_e = lane == 31;
    const long long pxy = L.pxy, px = L.px;

    Vec<T> zero;
#pragma unroll
    for (int i = 0; i < N; ++i) zero.a[i] = T(0);

    const long long base = loadable ? L.idx(x, y, zs) : L.idx(0, y, zs);
    const T* pu = u + base;
    const T* pv = v + base;
    Vec<T> um = loadable ? ld_vec(pu - pxy) : zero, vm = loadable ? ld_vec(pv - pxy) : zero;
    Vec<T> u0 = loadable ? ld_vec(pu) : zero, v0 = loadable ? ld_vec(pv) : zero;
    Vec<T> up = loadable ? ld_vec(pu + pxy) : zero, vp = loadable ? ld_vec(pv + pxy) : zero;

    for (int z = zs; z < ze; ++z) {
        // prefetch the centre of plane z+2 (needed next iteration); plane index <= sz exists
        Vec<T> upp = zero, vpp = zero;
        if (loadable && z + 1 < ze) { upp = ld_vec(pu + 2 * pxy); vpp = ld_vec(pv + 2 * pxy); }
        Vec<T> us = zero, un = zero, vs = zero, vn = zero;
        if (loadable) {
            us = ld_vec(pu - px); un = ld_vec(pu + px);
            vs = ld_vec(pv - px); vn = ld_vec(pv + px);
        }
        T uw = __shfl_up_sync(0xffffffffu, u0.a[N - 1], 1);
        T ue = __shfl_down_sync(0xffffffffu, u0.a[0], 1);
        T vw = __shfl_up_sync(0xffffffffu, v0.a[N - 1], 1);
        T ve = __shfl_down_sync(0xffffffffu, v0.a[0], 1);
        if (lane_w && loadable) { uw = pu[-1]; vw = pv[-1]; }
        if (lane_e && loadable) { ue = pu[N]; ve = pv[N]; }   // row padding guarantees this address
        if (loadable && x < box.x1 && x + N > box.x0) {
            Vec<T> ou, ov;
            update_vec<T, ARITH, NOISE>(um, u0, up, us, un, uw, ue, vm, v0, vp, vs, vn, vw, ve, p, x, y, z, ou, ov);
            store_masked(uo + (pu - u), vo + (pv - v), ou, ov, x, box.x0, box.x1);
        }
        um = u0; u0 = up; up = upp;
        vm = v0; v0 = vp; vp = vpp;
        pu += pxy; pv += pxy;
    }
}

// ------------------------------------------------------------------------------------------
// TMA variant.
//
// CTA = TY consumer warps (one tile row each) + 1 producer warp.  For every plane of its z chunk
// (plus the two halo planes) the producer issues two cp.async.bulk.tensor.3d loads -- the
// (32N + 2N) x (TY + 2) x 1 box of u and of v, x halo = one whole 16-byte vector each side so the
// centre stays 16-byte aligned in shared memory -- into a ring of STAGES slots, completion on a
// per-slot "full" mbarrier.  Consumers keep centre z-1/z/z+1 in registers, read the y+-1 rows and
// the two x edge scalars from the slot of plane z, compute, store with 16-byte coalesced st.global
// and release the slot on its "empty" mbarrier (one arrive per consumer warp).
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// One arrival per (converged) warp: elect.sync picks the lane, no lane-id test, no separate warp barrier.
__device__ __forceinline__ void mbar_arrive_one_per_warp(unsigned bar) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "@p mbarrier.arrive.shared::cta.b64 _, [%0];\n\t}" ::"r"(bar) : "memory");
}
// The suspend-time hint lets the hardware park the warp instead of re-issuing the poll: without it a
// waiting consumer warp spent 11-22 issue slots per plane on PHASECHK / BRA / YIELD (ncu source
// page, round 1), slots the computing warps of the instruction-bound variants need.
#ifndef GSB_MBAR_HINT_NS
#define GSB_MBAR_HINT_NS 20000u
#endif
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(bar), "r"(parity), "r"(GSB_MBAR_HINT_NS) : "memory");
}
__device__ __forceinline__ void tma_load_3d(unsigned dst, const CUtensorMap* map, int c0, int c1, int c2,
                                            unsigned bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
        : "memory");
}

// 1-D bulk copies (the x-face columns of the fused halo scheme): global -> shared completing on an
// mbarrier like the tensor loads, shared -> global in a bulk async-group.  16-byte aligned, n * 16 bytes.
__device__ __forceinline__ void bulk_load_1d(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void bulk_store_1d(void* dst, unsigned src, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
}

template <typename T, int TY>
struct TmaTile {
    static constexpr int N = Vec<T>::N;
    static constexpr int BW = 32 * N + 2 * N;                       // box width in elements (544 B)
    static constexpr int ROWS = TY + 2;
    static constexpr int BOX_BYTES = BW * ROWS * (int)sizeof(T);
    static constexpr int BOXR_BYTES = (BOX_BYTES + 127) / 128 * 128;   // TMA dst alignment
    static constexpr int BOXR = BOXR_BYTES / (int)sizeof(T);           // first element after the box
    // x-face side area behind each field's box (fused halo scheme only), TY elements each:
    //   [W ghosts in | E ghosts in | column x = 0 out | column x = sx-1 out]
    static constexpr int SIDE_BYTES = (4 * TY * (int)sizeof(T) + 127) / 128 * 128;
    __host__ __device__ static constexpr int field_bytes(bool xf) { return BOXR_BYTES + (xf ? SIDE_BYTES : 0); }
    __host__ __device__ static constexpr int stage_bytes(bool xf) { return 2 * field_bytes(xf); }
    __host__ __device__ static constexpr size_t smem_bytes(int stages, bool xf) { return (size_t)stages * stage_bytes(xf) + 128 + 16 * stages; }
};

// One consumer iteration: plane index i of the chunk has landed in slot `cur`; am/a0 (bm/b0) hold
// the centre vectors of planes i-2 / i-1, ap (bp) receives plane i; plane i-1 sits in slot `prev`.
// The caller rotates the register names (three-way unrolled loop), so no register moves.
// XF: an edge lane that owns a subdomain boundary column (own_x) finds its x ghost at `eoff` in the side
// area of the slot (put there by the producer) and leaves its new boundary value at eoff + 2 * TY.
template <typename T, int ARITH, bool NOISE, int TY, bool XF, bool UPD>
__device__ __forceinline__ void tma_consume(const unsigned char* cur, unsigned char* prev, int i,
                                            const Vec<T>& am, const Vec<T>& a0, Vec<T>& ap, const Vec<T>& bm,
                                            const Vec<T>& b0, Vec<T>& bp, const int roff, const int eoff,
                                            const bool lane_w, const bool lane_e, const bool vec_active,
                                            const Params& p, const Box& box, const int x, const int y, const int z,
                                            T*& pou, const long long vdelta, const long long pxy, const bool own_x,
                                            const bool pair_ok, unsigned& keep0, unsigned& keep1) {
    using Tile = TmaTile<T, TY>;
    constexpr int N = Tile::N;
    constexpr int FB = Tile::field_bytes(XF);
    const T* su_ = reinterpret_cast<const T*>(cur);
    const T* sv_ = reinterpret_cast<const T*>(cur + FB);
    ap = ld_vec(su_ + roff);
    bp = ld_vec(sv_ + roff);
    if constexpr (UPD) {                     // planes 0 and 1 of a chunk only fill the register pipeline
        T* pu_ = reinterpret_cast<T*>(prev);
        T* pv_ = reinterpret_cast<T*>(prev + FB);
        const Vec<T> us = ld_vec(pu_ + roff - Tile::BW), un = ld_vec(pu_ + roff + Tile::BW);
        const Vec<T> vs = ld_vec(pv_ + roff - Tile::BW), vn = ld_vec(pv_ + roff + Tile::BW);
        // x neighbours across vectors: shuffles inside the warp; lanes 0 and 31 take the scalar at eoff
        // (the box halo, or the side area when the column beyond is a neighbour's)
        // (every lane loads: for lanes 1..30 eoff is simply the next lane's first element, unused -- two
        // conflicted LDS cost less than the divergent region around them did)
        const T eu = pu_[eoff], ev = pv_[eoff];
        const T su = __shfl_up_sync(0xffffffffu, a0.a[N - 1], 1), sd = __shfl_down_sync(0xffffffffu, a0.a[0], 1);
        const T tu = __shfl_up_sync(0xffffffffu, b0.a[N - 1], 1), td = __shfl_down_sync(0xffffffffu, b0.a[0], 1);
        const T uw = lane_w ? eu : su, ue = lane_e ? eu : sd;
        const T vw = lane_w ? ev : tu, ve = lane_e ? ev : td;
        // Float64 with noise: a thread owns two cells but a Philox call yields four words.  Two lanes that
        // share a counter block (4-aligned global x) share the work across TWO planes: on even iterations
        // the even lane evaluates the block of this plane, the odd lane the block of the next one, and they
        // swap the halves the other needs; on odd iterations both use what they kept.  Same words as
        // noise_words (any alignment), half the generator work per cell.
        constexpr bool PAIR = NOISE && sizeof(T) == 8;
        unsigned nw[N];
        if constexpr (PAIR) {
            if (pair_ok) {
                const bool odd = (threadIdx.x & 1) != 0;
                if ((i & 1) == 0) {
                    const Philox4 blk = philox4x32_10((unsigned)(p.gx0 + x) >> 2, (unsigned)(p.gy0 + y),
                                                      (unsigned)(p.gz0 + z) + (odd ? 1u : 0u), p.step, p);
                    const unsigned r0 = __shfl_xor_sync(0xffffffffu, odd ? blk.w[0] : blk.w[2], 1);
                    const unsigned r1 = __shfl_xor_sync(0xffffffffu, odd ? blk.w[1] : blk.w[3], 1);
                    nw[0] = odd ? r0 : blk.w[0];
                    nw[1] = odd ? r1 : blk.w[1];
                    keep0 = odd ? blk.w[2] : r0;
                    keep1 = odd ? blk.w[3] : r1;
                } else {
                    nw[0] = keep0;
                    nw[1] = keep1;
                }
            } else if (vec_active) {
                noise_words<N>(p, (unsigned)(p.gx0 + x), (unsigned)(p.gy0 + y), (unsigned)(p.gz0 + z), nw);
            }
        }
        if (vec_active) {
            Vec<T> ou, ov;
            if constexpr (PAIR)
                update_vec_w<T, ARITH, NOISE>(am, a0, ap, us, un, uw, ue, bm, b0, bp, vs, vn, vw, ve, p, nw, ou, ov);
            else
                update_vec<T, ARITH, NOISE>(am, a0, ap, us, un, uw, ue, bm, b0, bp, vs, vn, vw, ve, p, x, y, z, ou, ov);
            // one running pointer; v's array lies a kernel-uniform distance away
            store_masked(pou, reinterpret_cast<T*>(reinterpret_cast<char*>(pou) + vdelta), ou, ov, x, box.x0, box.x1);
            pou += pxy;
            if (XF && own_x) {                                   // my boundary column, for the producer to send
                pu_[eoff + 2 * TY] = lane_w ? ou.a[0] : ou.a[N - 1];
                pv_[eoff + 2 * TY] = lane_w ? ov.a[0] : ov.a[N - 1];
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // a bulk copy will read it
            }
        }
    }
}

// XF = true: the x faces of the halo ride inside this kernel (XFaceArgs; pipelined scheme, sx a whole
// number of tiles, sy a whole number of 16-byte vectors).  Only the CTAs of the first / last x tile take
// part, and there the work is the producer's: per plane it fetches the TY ghost values of the W / E
// neighbour column from the dense receive buffer with a bulk copy into the slot's side area (same mbarrier
// as the tile), and when a slot comes back from the consumers it sends the TY new boundary values they
// left there to the dense message buffer with a bulk store.  The consumers pay two predicated shared
// stores per plane in one lane; the strided pack / unpack kernels (128 bytes of DRAM per 4-byte cell)
// disappear.
template <typename T, int ARITH, bool NOISE, int TY, int STAGES, int MINB, bool XF>
__global__ void __launch_bounds__(32 * (TY + 1), MINB)
stencil_tma_kernel(const __grid_constant__ CUtensorMap map_u, const __grid_constant__ CUtensorMap map_v,
                   T* __restrict__ uo, T* __restrict__ vo, const Layout L, const Params p, const Box box,
                   const int zchunk, const XFaceArgs xf) {
    using Tile = TmaTile<T, TY>;
    constexpr int N = Tile::N;
    constexpr int FB = Tile::field_bytes(XF), SB = Tile::stage_bytes(XF);
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    // carve: [stages][2 fields][box | side area] | full[STAGES] | empty[STAGES]; keep the shared address
    // space visible to the compiler (LDS, not generic LD): offset the array, do not round a pointer
    const unsigned pad = (128u - (smem_u32(smem_raw) & 127u)) & 127u;
    unsigned char* tiles = smem_raw + pad;
    const unsigned tiles_u32 = smem_u32(tiles);
    const unsigned bar0 = tiles_u32 + (unsigned)(STAGES * SB);
    auto full_bar = [&](int s) { return bar0 + 8u * (unsigned)s; };
    auto empty_bar = [&](int s) { return bar0 + 8u * (unsigned)(STAGES + s); };

    const int lane = threadIdx.x, w = threadIdx.y;
    const int xt0 = (box.x0 / N) * N;
    const int xtile = xt0 + blockIdx.x * 32 * N;         // first interior x of this tile
    const int ytile = box.y0 + blockIdx.y * TY;          // first interior y of this tile
    const int zs = box.z0 + blockIdx.z * zchunk;
    const int ze = min(zs + zchunk, box.z1);
    const int nplanes = ze - zs + 2;                     // zs-1 .. ze
    // does this tile hold a column whose x neighbour belongs to another subdomain?
    const bool tile_lo = XF && xtile == 0 && xf.send_lo != nullptr;
    const bool tile_hi = XF && xtile + 32 * N == L.sx && xf.send_hi != nullptr;

    if (w == 0 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), TY); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    if (w == TY) {
        // ------------------------------ producer warp ------------------------------
        if (lane == 0) {
            const int c0 = L.xoff + xtile - N;           // element index in row
            const int c1 = ytile;                        // row index of y = ytile-1
            // x faces: rows of this tile inside the subdomain, and where plane zs of the piece starts
            const unsigned xbytes = (unsigned)(min(TY, L.sy - ytile) * (int)sizeof(T));
            const long long f0 = (long long)ytile + (long long)L.sy * (zs - xf.zbase);   // element (ytile, zs)
            const long long fv = (long long)L.sy * xf.nz;                                // v part of a piece
            int s = 0;
            unsigned ph = 1;                             // fresh barriers: waiting on parity 1 passes
            // planes 1 .. nplanes-2 are the updated ones; a slot filled at iteration j comes back at j + STAGES
            const int last = (XF && (tile_lo || tile_hi)) ? nplanes - 2 + STAGES : nplanes - 1;
            for (int i = 0; i <= last; ++i) {
                mbar_wait(empty_bar(s), ph);
                const unsigned slot = tiles_u32 + (unsigned)(s * SB);
                if (XF && (tile_lo || tile_hi)) {
                    const int j = i - STAGES;            // the plane this slot held: its new columns are ready
                    if (j >= 1 && j <= nplanes - 2) {
                        const long long e = f0 + (long long)L.sy * (j - 1);
                        const unsigned out_lo = slot + (unsigned)((Tile::BOXR + 2 * TY) * (int)sizeof(T));
                        const unsigned out_hi = out_lo + (unsigned)(TY * (int)sizeof(T));
                        if (tile_lo) {
                            bulk_store_1d((T*)xf.send_lo + e, out_lo, xbytes);
                            bulk_store_1d((T*)xf.send_lo + e + fv, out_lo + FB, xbytes);
                        }
                        if (tile_hi) {
                            bulk_store_1d((T*)xf.send_hi + e, out_hi, xbytes);
                            bulk_store_1d((T*)xf.send_hi + e + fv, out_hi + FB, xbytes);
                        }
                        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                        // the side area is rewritten only after this slot has been refilled and consumed;
                        // still, do not refill before the copy has read it
                        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                    }
                }
                if (i < nplanes) {
                    const bool upd = XF && i >= 1 && i <= nplanes - 2;     // ghosts are needed for updated planes only
                    const unsigned extra = upd ? ((tile_lo ? 2u : 0u) + (tile_hi ? 2u : 0u)) * xbytes : 0u;
                    mbar_expect_tx(full_bar(s), 2u * (unsigned)Tile::BOX_BYTES + extra);
                    const int c2 = zs + i;               // plane index of z = zs-1+i
                    tma_load_3d(slot, &map_u, c0, c1, c2, full_bar(s));
                    tma_load_3d(slot + FB, &map_v, c0, c1, c2, full_bar(s));
                    if (upd) {
                        const long long e = f0 + (long long)L.sy * (i - 1);
                        const unsigned in_lo = slot + (unsigned)(Tile::BOXR * (int)sizeof(T));
                        const unsigned in_hi = in_lo + (unsigned)(TY * (int)sizeof(T));
                        if (tile_lo) {
                            bulk_load_1d(in_lo, (const T*)xf.ghost_lo + e, xbytes, full_bar(s));
                            bulk_load_1d(in_lo + FB, (const T*)xf.ghost_lo + e + fv, xbytes, full_bar(s));
                        }
                        if (tile_hi) {
                            bulk_load_1d(in_hi, (const T*)xf.ghost_hi + e, xbytes, full_bar(s));
                            bulk_load_1d(in_hi + FB, (const T*)xf.ghost_hi + e + fv, xbytes, full_bar(s));
                        }
                    }
                }
                if (++s == STAGES) { s = 0; ph ^= 1u; }
            }
            if (XF && (tile_lo || tile_hi)) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
        }
        return;
    }

    // ------------------------------ consumer warps ------------------------------
    const int x = xtile + lane * N;
    const int y = ytile + w;
    const bool row_active = y < box.y1;
    const bool vec_active = row_active && x < box.x1 && x + N > box.x0;
    const long long out0 = vec_active ? L.idx(x, y, zs) : 0;
    T* pou = uo + out0;
    const long long vdelta = reinterpret_cast<const char*>(vo) - reinterpret_cast<const char*>(uo);
    const int roff = (w + 1) * Tile::BW + N + lane * N;  // my centre vector in the box (elements)
    const bool lane_w = lane == 0, lane_e = lane == 31;
    // the lane that owns column 0 / sx-1 of a row whose x neighbour is another subdomain
    const bool own_x = XF && ((lane_w && tile_lo) || (lane_e && tile_hi));
    // the scalar an edge lane needs: the box halo, or the ghost the producer put into the side area
    const int eoff = own_x ? Tile::BOXR + (lane_w ? 0 : TY) + w : (lane_w ? roff - 1 : roff + N);

    Vec<T> ua, ub, uc, va, vb, vc;
#pragma unroll
    for (int q = 0; q < N; ++q) { ua.a[q] = ub.a[q] = uc.a[q] = va.a[q] = vb.a[q] = vc.a[q] = T(0); }

    // Float64 noise: lanes 2k, 2k+1 share one Philox counter block when the tile starts on a multiple of 4
    const bool pair_ok = ((p.gx0 + xtile) & 3) == 0;
    unsigned keep0 = 0u, keep1 = 0u;

    int s = 0, sprev = 0;
    unsigned ph = 0;
    // slot release: every lane has used what it loaded from the slot (the update is computed), one lane
    // arrives.  XF: the edge lane's boundary-column stores must be ordered before the arrival.
#ifdef GSB_NO_ELECT
#define GSB_RELEASE_SLOT(BAR) { __syncwarp(); if (lane == 0) mbar_arrive(BAR); }
#else
#define GSB_RELEASE_SLOT(BAR) { if (XF) __syncwarp(); mbar_arrive_one_per_warp(BAR); }
#endif
    // one step of the ring: wait for plane i, consume (plane i-1 is updated), release plane i-1
#define GSB_TMA_STEP(UPD, AM, A0, AP, BM, B0, BP)                                                         \
    {                                                                                                     \
        mbar_wait(full_bar(s), ph);                                                                       \
        tma_consume<T, ARITH, NOISE, TY, XF, UPD>(tiles + s * SB, tiles + sprev * SB, i, AM, A0, AP, BM, B0, BP, roff, eoff, \
                                             lane_w, lane_e, vec_active, p, box, x, y, zs + i - 2, pou, vdelta, L.pxy, \
                                             own_x, pair_ok, keep0, keep1);                               \
        if (i >= 1) GSB_RELEASE_SLOT(empty_bar(sprev))                                                     \
        sprev = s;                                                                                        \
        if (++s == STAGES) { s = 0; ph ^= 1u; }                                                           \
        ++i;                                                                                              \
    }
    int i = 0;
    GSB_TMA_STEP(false, ua, ub, uc, va, vb, vc)          // planes zs-1 and zs: nothing to update yet
    GSB_TMA_STEP(false, ub, uc, ua, vb, vc, va)
    while (i + 3 <= nplanes) {
        GSB_TMA_STEP(true, uc, ua, ub, vc, va, vb)
        GSB_TMA_STEP(true, ua, ub, uc, va, vb, vc)
        GSB_TMA_STEP(true, ub, uc, ua, vb, vc, va)
    }
    if (i < nplanes) {
        GSB_TMA_STEP(true, uc, ua, ub, vc, va, vb)
        if (i < nplanes) GSB_TMA_STEP(true, ua, ub, uc, va, vb, vc)
    }
#undef GSB_TMA_STEP
#undef GSB_RELEASE_SLOT
}

// ------------------------------------------------------------------------------------------
// Naive kernel on the reference layout: arrays are (sx+2)(sy+2)(sz+2), x fastest, no padding
// (what a CuArray-owning caller has).  One thread per cell, x <-> threadIdx.x (coalesced).
// ------------------------------------------------------------------------------------------
template <typename T, int ARITH, bool NOISE>
__global__ void __launch_bounds__(256)
naive_reflayout_kernel(const T* __restrict__ u, const T* __restrict__ v, T* __restrict__ uo, T* __restrict__ vo,
                       const int sx, const int sy, const int sz, const Params p) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    const int z = blockIdx.z;
    if (x >= sx || y >= sy || z >= sz) return;
    const long long nx = sx + 2, nxy = nx * (long long)(sy + 2);
    const long long c = (x + 1) + nx * (long long)(y + 1) + nxy * (long long)(z + 1);
    const T uc = u[c], vc = v[c];
    const T su = sum6(u[c - 1], u[c + 1], u[c - nx], u[c + nx], u[c - nxy], u[c + nxy]);
    const T sv = sum6(v[c - 1], v[c + 1], v[c - nx], v[c + nx], v[c - nxy], v[c + nxy]);
    unsigned word = 0;
    if (NOISE) {
        const unsigned gx = (unsigned)(p.gx0 + x);
        Philox4 ph = philox4x32_10(gx >> 2, (unsigned)(p.gy0 + y), (unsigned)(p.gz0 + z), p.step, p);
        const unsigned sel = gx & 3u;
        word = sel == 0 ? ph.w[0] : sel == 1 ? ph.w[1] : sel == 2 ? ph.w[2] : ph.w[3];
    }
    T un, vn;
    if constexpr (ARITH == ARITH_NATIVE) {
        cell_update_native<NOISE>(uc, vc, su, sv, NOISE ? noise_r32(word) : 0.f, p, un, vn);
    } else {
        cell_update_ref<NOISE, REF_GENERIC>(uc, vc, su, sv, NOISE ? noise_r64(word) : 0.0, p, un, vn);
    }
    uo[c] = un;
    vo[c] = vn;
}

}  // namespace gsb
