// gs_philox.cuh -- counter-based noise for the du term (Simulation_CPU.jl:103).
//
// The reference draws rand(Distributions.Uniform(-1,1)) = -1 + 2*rand(Float64) from Julia's
// unseeded task-local RNG, so its noise is not reproducible and parity is statistical only.
// Here the variate of a cell is a pure function of (seed, step, GLOBAL cell coordinate), so a
// run is reproducible and independent of the domain decomposition:
//
//   word(gx,gy,gz,step) = Philox4x32-10(counter = (gx>>2, gy, gz, step), key = seed)[gx & 3]
//   r  = -1 + 2 * word * 2^-32                       in [-1, 1)   (reference arithmetic, Float64)
//   r32 = float((int32)(word ^ 0x80000000) >> 8) * 2^-23  in [-1, 1)   (native Float32; the kernel
//         keeps the integer-valued float and folds the 2^-23 into the noise amplitude, which is exact)
//
// One Philox call serves four x-consecutive cells.  Philox4x32-10 is Salmon et al. (SC'11),
// constants as published in Random123; known-answer vectors are checked in tests/.
// oracle/gs_oracle.c holds the same definition on the CPU.
#pragma once
#include "gs_common.cuh"

namespace gsb {

struct Philox4 { unsigned w[4]; };

// The ten round keys (k0 + r * 0x9E3779B9, k1 + r * 0xBB67AE85) are prepared on the host
// (Params::pk): they reach the xor as constant-bank operands instead of 20 adds per call.
__device__ __forceinline__ Philox4 philox4x32_10(unsigned c0, unsigned c1, unsigned c2, unsigned c3,
                                                 const Params& p) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const unsigned hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const unsigned hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const unsigned n0 = hi1 ^ c1 ^ p.pk[r][0], n2 = hi0 ^ c3 ^ p.pk[r][1];
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    }
    Philox4 o;
    o.w[0] = c0; o.w[1] = c1; o.w[2] = c2; o.w[3] = c3;
    return o;
}

__device__ __forceinline__ double noise_r64(unsigned word) {
    // r = -1 + 2*(word * 2^-32) = ri * 2^-31 with ri = word - 2^31; returns ri (exact), the caller
    // applies the 2^-31 (gs_arith.cuh noise_term).  The integer becomes a double without a conversion
    // instruction: the bit pattern 0x43300000'word is exactly 2^52 + word, and subtracting 2^52 + 2^31 is exact
    // (an I2F.F64 costs ~5.5 dispatch clocks on B200, a DADD 2.2: profiles/r2_pipe_throughputs.txt).
    return __dadd_rn(__hiloint2double(0x43300000, (int)word), -4503601774854144.0);
}
__device__ __forceinline__ float noise_r32(unsigned word) {
    // the integer-valued variate r32 * 2^23 in [-2^23, 2^23); cell_update_native multiplies it by
    // Params::fnoise_s = noise * 2^-23 (scaling by a power of two commutes with the rounding)
    return (float)((int)(word ^ 0x80000000u) >> 8);
}

// Noise words for the N cells gx .. gx+N-1 of one row (N = 4 or 2), any alignment of gx.
template <int N>
__device__ __forceinline__ void noise_words(const Params& p, unsigned gx, unsigned gy, unsigned gz,
                                            unsigned (&w)[N]) {
    Philox4 a = philox4x32_10(gx >> 2, gy, gz, p.step, p);
    if (((gx & 3u) + N) <= 4u) {
#pragma unroll
        for (int i = 0; i < N; ++i) {
            // select a.w[(gx&3)+i] without dynamic register indexing
            const unsigned s = (gx & 3u) + i;
            w[i] = s == 0 ? a.w[0] : s == 1 ? a.w[1] : s == 2 ? a.w[2] : a.w[3];
        }
    } else {
        Philox4 b = philox4x32_10((gx >> 2) + 1, gy, gz, p.step, p);
#pragma unroll
        for (int i = 0; i < N; ++i) {
            const unsigned s = (gx & 3u) + i;  // 0..6
            w[i] = s == 0 ? a.w[0] : s == 1 ? a.w[1] : s == 2 ? a.w[2] : s == 3 ? a.w[3]
                 : s == 4 ? b.w[0] : s == 5 ? b.w[1] : b.w[2];
        }
    }
}

}  // namespace gsb
