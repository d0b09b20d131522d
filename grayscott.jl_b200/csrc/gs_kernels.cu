// gs_kernels.cu -- auxiliary kernels (init, halo pack/unpack, per-slice PDF, statistics),
// the TMA tensor-map encoder and the dtype dispatch of the stencil launchers.
#include <cfloat>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "gs_kernels.h"

namespace gsb {

template <typename T>
cudaError_t launch_stencil_T(const StencilCfg&, const void*, const void*, void*, void*, const CUtensorMap*,
                             const CUtensorMap*, const Layout&, const Params&, const Box&, const XFaceArgs&, cudaStream_t);
template <typename T>
cudaError_t launch_naive_T(int, int, const void*, const void*, void*, void*, int, int, int, const Params&,
                           cudaStream_t);

cudaError_t launch_stencil(const StencilCfg& c, const void* u, const void* v, void* uo, void* vo,
                           const CUtensorMap* mu, const CUtensorMap* mv, const Layout& L, const Params& p,
                           const Box& box, cudaStream_t st, const XFaceArgs* xface) {
    XFaceArgs xf;
    memset(&xf, 0, sizeof xf);
    xf.nz = 1;
    if (xface) xf = *xface;
    return c.dtype == DT_F32 ? launch_stencil_T<float>(c, u, v, uo, vo, mu, mv, L, p, box, xf, st)
                             : launch_stencil_T<double>(c, u, v, uo, vo, mu, mv, L, p, box, xf, st);
}

cudaError_t launch_naive_reflayout(int dtype, int arith, int noise, const void* u, const void* v, void* uo,
                                   void* vo, int sx, int sy, int sz, const Params& p, cudaStream_t st) {
    return dtype == DT_F32 ? launch_naive_T<float>(arith, noise, u, v, uo, vo, sx, sy, sz, p, st)
                           : launch_naive_T<double>(arith, noise, u, v, uo, vo, sx, sy, sz, p, st);
}

// ------------------------------------------------------------------ tensor map
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}

int make_field_tensor_map(CUtensorMap* out, int dtype, const void* base, const Layout& L, int ty, const char** err) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) { if (err) *err = "cuTensorMapEncodeTiled not available from the driver"; return 1; }
    const size_t es = dtype == DT_F32 ? 4 : 8;
    const int n = (int)(16 / es);
    cuuint64_t dims[3] = {(cuuint64_t)L.px, (cuuint64_t)(L.sy + 2), (cuuint64_t)(L.sz + 2)};
    cuuint64_t strides[2] = {(cuuint64_t)L.px * es, (cuuint64_t)L.pxy * es};
    cuuint32_t box[3] = {(cuuint32_t)(32 * n + 2 * n), (cuuint32_t)(ty + 2), 1};
    cuuint32_t estr[3] = {1, 1, 1};
    // L2 promotion (fetch granularity of the TMA reads): $GSB_TMA_L2PROMO = 0 | 64 | 128 | 256
    CUtensorMapL2promotion promo = CU_TENSOR_MAP_L2_PROMOTION_L2_128B;
    if (const char* e = getenv("GSB_TMA_L2PROMO")) {
        const int v = atoi(e);
        promo = v == 0 ? CU_TENSOR_MAP_L2_PROMOTION_NONE : v == 64 ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B
              : v == 256 ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : CU_TENSOR_MAP_L2_PROMOTION_L2_128B;
    }
    CUresult r = fn(out, dtype == DT_F32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3,
                    const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        static thread_local char msg[96];
        snprintf(msg, sizeof msg, "cuTensorMapEncodeTiled failed (CUresult %d)", (int)r);
        if (err) *err = msg;
        return 1;
    }
    return 0;
}

// ------------------------------------------------------------------ init_fields
template <typename T>
__global__ void fill_kernel(T* __restrict__ u, T* __restrict__ v, T* __restrict__ ut, T* __restrict__ vt,
                            long long n) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        u[i] = T(1); v[i] = T(0); ut[i] = T(0); vt[i] = T(0);
    }
}

// the 13^3 block [G/2-6, G/2+6]^3 in global coordinates (Simulation_CPU.jl:31-57)
template <typename T>
__global__ void seed_kernel(T* __restrict__ u, T* __restrict__ v, Layout L, long long mx, long long my,
                            long long mz, long long ox, long long oy, long long oz) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 13 * 13 * 13) return;
    const long long gx = mx + t % 13, gy = my + (t / 13) % 13, gz = mz + t / 169;
    const long long x = gx - ox, y = gy - oy, z = gz - oz;
    if (x < 0 || x >= L.sx || y < 0 || y >= L.sy || z < 0 || z >= L.sz) return;  // is_inside
    const long long i = L.idx((int)x, (int)y, (int)z);
    u[i] = T(0.25);
    v[i] = T(0.33);
}

template <typename T>
static cudaError_t init_T(void* u, void* v, void* ut, void* vt, const Layout& L, const long long g[3],
                          const long long off[3], cudaStream_t st) {
    fill_kernel<T><<<148 * 8, 256, 0, st>>>((T*)u, (T*)v, (T*)ut, (T*)vt, L.elems());
    seed_kernel<T><<<(13 * 13 * 13 + 255) / 256, 256, 0, st>>>((T*)u, (T*)v, L, g[0] / 2 - 6, g[1] / 2 - 6,
                                                                g[2] / 2 - 6, off[0], off[1], off[2]);
    return cudaGetLastError();
}

cudaError_t launch_init_fields(int dtype, void* u, void* v, void* ut, void* vt, const Layout& L,
                               const long long g[3], const long long off[3], cudaStream_t st) {
    return dtype == DT_F32 ? init_T<float>(u, v, ut, vt, L, g, off, st) : init_T<double>(u, v, ut, vt, L, g, off, st);
}

// ------------------------------------------------------------------ halo faces
long long face_elems(const Layout& L, int dir) {
    switch (dir >> 1) {
        case 0: return (long long)L.sy * L.sz;   // x faces (yz)
        case 1: return (long long)L.sx * L.sz;   // y faces (xz)
        default: return (long long)L.sx * L.sy;  // z faces (xy)
    }
}

long long face_piece_elems(const Layout& L, int dir, int z0, int z1) {
    switch (dir >> 1) {
        case 0: return (long long)L.sy * (z1 - z0);
        case 1: return (long long)L.sx * (z1 - z0);
        default: return (long long)L.sx * L.sy;
    }
}

// element e of the piece of face `dir` that starts at plane zb -> interior coordinates; `ghost`
// selects the ghost layer (unpack) or the boundary interior layer (pack)
__device__ __forceinline__ long long face_index(const Layout& L, int dir, long long e, int zb, bool ghost) {
    int x, y, z;
    const bool hi = dir & 1;
    switch (dir >> 1) {
        case 0:
            y = (int)(e % L.sy); z = zb + (int)(e / L.sy);
            x = hi ? (ghost ? L.sx : L.sx - 1) : (ghost ? -1 : 0);
            break;
        case 1:
            x = (int)(e % L.sx); z = zb + (int)(e / L.sx);
            y = hi ? (ghost ? L.sy : L.sy - 1) : (ghost ? -1 : 0);
            break;
        default:
            x = (int)(e % L.sx); y = (int)(e / L.sx);
            z = hi ? (ghost ? L.sz : L.sz - 1) : (ghost ? -1 : 0);
            break;
    }
    return L.idx(x, y, z);
}

struct FaceArgs {
    int n;
    int dir[6];
    void* buf[6];
    long long count[6];
    int zb[6];
};

template <typename T, bool PACK>
__global__ void faces_kernel(T* __restrict__ u, T* __restrict__ v, Layout L, FaceArgs fa) {
    const int f = blockIdx.y;
    if (f >= fa.n) return;
    const int dir = fa.dir[f];
    const long long n = fa.count[f];
    const int zb = fa.zb[f];
    T* buf = (T*)fa.buf[f];
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += stride) {
        const long long i = face_index(L, dir, e, zb, !PACK);
        if (PACK) { buf[e] = u[i]; buf[n + e] = v[i]; }
        else { u[i] = buf[e]; v[i] = buf[n + e]; }
    }
}

template <bool PACK>
static cudaError_t faces_launch(int dtype, void* u, void* v, const Layout& L, const FaceSet& fs, cudaStream_t st) {
    if (fs.n <= 0) return cudaSuccess;
    FaceArgs fa;
    fa.n = fs.n;
    long long maxn = 0;
    for (int i = 0; i < 6; ++i) { fa.dir[i] = 0; fa.buf[i] = nullptr; fa.count[i] = 0; fa.zb[i] = 0; }
    for (int i = 0; i < fs.n; ++i) {
        fa.dir[i] = fs.dir[i]; fa.buf[i] = fs.buf[i]; fa.zb[i] = fs.z0[i];
        fa.count[i] = face_piece_elems(L, fs.dir[i], fs.z0[i], fs.z1[i]);
        if (fa.count[i] > maxn) maxn = fa.count[i];
    }
    if (maxn == 0) return cudaSuccess;
    long long bx = (maxn + 255) / 256;
    if (bx > 148 * 4) bx = 148 * 4;
    dim3 grid((unsigned)bx, (unsigned)fs.n);
    if (dtype == DT_F32) faces_kernel<float, PACK><<<grid, 256, 0, st>>>((float*)u, (float*)v, L, fa);
    else faces_kernel<double, PACK><<<grid, 256, 0, st>>>((double*)u, (double*)v, L, fa);
    return cudaGetLastError();
}

cudaError_t launch_pack_faces(int dtype, const void* u, const void* v, const Layout& L, const FaceSet& fs,
                              cudaStream_t st) {
    return faces_launch<true>(dtype, const_cast<void*>(u), const_cast<void*>(v), L, fs, st);
}
cudaError_t launch_unpack_faces(int dtype, void* u, void* v, const Layout& L, const FaceSet& fs, cudaStream_t st) {
    return faces_launch<false>(dtype, u, v, L, fs, st);
}

// ------------------------------------------------------------------ per-slice PDF
// _compute_pdf (src/analysis/pdfcalc.jl:13-49, unfinished in the reference): for every z slice a
// histogram of `nbins` equal bins over [min, max]; bin = floor((value - min) / bin_width) evaluated in
// T, value == max falls into the last bin, values outside the range are not counted.
// grid = (z slices, row groups): a CTA histograms its share of the slice's rows in shared memory and
// adds the non-empty bins to the slice's histogram in global memory (zeroed by the launcher).
template <typename T>
__global__ void pdf_kernel(const T* __restrict__ f, Layout L, int nbins, T vmin, T vmax, T bin_width,
                           unsigned* __restrict__ pdf) {
    extern __shared__ unsigned hist[];
    for (int b = threadIdx.x; b < nbins; b += blockDim.x) hist[b] = 0;
    __syncthreads();
    const int z = blockIdx.x;
    for (int y = blockIdx.y; y < L.sy; y += gridDim.y) {
        const T* row = f + L.idx(0, y, z);
        for (int x = threadIdx.x; x < L.sx; x += blockDim.x) {
            const T v = row[x];
            if (!(v >= vmin && v <= vmax)) continue;       // outside the window, or NaN: not counted
            int bin = (int)floor((v - vmin) / bin_width);
            bin = bin < 0 ? 0 : (bin >= nbins ? nbins - 1 : bin);
            atomicAdd(&hist[bin], 1u);
        }
    }
    __syncthreads();
    for (int b = threadIdx.x; b < nbins; b += blockDim.x)
        if (hist[b]) atomicAdd(&pdf[(long long)z * nbins + b], hist[b]);
}

cudaError_t launch_pdf(int dtype, const void* f, const Layout& L, int nbins, double vmin, double vmax,
                       unsigned* pdf_dev, cudaStream_t st) {
    const size_t smem = sizeof(unsigned) * (size_t)nbins;
    cudaError_t e = cudaMemsetAsync(pdf_dev, 0, sizeof(unsigned) * (size_t)nbins * (size_t)L.sz, st);
    if (e != cudaSuccess) return e;
    // enough CTAs to fill the GPU (148 SMs x 8) without splitting small slices into crumbs
    int groups = (148 * 8 + L.sz - 1) / L.sz;
    if (groups > (L.sy + 7) / 8) groups = (L.sy + 7) / 8;
    if (groups < 1) groups = 1;
    dim3 grid((unsigned)L.sz, (unsigned)groups);
    if (dtype == DT_F32) {
        const float lo = (float)vmin, hi = (float)vmax, w = (hi - lo) / (float)nbins;
        pdf_kernel<float><<<grid, 256, smem, st>>>((const float*)f, L, nbins, lo, hi, w, pdf_dev);
    } else {
        const double w = (vmax - vmin) / (double)nbins;
        pdf_kernel<double><<<grid, 256, smem, st>>>((const double*)f, L, nbins, vmin, vmax, w, pdf_dev);
    }
    return cudaGetLastError();
}

// ------------------------------------------------------------------ statistics
static constexpr int STATS_BLOCKS = 148 * 4;
static constexpr int STATS_THREADS = 256;
int stats_scratch_doubles() { return STATS_BLOCKS * 5; }

template <typename T>
__global__ void stats_partial_kernel(const T* __restrict__ f, Layout L, double thresh, double* __restrict__ part) {
    double s = 0.0, s2 = 0.0, mn = DBL_MAX, mx = -DBL_MAX, cnt = 0.0;
    const long long rows = (long long)L.sy * L.sz;
    for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
        const int y = (int)(r % L.sy), z = (int)(r / L.sy);
        const T* row = f + L.idx(0, y, z);
        for (int x = threadIdx.x; x < L.sx; x += blockDim.x) {
            const double a = (double)row[x];
            s += a; s2 += a * a; mn = fmin(mn, a); mx = fmax(mx, a); cnt += a > thresh ? 1.0 : 0.0;
        }
    }
    __shared__ double sh[5][STATS_THREADS];
    sh[0][threadIdx.x] = s; sh[1][threadIdx.x] = s2; sh[2][threadIdx.x] = mn; sh[3][threadIdx.x] = mx;
    sh[4][threadIdx.x] = cnt;
    __syncthreads();
    for (int o = STATS_THREADS / 2; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) {
            sh[0][threadIdx.x] += sh[0][threadIdx.x + o];
            sh[1][threadIdx.x] += sh[1][threadIdx.x + o];
            sh[2][threadIdx.x] = fmin(sh[2][threadIdx.x], sh[2][threadIdx.x + o]);
            sh[3][threadIdx.x] = fmax(sh[3][threadIdx.x], sh[3][threadIdx.x + o]);
            sh[4][threadIdx.x] += sh[4][threadIdx.x + o];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0)
        for (int q = 0; q < 5; ++q) part[blockIdx.x * 5 + q] = sh[q][0];
}

__global__ void stats_final_kernel(const double* __restrict__ part, int nblocks, double* __restrict__ out) {
    // one thread per statistic, fixed order: deterministic
    const int q = threadIdx.x;
    if (q >= 5) return;
    double a = (q == 2) ? DBL_MAX : (q == 3 ? -DBL_MAX : 0.0);
    for (int b = 0; b < nblocks; ++b) {
        const double x = part[b * 5 + q];
        a = q == 2 ? fmin(a, x) : q == 3 ? fmax(a, x) : a + x;
    }
    out[q] = a;
}

cudaError_t launch_field_stats(int dtype, const void* f, const Layout& L, double thresh, double* partials,
                               double* out_dev, cudaStream_t st) {
    if (dtype == DT_F32) stats_partial_kernel<float><<<STATS_BLOCKS, STATS_THREADS, 0, st>>>((const float*)f, L, thresh, partials);
    else stats_partial_kernel<double><<<STATS_BLOCKS, STATS_THREADS, 0, st>>>((const double*)f, L, thresh, partials);
    stats_final_kernel<<<1, 32, 0, st>>>(partials, STATS_BLOCKS, out_dev);
    return cudaGetLastError();
}

}  // namespace gsb
